import sys, os; sys.path.insert(0,'tests'); sys.path.insert(0,'.')
import numpy as np, oracle_binding as ob
from robosoccer_motion_planning_b200 import scenes, capi
starts, goals, obs = scenes.batch_queries(400, 9, seed=23)
goals[::7] = starts[::7]; goals[3, :2] = obs[3, 0, :2]
for mode in (0,1):
  kw = dict(p_direct=0.1, seed=9, goal_mode=mode, goal_tol=0.7)
  for pre in (False, True):
    G = capi.Planner(capacity=1 << 12, **kw)
    if pre: G.plan_batch(starts, goals, obs, 1, 200, query_base=77)
    out = G.plan_batch(starts, goals, obs, 1, 30, query_base=77)
    ref = ob.plan_batch(ob.default_config(capacity=1 << 12, **kw), starts, goals, obs, 1, 30, query_base=77, nthreads=8)
    bad = np.nonzero(out['n_nodes'] != ref['n_nodes'])[0]
    print("mode", mode, "pre", pre, "bad queries", bad[:20], len(bad), "equal", out.tobytes()==ref.tobytes())
    for b in bad[:3]:
      o1 = G.plan_batch(starts[b:b+1], goals[b:b+1], obs[b:b+1], 1, 30, query_base=77+int(b))
      print("  single", b, o1['n_nodes'], ref['n_nodes'][b], G.batch_stats())
      P=ob.OraclePlanner(capacity=32,query_id=77+int(b),**kw); P.set_obstacles(obs[b],1); P.set_start(starts[b]); P.grow(goals[b],30); print("  oracle", P.stats())
