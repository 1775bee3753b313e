"""Regenerates the golden vectors in this directory from the CPU oracle (oracle/krrt_oracle.c).

The reference's own RRT* implementation is an absent submodule, so these are NOT reference
outputs: they pin the oracle against drift and give the GPU tests a fixture that does not need the oracle.
Run:  python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import oracle_binding as ob  # noqa: E402
from robosoccer_motion_planning_b200 import scenes  # noqa: E402

rng = np.random.default_rng(2024)
n = 2000
a = rng.uniform(scenes.FIELD_LO, scenes.FIELD_HI, (n, 4))
b = rng.uniform(scenes.FIELD_LO, scenes.FIELD_HI, (n, 4))
a[::5, 2:] *= 4
b[::5, 2:] *= 4
b[::5, :2] = a[::5, :2] + 0.3 * rng.uniform(-1, 1, (len(a[::5]), 2))
b[::9, :2] = a[::9, :2]
tau, cost = ob.steer(a, b, 0.5)
np.savez_compressed(os.path.join(HERE, "steer.npz"), a=a, b=b, r=0.5, tau=tau, cost=cost)

obs = scenes.dynamic_obstacles()
P = ob.OraclePlanner(capacity=2048, p_direct=0.1)
P.set_obstacles(obs, 1)
P.set_start(scenes.START)
solved, gi = P.grow(scenes.GOAL, 1500)
st, par, ctc, tim = P.tree()
np.savez_compressed(os.path.join(HERE, "tree_c2_small.npz"), capacity=2048, p_direct=0.1, obs=obs, start=scenes.START,
                    goal=scenes.GOAL, n_iter=1500, states=st, parent=par, ctc=ctc, time=tim, solved=solved,
                    goal_index=gi, result=np.frombuffer(P.result(solved, gi).tobytes(), dtype=np.uint8))

starts, goals, bobs = scenes.batch_queries(24, 9, seed=7)
out = ob.plan_batch(ob.default_config(p_direct=0.1, seed=5), starts, goals, bobs, 1, 120, query_base=500, nthreads=2)
tol = ob.plan_batch(ob.default_config(p_direct=0.1, seed=5, goal_mode=1), starts, goals, bobs, 1, 120, query_base=500,
                    nthreads=2)
np.savez_compressed(os.path.join(HERE, "batch_small.npz"), starts=starts, goals=goals, obs=bobs, n_iter=120,
                    query_base=500, results=out.view(np.uint8), results_tolerance=tol.view(np.uint8))
print("golden vectors written")
