"""Smallest C4 case for an ncu capture of k_plan_batch2: 4096 queries x 200 iterations, three launches."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from robosoccer_motion_planning_b200 import capi, scenes
bs, bg, bo = scenes.batch_queries(4096, 9, seed=0)
G = capi.Planner(p_direct=0.1)
for _ in range(3):
    out = G.plan_batch(bs, bg, bo, 1, 200)
print("solved", float(out["solved"].mean()))
