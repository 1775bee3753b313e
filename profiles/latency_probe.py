import sys, time
sys.path.insert(0, '/root/repo')
import numpy as np
from robosoccer_motion_planning_b200 import capi, scenes
G = capi.Planner(capacity=4096, p_direct=0.1)
G.set_obstacles(scenes.dynamic_obstacles(), 1)
for rep in range(3):
    G.set_start(scenes.START)
    t0 = time.perf_counter()
    for i in range(75):
        G.grow(scenes.GOAL, 1)
    t1 = time.perf_counter()
    G.set_start(scenes.START)
    t2 = time.perf_counter()
    G.grow(scenes.GOAL, 75)
    t3 = time.perf_counter()
    print(f"75 x grow(1): {1e3*(t1-t0):.3f} ms ({1e6*(t1-t0)/75:.1f} us per call);  grow(75): {1e3*(t3-t2):.3f} ms", flush=True)
