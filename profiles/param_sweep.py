#!/usr/bin/env python
"""Window size / sub-round threshold sweep of the single-tree planner (C2 20k iterations, C3 100k iterations).
Device time of the grow call, best of 3.  KRRT_WAVE_MIN_N is read at context creation."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from robosoccer_motion_planning_b200 import capi, scenes  # noqa: E402


def run(mode, batch, min_n):
    os.environ["KRRT_WAVE_MIN_N"] = str(min_n)
    n_iter = 20000 if mode == "c2" else 100000
    G = capi.Planner(capacity=n_iter + 2, p_direct=0.1 if mode == "c2" else 0.0, batch=batch)
    if mode == "c2":
        G.set_obstacles(scenes.dynamic_obstacles(), 1)
        goal = scenes.GOAL
    else:
        G.set_obstacles(scenes.static_obstacles(), 0)
        goal = None
    best = 1e9
    for _ in range(3):
        G.set_start(scenes.START)
        G.grow(goal, n_iter)
        best = min(best, G.last_timing()["total_ms"])
    w = G.wave_stats()
    G.close()
    return best, w["rounds_parallel"], w["rounds_sequential"]


for mode in ("c2", "c3"):
    for batch in (128, 256, 384, 512):
        for min_n in (64, 128, 256, 512):
            ms, rp, rs = run(mode, batch, min_n)
            print(f"{mode} batch={batch:4d} wave_min_n={min_n:4d}: {ms:8.3f} ms  rounds {rp}+{rs}", flush=True)
