#!/usr/bin/env python
"""Where the C2 plan's time goes as the tree grows: 20 calls of 1000 iterations, device ms and sub-round statistics per call."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from robosoccer_motion_planning_b200 import capi, scenes  # noqa: E402

chunk = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
G = capi.Planner(capacity=20002, p_direct=0.1)
G.set_obstacles(scenes.dynamic_obstacles(), 1)
for rep in range(2):
    G.set_start(scenes.START)
    prev = {"rounds_parallel": 0, "rounds_sequential": 0, "expansions_parallel": 0}
    tot = 0.0
    for k in range(20000 // chunk):
        G.grow(scenes.GOAL, chunk)
        t = G.last_timing()
        w = G.wave_stats()
        kt = G.kernel_times()
        ms = t["total_ms"]
        tot += ms
        if rep == 1:
            dr = w["rounds_parallel"] - prev["rounds_parallel"]
            ds = w["rounds_sequential"] - prev["rounds_sequential"]
            de = w["expansions_parallel"] - prev["expansions_parallel"]
            sh = {a: round(v["ms"], 3) for a, v in kt.items() if v["ms"] > 0}
            print(f"iterations {k * chunk:6d}+  n {G.size:6d}  {ms:7.3f} ms  sub-rounds {dr:4d} (+{ds} in order)  {de / max(dr, 1):6.1f} per sub-round  {sh}")
        prev = w
    print("total", round(tot, 3), "ms")
