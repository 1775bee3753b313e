#!/usr/bin/env python
"""Where CTA 0 spends the sub-rounds of k_grow (development build with -DKRRT_DEBUG_CLK in profiles/_dbg/): SM-clock sums
between the stamps inside decide / apply, per sub-round."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from robosoccer_motion_planning_b200 import capi  # noqa: E402

capi.LIB_PATH = os.path.join(ROOT, "profiles", "_dbg", "libkrrt.so")
capi.build_library = lambda *a, **k: capi.LIB_PATH
from robosoccer_motion_planning_b200 import scenes  # noqa: E402

NAMES = {1: "decide: candidate slots of the sample (round trip 1)", 13: "decide: map + node data issued", 0: "decide: count of earlier inserts (sync)",
         2: "decide: near flags, K (sync)", 3: "decide: forward edges collision", 4: "decide: parent arg-min", 5: "decide: candidates + hash (sync)",
         6: "decide: rewire edges collision", 7: "decide: children of the rewired nodes", 8: "decide: level walk", 9: "decide: outcome",
         12: "grid barriers (3 per sub-round) + validate", 10: "apply: insert + link", 11: "apply: detach, rewire, levels"}
for mode, n_iter in (("c2", 20000), ("c3", 100000)):
    G = capi.Planner(capacity=n_iter + 2, p_direct=0.1 if mode == "c2" else 0.0)
    G.set_obstacles(scenes.dynamic_obstacles() if mode == "c2" else scenes.static_obstacles(), 1 if mode == "c2" else 0)
    G.set_start(scenes.START)
    G.grow(scenes.GOAL if mode == "c2" else None, n_iter)
    ms = G.last_timing()["total_ms"]
    d = G.debug_clocks()
    w = G.wave_stats()
    rounds = w["rounds_parallel"] + w["rounds_sequential"] - G.tail_commits()
    tot = float(sum(int(d[k]) for k in NAMES))
    print(mode, f"{ms:.3f} ms", rounds, "sub-rounds; mean decide clocks over samples", int(d[14]) / max(1, int(d[15])))
    for k, nm in NAMES.items():
        print(f"  {nm:60s} {int(d[k]) / rounds:9.0f} clk/sub-round  {100.0 * int(d[k]) / tot:5.1f} %")
