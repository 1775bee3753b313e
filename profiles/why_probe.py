#!/usr/bin/env python
"""What ends the validated prefixes of k_grow's sub-rounds (diagnostic build with -DKRRT_WHY in profiles/_why/).
A failing sample may show several reasons at once."""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from robosoccer_motion_planning_b200 import capi  # noqa: E402

capi.LIB_PATH = os.path.join(ROOT, "profiles", "_why", "libkrrt.so")
capi.build_library = lambda *a, **k: capi.LIB_PATH
from robosoccer_motion_planning_b200 import scenes  # noqa: E402

NAMES = ["radius side", "reads a node an earlier sample writes", "newcomer: no free parent", "newcomer: duplicate",
         "newcomer would be tried as parent", "newcomer is a rewire candidate", "rewires a node an earlier sample read",
         "propagates into a node an earlier sample read", "out REJECT", "out DUP", "out NOFREE", "out INSERT",
         "out COMPLEX", "out CONFLICT/NEAREST"]
for mode, n_iter in (("c2", 20000), ("c3", 100000)):
    G = capi.Planner(capacity=n_iter + 2, p_direct=0.1 if mode == "c2" else 0.0)
    G.set_obstacles(scenes.dynamic_obstacles() if mode == "c2" else scenes.static_obstacles(), 1 if mode == "c2" else 0)
    G.set_start(scenes.START)
    G.grow(scenes.GOAL if mode == "c2" else None, n_iter)
    d = G.debug_clocks()
    print(mode, G.wave_stats())
    for k, nm in enumerate(NAMES):
        if d[k]:
            print(f"  {nm:50s} {int(d[k])}")
    print(f"  {'(of the reads: cost changed by propagation)':50s} {int(d[15])}")
