#!/usr/bin/env python
"""Decide time of every decided sample of a C3 plan (140k iterations) (development build, profiles/_dbg/) against what the sample had to do."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from robosoccer_motion_planning_b200 import capi  # noqa: E402

capi.LIB_PATH = os.path.join(ROOT, "profiles", "_dbg", "libkrrt.so")
capi.build_library = lambda *a, **k: capi.LIB_PATH
from robosoccer_motion_planning_b200 import scenes  # noqa: E402

path = os.path.join(ROOT, "gpurun_out", "dlog.bin")
os.environ["KRRT_DEBUG_LOG"] = path
G = capi.Planner(capacity=140002)
G.set_obstacles(scenes.static_obstacles(), 0)
G.set_start(scenes.START)
G.grow(None, 140000)
G.debug_clocks()
L = np.fromfile(path, dtype=np.int32).reshape(-1, 8)
os.remove(path)
cta, clk, out, K, cnt, nrew, nwl, nlvl = L.T
print("decided samples", len(L), "clk percentiles 50/90/99/max", np.percentile(clk, [50, 90, 99, 100]))
for name, col, edges in (("nlvl", nlvl, [0, 1, 2, 3, 4, 6, 8, 12, 100]), ("nrew", nrew, [0, 1, 2, 3, 5, 8, 100]),
                         ("nwl", nwl, [0, 1, 4, 16, 64, 256, 2000]), ("cnt", cnt, [0, 16, 32, 64, 128, 257]), ("K", K, [0, 8, 16, 32, 64, 257]),
                         ("outcome", out, [0, 1, 2, 3, 4, 5, 6, 7, 8])):
    print(name)
    for lo, hi in zip(edges[:-1], edges[1:]):
        m = (col >= lo) & (col < hi)
        if m.any():
            print(f"  [{lo:4d},{hi:4d})  n {m.sum():6d}  mean clk {clk[m].mean():8.0f}  p90 {np.percentile(clk[m], 90):8.0f}  max {clk[m].max():8d}")
# what the slowest sample of a sub-round looks like: the log is appended in completion order, so consecutive runs of ~100
# entries belong to one sub-round; take windows of 100 entries and report the max
W = 100
nwin = len(L) // W
mx = clk[:nwin * W].reshape(nwin, W)
arg = mx.argmax(axis=1)
rows = L[:nwin * W].reshape(nwin, W, 8)[np.arange(nwin), arg]
print("per window of 100 decided samples: median of max decide clk", np.median(mx.max(axis=1)), "median of median", np.median(np.median(mx, axis=1)))
print("the slowest sample of a window: nlvl percentiles 50/90", np.percentile(rows[:, 7], [50, 90]), "nwl 50/90", np.percentile(rows[:, 6], [50, 90]),
      "nrew 50/90", np.percentile(rows[:, 5], [50, 90]), "cnt 50/90", np.percentile(rows[:, 4], [50, 90]))
slow_has_walk = (rows[:, 7] > 0).mean()
print("slowest sample walks a subtree in", slow_has_walk, "of the windows")
