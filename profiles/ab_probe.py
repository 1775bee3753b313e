#!/usr/bin/env python
"""A/B of k_grow switches (environment variables given as NAME=VALUE arguments, each argument one variant; '-' = defaults):
C2 20k iterations and C3 100k iterations, device time of the grow call (best of 3), sub-round counts, hash of the tree."""
import hashlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from robosoccer_motion_planning_b200 import capi  # noqa: E402

if os.environ.get("KRRT_LIB"):  # a development build of the library (profiles/_*/libkrrt.so)
    capi.LIB_PATH = os.path.join(ROOT, os.environ["KRRT_LIB"])
    capi.build_library = lambda *a, **k: capi.LIB_PATH
from robosoccer_motion_planning_b200 import scenes  # noqa: E402

MODES = os.environ.get("AB_MODES", "c2,c3").split(",")


def run(mode):
    n_iter = {"c2": 20000, "c3": 100000, "c3m": 1000000}[mode]
    G = capi.Planner(capacity=n_iter + 2, p_direct=0.1 if mode == "c2" else 0.0)
    if mode == "c2":
        G.set_obstacles(scenes.dynamic_obstacles(), 1)
        goal = scenes.GOAL
    else:
        G.set_obstacles(scenes.static_obstacles(), 0)
        goal = None
    best = 1e9
    for _ in range(1 if mode == "c3m" else 3):
        G.set_start(scenes.START)
        G.grow(goal, n_iter)
        best = min(best, G.last_timing()["total_ms"])
    w = G.wave_stats()
    tc = G.tail_commits()
    t = G.tree()
    h = hashlib.sha256(b"".join(a.tobytes() for a in t)).hexdigest()[:12]
    G.close()
    return best, w, tc, h


for arg in sys.argv[1:] or ["-"]:
    kv = [] if arg == "-" else [a.split("=") for a in arg.split(",")]
    for k, v in kv:
        os.environ[k] = v
    for mode in MODES:
        ms, w, tc, h = run(mode)
        n = {"c2": 20000, "c3": 100000, "c3m": 1000000}[mode]
        print(f"{arg:28s} {mode}: {ms:8.3f} ms ({n / ms / 1e3:.3f} M/s) rounds {w['rounds_parallel']}+{w['rounds_sequential']} (tail {tc}) "
              f"ends {w['prefix_ended_by']} tree {h}", flush=True)
    for k, v in kv:
        os.environ.pop(k, None)
