#!/usr/bin/env python
"""Sub-round sizing sweep of the single-tree planner (C2 20k iterations, C3 100k iterations): samples decided per
sub-round = max(nb_min, nb_mul x mean prefix + nb_add), participants likewise.  Device time of the grow call, best of 3."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from robosoccer_motion_planning_b200 import capi, scenes  # noqa: E402


def run(mode, env):
    for k, v in env.items():
        os.environ[k] = str(v)
    n_iter = 20000 if mode == "c2" else 100000
    G = capi.Planner(capacity=n_iter + 2, p_direct=0.1 if mode == "c2" else 0.0)
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
    return best, w


base = dict(KRRT_NB_MUL=2, KRRT_NB_ADD=8, KRRT_NB_MIN=16, KRRT_PN_MUL=3, KRRT_PN_ADD=16, KRRT_PN_MIN=48)
variants = [dict(), dict(KRRT_NB_MUL=3), dict(KRRT_NB_MUL=4, KRRT_PN_MUL=5), dict(KRRT_NB_ADD=16), dict(KRRT_NB_MUL=3, KRRT_NB_ADD=16, KRRT_PN_MUL=4),
            dict(KRRT_NB_MUL=1, KRRT_NB_ADD=16), dict(KRRT_PN_MIN=32), dict(KRRT_PN_MIN=96, KRRT_PN_MUL=6), dict(KRRT_NB_MIN=32),
            dict(KRRT_NB_MUL=3, KRRT_NB_ADD=4), dict(KRRT_NB_MUL=2, KRRT_NB_ADD=0, KRRT_NB_MIN=8)]
for mode in ("c2", "c3"):
    for v in variants:
        env = dict(base, **v)
        ms, w = run(mode, env)
        print(f"{mode} {v}: {ms:8.3f} ms rounds {w['rounds_parallel']}+{w['rounds_sequential']} ends {w['prefix_ended_by']}", flush=True)
