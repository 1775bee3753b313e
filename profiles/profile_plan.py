#!/usr/bin/env python
"""One single-tree plan (or one batched plan) through the C ABI, for `ncu` launch lists and captures.

  python profiles/profile_plan.py c2 [n_iter]        C2 scene, 9 moving opponents, p_direct = 0.1
  python profiles/profile_plan.py c3 [n_iter]        C3 scene, 9 static obstacles, no goal
  python profiles/profile_plan.py batch [q] [iters]  C4: q independent queries
Numbers printed under a profiler are never bench values (bench.py is the only source of those)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from robosoccer_motion_planning_b200 import capi, scenes  # noqa: E402


def main():
    mode = sys.argv[1] if len(sys.argv) > 1 else "c2"
    if mode == "batch":
        q = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
        iters = int(sys.argv[3]) if len(sys.argv) > 3 else 200
        s, g, o = scenes.batch_queries(q, 9, seed=0)
        G = capi.Planner(p_direct=0.1)
        out = G.plan_batch(s, g, o, 1, iters)
        print("batch solved", int(out["solved"].sum()), "/", q, "launches", G.launch_count)
        return
    n_iter = int(sys.argv[2]) if len(sys.argv) > 2 else 20000
    if mode == "c2":
        G = capi.Planner(capacity=n_iter + 2, p_direct=0.1)
        G.set_obstacles(scenes.dynamic_obstacles(), 1)
        goal = scenes.GOAL
    else:
        G = capi.Planner(capacity=n_iter + 2)
        G.set_obstacles(scenes.static_obstacles(), 0)
        goal = None
    G.set_start(scenes.START)
    r = G.grow(goal, n_iter)
    print(mode, "n", G.size, "result", r, "ms", G.last_timing()["total_ms"], "launches", G.launch_count)
    print("wave", G.wave_stats())
    subs = max(1, sum(G.wave_stats()[k] for k in ("rounds_parallel", "rounds_sequential")))
    print("dbg cycles per sub-round", [int(v) // subs for v in G.debug_clocks()])
    print("stats", G.stats())
    if True:
        for k, v in G.kernel_times().items():
            print(f"  {k:20s} {v['ms']:9.3f} ms  {v['launches']:6d} launches  {1e3 * v['ms'] / max(1, v['launches']):8.2f} us avg")


if __name__ == "__main__":
    main()
