#!/usr/bin/env python
"""Round 2 probe of k_grow: C2 (20k iterations), C3 (140k iterations = 100k+ nodes), C3 (1M iterations); best of `reps` timings,
sub-round statistics and phase shares; `--parity` compares C2 and C3-140k with the CPU oracle bit for bit.
Environment switches (KRRT_GROW_PER_SM, KRRT_NB_*, KRRT_PN_*, ...) are read by the library: run once per setting."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402
from robosoccer_motion_planning_b200 import capi, scenes  # noqa: E402


F32 = "--f32" in sys.argv


def run(label, iters, obs, dynamic, goal, p_direct, reps, parity):
    if F32:
        return run_f32(label, iters, obs, dynamic, goal, p_direct, reps, parity)
    G = capi.Planner(capacity=iters + 2, p_direct=p_direct)
    G.set_obstacles(obs, dynamic)
    G.set_start(scenes.START)
    G.grow(goal, min(iters, 20000))
    best = None
    for _ in range(reps):
        G.set_start(scenes.START)
        G.flush_l2()
        G.grow(goal, iters)
        t = G.last_timing()["total_ms"]
        best = t if best is None else min(best, t)
    w = G.wave_stats()
    k = G.kernel_times()
    ks = sum(v["ms"] for v in k.values()) or 1.0
    subs = w["rounds_parallel"] + w["rounds_sequential"]
    print(f"{label}: {best:9.3f} ms  {iters / best / 1e3:7.3f} M exp/s  nodes {G.size}  sub-rounds {subs} "
          f"({w['expansions_parallel'] / max(1, w['rounds_parallel']):.1f} per parallel one)  ended_by {w['prefix_ended_by']}")
    print("   shares", {a: round(v["ms"] / ks, 3) for a, v in k.items() if v["ms"] > 0})
    if parity:
        import oracle_binding as ob
        O = ob.OraclePlanner(capacity=iters + 2, p_direct=p_direct)
        O.set_obstacles(obs, dynamic)
        O.set_start(scenes.START)
        t0 = time.perf_counter()
        O.grow(goal, iters)
        dt = time.perf_counter() - t0
        same = all(np.array_equal(a, b) for a, b in zip(G.tree(), O.tree())) and G.stats() == O.stats()
        print(f"   oracle {iters / dt / 1e3:.1f} k exp/s -> x{(iters / (best * 1e-3)) / (iters / dt):.1f}  parity {'bit-exact' if same else 'MISMATCH'}")
    G.close()


def run_f32(label, iters, obs, dynamic, goal, p_direct, reps, parity):
    """the same plan on a KRRT_F32 context (wall clock around the call: the diagnostics are binary64 only)"""
    G = capi.Planner(capacity=iters + 2, p_direct=p_direct, precision=capi.KRRT_F32)
    G.set_obstacles(obs, dynamic)
    G.set_start(scenes.START)
    G.grow(goal, min(iters, 20000))
    best = None
    for _ in range(reps):
        G.set_start(scenes.START)
        t0 = time.perf_counter()
        G.grow(goal, iters)
        t = 1e3 * (time.perf_counter() - t0)
        best = t if best is None else min(best, t)
    print(f"{label} (f32): {best:9.3f} ms wall  {iters / best / 1e3:7.3f} M exp/s  nodes {G.size}")
    if parity:
        import oracle_f32_binding as of
        O = of.OraclePlannerF32(capacity=iters + 2, p_direct=p_direct)
        O.set_obstacles(obs, dynamic)
        O.set_start(scenes.START)
        O.grow(goal, iters)
        same = all(np.array_equal(a, b) for a, b in zip(G.tree(), O.tree())) and G.stats() == O.stats()
        print(f"   parity with the float checker: {'bit-exact' if same else 'MISMATCH'}")
    G.close()


def main():
    parity = "--parity" in sys.argv
    reps = 3
    which = [a for a in sys.argv[1:] if not a.startswith("--")] or ["c2", "c3", "c3m"]
    print("env", {k: v for k, v in os.environ.items() if k.startswith("KRRT_")})
    if "c2" in which:
        run("C2  20k", 20000, scenes.dynamic_obstacles(), 1, scenes.GOAL, 0.1, reps, parity)
    if "c1" in which:
        run("C1  10k", 10000, scenes.static_obstacles(), 0, scenes.GOAL, 0.0, reps, parity)
    if "c3" in which:
        run("C3 140k", 140000, scenes.static_obstacles(), 0, None, 0.0, reps, parity)
    if "c3m" in which:
        run("C3   1M", 1000000, scenes.static_obstacles(), 0, None, 0.0, 2, False)


if __name__ == "__main__":
    main()
