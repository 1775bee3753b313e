#!/usr/bin/env python
"""bench.py — headline benchmark of the kinodynamic RRT* hot path on B200 (contract: see the task statement).

Workload (N = 1): BASELINE.json configs[1] — one RRT* tree on the soccer field with 9 moving opponents, time-space
collision checks, direct-goal sampling p = 0.1, 20 000 iterations.  A "step" is one complete plan.
  value  RRT* expansions/s, device time of the grow call only (inputs resident in HBM)
  e2e    the same metric through the C ABI with HOST buffers: obstacle upload + start + grow + result read-back,
         CUDA events on the context's stream around the whole sequence
N > 1: a single tree does not shard ("replicas only"): every rank plans the same scene with the same
Philox stream, so the per-GPU work is exactly the N = 1 work (with one stream per rank the plans differ in length by
+-30 % and the max over ranks measures the unluckiest seed: 4.88 M expansions/s at N = 8, profiles/r1_bench_n8_s25.json);
value = all ranks' expansions / max-over-ranks time.  The batched-query sweep, which
DOES shard (contiguous blocks + one NCCL all-gather of results), is reported in the same line under "batched".

--impl reference: the CPU oracle (oracle/krrt_oracle.c — the reference's own RRT* sources are an absent submodule,
so nothing of it can be built: DESIGN.md §3) timed on the host cores on the same config.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

N_ITER = 20000
P_DIRECT = 0.1
BATCH_Q = 4096          # config 3/4: queries per GPU
BATCH_ITERS = 200
METRIC = "rrtstar_expansions_per_sec"
UNIT = "expansions/s"
WORKLOAD = "C2: soccer field, 9 moving opponents, time-space collision, direct-goal p=0.1, 20000 iterations, fp64"


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def fp64_peak():
    """FP64 vector peak in TFLOP/s: measured on this pool's B200 by tests/native/fp_peak.cu (profiles/r1_fp_peak.json)."""
    try:
        with open(os.path.join(ROOT, "profiles", "r1_fp_peak.json")) as f:
            return float(json.load(f)["fp64_fma_tflops"]), "measured (profiles/r1_fp_peak.json, FMA-chain microbenchmark)"
    except Exception:
        return 37.0, "fallback (nominal B200 FP64 vector)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
            except ValueError:
                continue
            for k, nm in enumerate(names):
                if f[2 + k].lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


_REAL_STDOUT = None


def quiet_stdout():
    """Everything a library prints on stdout (NCCL's version banner, ...) goes to stderr: stdout carries ONE JSON line."""
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)


def emit(line):
    data = (json.dumps(line) + "\n").encode()
    sys.stdout.flush()
    if _REAL_STDOUT is None:
        os.write(1, data)
    else:
        os.write(_REAL_STDOUT, data)


def c2_scene():
    from robosoccer_motion_planning_b200 import scenes
    return scenes.dynamic_obstacles(), scenes.START, scenes.GOAL


# ---------------------------------------------------------------------------------------------- reference arm
def run_reference(args, rank, world):
    """CPU oracle on the host cores: `world` replicas of the C2 plan, one thread each (a single tree is sequential)."""
    if rank != 0:
        return
    import oracle_binding as ob
    obs, start, goal = c2_scene()
    n = max(1, world)
    cores = min(n, os.cpu_count() or 1)
    cfg = ob.default_config(p_direct=P_DIRECT, capacity=N_ITER + 2)
    starts = np.tile(start, (n, 1))
    goals = np.tile(goal, (n, 1))
    obss = np.tile(obs, (n, 1, 1))
    for _ in range(args.warmup):
        ob.plan_batch(cfg, starts, goals, obss, 1, N_ITER, nthreads=cores)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        out = ob.plan_batch(cfg, starts, goals, obss, 1, N_ITER, nthreads=cores)
    dt = time.perf_counter() - t0
    value = n * N_ITER * args.steps / dt
    # batched queries on all host cores, bounded sample
    from robosoccer_motion_planning_b200 import scenes
    allc = os.cpu_count() or 1
    qs = 16 * allc
    bs, bg, bo = scenes.batch_queries(qs, 9, seed=0)
    cfgb = ob.default_config(p_direct=P_DIRECT)
    tb = time.perf_counter()
    ob.plan_batch(cfgb, bs, bg, bo, 1, BATCH_ITERS, nthreads=allc)
    tb = time.perf_counter() - tb
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "replicas": n, "note": "CPU oracle port; the reference's RRT* sources "
                   "(submodule alifahrri/motion_planning@3b3183e) are absent, nothing of it builds"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": f"{args.steps} full plans x {n} replica(s), {N_ITER} iterations each, kd-tree near search"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "batched": {"queries_per_sec": qs / tb, "expansions_per_sec": qs * BATCH_ITERS / tb, "queries": qs,
                    "iters": BATCH_ITERS, "cores": allc},
        "solved": int(out["solved"].sum()),
    }
    emit(line)


# ---------------------------------------------------------------------------------------------------- our arm
def run_ours(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    from robosoccer_motion_planning_b200 import capi, scenes, sweep
    assert torch.cuda.is_available(), "bench.py needs a GPU: libkrrt has no CPU fallback"
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device=dev)   # > 126 MB L2

    def flush_l2():
        flush_buf.zero_()
        torch.cuda.synchronize()

    obs, start, goal = c2_scene()
    G = capi.Planner(device=local_rank, capacity=N_ITER + 2, p_direct=P_DIRECT, query_id=0, batch=args.batch)

    def plan_device_timed():
        G.set_obstacles(obs, 1)
        G.set_start(start)
        r = G.grow(goal, N_ITER)
        return G.last_timing()["total_ms"], r

    def plan_e2e():
        G.timer_begin()
        G.set_obstacles(obs, 1)
        G.set_start(start)
        r = G.grow(goal, N_ITER)
        res = G.result(*r)
        return G.timer_end(), res

    for _ in range(max(args.warmup, 3)):
        plan_device_timed()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    # ---- value: device time of the grow call, K steps
    barrier()
    l0 = G.launch_count
    tot_ms = 0.0
    for _ in range(args.steps):
        flush_l2()
        ms, r_last = plan_device_timed()
        tot_ms += ms
    launches = G.launch_count - l0
    barrier()
    tot_ms = max_over_ranks(tot_ms)
    value = world * N_ITER * args.steps / (tot_ms * 1e-3)
    # ---- e2e: host buffers, copies inside the timed region
    barrier()
    e2e_ms = 0.0
    for _ in range(args.steps):
        flush_l2()
        ms, res = plan_e2e()
        e2e_ms += ms
    barrier()
    e2e_ms = max_over_ranks(e2e_ms)
    e2e_value = world * N_ITER * args.steps / (e2e_ms * 1e-3)
    clocks = sampler.stop() if rank == 0 else None
    ctl_bytes = 224
    path_len = int(res["path_len"])
    h2d = obs.nbytes + 64
    d2h = 4 * ctl_bytes + min(path_len, 32) * 40 + 4 + 16

    # ---- phase shares and the roofline.  The whole grow call is ONE persistent kernel (k_grow); its phases are timed
    # by the kernel itself (SM clocks of its last CTA between grid barriers), scaled to the call's CUDA-event time
    peak, peak_src = peaks()
    flush_l2()
    plan_device_timed()
    lt = G.last_timing()
    kt = G.kernel_times()
    wave = G.wave_stats()
    ksum = sum(v["ms"] for v in kt.values()) or 1.0
    shares = {k: round(v["ms"] / ksum, 4) for k, v in kt.items() if v["ms"] > 0}
    scan_bytes = lt["scan_node_samples"] * 32.0
    achieved = scan_bytes / (lt["total_ms"] * 1e-3) / 1e9 if lt["total_ms"] > 0 else 0.0
    scan_only = scan_bytes / (lt["scan_ms"] * 1e-3) / 1e9 if lt["scan_ms"] > 0 else 0.0
    try:
        with open(os.path.join(ROOT, "profiles", "r1_k_grow_traffic.json")) as f:
            tr = json.load(f)
        traffic = float(tr["dram_bytes_read"] + tr["dram_bytes_write"])   # one ncu --set full capture of this kernel
    except Exception:
        traffic = None
    roofline = {"bound": "hbm", "kernel": "k_grow", "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": scan_bytes,
                "avg_launch_us": 1e3 * lt["total_ms"],
                "share_of_step": 1.0,
                "phase_shares": shares,
                "scan_phase_GBps_algorithmic": scan_only,
                "scan_index": dict(G.scan_stats(), visited_bytes=32.0 * G.scan_stats()["visited_pairs"]),
                "note": "one persistent kernel per grow call; algorithmic bytes = 32 B x tree nodes x samples scanned "
                        "(brute-force equivalent; scan_index.visited_bytes is what the grid-indexed scan really read) "
                        "(SURVEY 8d: N x 32 B per expansion) over the kernel's whole duration.  A <= 20k-node tree is "
                        "640 KB and L2-resident: this config is bound by the latency chain of the commit sub-rounds "
                        "(decide / validate / apply, three grid barriers each), not by HBM; scan_phase_GBps is the "
                        "same byte count over the scan phase alone (see c3_100k / c3_1m for larger trees)"}
    # ---- batched queries (configs 3/4): one warp per query, sharded over ranks, NCCL all-gather of the results
    q_total = BATCH_Q * world
    lo, hi = sweep.shard_bounds(q_total, rank, world)
    bs, bg, bo = scenes.batch_queries(hi - lo, 9, seed=0, first=lo)
    Gb = capi.Planner(device=local_rank, p_direct=P_DIRECT)
    d_s = torch.from_numpy(bs).to(dev)
    d_g = torch.from_numpy(bg).to(dev)
    d_o = torch.from_numpy(bo).to(dev)
    item = capi.RESULT_DTYPE.itemsize
    d_out = torch.zeros((hi - lo) * item, dtype=torch.uint8, device=dev)
    d_all = torch.zeros(q_total * item, dtype=torch.uint8, device=dev)
    # a dedicated torch stream (non-zero handle): the kernel, the NCCL gather and the timing events all go on it
    tstream = torch.cuda.Stream(device=dev)
    assert tstream.cuda_stream != 0

    def batched_step():
        Gb.plan_batch_device(d_s.data_ptr(), d_g.data_ptr(), d_o.data_ptr(), hi - lo, 9, 1, BATCH_ITERS, lo,
                             d_out.data_ptr(), tstream.cuda_stream)
        if world > 1:
            dist.all_gather_into_tensor(d_all, d_out)

    torch.cuda.synchronize()
    with torch.cuda.stream(tstream):
        for _ in range(3):
            batched_step()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    bsteps = max(3, args.steps)
    b_ms = 0.0
    for _ in range(bsteps):
        flush_l2()
        with torch.cuda.stream(tstream):
            e0.record(tstream)
            batched_step()
            e1.record(tstream)
        e1.synchronize()
        b_ms += e0.elapsed_time(e1)
    barrier()
    b_ms = max_over_ranks(b_ms)
    # e2e batched: pinned host buffers through krrt_plan_batch
    p_s = torch.from_numpy(bs).pin_memory().numpy()
    p_g = torch.from_numpy(bg).pin_memory().numpy()
    p_o = torch.from_numpy(bo).pin_memory().numpy()
    p_out = torch.zeros((hi - lo) * item, dtype=torch.uint8).pin_memory().numpy().view(capi.RESULT_DTYPE)
    Gb.plan_batch(p_s, p_g, p_o, 1, BATCH_ITERS, query_base=lo, out=p_out)
    be_ms = 0.0
    for _ in range(bsteps):
        flush_l2()
        Gb.timer_begin()
        out_h = Gb.plan_batch(p_s, p_g, p_o, 1, BATCH_ITERS, query_base=lo, out=p_out)
        be_ms += Gb.timer_end()
    barrier()
    be_ms = max_over_ranks(be_ms)
    bst = Gb.batch_stats()          # counters of the last C4 launch (one rank's shard)
    # ---- config 4 (C5): ONE Monte-Carlo sweep of 65536 queries sharded over the ranks (strong scaling: the sweep is fixed,
    # the shard shrinks with N) + NCCL all-gather of all 65536 result records.  Scenes: the 4096 C4 scenes, each planned
    # with 16 different Philox streams (global query id = stream of the query)
    SWEEP_Q = 65536
    slo, shi = sweep.shard_bounds(SWEEP_Q, rank, world)
    base_s, base_g, base_o = scenes.batch_queries(BATCH_Q, 9, seed=0)
    sel = np.arange(slo, shi) % BATCH_Q
    d_ss = torch.from_numpy(np.ascontiguousarray(base_s[sel])).to(dev)
    d_sg = torch.from_numpy(np.ascontiguousarray(base_g[sel])).to(dev)
    d_so = torch.from_numpy(np.ascontiguousarray(base_o[sel])).to(dev)
    spad = sweep.max_shard(SWEEP_Q, world)
    d_sout = torch.zeros(spad * item, dtype=torch.uint8, device=dev)
    d_sall = torch.zeros(world * spad * item, dtype=torch.uint8, device=dev)

    def sweep_step():
        Gb.plan_batch_device(d_ss.data_ptr(), d_sg.data_ptr(), d_so.data_ptr(), shi - slo, 9, 1, BATCH_ITERS, slo,
                             d_sout.data_ptr(), tstream.cuda_stream)
        if world > 1:
            dist.all_gather_into_tensor(d_sall, d_sout)

    torch.cuda.synchronize()
    with torch.cuda.stream(tstream):
        sweep_step()
    barrier()
    s_ms = 0.0
    for _ in range(3):
        flush_l2()
        with torch.cuda.stream(tstream):
            e0.record(tstream)
            sweep_step()
            e1.record(tstream)
        e1.synchronize()
        s_ms += e0.elapsed_time(e1)
    barrier()
    s_ms = max_over_ranks(s_ms)
    sweep_line = {"workload": f"C5: one sweep of {SWEEP_Q} queries x {BATCH_ITERS} iterations sharded over {world} GPU(s) "
                              "(strong scaling), results all-gathered" + (" over NCCL" if world > 1 else " (1 GPU: no gather)"),
                  "queries_per_sec": SWEEP_Q * 3 / (s_ms * 1e-3), "ms_per_sweep": s_ms / 3,
                  "gather_bytes": int(SWEEP_Q * item) if world > 1 else 0,
                  "solved_frac": float((d_sout.cpu().numpy()[: (shi - slo) * item].view(capi.RESULT_DTYPE)["solved"] != 0).mean())}
    b_flop = 134.0 * bst["steer_calls"] + 1180.0 * bst["collide_calls"]      # SURVEY 8(d) accounting, one rank's shard
    f64_peak, f64_src = fp64_peak()
    b_tflops = b_flop * bsteps / (b_ms * 1e-3) / 1e12
    batched = {
        "roofline": {"bound": "fp64 pipe / latency", "kernel": "k_plan_batch2", "achieved": b_tflops, "peak": f64_peak,
                     "unit": "TFLOP/s", "frac": b_tflops / f64_peak, "peak_source": f64_src,
                     "algorithmic_flop_per_launch": b_flop,
                     "note": "134 flop per steering pair + 1180 flop per edge test (SURVEY 8d), counted from the "
                             "kernel's own steer_calls / collide_calls; trees live in shared memory, HBM traffic is the "
                             "inputs and the 1304-byte result records only"},
        "workload": f"{BATCH_Q} queries/GPU x {BATCH_ITERS} iterations, 9 moving obstacles, p_direct=0.1 (C4; sharded sweep C5 for N>1)",
        "queries_per_sec": q_total * bsteps / (b_ms * 1e-3),
        "expansions_per_sec": q_total * BATCH_ITERS * bsteps / (b_ms * 1e-3),
        "ms_per_step": b_ms / bsteps,
        "e2e_queries_per_sec": q_total * bsteps / (be_ms * 1e-3),
        "h2d_bytes_per_step": int(bs.nbytes + bg.nbytes + bo.nbytes), "d2h_bytes_per_step": int((hi - lo) * item),
        "gather": "nccl all_gather_into_tensor of result records" if world > 1 else "none (1 GPU)",
        "solved_frac": float(out_h["solved"].mean()),
    }

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": tot_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "replicas": world, "window": int(G.cfg.batch),
                   "l2": "flushed between timed steps (256 MiB write)", "parallelism": "replicas only (single tree; every rank grows the same plan, same seed); "
                   "batched queries shard by contiguous blocks"},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                "ms_per_step": e2e_ms / args.steps},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": roofline,
        "batched": batched,
        "sweep_c5": sweep_line,
        "tree_nodes": int(res["n_nodes"]), "solved": bool(res["solved"]), "cost": float(res["cost"]),
        "rounds": wave,
    }

    if rank == 0 and world == 1:
        import oracle_binding as ob
        O = ob.OraclePlanner(capacity=N_ITER + 2, p_direct=P_DIRECT, query_id=0)
        reps = 3
        t0 = time.perf_counter()
        for _ in range(reps):
            O.set_obstacles(obs, 1)
            O.set_start(start)
            ro = O.grow(goal, N_ITER)
        dt = time.perf_counter() - t0
        line["cpu_baseline"] = {"value": reps * N_ITER / dt, "unit": UNIT, "cores": 1, "kind": "port",
                                "sample": f"{reps} full C2 plans ({N_ITER} iterations each), kd-tree near search, 1 thread "
                                          "(a single RRT* tree is sequential)"}
        same = all(np.array_equal(a, b) for a, b in zip(G.tree(), O.tree())) and r_last == ro
        line["parity_vs_oracle"] = "bit-exact" if same else "MISMATCH"
        # batched CPU baseline on all host cores, bounded sample
        allc = os.cpu_count() or 1
        qs = min(hi - lo, 16 * allc)
        tb = time.perf_counter()
        refb = ob.plan_batch(ob.default_config(p_direct=P_DIRECT), bs[:qs], bg[:qs], bo[:qs], 1, BATCH_ITERS,
                             query_base=lo, nthreads=allc)
        tb = time.perf_counter() - tb
        line["batched"]["cpu_baseline"] = {"queries_per_sec": qs / tb, "cores": allc, "kind": "port",
                                           "sample": f"{qs} of the {BATCH_Q} queries"}
        line["batched"]["parity_vs_oracle"] = ("bit-exact" if np.array_equal(out_h[:qs].view(np.uint8), refb.view(np.uint8))
                                               else "MISMATCH")
        # config 0 (rrtstar_benchmark, C1): static obstacles, fixed seed, 10k iterations — GPU and the CPU port, same plan
        G1 = capi.Planner(device=local_rank, capacity=10002, batch=args.batch)
        O1 = ob.OraclePlanner(capacity=10002, query_id=0)
        for P_ in (G1, O1):
            P_.set_obstacles(scenes.static_obstacles(), 0)
            P_.set_start(start)
        G1.grow(goal, 10000)                       # warm-up
        G1.set_start(start)
        flush_l2()
        r1 = G1.grow(goal, 10000)
        ms1 = G1.last_timing()["total_ms"]
        t0 = time.perf_counter()
        ro1 = O1.grow(goal, 10000)
        dt1 = time.perf_counter() - t0
        same1 = r1 == ro1 and all(np.array_equal(a, b) for a, b in zip(G1.tree(), O1.tree()))
        line["c1_rrtstar_benchmark"] = {
            "workload": "C1: 9 static obstacles, fixed seed, 10000 iterations, p_direct = 0, goal connected (O8)",
            "cost": float(G1.result(*r1)["cost"]) if r1[0] else None,
            "expansions_per_sec": 10000 / (ms1 * 1e-3), "ms": ms1, "nodes": G1.size, "solved": bool(r1[0]),
            "cpu_baseline": {"value": 10000 / dt1, "unit": UNIT, "cores": 1, "kind": "port", "sample": "the full plan, once"},
            "parity_vs_oracle": "bit-exact" if same1 else "MISMATCH"}
        G1.close()
        # the README's curves: computation time and solution cost against node size on the C1 scene,
        # 1024 seeds per size through the batched planner; CPU port on all host cores beside it (context: the README's
        # numbers are an i3 at 2.0 GHz, scene and trial count not stated)
        curve = []
        qn = 1024
        c_s = np.tile(start, (qn, 1))
        c_g = np.tile(goal, (qn, 1))
        c_o = np.tile(scenes.static_obstacles(), (qn, 1, 1))
        for p_dir in (0.0, 0.1):
            Gc = capi.Planner(device=local_rank, p_direct=p_dir)
            for nodes in (10, 25, 50, 75, 100, 150, 200):
                Gc.plan_batch(c_s, c_g, c_o, 0, nodes)
                Gc.timer_begin()
                oc = Gc.plan_batch(c_s, c_g, c_o, 0, nodes)
                msb = Gc.timer_end()
                tcp = time.perf_counter()
                rc_ = ob.plan_batch(ob.default_config(p_direct=p_dir), c_s[:4 * allc], c_g[:4 * allc], c_o[:4 * allc], 0,
                                    nodes, nthreads=allc)
                tcp = time.perf_counter() - tcp
                sv = oc["solved"] != 0
                curve.append({"p_direct": p_dir, "node_size": nodes, "gpu_ms_per_plan_batched": msb / qn,
                              "cpu_port_ms_per_plan_1core": 1e3 * tcp / 4.0,   # 4 plans per thread, all threads busy
                              "solved_frac": float(sv.mean()),
                              "mean_cost": float(oc["cost"][sv].mean()) if sv.any() else None,
                              "parity": bool(np.array_equal(oc[:4 * allc].view(np.uint8), rc_.view(np.uint8)))})
            Gc.close()
        line["readme_curve"] = {"scene": "C1 start/goal, 9 static obstacles, goal connected to the cheapest collision-free node (O8), 1024 seeds per point (query_id = seed)",
                                "points": curve}
        # config 3: single large tree, 100k iterations, static obstacles (+ the 1M-iteration tree with --c3-1m)
        if not args.skip_c3:
            for label, iters in (("c3_100k", 100000),) + ((("c3_1m", 1000000),) if args.c3_1m else ()):
                G3 = capi.Planner(device=local_rank, capacity=iters + 2, batch=args.batch)
                G3.set_obstacles(scenes.static_obstacles(), 0)
                G3.set_start(start)
                G3.grow(None, min(iters, 20000))     # warm-up
                G3.set_start(start)
                flush_l2()
                G3.grow(None, iters)
                t3 = G3.last_timing()
                t_plain = t3["total_ms"]
                k3 = G3.kernel_times()
                ks = sum(v["ms"] for v in k3.values()) or 1.0
                line[label] = {"expansions_per_sec": iters / (t_plain * 1e-3), "ms": t_plain, "nodes": G3.size,
                               "kernel_GBps_algorithmic": t3["scan_node_samples"] * 32 / (t_plain * 1e-3) / 1e9,
                               "kernel_frac_of_hbm_peak": t3["scan_node_samples"] * 32 / (t_plain * 1e-3) / 1e9 / peak,
                               "scan_phase_GBps_algorithmic": t3["scan_node_samples"] * 32 / (t3["scan_ms"] * 1e-3) / 1e9,
                               "phase_shares": {k: round(v["ms"] / ks, 4) for k, v in k3.items() if v["ms"] > 0},
                               "scan_index": dict(G3.scan_stats(), visited_bytes=32.0 * G3.scan_stats()["visited_pairs"]),
                               "rounds": G3.wave_stats()}
                if label == "c3_100k":   # the same plan on one host core (kd-tree near search), once
                    O3 = ob.OraclePlanner(capacity=iters + 2, query_id=0)
                    O3.set_obstacles(scenes.static_obstacles(), 0)
                    O3.set_start(start)
                    t0 = time.perf_counter()
                    O3.grow(None, iters)
                    dt3 = time.perf_counter() - t0
                    same3 = all(np.array_equal(a, b) for a, b in zip(G3.tree(), O3.tree()))
                    line[label]["cpu_baseline"] = {"value": iters / dt3, "unit": UNIT, "cores": 1, "kind": "port",
                                                   "sample": f"the full {iters}-iteration plan, once, kd-tree near search"}
                    line[label]["gpu_over_cpu"] = line[label]["expansions_per_sec"] / (iters / dt3)
                    line[label]["parity_vs_oracle"] = "bit-exact" if same3 else "MISMATCH"
                G3.close()
    if rank == 0:
        emit(line)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=0, help="window size: samples scanned + steered at once (0 = library default)")
    ap.add_argument("--skip-c3", action="store_true")
    ap.add_argument("--c3-1m", action="store_true", help="also grow the 1M-iteration single tree (a few seconds)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    quiet_stdout()
    if args.impl == "reference":
        run_reference(args, rank, max(world, args.gpus if world == 1 else world))
    else:
        run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
