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
DOES shard (contiguous blocks + NCCL gather of the result records inside libkrrt_mgpu.so), is reported in the same line
under "batched" (C4, weak scaling) and "sweep_c5" (C5, strong scaling, with the sha256 of the gathered records).
No torch anywhere on this path: processes are launched by torch.distributed.run, everything else goes through the C ABIs.

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
                                          "--format=csv,noheader,nounits", "-lms", "20"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append((time.monotonic(), ln.strip()))

    def mark(self):
        """start of the timed region (the sampler itself is started before the warm-up: nvidia-smi needs ~0.1 s to come up)"""
        self.t0 = time.monotonic()

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        t1 = time.monotonic()
        t0 = getattr(self, "t0", 0.0)
        time.sleep(0.03)    # (a sample that was being taken when the region ended)
        raw = list(self.lines)    # (the reader thread keeps appending until the process is terminated below)
        inside = [ln for (ts, ln) in raw if t0 <= ts <= t1 + 0.03]
        lines = inside if inside else [ln for (_, ln) in raw[-3:]]
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in lines:
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
        "config": bench_config(n),
        "note": "CPU oracle port; the reference's RRT* sources (submodule alifahrri/motion_planning@3b3183e) are absent, "
                "nothing of it builds",
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": f"{args.steps} full plans x {n} replica(s), {N_ITER} iterations each, kd-tree near search"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "batched": {"queries_per_sec": qs / tb, "expansions_per_sec": qs * BATCH_ITERS / tb, "queries": qs,
                    "iters": BATCH_ITERS, "cores": allc},
        "solved": int(out["solved"].sum()),
    }
    emit(line)


def bench_config(replicas):
    """`config` of BOTH arms, key for key (the driver compares them): what is planned, how often, and how the timed steps
    are kept honest."""
    return {"workload": WORKLOAD, "replicas": int(replicas),
            "l2": "GPU arm: flushed between timed steps (write of 2 x L2); CPU arm: not applicable",
            "parallelism": "replicas only (single tree; every rank grows the same plan, same seed); "
                           "batched queries shard by contiguous blocks"}


# ---------------------------------------------------------------------------------------------------- our arm
def profile_json(name):
    """a number taken from an ncu capture / microbenchmark committed under profiles/ (None when the file is absent)"""
    try:
        with open(os.path.join(ROOT, "profiles", name)) as f:
            return json.load(f)
    except Exception:
        return None


def run_ours(args, rank, world, local_rank):
    """No torch on this path: device memory, streams, NCCL and timing all live behind the C ABIs of libkrrt.so
    (include/krrt.h) and libkrrt_mgpu.so (include/krrt_mgpu.h)."""
    import hashlib
    from robosoccer_motion_planning_b200 import capi, scenes, sweep
    assert capi.lib().krrt_device_count() > local_rank, "bench.py needs a GPU: libkrrt has no CPU fallback"
    # the batched-query handle also carries the job's NCCL communicator (barrier / max over ranks of the headline)
    S = sweep.Sweep.from_env(p_direct=P_DIRECT)
    barrier = S.barrier
    max_over_ranks = S.max_over_ranks

    obs, start, goal = c2_scene()
    G = capi.Planner(device=local_rank, capacity=N_ITER + 2, p_direct=P_DIRECT, query_id=0, batch=args.batch)
    flush_l2 = G.flush_l2          # overwrites 2 x L2 on the context's stream and waits

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

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    for _ in range(max(args.warmup, 3)):
        plan_device_timed()
    # ---- value: device time of the grow call, K steps
    barrier()
    sampler.mark()
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
    ncu_k = profile_json("r2_k_grow_ncu.json") or profile_json("r1_k_grow_traffic.json") or {}
    traffic = (float(ncu_k["dram_bytes_read"] + ncu_k["dram_bytes_write"])
               if "dram_bytes_read" in ncu_k else None)   # one ncu --set full capture of this kernel
    sst = G.scan_stats()
    sub_rounds = wave["rounds_parallel"] + wave["rounds_sequential"]
    lat = profile_json("r2_latency_model.json") or {}
    # what really bounds the kernel: the chain of sub-rounds.  floor = sub-rounds x (grid barriers x their
    # measured cost + the decide / validate / apply critical path of an undisturbed CTA) + windows x (scan + steer of one)
    commit_ms = sum(kt[k]["ms"] for k in ("decide", "validate", "apply"))
    latency = {"sub_rounds": int(sub_rounds), "windows": int(kt["scan"]["launches"]),
               "us_per_sub_round": 1e3 * commit_ms / max(sub_rounds, 1),
               "expansions_per_sub_round": N_ITER / max(sub_rounds, 1)}
    if lat:
        floor_us = lat["barriers_per_sub_round"] * lat["barrier_us"] + lat["cta0_critical_path_us"]
        latency.update({"barrier_us": lat["barrier_us"], "barriers_per_sub_round": lat["barriers_per_sub_round"],
                        "cta0_critical_path_us": lat["cta0_critical_path_us"],
                        "latency_floor_ms": 1e-3 * floor_us * sub_rounds,
                        "frac_of_latency_floor": (1e-3 * floor_us * sub_rounds) / commit_ms if commit_ms > 0 else None,
                        "source": lat.get("source")})
    roofline = {"bound": "hbm", "kernel": "k_grow", "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": scan_bytes,
                "avg_launch_us": 1e3 * lt["total_ms"],
                "share_of_step": 1.0,
                "phase_shares": shares,
                "scan_phase_brute_force_equivalent_GBps": scan_only,
                "scan_index": dict(sst, visited_bytes=32.0 * sst["visited_pairs"]),
                "visited_bytes": 32.0 * sst["visited_pairs"],
                "l2_bytes": ncu_k.get("lts_t_bytes"),
                "scan_phase_GBps_visited": (32.0 * sst["visited_pairs"] / (lt["scan_ms"] * 1e-3) / 1e9
                                            if lt["scan_ms"] > 0 else None),
                "lts_pct_of_peak": ncu_k.get("lts_pct_of_peak"), "issue_active_pct": ncu_k.get("issue_active_pct"),
                "stall_barrier_per_inst": ncu_k.get("stall_barrier_per_inst"),
                "stall_cycles_per_inst": ncu_k.get("stall_cycles_per_inst"),
                "l2_peak_GBps": (profile_json("r2_l2_peak.json") or {}).get("l2_read_GBps"),
                "fp64_pipe_pct": ncu_k.get("fp64_pipe_pct"),
                "latency_model": latency,
                "note": "one persistent kernel per grow call; algorithmic bytes = 32 B x tree nodes x samples scanned "
                        "(brute-force equivalent, SURVEY 8d: N x 32 B per expansion) over the kernel's whole duration; "
                        "visited_bytes is what the grid-indexed scan really read (L2-resident: `traffic` is the DRAM "
                        "bytes of one ncu capture).  A <= 20k-node tree is 640 KB: this config is bound by the latency "
                        "chain of the commit sub-rounds (latency_model), not by HBM"}

    # ---- batched queries (config 3, C4): BATCH_Q queries per GPU (weak scaling), blocks + chunked NCCL gather inside
    # libkrrt_mgpu (krrt_sweep_load / run: inputs resident in HBM, device time of plan + gather, max over ranks)
    q_total = BATCH_Q * world
    bs, bg, bo = scenes.batch_queries(q_total, 9, seed=0)
    S.load(bs, bg, bo)
    for _ in range(3):
        S.run(1, BATCH_ITERS)
    barrier()
    bsteps = max(3, args.steps)
    l0b = S.launch_count
    b_ms = 0.0
    for _ in range(bsteps):
        flush_l2()
        b_ms += S.run(1, BATCH_ITERS)
    b_launches = S.launch_count - l0b
    barrier()
    b_ms = max_over_ranks(b_ms)
    bst = S.stats()                 # counters of the last C4 run, all ranks
    # e2e batched: page-locked host buffers through krrt_sweep_multi_gpu (H2D of the block, plan, gather, D2H of all records)
    p_s, p_g, p_o = capi.PinnedArray.like(bs), capi.PinnedArray.like(bg), capi.PinnedArray.like(bo)
    p_out = capi.PinnedArray((q_total,), capi.RESULT_DTYPE)
    S.sweep(p_s.array, p_g.array, p_o.array, 1, BATCH_ITERS, out=p_out.array)
    barrier()
    be_ms = 0.0
    for _ in range(bsteps):
        flush_l2()
        out_h = S.sweep(p_s.array, p_g.array, p_o.array, 1, BATCH_ITERS, out=p_out.array)
        be_ms += S.last_timing()["e2e_ms"]
    barrier()
    be_ms = max_over_ranks(be_ms)
    out_h = out_h.copy()
    lo, hi = sweep.shard_bounds(q_total, rank, world)

    # ---- config 4 (C5): ONE Monte-Carlo sweep of 65536 queries sharded over the ranks (strong scaling: the sweep is fixed,
    # the block shrinks with N) + gather of all 65536 result records.  Scenes: the 4096 C4 scenes, each planned with 16
    # different Philox streams (global query id = stream of the query)
    SWEEP_Q = 65536
    base_s, base_g, base_o = scenes.batch_queries(BATCH_Q, 9, seed=0)
    sel = np.arange(SWEEP_Q) % BATCH_Q
    sw_s, sw_g, sw_o = (np.ascontiguousarray(x[sel]) for x in (base_s, base_g, base_o))
    S.load(sw_s, sw_g, sw_o)
    S.run(1, BATCH_ITERS)
    barrier()
    s_ms = 0.0
    for _ in range(3):
        flush_l2()
        s_ms += S.run(1, BATCH_ITERS)
    barrier()
    s_ms = max_over_ranks(s_ms)
    ov = S.overlap()
    info = S.info()
    sweep_line = {"workload": f"C5: one sweep of {SWEEP_Q} queries x {BATCH_ITERS} iterations sharded over {world} GPU(s) "
                              "(strong scaling), results gathered" + (" over NCCL, chunk by chunk while later chunks are planned"
                                                                      if world > 1 else " (1 GPU: no gather)"),
                  "queries_per_sec": SWEEP_Q * 3 / (s_ms * 1e-3), "ms_per_sweep": s_ms / 3,
                  "gather_bytes": int(SWEEP_Q * capi.RESULT_DTYPE.itemsize) if world > 1 else 0,
                  "gather": {"nccl": bool(info["nccl"]), "nccl_version": info["nccl_version"], "ranks": info["world"],
                             "overlapped": ov["overlapped"], "chunks_per_block": ov["chunks"]}}
    if rank == 0:
        recs = S.results()                     # all 65536 records, global order, from the gathered device array
        sweep_line["records_sha256"] = hashlib.sha256(recs.tobytes()).hexdigest()
        sweep_line["solved_frac"] = float((recs["solved"] != 0).mean())
        import oracle_binding as ob
        # EVERY record against the CPU checker (all host cores; ~2 s for 65536 plans of 200 iterations) — outside any timed region
        cfgo = ob.default_config(p_direct=P_DIRECT)
        ref_all = ob.plan_batch(cfgo, sw_s, sw_g, sw_o, 1, BATCH_ITERS, query_base=0, nthreads=os.cpu_count() or 1)
        bad = int((ref_all.view(np.uint8).reshape(SWEEP_Q, -1) != recs.view(np.uint8).reshape(SWEEP_Q, -1)).any(axis=1).sum())
        sweep_line["parity_vs_oracle"] = {"sample": f"all {SWEEP_Q} gathered records",
                                          "result": "bit-exact" if bad == 0 else f"MISMATCH in {bad} records"}
    if world > 1:
        # the same sweep on ONE GPU, timed in the same run (rank 0; the others wait at the barrier): strong-scaling efficiency
        t1 = 0.0
        if rank == 0:
            S1 = sweep.Sweep.single_process(1, devices=[local_rank], p_direct=P_DIRECT)
            S1.load(sw_s, sw_g, sw_o)
            S1.run(1, BATCH_ITERS)
            for _ in range(2):
                flush_l2()
                t1 += S1.run(1, BATCH_ITERS) / 2
            sweep_line["one_gpu_records_sha256"] = hashlib.sha256(S1.results().tobytes()).hexdigest()
            S1.close()
        barrier()
        t1 = max_over_ranks(t1)
        sweep_line["one_gpu_ms_per_sweep"] = t1
        sweep_line["efficiency"] = t1 / (world * (s_ms / 3))
    b_flop = 134.0 * bst["steer_calls"] + 1180.0 * bst["collide_calls"]      # SURVEY 8(d) accounting, all ranks
    f64_peak, f64_src = fp64_peak()
    b_tflops = b_flop * bsteps / (b_ms * 1e-3) / 1e12 / world
    batched = {
        "roofline": {"bound": "fp64 pipe / latency", "kernel": "k_plan_batch2", "achieved": b_tflops, "peak": f64_peak,
                     "unit": "TFLOP/s", "frac": b_tflops / f64_peak, "peak_source": f64_src,
                     "algorithmic_flop_per_launch": b_flop / world,
                     "note": "per GPU; 134 flop per steering pair + 1180 flop per edge test (SURVEY 8d), counted from the "
                             "kernel's own steer_calls / collide_calls; trees live in shared memory, HBM traffic is the "
                             "inputs and the 1312-byte result records only"},
        "workload": f"{BATCH_Q} queries/GPU x {BATCH_ITERS} iterations, 9 moving obstacles, p_direct=0.1 (C4; sharded sweep C5 for N>1)",
        "queries_per_sec": q_total * bsteps / (b_ms * 1e-3),
        "expansions_per_sec": q_total * BATCH_ITERS * bsteps / (b_ms * 1e-3),
        "ms_per_step": b_ms / bsteps,
        "e2e_queries_per_sec": q_total * bsteps / (be_ms * 1e-3),
        "h2d_bytes_per_step": int((bs.nbytes + bg.nbytes + bo.nbytes) // world),
        "d2h_bytes_per_step": int(q_total * capi.RESULT_DTYPE.itemsize),
        "gpu_launches": int(b_launches),
        "gather": "NCCL all-gather of result records inside libkrrt_mgpu, overlapped chunk by chunk" if world > 1 else "none (1 GPU)",
        "solved_frac": float(out_h["solved"].mean()),
    }

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": tot_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": bench_config(world),
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                "ms_per_step": e2e_ms / args.steps},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": roofline,
        "batched": batched,
        "sweep_c5": sweep_line,
        "tree_nodes": int(res["n_nodes"]), "solved": bool(res["solved"]), "cost": float(res["cost"]),
        "rounds": wave,
        "cpu_flags": "oracle built with -O3 -mavx2 -mfma -ffp-contract=off (x86-64-v3, not -march=native: the .so built in the "
                     "build container must run on the GPU box's host)",
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
        same = same and G.result(*r_last).tobytes() == O.result(*ro).tobytes()
        line["parity_vs_oracle"] = "bit-exact" if same else "MISMATCH"
        # batched CPU baseline on all host cores, bounded sample
        allc = os.cpu_count() or 1
        qs = min(hi - lo, 32 * allc)
        tb = time.perf_counter()
        refb = ob.plan_batch(ob.default_config(p_direct=P_DIRECT), bs[:qs], bg[:qs], bo[:qs], 1, BATCH_ITERS,
                             query_base=lo, nthreads=allc)
        tb = time.perf_counter() - tb
        line["batched"]["cpu_baseline"] = {"queries_per_sec": qs / tb, "cores": allc, "kind": "port",
                                           "sample": f"{qs} of the {BATCH_Q} queries, one query per thread at a time"}
        line["batched"]["parity_vs_oracle"] = ("bit-exact" if np.array_equal(out_h[:qs].view(np.uint8), refb.view(np.uint8))
                                               else "MISMATCH")
        # fp32 mode (KRRT_F32 contexts): the same C4 workload through the binary32 build of the batched planner, its own line
        # (the headline stays fp64); checked against the CPU checker's float build, and against the fp64 records within
        # north_star's 1e-4
        import oracle_f32_binding as of
        S32 = sweep.Sweep.single_process(1, devices=[local_rank], p_direct=P_DIRECT, precision=capi.KRRT_F32)
        S32.load(bs[:BATCH_Q], bg[:BATCH_Q], bo[:BATCH_Q])
        for _ in range(3):
            S32.run(1, BATCH_ITERS)
        f_ms = 0.0
        for _ in range(bsteps):
            flush_l2()
            f_ms += S32.run(1, BATCH_ITERS)
        rec32 = S32.results()
        st32 = S32.stats()
        ref32 = of.plan_batch(of.default_config(p_direct=P_DIRECT), bs[:qs], bg[:qs], bo[:qs], 1, BATCH_ITERS, nthreads=allc)
        rec64 = out_h[:BATCH_Q]
        same_tree = ((rec32["n_nodes"] == rec64["n_nodes"]) & (rec32["goal_index"] == rec64["goal_index"])
                     & (rec32["path_len"] == rec64["path_len"]))
        okc = same_tree & (rec64["solved"] != 0)
        relc = np.abs(rec32["cost"][okc] - rec64["cost"][okc]) / rec64["cost"][okc]
        try:
            with open(os.path.join(ROOT, "profiles", "r1_fp_peak.json")) as f:
                f32_peak = float(json.load(f)["fp32_fma_tflops"])
        except Exception:
            f32_peak = 75.0
        f_flop = 134.0 * st32["steer_calls"] + 1180.0 * st32["collide_calls"]
        f_tflops = f_flop * bsteps / (f_ms * 1e-3) / 1e12
        line["batched_f32"] = {
            "dtype": "f32", "workload": line["batched"]["workload"] + ", KRRT_F32 context",
            "queries_per_sec": BATCH_Q * bsteps / (f_ms * 1e-3), "ms_per_step": f_ms / bsteps,
            "speedup_over_f64": (b_ms / bsteps) / (f_ms / bsteps),
            "solved_frac": float((rec32["solved"] != 0).mean()),
            "parity_vs_float_checker": ("bit-exact" if np.array_equal(rec32[:qs].view(np.uint8), ref32.view(np.uint8)) else "MISMATCH"),
            "vs_f64_records": {"same_tree_size_goal_and_path_length": float(same_tree.mean()),
                               "max_rel_cost_diff": float(relc.max()) if relc.size else None, "tolerance": 1e-4},
            "roofline": {"bound": "fp32 pipe / latency", "kernel": "krrt_f32::k_plan_batch2", "achieved": f_tflops,
                         "peak": f32_peak, "unit": "TFLOP/s", "frac": f_tflops / f32_peak,
                         "peak_source": "measured (profiles/r1_fp_peak.json, FMA-chain microbenchmark)"}}
        # the C5 sweep (65536 queries, this GPU alone) in fp32: 28 rounds of the resident warps hide the tail that the 1.4 rounds of
        # C4 expose, so this is what the float kernel sustains; every record against the float checker
        S32.load(sw_s, sw_g, sw_o)
        S32.run(1, BATCH_ITERS)
        f5_ms = 0.0
        for _ in range(2):
            flush_l2()
            f5_ms += S32.run(1, BATCH_ITERS)
        rec5 = S32.results()
        ref5 = of.plan_batch(of.default_config(p_direct=P_DIRECT), sw_s, sw_g, sw_o, 1, BATCH_ITERS, query_base=0,
                             nthreads=os.cpu_count() or 1)       # every record, all host cores, outside the timed region
        bad5 = int((rec5.view(np.uint8).reshape(SWEEP_Q, -1) != ref5.view(np.uint8).reshape(SWEEP_Q, -1)).any(axis=1).sum())
        line["batched_f32"]["sweep_c5_one_gpu"] = {
            "workload": f"{SWEEP_Q} queries x {BATCH_ITERS} iterations on one GPU, KRRT_F32 context",
            "queries_per_sec": SWEEP_Q * 2 / (f5_ms * 1e-3), "ms_per_sweep": f5_ms / 2,
            "speedup_over_f64_one_gpu_sweep": ((s_ms / 3) / (f5_ms / 2)) if world == 1 else None,
            "parity_vs_float_checker": {"sample": f"all {SWEEP_Q} records",
                                        "result": "bit-exact" if bad5 == 0 else f"MISMATCH in {bad5} records"}}
        S32.close()
        # ... and the headline plan on a KRRT_F32 context (k_grow compiled in float), against the float checker
        G32 = capi.Planner(device=local_rank, capacity=N_ITER + 2, p_direct=P_DIRECT, precision=capi.KRRT_F32)
        O32 = of.OraclePlannerF32(capacity=N_ITER + 2, p_direct=P_DIRECT)
        for P_ in (G32, O32):
            P_.set_obstacles(obs, 1)
            P_.set_start(start)
        r32 = G32.grow(goal, N_ITER)
        ro32 = O32.grow(goal, N_ITER)
        same32 = r32 == ro32 and all(np.array_equal(a, b) for a, b in zip(G32.tree(), O32.tree())) and G32.stats() == O32.stats()
        ms32 = 0.0
        for _ in range(args.steps):
            G32.set_start(start)
            flush_l2()
            G32.grow(goal, N_ITER)
            ms32 += G32.last_timing()["total_ms"]
        line["single_tree_f32"] = {"dtype": "f32", "workload": WORKLOAD.replace("fp64", "KRRT_F32 context"),
                                   "expansions_per_sec": N_ITER * args.steps / (ms32 * 1e-3), "ms_per_step": ms32 / args.steps,
                                   "speedup_over_f64": (tot_ms / args.steps) / (ms32 / args.steps),
                                   "tree_nodes": int(G32.size), "cost": float(G32.result(*r32)["cost"]),
                                   "cost_rel_diff_vs_f64": abs(float(G32.result(*r32)["cost"]) - float(res["cost"])) / float(res["cost"]),
                                   "parity_vs_float_checker": "bit-exact" if same32 else "MISMATCH"}
        G32.close()
        O32.close()
        # config 0 (rrtstar_benchmark, C1): static obstacles, fixed seed, 10k iterations — GPU and the CPU port, same plan
        G1 = capi.Planner(device=local_rank, capacity=10002, batch=args.batch)
        O1 = ob.OraclePlanner(capacity=10002, query_id=0)
        for P_ in (G1, O1):
            P_.set_obstacles(scenes.static_obstacles(), 0)
            P_.set_start(start)
        G1.grow(goal, 10000)                       # warm-up
        ms1 = 0.0
        for _ in range(args.steps):
            G1.set_start(start)
            flush_l2()
            r1 = G1.grow(goal, 10000)
            ms1 += G1.last_timing()["total_ms"] / args.steps
        t0 = time.perf_counter()
        ro1 = O1.grow(goal, 10000)
        dt1 = time.perf_counter() - t0
        same1 = r1 == ro1 and all(np.array_equal(a, b) for a, b in zip(G1.tree(), O1.tree()))
        line["c1_rrtstar_benchmark"] = {
            "workload": "C1: 9 static obstacles, fixed seed, 10000 iterations, p_direct = 0, goal connected (O8)",
            "cost": float(G1.result(*r1)["cost"]) if r1[0] else None,
            "expansions_per_sec": 10000 / (ms1 * 1e-3), "ms": ms1, "steps": args.steps, "nodes": G1.size, "solved": bool(r1[0]),
            "cpu_baseline": {"value": 10000 / dt1, "unit": UNIT, "cores": 1, "kind": "port", "sample": "the full plan, once"},
            "parity_vs_oracle": "bit-exact" if same1 else "MISMATCH"}
        G1.close()
        # the README's curves: computation time and solution cost against node size, 1024 seeds per size
        # through the batched planner, on the obstacle-free field (where the README's cost band is reproduced) and on the
        # C1 scene; CPU port on all host cores beside it (the README's numbers are an i3 at 2.0 GHz, scene not stated)
        curve = []
        qn = 1024
        c_s = np.tile(start, (qn, 1))
        c_g = np.tile(goal, (qn, 1))
        for scene_name, c_o in (("free field", np.zeros((qn, 0, 4))), ("C1, 9 static robots", np.tile(scenes.static_obstacles(), (qn, 1, 1)))):
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
                    curve.append({"scene": scene_name, "p_direct": p_dir, "node_size": nodes, "gpu_ms_per_plan_batched": msb / qn,
                                  "cpu_port_ms_per_plan_1core": 1e3 * tcp / 4.0,   # 4 plans per thread, all threads busy
                                  "solved_frac": float(sv.mean()),
                                  "mean_cost": float(oc["cost"][sv].mean()) if sv.any() else None,
                                  "parity": bool(np.array_equal(oc[:4 * allc].view(np.uint8), rc_.view(np.uint8)))})
                Gc.close()
        line["readme_curve"] = {"note": "C1 start/goal, goal connected to the cheapest collision-free node (O8), 1024 seeds per "
                                        "point (query_id = seed); README cost_plot.png: 16.26 @ 10 -> 12.22 @ 200 without direct sampling",
                                "points": curve}
        # config 2 (C3): single large trees on static obstacles — 140k iterations (>= 100k NODES: the tree north_star's 100x
        # target is quoted on) beside the CPU port on one host core, and the 1M-iteration tree
        if not args.skip_c3:
            l2_peak = (profile_json("r2_l2_peak.json") or {}).get("l2_read_GBps")
            for label, iters in (("c3_100k_nodes", 140000), ("c3_1m", 1000000)):
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
                s3 = G3.scan_stats()
                line[label] = {"iterations": iters, "expansions_per_sec": iters / (t_plain * 1e-3), "ms": t_plain, "nodes": G3.size,
                               "brute_force_equivalent_GBps": t3["scan_node_samples"] * 32 / (t_plain * 1e-3) / 1e9,
                               "scan_phase_GBps_visited": 32.0 * s3["visited_pairs"] / (t3["scan_ms"] * 1e-3) / 1e9,
                               "scan_phase_frac_of_l2_peak": ((32.0 * s3["visited_pairs"] / (t3["scan_ms"] * 1e-3) / 1e9) / l2_peak
                                                              if l2_peak else None),
                               "phase_shares": {k: round(v["ms"] / ks, 4) for k, v in k3.items() if v["ms"] > 0},
                               "scan_index": dict(s3, visited_bytes=32.0 * s3["visited_pairs"]),
                               "rounds": G3.wave_stats(),
                               "note": "brute_force_equivalent_GBps = SURVEY 8d's 32 B x tree nodes x samples over the call's duration: what a "
                                       "brute-force scan would have to sustain to be as fast — an EQUIVALENCE, not a bandwidth (the "
                                       "grid-indexed scan reads scan_index.visited_bytes, all of them L2 hits: scan_phase_frac_of_l2_peak "
                                       "holds them against the measured L2 read peak, profiles/r2_l2_peak.json); the call is bound by "
                                       "the latency of its commit sub-rounds (roofline.latency_model)"}
                if label == "c3_100k_nodes":   # the same plan on one host core (kd-tree near search), once
                    O3 = ob.OraclePlanner(capacity=iters + 2, query_id=0)
                    O3.set_obstacles(scenes.static_obstacles(), 0)
                    O3.set_start(start)
                    t0 = time.perf_counter()
                    O3.grow(None, iters)
                    dt3 = time.perf_counter() - t0
                    same3 = all(np.array_equal(a, b) for a, b in zip(G3.tree(), O3.tree())) and G3.stats() == O3.stats()
                    line[label]["cpu_baseline"] = {"value": iters / dt3, "unit": UNIT, "cores": 1, "kind": "port",
                                                   "sample": f"the full {iters}-iteration plan, once, kd-tree near search"}
                    line[label]["gpu_over_cpu"] = line[label]["expansions_per_sec"] / (iters / dt3)
                    line[label]["parity_vs_oracle"] = "bit-exact" if same3 else "MISMATCH"
                G3.close()
    if rank == 0:
        emit(line)
    S.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=0, help="window size: samples scanned + steered at once (0 = library default)")
    ap.add_argument("--skip-c3", action="store_true")
    ap.add_argument("--c3-1m", action="store_true", help="(kept for compatibility: the 1M-iteration tree is part of the default line)")
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
