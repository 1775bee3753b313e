#!/usr/bin/env python
"""Summaries of ncu outputs brought back in gpurun_out/ (run here, no GPU needed).
  python profiles/summarize.py launches <launches.csv>         per-kernel count / total / average / share
  python profiles/summarize.py raw <file.ncu-rep> [regex]      selected raw metrics of every captured launch
  python profiles/summarize.py json <file.ncu-rep>             the numbers bench.py's roofline object quotes, as JSON
  python profiles/summarize.py sass <lib.so> <kernel regex>    instruction mix of a kernel from cuobjdump -sass"""
import collections
import csv
import re
import subprocess
import sys

DEFAULT = (r"^(gpu__time_duration.sum|sm__cycles_elapsed.max|smsp__inst_executed.sum|dram__bytes_(read|write).sum|"
           r"launch__(registers_per_thread|grid_size|block_size|shared_mem_per_block|occupancy_limit_\w+)|"
           r"sm__warps_active.avg.pct_of_peak_sustained_active|sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_elapsed|"
           r"smsp__issue_active.avg.pct_of_peak_sustained_active|smsp__thread_inst_executed_per_inst_executed.ratio|"
           r"smsp__average_warp_latency_per_inst_issued.ratio|lts__t_bytes.sum|lts__t_sector_hit_rate.pct|"
           r"smsp__average_warps_issue_stalled_\w+_per_issue_active.ratio)$")


def launches(path):
    rows = [r for r in csv.reader(open(path)) if len(r) > 10]
    h = rows[0]
    ki, vi, ui = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Unit")
    agg = collections.OrderedDict()
    for r in rows[1:]:
        v = float(r[vi].replace(",", ""))
        v = {"ns": v / 1e3, "us": v, "ms": v * 1e3, "s": v * 1e6}[r[ui]]
        a = agg.setdefault(r[ki].split("(")[0], [0, 0.0])
        a[0] += 1
        a[1] += v
    tot = sum(v[1] for v in agg.values())
    print("| kernel | launches | total us | avg us | share |\n|---|---|---|---|---|")
    for k, v in sorted(agg.items(), key=lambda x: -x[1][1]):
        print(f"| `{k}` | {v[0]} | {v[1]:.1f} | {v[1] / v[0]:.2f} | {v[1] / tot:.3f} |")


def raw(path, pattern):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    h = rows[0]
    pat = re.compile(pattern)
    print("| metric | unit | " + " | ".join(f"launch {i}" for i in range(len(rows) - 2)) + " |")
    print("|---|---|" + "---|" * (len(rows) - 2))
    for i, c in enumerate(h):
        if pat.search(c):
            vals = [r[i] for r in rows[2:]]
            if all(v in ("0", "") for v in vals) and "stalled" in c:
                continue
            print(f"| {c} | {rows[1][i]} | " + " | ".join(vals) + " |")


def as_json(path):
    """dram bytes, L2 bytes, pipe / issue utilisation, stall split of the first captured launch (bench.py reads this)"""
    import json
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    h, units, v = rows[0], rows[1], rows[2]

    def get(name, scale=True):
        if name not in h:
            return None
        i = h.index(name)
        x = float(v[i].replace(",", ""))
        u = units[i]
        if scale:
            x *= {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12, "byte": 1.0, "us": 1e-6, "ms": 1e-3, "ns": 1e-9,
                  "s": 1.0}.get(u, 1.0)
        return x
    d = {"source": path.split("/")[-1], "kernel": v[h.index("Kernel Name")],
         "duration_s_under_ncu": get("gpu__time_duration.sum"),
         "dram_bytes_read": get("dram__bytes_read.sum"), "dram_bytes_write": get("dram__bytes_write.sum"),
         "lts_t_bytes": (get("lts__t_bytes.sum") if "lts__t_bytes.sum" in h
                         else (32.0 * get("lts__t_sectors.sum") if get("lts__t_sectors.sum") is not None else None)),
         "lts_pct_of_peak": get("lts__t_sectors.sum.pct_of_peak_sustained_elapsed"),
         "xbar2l1tex_read_bytes": get("l1tex__m_xbar2l1tex_read_bytes.sum"), "lts_hit_rate_pct": get("lts__t_sector_hit_rate.pct"),
         "fp64_pipe_pct": get("sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_elapsed"),
         "issue_active_pct": get("smsp__issue_active.avg.pct_of_peak_sustained_active"),
         "warps_active_pct": get("sm__warps_active.avg.pct_of_peak_sustained_active"),
         "threads_per_inst": get("smsp__thread_inst_executed_per_inst_executed.ratio"),
         "registers_per_thread": get("launch__registers_per_thread"),
         "stall_cycles_per_inst": get("smsp__average_warp_latency_per_inst_issued.ratio"),
         "stall_barrier_per_inst": get("smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio"),
         "stall_long_scoreboard_per_inst": get("smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio"),
         "inst_executed": get("smsp__inst_executed.sum")}
    print(json.dumps(d, indent=1))


def sass(lib, pattern):
    """instruction mix of every kernel of `lib` whose mangled name matches `pattern`"""
    out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
    pat = re.compile(pattern)
    fn, mix = None, collections.OrderedDict()
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            fn = m.group(1) if pat.search(m.group(1)) else None
            if fn:
                mix[fn] = collections.Counter()
            continue
        if fn:
            m = re.search(r"/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d\s+)?([A-Z][A-Z0-9_]*)", line)
            if m:
                mix[fn][m.group(1)] += 1
    for fn, c in mix.items():
        tot = sum(c.values())
        print(f"### `{fn}` — {tot} SASS instructions")
        print("| opcode | count | share |\n|---|---|---|")
        for op, k in c.most_common(18):
            print(f"| {op} | {k} | {k / tot:.3f} |")
        groups = {"fp64 (DADD/DMUL/DFMA/DSETP/MUFU.*64)": sum(k for op, k in c.items() if op.startswith("D") and op not in ("DEPBAR",)),
                  "fp32 (FADD/FMUL/FFMA/FSETP/FMNMX)": sum(k for op, k in c.items() if op.startswith("F") and op[:3] in ("FAD", "FMU", "FFM", "FSE", "FMN", "FCH", "FSE")),
                  "global/local memory (LDG/STG/LDL/STL/ATOMG/RED)": sum(k for op, k in c.items() if op in ("LDG", "STG", "LDL", "STL", "ATOMG", "RED", "ATOM")),
                  "shared memory (LDS/STS/ATOMS)": sum(k for op, k in c.items() if op in ("LDS", "STS", "ATOMS", "LDSM")),
                  "warp collectives (SHFL/VOTE/REDUX/MATCH)": sum(k for op, k in c.items() if op in ("SHFL", "VOTE", "REDUX", "MATCH")),
                  "barriers (BAR/MEMBAR/ERRBAR)": sum(k for op, k in c.items() if op in ("BAR", "MEMBAR", "ERRBAR")),
                  "tensor / TMA (HMMA/UTC*/UBLKCP/UTMA*)": sum(k for op, k in c.items() if op.startswith(("HMMA", "UTC", "UBLK", "UTMA", "QGMMA", "IMMA")))}
        for g, k in groups.items():
            print(f"- {g}: {k}")
        print()


if __name__ == "__main__":
    if sys.argv[1] == "json":
        as_json(sys.argv[2])
        sys.exit(0)
    if sys.argv[1] == "sass":
        sass(sys.argv[2], sys.argv[3])
        sys.exit(0)
    if sys.argv[1] == "launches":
        launches(sys.argv[2])
    else:
        raw(sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else DEFAULT)
