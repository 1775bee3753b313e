#!/usr/bin/env python
"""Summaries of ncu outputs brought back in gpurun_out/ (run here, no GPU needed).
  python profiles/summarize.py launches <launches.csv>         per-kernel count / total / average / share
  python profiles/summarize.py raw <file.ncu-rep> [regex]      selected raw metrics of every captured launch"""
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


if __name__ == "__main__":
    if sys.argv[1] == "launches":
        launches(sys.argv[2])
    else:
        raw(sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else DEFAULT)
