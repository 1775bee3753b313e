#!/bin/bash
# Round 2, GPU session 9: how large the un-indexed tail of the neighbour scan may grow before the index is rebuilt.
set -u
mkdir -p gpurun_out
{ for tail in 1073741824 32768 16384 8192 4096 2048; do KRRT_GRID_TAIL=$tail timeout 300 python profiles/r2_grow_probe.py c3 c3m; done; } > gpurun_out/grid_tail.log 2>&1
grep -E 'env|C3' gpurun_out/grid_tail.log
timeout 300 python profiles/r2_dlog_c3.py > gpurun_out/dlog_c3.txt 2>&1; tail -3 gpurun_out/dlog_c3.txt
