#!/bin/bash
# Round 2, GPU session 10: the next window evaluated (scan + steer) by the CTAs that sit out the current window's sub-rounds.
set -u
mkdir -p gpurun_out
{
KRRT_AHEAD=0 timeout 300 python profiles/r2_grow_probe.py c2 c3
timeout 300 python profiles/r2_grow_probe.py c2 c3 --parity
for pn in "2 8 16" "2 0 16" "2 -8 16" "1 24 24" "1 40 24"; do set -- $pn; KRRT_PN_MUL=$1 KRRT_PN_ADD=$2 KRRT_AHEAD_MIN=$3 timeout 300 python profiles/r2_grow_probe.py c2 c3 --parity; done
} > gpurun_out/ahead_probe.log 2>&1
grep -v shares gpurun_out/ahead_probe.log | cut -c1-200
