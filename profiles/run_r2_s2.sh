#!/bin/bash
# Round 2, GPU session 2: two k_grow CTAs per SM (128 registers) against one; the reference's own node on the GPU.
set -u
mkdir -p gpurun_out
{
KRRT_GROW_PER_SM=1 timeout 300 python profiles/r2_grow_probe.py
timeout 300 python profiles/r2_grow_probe.py --parity
for nb in "3 16" "4 16"; do set -- $nb; KRRT_NB_MUL=$1 KRRT_NB_ADD=$2 KRRT_PN_MUL=$(( $1 + 1 )) KRRT_PN_ADD=32 timeout 300 python profiles/r2_grow_probe.py c2 c3; done
} > gpurun_out/grow_probe.log 2>&1
tail -40 gpurun_out/grow_probe.log
timeout 600 python -m pytest tests/test_reference_node.py tests/test_host_surface.py tests/test_gpu_parity.py -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; tail -5 gpurun_out/pytest_gpu.log
