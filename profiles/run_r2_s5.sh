#!/bin/bash
# Round 2, GPU session 5: lane-parallel goal connection in the batched kernels: parity, C4 timing, full ncu capture of k_plan_batch2.
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_f32.py tests/test_gpu_parity.py tests/test_sweep_gpu.py -m gpu -q -x -k "f32 or batch or sweep or goal" > gpurun_out/pytest_batch.log 2>&1; tail -5 gpurun_out/pytest_batch.log
timeout 300 python profiles/r2_c4_probe.py > gpurun_out/c4_probe.log 2>&1; tail -3 gpurun_out/c4_probe.log
timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_plan_batch2 --launch-count 1 -f -o gpurun_out/r2_plan_batch2 python profiles/profile_plan.py batch 4096 200 > gpurun_out/ncu_batch2.log 2>&1; echo "ncu rc=$?"
