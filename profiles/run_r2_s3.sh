#!/bin/bash
# Round 2, GPU session 3: fp32 mode (KRRT_F32) on hardware: parity tests against the float checker, C4 timing per precision.
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_f32.py tests/test_gpu_parity.py -m gpu -q -x -k "f32 or batch" > gpurun_out/pytest_f32.log 2>&1; tail -15 gpurun_out/pytest_f32.log
timeout 300 python profiles/r2_c4_probe.py > gpurun_out/c4_probe.log 2>&1; cat gpurun_out/c4_probe.log | tail -5
