#!/bin/bash
# Session GPU run: parity tests, fp peaks, bench (both arms), ncu full capture of k_plan_batch.
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
nproc >> gpurun_out/smi.txt
tests/native/_build/fp_peak > gpurun_out/fp_peak.json 2>&1; cat gpurun_out/fp_peak.json
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_plan_batch --launch-count 1 \
  -f -o gpurun_out/s20_k_plan_batch python profiles/profile_plan.py batch 4096 200 > gpurun_out/ncu_batch.log 2>&1; echo "ncu batch rc=$?"
ls -la gpurun_out
