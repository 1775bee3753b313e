#!/bin/bash
# GPU session: hang check of the window-barrier fix (repeated small-window plans under a short timeout), full parity
# suite, bench (both arms), ncu full capture of k_plan_batch.  Every step has its own timeout.
set -u
mkdir -p gpurun_out
for i in 1 2 3 4 5 6; do
  timeout 120 python -m pytest "tests/test_gpu_parity.py::test_grow_static_c1" -x -q 2>&1 | tail -1
  rc=${PIPESTATUS[0]}
  if [ "$rc" != "0" ]; then echo "stress iteration $i rc=$rc"; break; fi
done | tee gpurun_out/stress.log
if grep -q "rc=" gpurun_out/stress.log; then echo "STRESS FAILED"; exit 1; fi
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
timeout 400 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "bench ref rc=$?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_plan_batch --launch-count 1 \
  -f -o gpurun_out/s21_k_plan_batch python profiles/profile_plan.py batch 4096 200 > gpurun_out/ncu_batch.log 2>&1; echo "ncu batch rc=$?"
ls -la gpurun_out
