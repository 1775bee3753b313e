#!/bin/bash
# GPU session: k_plan_batch2 parity, full suite, bench, ncu of k_plan_batch2.  Every step has its own timeout.
set -u
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_parity.py -k "plan_batch" -x -q > gpurun_out/pytest_batch.log 2>&1; echo "batch rc=$?" >> gpurun_out/pytest_batch.log
tail -25 gpurun_out/pytest_batch.log
timeout 600 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
timeout 400 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
KRRT_BATCH_V1=1 timeout 400 python bench.py --skip-c3 > gpurun_out/bench_v1.json 2> gpurun_out/bench_v1.err; echo "bench v1 rc=$?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_plan_batch2 --launch-count 1 \
  -f -o gpurun_out/s22_k_plan_batch2 python profiles/profile_plan.py batch 4096 200 > gpurun_out/ncu_batch.log 2>&1; echo "ncu batch rc=$?"
python -c "
import json
for f in ('bench','bench_v1'):
    d=json.load(open('gpurun_out/%s.json'%f)); print(f, d['value'], d['batched']['queries_per_sec'], d['batched']['e2e_queries_per_sec'], d['batched'].get('parity_vs_oracle'))
"
