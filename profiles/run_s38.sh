#!/bin/bash
# GPU session 38: final-state evidence after the parallel-commit widenings — tests, bench (both arms), launch list of the bench command, full capture of k_grow.
set -u
mkdir -p gpurun_out
timeout 400 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -3 gpurun_out/pytest_gpu.log
timeout 500 python bench.py --c3-1m > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -3 gpurun_out/bench.err
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_bench.csv \
  python bench.py --steps 2 --warmup 1 --skip-c3 > gpurun_out/ncu_launches.log 2>&1; echo "ncu launches rc=$?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_grow --launch-count 1 \
  -f -o gpurun_out/s38_k_grow python profiles/profile_plan.py c2 20000 > gpurun_out/ncu_kgrow.log 2>&1; echo "ncu k_grow rc=$?"
python -c "
import json
d=json.load(open('gpurun_out/bench.json'))
print('value',d['value'],'e2e',d['e2e']['value'],'cpu',d['cpu_baseline']['value'], d['roofline']['achieved'], d['roofline']['frac'])
b=d['batched']; print('batched',b['queries_per_sec'],b['e2e_queries_per_sec'],b['roofline']['frac'],b['cpu_baseline']['queries_per_sec'])
print('sweep', d['sweep_c5']['queries_per_sec'])
for k in ('c1_rrtstar_benchmark','c3_100k','c3_1m'): print(k, d[k]['expansions_per_sec'], d[k].get('gpu_over_cpu'), d[k].get('cpu_baseline'))
"
