#!/bin/bash
# GPU session: bench line with the batched roofline / C3 CPU baseline, launch list of the bench command, full capture of k_grow.
set -u
mkdir -p gpurun_out
timeout 300 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
timeout 500 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -3 gpurun_out/bench.err
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_bench.csv \
  python bench.py --steps 2 --warmup 1 --skip-c3 > gpurun_out/ncu_launches.log 2>&1; echo "ncu launches rc=$?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_grow --launch-count 1 \
  -f -o gpurun_out/s23_k_grow python profiles/profile_plan.py c2 20000 > gpurun_out/ncu_kgrow.log 2>&1; echo "ncu k_grow rc=$?"
python -c "
import json
d=json.load(open('gpurun_out/bench.json'))
print('value',d['value'],'e2e',d['e2e']['value'],'cpu',d['cpu_baseline']['value'])
b=d['batched']; print('batched',b['queries_per_sec'],b['e2e_queries_per_sec'],b['roofline']['achieved'],b['roofline']['frac'],b['cpu_baseline'])
print('c3',{k:d['c3_100k'].get(k) for k in ('expansions_per_sec','cpu_baseline','gpu_over_cpu','parity_vs_oracle')})
"
