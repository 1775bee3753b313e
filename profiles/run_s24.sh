#!/bin/bash
# GPU session: grid-indexed neighbour scan of k_grow — parity, then bench with the 1M-iteration tree.
set -u
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_parity.py -k "grid_index or large_tree or grow" -x -q > gpurun_out/pytest_grid.log 2>&1; echo "grid rc=$?" >> gpurun_out/pytest_grid.log
tail -30 gpurun_out/pytest_grid.log
timeout 400 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --c3-1m > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -3 gpurun_out/bench.err
python -c "
import json
d=json.load(open('gpurun_out/bench.json'))
print('value',d['value'],'e2e',d['e2e']['value'], d['parity_vs_oracle'], d['roofline']['phase_shares'], d['roofline']['scan_index'])
for k in ('c3_100k','c3_1m'):
    c=d[k]; print(k, c['expansions_per_sec'], c['ms'], c['phase_shares'], c['scan_index'], c.get('gpu_over_cpu'), c.get('parity_vs_oracle'))
"
