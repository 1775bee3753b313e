#!/bin/bash
# GPU session: candidates below rewired nodes settled in decide — the new tests, the full suite, bench, stress repeats.
set -u
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_gpu_parity.py -k "candidates_below or other_seeds" -x -q 2>&1 | tail -15
timeout 500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --c3-1m > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -3 gpurun_out/bench.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench.json'))
print('value',d['value'],'e2e',d['e2e']['value'], d['parity_vs_oracle'], d['roofline']['phase_shares'])
for k in ('c1','c3_100k','c3_1m'):
    c=d.get(k)
    if c: print(k, c['expansions_per_sec'], c['ms'], c.get('gpu_over_cpu'), c.get('parity_vs_oracle'))
print({k: d['batched'][k] for k in d['batched'] if k in ('value','parity_vs_oracle')})
PY
