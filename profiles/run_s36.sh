#!/bin/bash
# GPU session: in-order samples committed behind the prefix in the same sub-round (tail commit) — parity, bench, switch off for comparison.
set -u
mkdir -p gpurun_out
timeout 500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --c3-1m > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -3 gpurun_out/bench.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench.json'))
print('value',d['value'],'e2e',d['e2e']['value'], d['parity_vs_oracle'], d['roofline']['phase_shares'])
print(d.get('wave'), d['batched']['value'] if 'batched' in d else None)
for k in ('c1','c3_100k','c3_1m'):
    c=d.get(k)
    if c: print(k, c['expansions_per_sec'], c['ms'], c.get('gpu_over_cpu'), c.get('parity_vs_oracle'))
PY
timeout 400 python profiles/ab_probe.py - KRRT_INTRA_FIX=0 KRRT_INTRA_FIX=0,KRRT_TAIL_SEQ=0 - 2>&1 | tee gpurun_out/ab_tail.txt
