#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 400 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python profiles/latency_probe.py
timeout 300 python bench.py --skip-c3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
python -c "
import json
d=json.load(open('gpurun_out/bench.json')); print(d['value'], d['e2e']['value'], d['gpu_launches'], d['parity_vs_oracle'])"
timeout 100 tests/native/_build/rrtstar_custom_benchmark 20 | head -12
