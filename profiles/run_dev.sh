set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 900 python bench.py --c3-1m > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -3 gpurun_out/bench.err
python -c "
import json; d=json.load(open('gpurun_out/bench.json'))
print('value',d['value'],'ms',d['ms_per_step'],'e2e',d['e2e']['value'])
print('roofline',d['roofline']['achieved'],d['roofline']['phase_shares'])
print('batched',d['batched']['queries_per_sec'], d['batched'].get('cpu_baseline'))
print('cpu',d['cpu_baseline'])
for k in ('c3_100k','c3_1m'):
    print(k,d.get(k))
"
