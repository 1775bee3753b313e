#!/bin/bash
# Round-1 GPU session: parity tests, bench (both arms), ncu launch list + full captures of the round kernels.
# Run under gpurun from the repo root; everything lands in gpurun_out/.
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
nproc >> gpurun_out/smi.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "bench ref rc=$?"
KRRT_PROFILE2=1 timeout 300 python profiles/profile_plan.py c2 20000 > gpurun_out/plan_c2.log 2>&1; echo "plan rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/launches_c2.csv \
  python profiles/profile_plan.py c2 20000 > gpurun_out/ncu_launches.log 2>&1; echo "ncu launches rc=$?"
for k in k_round_scan k_round_decide k_round_validate k_round_apply; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$k --launch-skip 600 --launch-count 2 \
    -f -o gpurun_out/r1_$k python profiles/profile_plan.py c2 20000 > gpurun_out/ncu_$k.log 2>&1; echo "ncu $k rc=$?"
done
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_plan_batch --launch-count 1 \
  -f -o gpurun_out/r1_k_plan_batch python profiles/profile_plan.py batch 4096 200 > gpurun_out/ncu_batch.log 2>&1; echo "ncu batch rc=$?"
ls -la gpurun_out
