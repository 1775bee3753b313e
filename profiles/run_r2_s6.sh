#!/bin/bash
# Round 2, GPU session 6: evidence run — full GPU suite, bench (both arms), microbenchmarks behind the roofline line (grid barrier,
# L2 read bandwidth, CTA-0 clock profile of a sub-round), launch list of the bench command, full ncu captures of k_grow (C2, C3 at
# 100k+ nodes, C3 at 1 M iterations) — each capture after the same command has exited 0 without ncu.
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -4 gpurun_out/pytest_gpu.log
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -3 gpurun_out/bench.err
timeout 400 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
tests/native/_build/barrier_bench > gpurun_out/barrier_bench.txt 2>&1; head -4 gpurun_out/barrier_bench.txt
tests/native/_build/l2_peak > gpurun_out/l2_peak.json 2>&1; cat gpurun_out/l2_peak.json
timeout 300 python profiles/clk_probe.py > gpurun_out/clk_probe.txt 2>&1; head -3 gpurun_out/clk_probe.txt
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_bench.csv \
  python bench.py --steps 2 --warmup 1 --skip-c3 > gpurun_out/ncu_launches.log 2>&1; echo "ncu launches rc=$?"
for cfg in "c2 20000" "c3 140000" "c3 1000000"; do
  set -- $cfg
  timeout 300 python profiles/profile_plan.py $1 $2 > gpurun_out/plan_$1_$2.log 2>&1 || { echo "plan $cfg failed"; continue; }
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_grow --launch-count 1 -f -o gpurun_out/r2_k_grow_$1_$2 \
    python profiles/profile_plan.py $1 $2 > gpurun_out/ncu_$1_$2.log 2>&1; echo "ncu $cfg rc=$?"
done
ls -la gpurun_out | tail -20
