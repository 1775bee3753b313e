#!/bin/bash
# Round 2, GPU session 12: evidence of the final state — full GPU suite, bench (both arms), smoke, launch list of the bench command,
# one full capture of k_grow (C2) and of k_plan_batch2, the native benchmark programs.
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -4 gpurun_out/pytest_gpu.log
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -3 gpurun_out/bench.err
timeout 400 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log
tests/native/_build/rrtstar_benchmark --benchmark_min_time=0.3 > gpurun_out/rrtstar_benchmark.txt 2>&1
KRRT_RUN_AHEAD=0 tests/native/_build/rrtstar_benchmark --benchmark_min_time=0.3 --benchmark_filter=PerCall 2>&1 | tail -1 | sed 's/$/   <- KRRT_RUN_AHEAD=0 (one device call per iteration)/' >> gpurun_out/rrtstar_benchmark.txt
tail -3 gpurun_out/rrtstar_benchmark.txt
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/launches_bench.csv \
  python bench.py --steps 2 --warmup 1 --skip-c3 > gpurun_out/ncu_launches.log 2>&1; echo "ncu launches rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_grow --launch-count 1 -f -o gpurun_out/r2_k_grow_c2_final \
  python profiles/profile_plan.py c2 20000 > gpurun_out/ncu_c2.log 2>&1; echo "ncu k_grow rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_plan_batch2 --launch-count 1 -f -o gpurun_out/r2_plan_batch2_final \
  python profiles/profile_plan.py batch 4096 200 > gpurun_out/ncu_batch2.log 2>&1; echo "ncu batch2 rc=$?"
