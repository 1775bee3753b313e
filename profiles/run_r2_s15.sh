#!/bin/bash
# Round 2, GPU session 15: the final state once more — full GPU suite, bench (both arms), smoke, the native benchmark program,
# the reference's unmodified node in both precisions, launch list of the bench command.
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -4 gpurun_out/pytest_gpu.log
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -3 gpurun_out/bench.err
timeout 400 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log
tests/native/_build/rrtstar_benchmark --benchmark_min_time=0.3 > gpurun_out/rrtstar_benchmark.txt 2>&1
KRRT_RUN_AHEAD=0 tests/native/_build/rrtstar_benchmark --benchmark_min_time=0.3 --benchmark_filter=PerCall 2>&1 | tail -1 | sed 's/$/   <- KRRT_RUN_AHEAD=0 (one device call per iteration)/' >> gpurun_out/rrtstar_benchmark.txt
KRRT_PRECISION=f32 tests/native/_build/rrtstar_benchmark --benchmark_min_time=0.3 --benchmark_filter=PerCall 2>&1 | tail -1 | sed 's/$/   <- KRRT_PRECISION=f32/' >> gpurun_out/rrtstar_benchmark.txt
tail -4 gpurun_out/rrtstar_benchmark.txt
for prec in f64 f32; do for e in static dynamic; do
  echo "== reference_node precision=$prec environment=$e (10 cycles of 75 iterations)"
  KRRT_PRECISION=$prec KRRT_STUB_CYCLES=10 KRRT_STUB_GOAL=7,0 ROSPARAM_environment=$e ROSPARAM_target_size=75 ROSPARAM_direct_sampling=true \
    ROSPARAM_direct_sampling_prob=0.1 tests/native/_build/reference_node 2>&1 | grep -E "solution" | tail -3
done; done > gpurun_out/reference_node.txt 2>&1; tail -3 gpurun_out/reference_node.txt
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/launches_bench.csv \
  python bench.py --steps 2 --warmup 1 --skip-c3 > gpurun_out/ncu_launches.log 2>&1; echo "ncu launches rc=$?"
