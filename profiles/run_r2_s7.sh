#!/bin/bash
# Round 2, GPU session 7: krrt_grow_logged + the run-ahead of the per-iteration loop (include/rrtstar.hpp).
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_host_surface.py tests/test_reference_node.py -m gpu -q -x > gpurun_out/pytest_runahead.log 2>&1; tail -8 gpurun_out/pytest_runahead.log
tests/native/_build/rrtstar_benchmark --benchmark_min_time=0.2 > gpurun_out/rrtstar_benchmark.txt 2>&1; cat gpurun_out/rrtstar_benchmark.txt
KRRT_RUN_AHEAD=0 tests/native/_build/rrtstar_benchmark --benchmark_min_time=0.2 --benchmark_filter=PerCall > gpurun_out/rrtstar_benchmark_norunahead.txt 2>&1; cat gpurun_out/rrtstar_benchmark_norunahead.txt
tests/native/_build/l2_peak > gpurun_out/l2_peak.json 2>&1; cat gpurun_out/l2_peak.json
for e in static dynamic; do KRRT_STUB_CYCLES=20 KRRT_STUB_GOAL=7,0 ROSPARAM_environment=$e ROSPARAM_target_size=75 ROSPARAM_direct_sampling=true ROSPARAM_direct_sampling_prob=0.1 tests/native/_build/reference_node 2>&1 | grep -E 'solution|cycle' | tail -4; done
