#!/bin/bash
# Hang / flake check: the GPU suite ten times and the bench three times, each under its own timeout.
set -u
mkdir -p gpurun_out
: > gpurun_out/stress.log
for i in 1 2 3 4 5 6 7 8 9 10; do
  timeout 200 python -m pytest tests -m gpu -x -q 2>&1 | tail -1 >> gpurun_out/stress.log
  echo "suite $i rc=${PIPESTATUS[0]}" >> gpurun_out/stress.log
done
for i in 1 2 3; do
  timeout 300 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_stress_$i.json 2>> gpurun_out/stress.log; echo "bench $i rc=$?" >> gpurun_out/stress.log
done
cat gpurun_out/stress.log
