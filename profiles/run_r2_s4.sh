#!/bin/bash
# Round 2, GPU session 4: warps per CTA of the binary32 batched planner (compiled on the box per setting).
set -u
mkdir -p gpurun_out
: > gpurun_out/c4_warps.log
for w in 16 20 24 32; do
  touch robosoccer_motion_planning_b200/csrc/krrt_f32.cu
  make -C robosoccer_motion_planning_b200 -s all EXTRA=-DKRRT_B2_MAX_WARPS_F32=$w >> gpurun_out/c4_warps.log 2>&1
  echo "== KRRT_B2_MAX_WARPS_F32=$w" >> gpurun_out/c4_warps.log
  timeout 300 python profiles/r2_c4_probe.py 2>&1 | grep f32 >> gpurun_out/c4_warps.log
  ITERS=75 timeout 300 python profiles/r2_c4_probe.py 2>&1 | grep f32 | sed 's/^/   75 iterations: /' >> gpurun_out/c4_warps.log
done
cat gpurun_out/c4_warps.log
