set -u
mkdir -p gpurun_out
KRRT_DEBUG_LOG=gpurun_out/declog.bin KRRT_DEBUG_DUMP=1 timeout 30 python profiles/profile_plan.py c2 20000 > gpurun_out/p_c2.log 2>&1; echo "c2 rc=$?"; grep -v "^cta" gpurun_out/p_c2.log; grep "cta   [01] \|cta 147" gpurun_out/p_c2.log
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -5
