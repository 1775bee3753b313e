set -u
mkdir -p gpurun_out
timeout 120 python profiles/profile_plan.py c2 2000 > gpurun_out/p_c2_2k.log 2>&1; echo "c2 2k rc=$?"; tail -12 gpurun_out/p_c2_2k.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu.log
timeout 120 python profiles/profile_plan.py c2 20000 > gpurun_out/p_c2.log 2>&1; echo "c2 rc=$?"; cat gpurun_out/p_c2.log
timeout 120 python profiles/profile_plan.py c3 100000 > gpurun_out/p_c3.log 2>&1; echo "c3 rc=$?"; cat gpurun_out/p_c3.log
