set -u
mkdir -p gpurun_out
timeout 60 python profiles/profile_plan.py c2 20000 > gpurun_out/p_c2.log 2>&1; echo "c2 rc=$?"; head -4 gpurun_out/p_c2.log
