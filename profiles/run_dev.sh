set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_yaw.py -x -q 2>&1 | tail -5
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
