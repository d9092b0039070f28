set -x
mkdir -p gpurun_out
timeout 60 python tools/icp_time.py 2>&1 | tail -4
timeout 250 python -m pytest tests/test_icp_gpu.py -x -q -m gpu --timeout 60 > gpurun_out/r02_51_pytest_icp.log 2>&1; echo "pytest rc $?"; tail -3 gpurun_out/r02_51_pytest_icp.log
timeout 60 python tools/icp_time.py 2>&1 | tail -4
