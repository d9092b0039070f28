set -x
mkdir -p gpurun_out
timeout 400 python -m pytest tests -x -q -m gpu --timeout 120 > gpurun_out/r02_pytest_gpu.log 2>&1; echo "pytest rc $?"; tail -2 gpurun_out/r02_pytest_gpu.log
