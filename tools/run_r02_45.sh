set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_forward_gpu.py -x -q -m gpu > gpurun_out/r02_45_pytest_forward.log 2>&1; echo "pytest rc $?"; tail -3 gpurun_out/r02_45_pytest_forward.log
for i in 1 2; do timeout 120 python tools/prof.py forward 65536 f32; done 2>&1 | tail -4
timeout 120 python tools/prof.py forward 16384 f32 2>&1 | tail -1
