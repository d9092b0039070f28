set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_icp_gpu.py -m gpu -q -x -s -k "solve" 2>&1 | tail -30
