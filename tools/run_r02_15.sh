set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_icp_gpu.py -m gpu -q -x -s 2>&1 | tail -30
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
