timeout 250 python -m pytest tests/test_forward_gpu.py tests/test_icp_gpu.py -x -q -m gpu --timeout 60 -k "random_rigs or block_rigged" 2>&1 | tail -8
