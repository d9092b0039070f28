timeout 200 python tools/blocks_check.py 2>&1 | tail -10; timeout 200 python -m pytest tests/test_forward_gpu.py -x -q -m gpu --timeout 60 2>&1 | tail -2
