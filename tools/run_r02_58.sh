timeout 250 python -m pytest tests/test_forward_gpu.py -x -q -m gpu --timeout 60 -k "more_than_four" 2>&1 | tail -12
