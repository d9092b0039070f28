set -x
timeout 900 python -m pytest tests/test_fit_window_gpu.py -m gpu -q -x 2>&1 | tail -3
