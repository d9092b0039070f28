set -x
timeout 1500 python -m pytest tests/test_fit_single_gpu.py tests/test_fit_window_gpu.py -m gpu -q -x 2>&1 | tail -6
