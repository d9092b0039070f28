set -x
timeout 1200 python -m pytest tests/test_fit_window_gpu.py tests/test_drivers_gpu.py -m gpu -q -x 2>&1 | tail -3
python tools/c5_time.py 100000 32 2>&1 | tail -1 | cut -c1-400
python tools/c5_time.py 100000 74 2>&1 | tail -1 | cut -c1-400
