set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_fit_window_gpu.py tests/test_drivers_gpu.py -m gpu -q -x -s 2>&1 | tail -12
python tools/c3_time.py 2000
python tools/c5_time.py 100000 32 2>&1 | tail -5
