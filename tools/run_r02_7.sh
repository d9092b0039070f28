set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_fit_window_gpu.py tests/test_drivers_gpu.py -m gpu -q -s > gpurun_out/r02_pytest_win_v4.log 2>&1; echo "pytest rc $?"
grep -E "sharded vs serial|passed|failed|^E " gpurun_out/r02_pytest_win_v4.log | cut -c1-300
python tools/c5_time.py 20000 32
python tools/c5_time.py 20000 96
