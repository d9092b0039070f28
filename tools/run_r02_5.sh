set -x
mkdir -p gpurun_out
timeout 300 python tools/fit_small.py 3 2>&1 | tail -2
timeout 900 python -m pytest tests/test_fit_window_gpu.py tests/test_drivers_gpu.py tests/test_shim_gpu.py -m gpu -q -x > gpurun_out/r02_pytest_win_v5.log 2>&1; echo "pytest rc $?"
grep -E "passed|failed|^E " gpurun_out/r02_pytest_win_v5.log | cut -c1-300
timeout 300 python tools/c3_time.py 2000
timeout 300 python tools/prof.py winphases 20 | head -3
