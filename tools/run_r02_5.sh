set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_fit_window_gpu.py tests/test_drivers_gpu.py tests/test_shim_gpu.py -m gpu -q > gpurun_out/r02_pytest_win_v2.log 2>&1; echo "pytest rc $?"
grep -E "passed|failed|^E " gpurun_out/r02_pytest_win_v2.log | cut -c1-300
python tools/c3_time.py 2000
python tools/prof.py winphases 200 | head -3
