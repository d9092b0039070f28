set -x
mkdir -p gpurun_out
python -m pytest tests/test_fit_window_gpu.py tests/test_drivers_gpu.py -m gpu -q -x 2>&1 | tail -8
python tools/prof.py winphases > gpurun_out/r02_win_phases_b.txt 2>&1; tail -20 gpurun_out/r02_win_phases_b.txt
python tools/c3_time.py 2000
