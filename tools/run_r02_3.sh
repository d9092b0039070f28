set -x
mkdir -p gpurun_out
python -m pytest tests/test_fit_single_gpu.py tests/test_drivers_gpu.py -m gpu -q -s > gpurun_out/r02_pytest_fit_v3.log 2>&1; echo "pytest rc $?"
grep -E "x 100|x 1000|worst|passed|failed|Error" gpurun_out/r02_pytest_fit_v3.log | head -30
python tools/prof.py phases 1000 > gpurun_out/r02_fit_phases_v3.txt 2>&1; head -36 gpurun_out/r02_fit_phases_v3.txt
python tools/prof.py fit 1000 C1
