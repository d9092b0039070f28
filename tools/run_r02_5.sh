set -x
mkdir -p gpurun_out
python -m pytest tests/test_fit_single_gpu.py -m gpu -q > gpurun_out/r02_pytest_fit_v5.log 2>&1; echo "pytest rc $?"
tail -3 gpurun_out/r02_pytest_fit_v5.log
python tools/prof.py fit 1000 C2
python tools/prof.py phases 1000 > gpurun_out/r02_fit_phases_v5.txt 2>&1; head -26 gpurun_out/r02_fit_phases_v5.txt
