set -x
mkdir -p gpurun_out
python -m pytest tests/test_fit_single_gpu.py tests/test_drivers_gpu.py -m gpu -x -q -s > gpurun_out/r02_pytest_fit_v2.log 2>&1; echo "pytest rc $?"
tail -15 gpurun_out/r02_pytest_fit_v2.log
python tools/prof.py phases 1000 > gpurun_out/r02_fit_phases_v2.txt 2>&1; head -30 gpurun_out/r02_fit_phases_v2.txt
timeout 600 compute-sanitizer --tool racecheck python tools/fit_small.py 3 > gpurun_out/r02_racecheck_v2.log 2>&1; tail -5 gpurun_out/r02_racecheck_v2.log
timeout 600 compute-sanitizer --tool memcheck python tools/fit_small.py 3 > gpurun_out/r02_memcheck_v2.log 2>&1; tail -5 gpurun_out/r02_memcheck_v2.log
