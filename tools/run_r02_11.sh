set -x
mkdir -p gpurun_out
python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/r02_noalloc_alone.log 2>&1; echo "alone rc $?"; tail -3 gpurun_out/r02_noalloc_alone.log
python -m pytest tests -m gpu -q > gpurun_out/r02_pytest_gpu.log 2>&1; echo "pytest rc $?"
tail -5 gpurun_out/r02_pytest_gpu.log
python tools/c3_time.py 400 > gpurun_out/plain_c3.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:fit_window_kernel -c 1 -o gpurun_out/r02_fit_window_final python tools/c3_time.py 400 > gpurun_out/ncu_win.log 2>&1
tail -2 gpurun_out/ncu_win.log; cat gpurun_out/plain_c3.log
