set -x
mkdir -p gpurun_out
python tools/c3_time.py 400 > gpurun_out/plain_c3.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:fit_window_kernel -s 1 -c 1 -o gpurun_out/r02_fit_window_final python tools/c3_time.py 400 > gpurun_out/ncu_win.log 2>&1
tail -1 gpurun_out/ncu_win.log; cat gpurun_out/plain_c3.log
