set -x
mkdir -p gpurun_out
python tools/prof.py fit 1000 C2 > gpurun_out/plain_fit.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:fit_single_kernel -s 3 -c 1 -o gpurun_out/r02_fit_single_v3 python tools/prof.py fit 1000 C2 > gpurun_out/ncu_fit.log 2>&1
tail -3 gpurun_out/ncu_fit.log; cat gpurun_out/plain_fit.log
