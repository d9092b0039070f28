set -x
mkdir -p gpurun_out
python tools/prof.py fit 1000 C2 > gpurun_out/r02_plain_fit.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:fit_single_kernel -s 3 -c 1 -o gpurun_out/r02_fit_single_final python tools/prof.py fit 1000 C2 > gpurun_out/ncu_fit.log 2>&1
tail -1 gpurun_out/ncu_fit.log; cat gpurun_out/r02_plain_fit.log
python tools/prof.py phases 1000 C2 > gpurun_out/r02_fit_phases.txt 2>&1
python tools/prof.py winphases > gpurun_out/r02_win_phases.txt 2>&1; tail -17 gpurun_out/r02_win_phases.txt
python tools/prof.py chainphases 200 > gpurun_out/r02_chain_phases_200.txt 2>&1
python tools/prof.py chainphases 4000 > gpurun_out/r02_chain_phases_4000.txt 2>&1
