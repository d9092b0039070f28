set -x
mkdir -p gpurun_out
python tools/prof.py phases 1000 > gpurun_out/r02_fit_phases_v9.txt 2>&1; tail -22 gpurun_out/r02_fit_phases_v9.txt
