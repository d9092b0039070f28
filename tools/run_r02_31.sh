set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/r02_pytest_gpu.log 2>&1; echo "pytest rc $?"; tail -2 gpurun_out/r02_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02_smoke.log 2>&1; echo "smoke rc $?"
python bench.py --steps 10 --warmup 3 > gpurun_out/r02_bench_n1.json 2> gpurun_out/r02_bench_n1.err; echo "bench rc $?"
python tools/prof.py fit 1000 C2 > gpurun_out/r02_plain_fit.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:fit_single_kernel -s 3 -c 1 -o gpurun_out/r02_fit_single_final python tools/prof.py fit 1000 C2 > gpurun_out/ncu_fit.log 2>&1
tail -2 gpurun_out/ncu_fit.log; cat gpurun_out/r02_plain_fit.log
python tools/prof.py phases 1000 C2 > gpurun_out/r02_fit_phases.txt 2>&1
