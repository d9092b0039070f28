set -x
mkdir -p gpurun_out
timeout 400 python -m pytest tests -x -q -m gpu --timeout 120 > gpurun_out/r02_pytest_gpu.log 2>&1; echo "pytest rc $?"; tail -2 gpurun_out/r02_pytest_gpu.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02_smoke.log 2>&1; echo "smoke rc $?"; tail -1 gpurun_out/r02_smoke.log | cut -c1-300
timeout 400 python bench.py --steps 10 --warmup 3 > gpurun_out/r02_bench_n1.json 2> gpurun_out/r02_bench_n1.err; echo "bench rc $?"
