set -x
mkdir -p gpurun_out
timeout 400 python -m pytest tests -x -q -m gpu --timeout 120 > gpurun_out/r02_pytest_gpu.log 2>&1; echo "pytest rc $?"; tail -2 gpurun_out/r02_pytest_gpu.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02_smoke.log 2>&1; echo "smoke rc $?"; tail -1 gpurun_out/r02_smoke.log | cut -c1-200
timeout 400 python bench.py --steps 10 --warmup 3 > gpurun_out/r02_bench_n1.json 2> gpurun_out/r02_bench_n1.err; echo "bench rc $?"
timeout 200 python bench.py --steps 2 --warmup 3 --no-extra-legs > gpurun_out/plain_small.json 2>/dev/null && timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches_bench.csv python bench.py --steps 2 --warmup 3 --no-extra-legs > gpurun_out/ncu_launches.log 2>&1
echo "launch list rc $?"
