set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/r02_pytest_gpu_v2.log 2>&1; echo "pytest rc $?"
tail -3 gpurun_out/r02_pytest_gpu_v2.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02_smoke_v1.log 2>&1; echo "smoke rc $?"; tail -2 gpurun_out/r02_smoke_v1.log
python bench.py --steps 10 --warmup 3 > gpurun_out/r02_bench_n1_v2.json 2> gpurun_out/r02_bench_n1_v2.err; echo "bench rc $?"
tail -c 300 gpurun_out/r02_bench_n1_v2.err
