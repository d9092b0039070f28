set -x
mkdir -p gpurun_out
./tools/microbench > gpurun_out/r02_microbench.log 2>&1
python -m pytest tests -m gpu -x -q -s > gpurun_out/r02_pytest_gpu_v1.log 2>&1; echo "pytest rc $?"
tail -5 gpurun_out/r02_pytest_gpu_v1.log
python bench.py --steps 10 --warmup 3 > gpurun_out/r02_bench_n1_v1.json 2> gpurun_out/r02_bench_n1_v1.err; echo "bench rc $?"
tail -c 600 gpurun_out/r02_bench_n1_v1.err
python tools/prof.py phases 1000 > gpurun_out/r02_fit_phases_v1.txt 2>&1
