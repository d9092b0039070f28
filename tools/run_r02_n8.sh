set -x
mkdir -p gpurun_out
N=${NG:-8}
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/r02_bench_n$N.json 2> gpurun_out/r02_bench_n$N.err; echo "bench rc $?"
tail -c 600 gpurun_out/r02_bench_n$N.err
head -c 1500 gpurun_out/r02_bench_n$N.json
