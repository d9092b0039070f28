set -x
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r02_bench_n2_v1.json 2> gpurun_out/r02_bench_n2_v1.err; echo "bench rc $?"
tail -c 400 gpurun_out/r02_bench_n2_v1.err
python -m pytest tests/test_fit_window_gpu.py -m gpu -q -s -k "deviation" 2>&1 | grep -E "sharded vs|passed|failed"
