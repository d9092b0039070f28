set -x
N=${NG:-2}
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519 tools/c5_time.py 100000 74 2>&1 | grep frames_per_sec | cut -c1-420
