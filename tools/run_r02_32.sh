set -x
python tools/c3_sharded_time.py 2000 16 2>&1 | tail -6
