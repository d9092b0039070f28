set -x
python tools/c5_time.py 100000 74 2>&1 | tail -2
