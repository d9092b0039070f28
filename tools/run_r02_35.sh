set -x
timeout 900 python tools/chainmode_check.py 2000 2>&1 | tail -5
