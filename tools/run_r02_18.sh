set -x
timeout 900 python tools/chain_check.py 1000 4000 2>&1 | tail -12
