set -x
timeout 600 python tools/chain_check.py 48 61 200 2>&1 | tail -12
