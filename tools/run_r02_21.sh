set -x
timeout 600 python tools/chain_check.py 48 97 200 1000 2>&1 | tail -6
python tools/prof.py chainphases 200 2>&1 | tail -18
