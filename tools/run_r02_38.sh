set -x
timeout 600 python tools/chain_check.py 97 200 4000 2>&1 | tail -4
python tools/prof.py chainphases 4000 2>&1 | tail -18
