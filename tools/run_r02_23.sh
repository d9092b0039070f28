set -x
python tools/prof.py chainphases 4000 2>&1 | tail -18
