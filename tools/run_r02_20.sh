set -x
python tools/prof.py chainphases 20 2>&1 | tail -18
python tools/prof.py chainphases 200 2>&1 | tail -18
python tools/prof.py chainphases 4000 2>&1 | tail -18
