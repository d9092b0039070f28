set -x
python tools/prof.py passes 1000 2>&1 | tail -12
