set -x
python tools/prof.py phases 1000 C2 2>&1 | tail -22
