set -x
python tools/prof.py fit 1000 C2 2>&1 | tail -2
timeout 1200 python -m pytest tests/test_fit_single_gpu.py -m gpu -q -x 2>&1 | tail -4
python tools/prof.py phases 1000 C2 2>&1 | grep -A2 "line-search fall-backs"
