set -x
timeout 1500 python -m pytest tests/test_fit_single_gpu.py tests/test_drivers_gpu.py -m gpu -q -x 2>&1 | tail -12
python tools/prof.py fit 1000 C2 2>&1 | tail -1
python tools/prof.py phases 1000 C2 2>&1 | grep -A4 "line-search fall-backs"
