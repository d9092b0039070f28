set -x
timeout 900 python -m pytest tests/test_fit_window_gpu.py -m gpu -q -x 2>&1 | tail -4
timeout 600 python tools/chain_check.py 200 4000 2>&1 | tail -3
python tools/prof.py chainphases 200 2>&1 | tail -18
python tools/c3_time.py 2000
