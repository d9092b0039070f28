set -x
timeout 900 python -m pytest tests/test_fit_single_gpu.py -m gpu -q -x -k "polynomial or queue or screening" 2>&1 | tail -40
