timeout 900 python -m pytest tests/test_drivers_gpu.py -m gpu -q -x -k every_solve 2>&1 | grep -E "^E  |assert" | cut -c1-250 | head -12
