python -m pytest tests/test_forward_gpu.py -m gpu -q -x 2>&1 | tail -2
python tools/prof.py forward 65536 f32
python tools/prof.py forward 65536 f32
