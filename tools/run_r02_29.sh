set -x
python tools/prof.py fit 1000 C2 2>&1 | tail -1
python tools/prof.py phases 1000 C2 2>&1 | grep -A3 "line-search fall-backs"
timeout 1500 python -m pytest tests/test_fit_single_gpu.py tests/test_drivers_gpu.py tests/test_shim_gpu.py -m gpu -q -x 2>&1 | tail -2
python bench.py --steps 10 --warmup 3 --no-extra-legs 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().splitlines()[-1]); print(d['value'], d['ms_per_step'], d['e2e']['value'])"
