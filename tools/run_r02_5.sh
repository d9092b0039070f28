timeout 600 python -m pytest tests/test_fit_window_gpu.py tests/test_fit_single_gpu.py -m gpu -q -x 2>&1 | tail -2
python tools/prof.py phases 1000 > gpurun_out/r02_fit_phases_v11.txt 2>&1; head -6 gpurun_out/r02_fit_phases_v11.txt
python tools/prof.py fit 1000 C2
python tools/prof.py winphases 20 | grep -E "cycles per|back sweep"
