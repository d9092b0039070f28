mkdir -p gpurun_out
timeout 120 python tools/prof.py tctrace 65536 2>&1 | tail -12
