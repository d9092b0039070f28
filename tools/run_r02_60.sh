timeout 80 python tools/rig_fuzz.py 300 2>&1 | tail -12
