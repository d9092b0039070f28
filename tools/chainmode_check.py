"""Stage 2 as one chain-mode launch (cluster = shard, windows back to back on the device) against one launch per wave:
   python tools/chainmode_check.py [frames]"""
import importlib, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = lambda n: importlib.import_module("3dbodyanimation_b200." + n)
synth, api = pkg("synth"), pkg("api")
import torch
torch.cuda.init()
lib = api.lib()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
model_np = synth.make_model()
seq = synth.make_sequence(model_np, n, 31)
gm = api.AvatarModel(model_np)
base = api.fit_opts(cam=(synth.FX, synth.FY, synth.CX, synth.CY))
for shards in (1, 3, 16):
    so = api.seq_opts(skip=10, wsize=20, overlap=5, it1=1000, it2=60, num_shards=shards, beta_pose=20.0, beta_shape=30.0, lambda_temp=3.0)
    out = {}
    for mode in (1, 0, 1, 0):
        lib.b3d_debug_seq_waves(mode)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        g = api.fit_sequences(gm, base, so, seq["kps"], seq["nkp"], seq["poses"], seq["R0"], seq["W"])
        dt = time.perf_counter() - t0
        out[mode] = (g, dt)
    lib.b3d_debug_seq_waves(0)
    a, b = out[1][0], out[0][0]
    same = all(np.array_equal(a[k], b[k]) for k in ("poses", "R0", "W", "r", "p", "solve_iters", "solve_term"))
    dmax = max(float(np.max(np.abs(a[k] - b[k]))) for k in ("poses", "R0", "W", "r", "p"))
    print(f"num_shards {shards}: waves {out[1][1] * 1e3:.1f} ms, chains {out[0][1] * 1e3:.1f} ms, bit-identical {same}, max |diff| {dmax:.2e}, failed {a['nfail']} / {b['nfail']}")
