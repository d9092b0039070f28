"""Exact and sharded C3 schedules timed several times in one process: python tools/c3_sharded_time.py [frames] [shards]"""
import importlib, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = lambda n: importlib.import_module("3dbodyanimation_b200." + n)
synth, api = pkg("synth"), pkg("api")
import torch
torch.cuda.init()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
G = int(sys.argv[2]) if len(sys.argv) > 2 else 16
model_np = synth.make_model()
seq = synth.make_sequence(model_np, n, 31)
gm = api.AvatarModel(model_np)
base = api.fit_opts(cam=(synth.FX, synth.FY, synth.CX, synth.CY))
for shards in (1, G, G, 1, G):
    so = api.seq_opts(skip=10, wsize=20, overlap=5, it1=1000, it2=60, num_shards=shards, beta_pose=20.0, beta_shape=30.0, lambda_temp=3.0)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    g = api.fit_sequences(gm, base, so, seq["kps"], seq["nkp"], seq["poses"], seq["R0"], seq["W"])
    dt = time.perf_counter() - t0
    print(f"num_shards {shards}: {dt * 1e3:.1f} ms, {n / dt:.0f} frames/s, solves {len(g['solve_iters'])}, LM iterations total {int(np.sum(g['solve_iters']))}, failed {g['nfail']}")
