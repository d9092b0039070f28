import os, sys, time, importlib
import numpy as np, torch
sys.path.insert(0, "/root/repo")
pkg = lambda n="": importlib.import_module("3dbodyanimation_b200" + ("." + n if n else ""))
synth, api, capi = pkg("synth"), pkg("api"), pkg("_capi")
model_np = synth.make_model()
prior = synth.load_pose_prior("/root/repo/tests/golden/pose_prior.txt")
gm, gg = api.AvatarModel(model_np), api.GaussianMixture(prior=prior)
n = 1000
fr = synth.make_frames(model_np, n, 1000, shape=True, prior=prior)
opts = api.fit_opts(cam=(synth.FX, synth.FY, synth.CX, synth.CY), mode=1, max_iters=100, beta_pose=20.0, beta_shape=30.0)
ws = api.Workspace()
s3, w = fr["sim3"].copy(), fr["w"].copy()
for i in range(6):
    s3[...] = fr["sim3"]; w[...] = fr["w"]
    t0 = time.perf_counter()
    api.fit_frames(ws, gm, gg, opts, fr["kps"], fr["nkp"], fr["R0"], s3, w, cloud_dtype=np.float32, want_cloud=False)
    print("fit_frames call %d: %.2f ms" % (i, (time.perf_counter() - t0) * 1e3), flush=True)
