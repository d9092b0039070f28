"""A few single-frame and window fits through the C-ABI (for compute-sanitizer runs): `python tools/fit_small.py [frames]`."""
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = lambda n: importlib.import_module("3dbodyanimation_b200." + n)
synth, api = pkg("synth"), pkg("api")

n = int(sys.argv[1]) if len(sys.argv) > 1 else 6
model_np = synth.make_model()
prior = synth.load_pose_prior(os.path.join(ROOT, "tests", "golden", "pose_prior.txt"))
gm, gg = api.AvatarModel(model_np), api.GaussianMixture(prior=prior)
cam = (synth.FX, synth.FY, synth.CX, synth.CY)
for name, shape, g, kw in (("C2", True, gg, dict(mode=1, max_iters=100, beta_pose=20.0, beta_shape=30.0)),
                           ("C1", False, None, dict(mode=0, max_iters=100, beta_pose=20.0, beta_shape=30.0))):
    fr = synth.make_frames(model_np, n, 123, shape=shape, prior=prior if shape else None, drop_prob=0.2)
    r = api.fit_single_batch(gm, g, api.fit_opts(cam=cam, **kw), fr["kps"], fr["nkp"], fr["R0"], fr["sim3"], fr["w"])
    print(name, "iters", r["niters"].tolist(), "term", r["term"].tolist(), "final cost", [float(r["trace"][f]["cost"][r["niters"][f] - 1]) for f in range(n)])
seq = synth.make_sequence(model_np, 8, 5, drop_prob=0.1)
w = api.fit_window(gm, api.fit_opts(cam=cam, max_iters=12, beta_pose=20.0, beta_shape=30.0, lambda_temp=3.0), seq["kps"], seq["nkp"], seq["R0"], seq["poses"], seq["W"][0])
print("window iters", w["niters"], "term", w["term"], "final cost", float(w["trace"]["cost"][-1]))
