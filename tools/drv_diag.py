"""Diagnostic: driver vs oracle on the fixture frames, per-frame differences."""
import glob, json, os, sys, importlib
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch; torch.cuda.init()
api = importlib.import_module("3dbodyanimation_b200.api"); drv = importlib.import_module("3dbodyanimation_b200.drivers"); synth = importlib.import_module("3dbodyanimation_b200.synth")
from oracle import oracle_api as orc
G = "tests/golden"
m = synth.make_model(with_posedirs=True); prior = synth.load_pose_prior(G + "/pose_prior.txt")
gm = api.AvatarModel(m); gm.posePrior = api.GaussianMixture(prior=prior); om = orc.Model(m); og = orc.GMM(prior=prior)
files = sorted(glob.glob(G + "/keypoints_video1/*.json")); frames = [drv.load_mp_json(f, 480, 270) for f in files]
keep = [i for i, k in enumerate(frames) if k]
kps, nkp = drv.pack_keypoints([frames[i] for i in keep]); cam = drv.intrinsics(480, 270)
for shape in (False, True):
    for iters in (60, 100):
        z = drv.run_single(gm, frames, 480, 270, iters, 20.0, 30.0, opt_shape=shape, use_gmm=shape)
        R0 = np.tile(drv.initial_root_rotation(), (33, 1)); sim3 = np.tile(np.array([1.0, 0, 0, 0, 0, 0, 3.0]), (33, 1))
        for ls, jets in ((orc.LS_DENSE_CHOLESKY, 0), (orc.LS_DENSE_QR, 1)):
            o = om.fit_single(og if shape else None, orc.fit_options(mode=1 if shape else 0, max_iters=iters, beta_pose=20.0, beta_shape=30.0, cam=cam, linear_solver=ls, use_jets=jets), kps, nkp, R0, sim3, np.zeros((33, 10)), nthreads=8)
            print(f"shape={shape} iters={iters} ls={ls}")
            for f in range(33):
                dj = np.max(np.abs(z["jointAA"][f] - o["jointAA"][f])); dr = np.max(np.abs(z["sim3"][f, 1:4] - o["sim3"][f, 1:4]))
                dt = np.max(np.abs(z["sim3"][f, 4:7] / z["sim3"][f, 0] - o["sim3"][f, 4:7] / o["sim3"][f, 0])); ds = abs(z["sim3"][f, 0] - o["sim3"][f, 0])
                n = min(z["niters"][f], o["niters"][f]); rad = o["trace"][f]["radius"][:n].max()
                if dj > 1e-4 or z["niters"][f] != o["niters"][f]:
                    print(f"  f{f:2d} it {z['niters'][f]:3d}/{o['niters'][f]:3d} term {z['term'][f]}/{o['term'][f]} dJ {dj:.2e} dR {dr:.2e} dT/s {dt:.2e} ds {ds:.2e} s {z['sim3'][f,0]:.3f} maxrad {rad:.1e} cost {o['trace'][f]['cost'][o['niters'][f]-1]:.4e}")
