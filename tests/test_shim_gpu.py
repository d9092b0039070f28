"""The reference-signature C++ shim (include/b3d_shim.hpp) on a GPU: OptimizePoseReprojection, OptimizePoseShapeReprojection,
OptimizeMultiFrame and mean_pixel_error called from a C++ program give the numbers the ctypes path gives on the same inputs
(both sit on the same C-ABI, so agreement is to rounding of the decimal round trip: 1e-12), and the public GMM entry points
b3d_gmm_get / b3d_gmm_residual (include/b3d.h:112-124) agree with the oracle on the reference's real pose prior."""
import os
import shutil
import subprocess

import numpy as np
import pytest

from conftest import PRIOR_TXT, ROOT

pytestmark = pytest.mark.gpu


def _parse(stdout):
    out = {}
    for line in stdout.splitlines():
        tok = line.split()
        if len(tok) >= 3 and tok[1] in ("sim3", "w", "r", "px_err", "pose", "jointPos"):
            out.setdefault(tok[0] + " " + tok[1], []).append(np.array([float(x) for x in tok[2:]]))
        elif len(tok) >= 3 and tok[1] == "usable":
            out[tok[0] + " usable"] = int(tok[2])
    return out


def test_shim_calls_equal_ctypes_path(api, gmodel, ggmm, synth, model_np, prior_np, tmp_path):
    cxx = shutil.which("g++")
    if cxx is None:
        pytest.skip("no g++")
    nf = 6
    seq = synth.make_sequence(model_np, nf, 12)
    R0 = seq["R0"][0]
    blob = tmp_path / "in.bin"
    with open(blob, "wb") as f:
        f.write(np.array([synth.V, nf], dtype=np.int32).tobytes())
        for k in ("v_template", "shapedirs", "J_regressor", "weights"):
            f.write(np.ascontiguousarray(model_np[k], dtype=np.float32).tobytes())
        f.write(np.ascontiguousarray(model_np["kintree_table"][0], dtype=np.uint32).tobytes())
        for i in range(nf):
            f.write(np.array([seq["nkp"][i]], dtype=np.int32).tobytes())
            f.write(np.ascontiguousarray(seq["kps"][i]).tobytes())
        f.write(np.ascontiguousarray(R0, dtype=np.float64).tobytes())
        f.write(PRIOR_TXT.encode() + b"\0")
    libdir = os.path.join(ROOT, "3dbodyanimation_b200")
    exe = str(tmp_path / "shim_parity")
    res = subprocess.run([cxx, "-std=c++14", "-O1", "-Wall", "-I" + os.path.join(ROOT, "include"), os.path.join(ROOT, "tests", "shim_parity.cpp"), "-o", exe,
                          "-L" + libdir, "-l:lib3dba_b200.so", "-Wl,-rpath," + libdir], capture_output=True, text=True)
    assert res.returncode == 0, res.stderr
    run = subprocess.run([exe, str(blob)], capture_output=True, text=True)
    assert run.returncode == 0, run.stdout + run.stderr
    got = _parse(run.stdout)

    cam = (432.0, 432.0, 240.0, 135.0)
    n0 = int(seq["nkp"][0])
    kps0 = [api.PixelKP(int(k["jid"]), float(k["u"]), float(k["v"])) for k in seq["kps"][0][:n0]]
    for mode, fn, gmm in ((0, api.OptimizePoseReprojection, None), (1, api.OptimizePoseShapeReprojection, ggmm)):
        av = api.Avatar(gmodel)
        av.r[0] = R0.copy()
        av.p = np.array([0.0, 0.0, 3.0])
        av.update()
        s3 = api.Sim3Params(np.array([1.0, 0, 0, 0, 0, 0, 3.0]))
        ok, _ = fn(gmodel, av, kps0, *cam, [k.jid for k in kps0], s3, 100, 20.0, 30.0, gmm)
        av.update()
        err = api.mean_pixel_error(kps0, av, *cam)
        assert got[f"single{mode} usable"] == int(ok)
        assert np.max(np.abs(got[f"single{mode} sim3"][0] - s3.data)) < 1e-12
        assert np.max(np.abs(got[f"single{mode} w"][0] - av.w)) < 1e-12
        assert np.max(np.abs(got[f"single{mode} r"][0] - av.r.ravel())) < 1e-12
        assert abs(got[f"single{mode} px_err"][0][0] - err) < 1e-12 * max(1.0, err)
    # OptimizeMultiFrame: same solve through api.fit_window + write-back + update
    kw = dict(max_iters=40, beta_pose=20.0, beta_shape=30.0, lambda_temp=3.0)
    g = api.fit_window(gmodel, api.fit_opts(mode=1, cam=cam, **kw), seq["kps"], seq["nkp"], seq["R0"], seq["poses"], np.zeros(10))
    assert got["multi usable"] == int(g["usable"])
    assert np.max(np.abs(got["multi w"][0] - g["w"])) < 1e-12
    assert np.max(np.abs(np.stack(got["multi pose"]) - g["poses"])) < 1e-12
    _, r, p = api.writeback(g["poses"], seq["R0"])
    _, jp, _ = api.smpl_forward(gmodel, np.tile(np.zeros(10), (nf, 1)), p, r, dtype=np.float64)
    # the shim's avatars keep their own w (only the front avatar receives the optimised block, include/MultiFrameBA.h:67)
    _, jp0, _ = api.smpl_forward(gmodel, g["w"][None], p[:1], r[:1], dtype=np.float64)
    jp[0] = jp0[0]
    assert np.max(np.abs(np.stack(got["multi jointPos"]) - jp.reshape(nf, 72))) < 1e-12


def test_gmm_get_and_residual_vs_oracle(api, ggmm, ogmm, prior_np):
    """b3d_gmm_get (prec_cho = chol(cov^-1) lower, consts_log) and b3d_gmm_residual (GaussianMixture.cpp:94-113) on the GPU against
    the oracle on the reference's real pose_prior.txt: 8 components x 69 dims."""
    assert (ggmm.nComps, ggmm.nDims) == (8, 69)
    ga, oa = ggmm.arrays(), ogmm.arrays()
    scale = np.max(np.abs(oa["prec_cho"]))
    assert np.max(np.abs(ga["prec_cho"] - oa["prec_cho"])) < 1e-9 * scale
    assert np.max(np.abs(ga["consts_log"] - oa["consts_log"])) < 1e-9 * np.max(np.abs(oa["consts_log"]))
    rng = np.random.default_rng(3)
    # poses near every component mean (each component wins somewhere) plus far-away and zero poses
    xs = [prior_np["means"][c] + 0.05 * rng.normal(size=69) for c in range(8)] + [rng.normal(0, 0.4, size=69) for _ in range(24)] + [np.zeros(69)]
    x = np.stack(xs)
    out, comp = ggmm.residual(x)
    seen = set()
    for i in range(x.shape[0]):
        o, c = ogmm.residual(x[i])
        assert comp[i] == c
        seen.add(int(c))
        assert np.max(np.abs(out[i] - o)) < 1e-9 * max(1.0, np.max(np.abs(o)))
    assert len(seen) >= 4
    # single-vector form of the mirrored interface
    o1, c1 = ggmm.residual(x[0])
    assert o1.shape == (1, 70) and c1[0] == comp[0]
