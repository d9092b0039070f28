"""Parity of the batched single-frame LM driver (b3d_fit_single through the C-ABI) with the oracle.

Tolerances (BASELINE.json north_star): per-iteration residual norms (cost) 1e-4 relative, fitted parameters 1e-3.
Two oracle variants are checked:
  * Cholesky + analytic Jacobians  — the same algebra as the GPU kernel: trajectories must agree tightly;
  * dense Householder QR + Jet autodiff — what the reference's ceres::Solve does (DENSE_QR, DynamicAutoDiffCostFunction).
The reprojection cost is exactly invariant under (s, t) -> a (s, t) (X = s q + t, pixels depend on X/X_z), so J^T J is
singular along that gauge direction and only the LM damping D^2 / radius regularises it.  Once the radius has grown past
~1e12 the step along the gauge is rounding noise in ANY implementation (QR-vs-Cholesky in the oracle itself differ there,
see DESIGN.md "gauge freedom"), so per-iteration checks are made on the prefix with radius <= 1e12, and scale/translation are
compared through the gauge-invariant t / s unless the whole solve stayed in the stable regime."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

COST_TOL = 1e-4     # north_star: per-iteration residual norms to 1e-4
PARAM_TOL = 1e-3    # north_star: fitted pose/shape/Sim3 parameters to 1e-3
STABLE_RADIUS = 1e12


def _cfg(name, synth, model_np, prior_np, n, seed, drop=0.0):
    if name == "C1":
        fr = synth.make_frames(model_np, n, seed, shape=False, drop_prob=drop)
        return fr, dict(mode=0, max_iters=100, beta_pose=20.0, beta_shape=30.0), False
    fr = synth.make_frames(model_np, n, seed, shape=True, prior=prior_np, drop_prob=drop)
    return fr, dict(mode=1, max_iters=100, beta_pose=20.0, beta_shape=30.0), True


def _compare(g, o, tight):
    B = g["niters"].shape[0]
    assert g["usable"].all() and o["usable"].all()
    n_full = 0
    for f in range(B):
        ng, no = int(g["niters"][f]), int(o["niters"][f])
        n = min(ng, no)
        rad = o["trace"][f]["radius"][:n]
        stable = np.cumprod(rad <= STABLE_RADIUS).astype(bool)
        k = int(stable.sum())
        assert k >= 3
        cg, co = g["trace"][f]["cost"][:k], o["trace"][f]["cost"][:k]
        assert np.max(np.abs(cg - co) / np.abs(co)) < COST_TOL, f"frame {f}"
        if tight:
            assert np.array_equal(g["trace"][f]["flags"][:k], o["trace"][f]["flags"][:k])
            assert np.allclose(g["trace"][f]["radius"][:k], o["trace"][f]["radius"][:k], rtol=1e-4)
            assert np.array_equal(g["trace"][f]["aux"][:k], o["trace"][f]["aux"][:k])
        # final cost
        fg, fo = g["trace"][f]["cost"][:ng], o["trace"][f]["cost"][:no]
        assert abs(fg.min() - fo.min()) / fo.min() < COST_TOL
        # gauge-invariant parameters
        assert np.max(np.abs(g["jointAA"][f] - o["jointAA"][f])) < PARAM_TOL
        assert np.max(np.abs(g["sim3"][f, 1:4] - o["sim3"][f, 1:4])) < PARAM_TOL
        assert np.max(np.abs(g["w"][f] - o["w"][f])) < PARAM_TOL
        # translation as t / s, the gauge-invariant combination (DESIGN "Gauge freedom"); it is a length of ~3-4 units (the camera
        # distance), so 1e-3 is taken relative to it
        ts_g, ts_o = g["sim3"][f, 4:7] / g["sim3"][f, 0], o["sim3"][f, 4:7] / o["sim3"][f, 0]
        assert np.max(np.abs(ts_g - ts_o)) < PARAM_TOL * max(1.0, float(np.max(np.abs(ts_o))))
        if k == n and ng == no:     # the whole solve stayed in the stable regime: scale and translation themselves
            n_full += 1
            assert np.max(np.abs(g["sim3"][f] - o["sim3"][f])) < PARAM_TOL
            assert g["term"][f] == o["term"][f]
    return n_full


@pytest.mark.parametrize("name", ["C1", "C2"])
def test_fit_single_vs_oracle(api, orc, gmodel, ggmm, omodel, ogmm, synth, model_np, prior_np, name):
    fr, kw, use_gmm = _cfg(name, synth, model_np, prior_np, 48, 21)
    cam = (synth.FX, synth.FY, synth.CX, synth.CY)
    g = api.fit_single_batch(gmodel, ggmm if use_gmm else None, api.fit_opts(cam=cam, **kw), fr["kps"], fr["nkp"], fr["R0"], fr["sim3"], fr["w"])
    o_ch = omodel.fit_single(ogmm if use_gmm else None, orc.fit_options(cam=cam, linear_solver=orc.LS_DENSE_CHOLESKY, use_jets=0, **kw),
                             fr["kps"], fr["nkp"], fr["R0"], fr["sim3"], fr["w"], nthreads=8)
    o_qr = omodel.fit_single(ogmm if use_gmm else None, orc.fit_options(cam=cam, linear_solver=orc.LS_DENSE_QR, use_jets=1, **kw),
                             fr["kps"], fr["nkp"], fr["R0"], fr["sim3"], fr["w"], nthreads=8)
    n1 = _compare(g, o_ch, tight=True)
    n2 = _compare(g, o_qr, tight=False)
    assert n1 >= 24 and n2 >= 24      # most frames converge before the gauge noise regime
    # recovered pose is close to the ground truth pose that generated the keypoints (sanity of the whole pipeline)
    gt = fr["gt"]
    if name == "C2":
        assert np.median(np.abs(g["sim3"][:, 1:4] - gt[:, 1:4])) < 0.2


def test_fit_single_ragged_and_degenerate(api, orc, gmodel, ggmm, omodel, ogmm, synth, model_np, prior_np):
    """Dropped keypoints (ragged nkp), a frame with a single keypoint, scale starting outside its bounds."""
    fr, kw, _ = _cfg("C2", synth, model_np, prior_np, 12, 5, drop=0.3)
    fr["nkp"][3] = 1
    fr["sim3"][4, 0] = 5.0     # projected onto [0.3, 3] at iteration 0
    cam = (synth.FX, synth.FY, synth.CX, synth.CY)
    g = api.fit_single_batch(gmodel, ggmm, api.fit_opts(cam=cam, **kw), fr["kps"], fr["nkp"], fr["R0"], fr["sim3"], fr["w"])
    o = omodel.fit_single(ogmm, orc.fit_options(cam=cam, linear_solver=orc.LS_DENSE_CHOLESKY, use_jets=0, **kw),
                          fr["kps"], fr["nkp"], fr["R0"], fr["sim3"], fr["w"], nthreads=8)
    _compare(g, o, tight=True)
    assert np.all(g["sim3"][:, 0] >= 0.3) and np.all(g["sim3"][:, 0] <= 3.0)


def test_fit_single_l2_prior_and_no_gmm_degrade(api, orc, gmodel, omodel, synth, model_np, prior_np):
    """gmm == NULL with --use-gmm semantics degrades silently to the L2 prior (Sim3BA.h:266); max_iters cap is honoured."""
    fr, kw, _ = _cfg("C2", synth, model_np, prior_np, 8, 9)
    kw["max_iters"] = 7
    cam = (synth.FX, synth.FY, synth.CX, synth.CY)
    g = api.fit_single_batch(gmodel, None, api.fit_opts(cam=cam, **kw), fr["kps"], fr["nkp"], fr["R0"], fr["sim3"], fr["w"])
    o = omodel.fit_single(None, orc.fit_options(cam=cam, linear_solver=orc.LS_DENSE_CHOLESKY, use_jets=0, **kw),
                          fr["kps"], fr["nkp"], fr["R0"], fr["sim3"], fr["w"], nthreads=8)
    assert np.array_equal(g["niters"], o["niters"]) and np.array_equal(g["term"], o["term"])
    assert (g["term"] == 4).any()
    for f in range(8):
        n = g["niters"][f]
        assert np.max(np.abs(g["trace"][f]["cost"][:n] - o["trace"][f]["cost"][:n]) / o["trace"][f]["cost"][:n]) < 1e-8
        assert np.all(g["trace"][f]["aux"][:n] == -1)


def test_reference_style_call(api, orc, gmodel, ggmm, omodel, ogmm, synth, model_np, prior_np):
    """The per-frame call of src/main_single_frame.cpp:205-255 through the mirrored interface."""
    fr, kw, _ = _cfg("C2", synth, model_np, prior_np, 1, 33)
    av = api.Avatar(gmodel)
    av.p = np.array([0.0, 0.0, 3.0])
    av.r[0] = synth.initial_root_rotation().T.reshape(9)
    av.update()
    kps = [api.PixelKP(int(k["jid"]), float(k["u"]), float(k["v"])) for k in fr["kps"][0][:fr["nkp"][0]]]
    sim3 = api.Sim3Params(np.array([1.0, 0, 0, 0, 0, 0, 3.0]))
    ok, report = api.OptimizePoseShapeReprojection(gmodel, av, kps, synth.FX, synth.FY, synth.CX, synth.CY, [k.jid for k in kps], sim3,
                                                   100, 20.0, 30.0, ggmm)
    assert ok and "termination" in report
    av.update()
    err = api.mean_pixel_error(kps, av, synth.FX, synth.FY, synth.CX, synth.CY)
    # oracle: same solve, write-back, update, pixel error in numpy
    o = omodel.fit_single(ogmm, orc.fit_options(cam=(synth.FX, synth.FY, synth.CX, synth.CY), linear_solver=orc.LS_DENSE_CHOLESKY, use_jets=0, **kw),
                          fr["kps"], fr["nkp"], fr["R0"], fr["sim3"], fr["w"])
    pose = np.concatenate([o["sim3"][0], o["jointAA"][0, 1:].ravel()])
    R0, r, p = orc.writeback(pose[None], fr["R0"][:1])
    _, jp, _ = omodel.forward(o["w"], p, r)
    e0 = np.mean([np.hypot(synth.FX * jp[0, k.jid, 0] / jp[0, k.jid, 2] + synth.CX - k.u, synth.FY * jp[0, k.jid, 1] / jp[0, k.jid, 2] + synth.CY - k.v) for k in kps])
    assert abs(err - e0) < 1e-3 * max(1.0, e0)
    assert np.max(np.abs(av.r - r[0])) < 1e-3 and np.max(np.abs(sim3.data[1:4] - o["sim3"][0, 1:4])) < 1e-3


def test_fit_frames_pipeline(api, orc, gmodel, ggmm, omodel, ogmm, synth, model_np, prior_np):
    """b3d_fit_frames = fit -> write-back -> update -> pixel error in one call equals the staged calls."""
    fr, kw, _ = _cfg("C2", synth, model_np, prior_np, 20, 17)
    cam = (synth.FX, synth.FY, synth.CX, synth.CY)
    o = api.fit_opts(cam=cam, **kw)
    ws = api.Workspace()
    res = api.fit_frames(ws, gmodel, ggmm, o, fr["kps"], fr["nkp"], fr["R0"], fr["sim3"].copy(), fr["w"].copy(), want_cloud=True,
                         cloud_dtype=np.float64)
    g = api.fit_single_batch(gmodel, ggmm, o, fr["kps"], fr["nkp"], fr["R0"], fr["sim3"], fr["w"])
    assert np.array_equal(res["sim3"], g["sim3"]) and np.array_equal(res["jointAA"], g["jointAA"]) and np.array_equal(res["w"], g["w"])
    assert np.array_equal(res["niters"], g["niters"]) and res["usable"].all()
    poses = np.concatenate([g["sim3"], g["jointAA"][:, 1:].reshape(20, 69)], axis=1)
    R0, r, p = orc.writeback(poses, fr["R0"])
    c0, jp0, _ = omodel.forward(g["w"], p, r)
    assert np.max(np.abs(res["r"] - r)) < 1e-13 and np.max(np.abs(res["p"] - p)) == 0.0
    assert np.max(np.abs(res["cloud"] - c0)) / np.max(np.abs(c0)) < 1e-12
    e0 = np.array([np.mean([np.hypot(synth.FX * jp0[f, k["jid"], 0] / jp0[f, k["jid"], 2] + synth.CX - k["u"],
                                     synth.FY * jp0[f, k["jid"], 1] / jp0[f, k["jid"], 2] + synth.CY - k["v"]) for k in fr["kps"][f][:fr["nkp"][f]]])
                   for f in range(20)])
    assert np.max(np.abs(res["px_err"] - e0)) < 1e-9
    # float32 cloud variant and a second call on the same workspace (buffers reused)
    res32 = api.fit_frames(ws, gmodel, ggmm, o, fr["kps"], fr["nkp"], fr["R0"], fr["sim3"].copy(), fr["w"].copy(), want_cloud=True)
    assert res32["cloud"].dtype == np.float32 and np.max(np.abs(res32["cloud"] - c0)) / np.max(np.abs(c0)) < 1e-5


GAUGE_COST_TOL = 2e-3    # final cost of frames that left the stable regime (see _regimes)


def _regimes(g, o):
    """Per-frame comparison of two solves of the same frames.  A frame is 'stable' when the second solve's whole trace stayed at
    radius <= 1e12 and both took the same number of iterations; otherwise it is in the 'gauge' regime: the trust region has opened so
    far that the step along the (s, t) -> a (s, t) null direction of the reprojection term — and along any other direction the 17
    keypoints barely constrain — is rounding noise in ANY implementation (DESIGN.md "gauge freedom"); such trajectories separate
    after the stable prefix and can end in different points of the same flat valley.  Returns boolean stable[], and per frame the
    worst relative cost deviation on the stable prefix, the relative final-cost deviation, the worst deviation of the gauge-invariant
    parameters (jointAA, rootAA, w) and of the whole Sim3 block."""
    B = g["niters"].shape[0]
    stable = np.zeros(B, dtype=bool)
    prefix, final, pose, sim3 = np.zeros(B), np.zeros(B), np.zeros(B), np.zeros(B)
    for f in range(B):
        ng, no = int(g["niters"][f]), int(o["niters"][f])
        n = min(ng, no)
        k = int(np.cumprod(o["trace"][f]["radius"][:n] <= STABLE_RADIUS).sum())
        stable[f] = (k == n and ng == no)
        cg, co = g["trace"][f]["cost"][:k], o["trace"][f]["cost"][:k]
        prefix[f] = np.max(np.abs(cg - co) / np.abs(co))
        fg, fo = g["trace"][f]["cost"][:ng].min(), o["trace"][f]["cost"][:no].min()
        final[f] = abs(fg - fo) / fo
        pose[f] = max(np.max(np.abs(g["jointAA"][f] - o["jointAA"][f])), np.max(np.abs(g["sim3"][f, 1:4] - o["sim3"][f, 1:4])),
                      np.max(np.abs(g["w"][f] - o["w"][f])))
        sim3[f] = np.max(np.abs(g["sim3"][f] - o["sim3"][f]))
    return dict(stable=stable, prefix=prefix, final=final, pose=pose, sim3=sim3)


def _report(label, d):
    st, ga = d["stable"], ~d["stable"]
    q = lambda a: "-" if a.size == 0 else f"median {np.median(a):.1e} p90 {np.quantile(a, 0.9):.1e} max {a.max():.1e}"
    return (f"{label}: stable {int(st.sum())} gauge {int(ga.sum())}\n"
            f"      stable frames: cost prefix max {d['prefix'][st].max():.1e}, final cost max {d['final'][st].max():.1e}, pose {q(d['pose'][st])}, sim3 max {d['sim3'][st].max():.1e}\n"
            f"      gauge  frames: cost prefix max {d['prefix'][ga].max() if ga.any() else 0:.1e}, final cost {q(d['final'][ga])}, pose {q(d['pose'][ga])}; "
            f"frames with pose deviation > 1e-3: {int((d['pose'][ga] > PARAM_TOL).sum())}")


@pytest.mark.parametrize("name,nframes", [("C1", 100), ("C2", 1000)])
def test_fit_single_baseline_config_at_stated_size(api, orc, gmodel, ggmm, omodel, ogmm, synth, model_np, prior_np, name, nframes, capsys):
    """BASELINE configs[0] (100 frames, pose-only) and configs[1] (1,000 frames, --opt-shape --use-gmm) at their stated sizes against
    both oracle variants (same seeds as bench.py: 2000 / 1000).  Every frame: per-iteration costs to 1e-4 on the stable prefix of the
    trace.  Stable frames: final cost 1e-4 and every fitted parameter to 1e-3 (the north_star tolerances).  Gauge frames: final cost
    within GAUGE_COST_TOL, and the number of frames whose gauge-invariant parameters differ by more than 1e-3 is judged against the same
    count between the oracle's OWN two variants (dense QR + Jets vs Cholesky + analytic Jacobians) — the implementation noise of
    the reference's algorithm on these frames.  All counts are printed."""
    fr, kw, use_gmm = _cfg(name, synth, model_np, prior_np, nframes, 2000 if name == "C1" else 1000)
    cam = (synth.FX, synth.FY, synth.CX, synth.CY)
    g = api.fit_single_batch(gmodel, ggmm if use_gmm else None, api.fit_opts(cam=cam, **kw), fr["kps"], fr["nkp"], fr["R0"], fr["sim3"], fr["w"])
    th = len(os.sched_getaffinity(0))
    o_ch = omodel.fit_single(ogmm if use_gmm else None, orc.fit_options(cam=cam, linear_solver=orc.LS_DENSE_CHOLESKY, use_jets=0, **kw),
                             fr["kps"], fr["nkp"], fr["R0"], fr["sim3"], fr["w"], nthreads=th)
    o_qr = omodel.fit_single(ogmm if use_gmm else None, orc.fit_options(cam=cam, linear_solver=orc.LS_DENSE_QR, use_jets=1, **kw),
                             fr["kps"], fr["nkp"], fr["R0"], fr["sim3"], fr["w"], nthreads=th)
    assert g["usable"].all() and o_ch["usable"].all() and o_qr["usable"].all()
    d_ch, d_qr, d_oo = _regimes(g, o_ch), _regimes(g, o_qr), _regimes(o_ch, o_qr)
    with capsys.disabled():
        print(f"\n[{name} x {nframes}] iterations: gpu mean {g['niters'].mean():.1f} max {g['niters'].max()}, at the cap: {int((g['term'] == 4).sum())}")
        print("   " + _report("GPU vs oracle Cholesky/analytic", d_ch))
        print("   " + _report("GPU vs oracle dense-QR/Jets    ", d_qr))
        print("   " + _report("oracle Cholesky vs oracle QR   ", d_oo))
    noise = int(((d_oo["pose"] > PARAM_TOL) & ~d_oo["stable"]).sum())
    for d in (d_ch, d_qr):
        st, ga = d["stable"], ~d["stable"]
        assert st.sum() >= nframes // 2                      # most frames converge before the trust region opens that far
        assert d["prefix"].max() < COST_TOL                  # per-iteration residual norms, every frame, stable prefix
        assert d["final"][st].max() < COST_TOL and d["pose"][st].max() < PARAM_TOL and d["sim3"][st].max() < PARAM_TOL
        if ga.any():
            assert d["final"][ga].max() < GAUGE_COST_TOL
            assert int((d["pose"][ga] > PARAM_TOL).sum()) <= 2 * noise + 5


@pytest.mark.parametrize("name", ["C1", "C2"])
def test_fit_single_queue_schedule_is_bit_identical(api, gmodel, ggmm, synth, model_np, prior_np, name):
    """Batches larger than the machine's CTA slots run through a persistent kernel: a fresh frame checkpoints itself after a capped
    number of iterations and is resumed from a device-side queue (csrc/fit_single.cu FitCkpt).  Whatever the cap, every output and
    every trace record must equal the one-CTA-per-frame solve bit for bit."""
    nfr = 700          # more frames than 2 CTAs x 148 SMs
    fr, kw, use_gmm = _cfg(name, synth, model_np, prior_np, nfr, 77, drop=0.1)
    cam = (synth.FX, synth.FY, synth.CX, synth.CY)
    lib = api.lib()
    run = lambda: api.fit_single_batch(gmodel, ggmm if use_gmm else None, api.fit_opts(cam=cam, **kw), fr["kps"], fr["nkp"], fr["R0"], fr["sim3"], fr["w"])
    try:
        lib.b3d_debug_fit_stop_at(0)
        ref = run()
        assert ref["niters"].max() > 40          # some frames run long past the caps below
        for cap in (1, 7, 20, 33):
            lib.b3d_debug_fit_stop_at(cap)
            got = run()
            for k in ("sim3", "w", "jointAA", "usable", "niters", "term"):
                assert np.array_equal(got[k], ref[k]), (cap, k)
            assert got["trace"].tobytes() == ref["trace"].tobytes(), cap
    finally:
        lib.b3d_debug_fit_stop_at(-1)


def test_gmm_component_screening_is_bit_identical(api, gmodel, ggmm, synth, model_np, prior_np):
    """The kernel skips GMM components that provably cannot be the argmin of GaussianMixture::residual (norm bounds from the last
    exact evaluation of each component).  Switching the screening off must not change a single bit of any output or trace record."""
    fr, kw, _ = _cfg("C2", synth, model_np, prior_np, 96, 314, drop=0.05)
    cam = (synth.FX, synth.FY, synth.CX, synth.CY)
    lib = api.lib()
    run = lambda: api.fit_single_batch(gmodel, ggmm, api.fit_opts(cam=cam, **kw), fr["kps"], fr["nkp"], fr["R0"], fr["sim3"], fr["w"])
    try:
        assert lib.b3d_debug_gmm_screen(0) == 0
        ref = run()
        assert lib.b3d_debug_gmm_screen(1) == 0
        got = run()
    finally:
        lib.b3d_debug_gmm_screen(1)
    for k in ("sim3", "w", "jointAA", "usable", "niters", "term"):
        assert np.array_equal(got[k], ref[k]), k
    assert got["trace"].tobytes() == ref["trace"].tobytes()
    assert len(set(int(a) for f in range(96) for a in ref["trace"][f]["aux"][:ref["niters"][f]])) >= 2      # more than one component is active


def test_line_search_polynomial_minimiser_vs_numpy(api):
    """ceres::MinimizeInterpolatingPolynomial as the Armijo search runs it on the device (ls_poly.cuh: Hermite interpolant through two or
    three samples with gradients, real roots of its derivative inside [x_min, x_max] by exact isolation + whole-warp refinement) against
    numpy on random samples: the chosen step must give the interpolant's minimum over the candidates Ceres tries (mid point, end
    points, real parts of ALL roots of the derivative that fall inside the interval) to 1e-6 of the polynomial's range, and the step
    itself must agree to 1e-9 wherever that minimum is not a near-tie."""
    import ctypes as C
    rng = np.random.default_rng(5)
    n = 4000
    rows = np.zeros((n, 12))
    for i in range(n):
        ns = 2 if i % 3 == 0 else 3
        xs = np.sort(rng.uniform(0.0, 1.0, 3)) * rng.choice([1.0, 1e-3, 10.0])
        xs[0] = 0.0 if i % 2 == 0 else xs[0]
        xs += np.arange(3) * 1e-3 * max(xs[-1], 1e-3)           # distinct nodes
        vals = rng.normal(0, 1, 3) * rng.choice([1.0, 1e3, 1e-3])
        grads = rng.normal(0, 1, 3) * np.abs(vals).max() / max(xs[-1] - xs[0], 1e-6)
        lo, hi = sorted(rng.uniform(xs[0], xs[-1] * 1.2 + 1e-6, 2))
        rows[i, 0] = ns
        for k in range(3):
            rows[i, 1 + 3 * k:4 + 3 * k] = (xs[k], vals[k], grads[k])
        rows[i, 10], rows[i, 11] = lo, hi + 1e-9
    out = np.zeros(n)
    assert api.lib().b3d_debug_poly_min(n, rows.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p)) == 0
    worst_val, nclose = 0.0, 0
    for i in range(n):
        ns = int(rows[i, 0])
        xs, vals, grads = rows[i, 1:1 + 3 * ns:3], rows[i, 2:2 + 3 * ns:3], rows[i, 3:3 + 3 * ns:3]
        deg = 2 * ns - 1
        A, b = [], []
        for x, v, g in zip(xs, vals, grads):
            A.append([x ** (deg - j) for j in range(deg + 1)]); b.append(v)
            A.append([(deg - j) * x ** (deg - j - 1) if deg - j > 0 else 0.0 for j in range(deg + 1)]); b.append(g)
        coef = np.linalg.solve(np.array(A), np.array(b))
        lo, hi = rows[i, 10], rows[i, 11]
        cand = [0.5 * (lo + hi), lo, hi] + [float(z.real) for z in np.roots(np.polyder(coef)) if lo <= z.real <= hi]
        pv = np.polyval(coef, cand)
        best = float(np.min(pv))
        scale = max(float(np.ptp(np.polyval(coef, np.linspace(lo, hi, 64)))), 1e-300)
        got_v = float(np.polyval(coef, out[i]))
        assert lo <= out[i] <= hi
        worst_val = max(worst_val, (got_v - best) / scale)
        second = np.sort(pv)[1] if len(pv) > 1 else best
        if (second - best) / scale > 1e-6:        # a clear winner: the step itself must match
            nclose += 1
            assert abs(out[i] - cand[int(np.argmin(pv))]) <= 1e-9 * max(abs(hi), 1e-300) + 1e-7 * (hi - lo), (i, out[i], cand, pv)
    assert worst_val < 1e-6, worst_val      # (both sides evaluate an ill-conditioned monomial form: coefficients up to 1e15 at x ~ 1e-3)
    assert nclose > n // 2
