"""Oracle self-checks for the vertex-level ICP term (oracle/icp.cpp restating
external/avatar/AvatarOptimizer.cpp:161-641).  No reference golden vectors exist for this path (parity unpinned): the analytic
Jacobians are pinned by central differences through the reference's own Plus() and the residuals by Avatar::update."""
import numpy as np


def _one(synth, model_np, seed=5, nd=600):
    return synth.make_icp_inputs(model_np, 1, seed, ndata=nd)


def test_icp_residual_is_cloud_minus_data(synth, orc, model_np):
    om = orc.Model(model_np)
    x = _one(synth, model_np)
    out = om.icp_eval(x["w"], x["p"], x["q"], x["cptr"], x["mpt"], x["dpt"], x["dptr"], x["data"], want_jac=True)
    cloud, _, _ = om.forward(x["w"], x["p"], synth.quat_to_r(x["q"]))
    ref = cloud[0][x["mpt"]] - x["data"][x["dpt"]]
    assert np.max(np.abs(out["resid"] - ref)) < 1e-12
    assert abs(out["cost"][0] - 0.5 * np.sum(ref ** 2)) < 1e-10 * out["cost"][0]
    g = np.einsum("erc,er->c", out["jac"], out["resid"])
    assert np.max(np.abs(g - out["grad"][0])) < 1e-9 * np.max(np.abs(g))


def test_icp_jacobian_matches_central_differences_through_plus(synth, orc, model_np):
    om = orc.Model(model_np)
    x = _one(synth, model_np, seed=8, nd=300)
    args = (x["cptr"], x["mpt"], x["dpt"], x["dptr"], x["data"])
    base = om.icp_eval(x["w"], x["p"], x["q"], *args, want_jac=True)
    jac = base["jac"]
    h = 1e-6

    def resid(delta):
        p = x["p"] + delta[:3]
        q = orc.quat_plus(x["q"][0], delta[3:75].reshape(24, 3))[None]
        w = x["w"] + delta[75:]
        return om.icp_eval(w, p, q, *args, want_jac=True)["resid"]

    worst = 0.0
    for k in range(85):
        d = np.zeros(85); d[k] = h
        num = (resid(d) - resid(-d)) / (2 * h)
        worst = max(worst, float(np.max(np.abs(num - jac[:, :, k]))))
    assert worst < 1e-7 * max(1.0, float(np.max(np.abs(jac)))), worst
    # the quaternions stay unit under Plus
    assert np.allclose(np.linalg.norm(orc.quat_plus(x["q"][0], np.full((24, 3), 0.3)), axis=1), 1.0)


def test_icp_pose_only_has_no_shape_columns(synth, orc, model_np):
    om = orc.Model(model_np)
    x = _one(synth, model_np, seed=9, nd=200)
    a = om.icp_eval(x["w"], x["p"], x["q"], x["cptr"], x["mpt"], x["dpt"], x["dptr"], x["data"], shape=False)
    b = om.icp_eval(x["w"], x["p"], x["q"], x["cptr"], x["mpt"], x["dpt"], x["dptr"], x["data"], shape=True)
    assert np.all(a["grad"][:, 75:] == 0.0) and np.allclose(a["grad"][:, :75], b["grad"][:, :75], rtol=0, atol=1e-12)
    assert np.any(b["grad"][:, 75:] != 0.0)


def test_icp_nearest_oracle_matches_numpy(synth, orc, model_np):
    """The oracle's findNN restatement (exhaustive search per part, visible points only) against numpy."""
    rng = np.random.default_rng(2)
    V = model_np["v_template"].shape[0]
    fin = synth.make_forward_inputs(1, 12)
    cloud = synth.forward_numpy(model_np, fin["w"][0], fin["p"][0], fin["r"][0])[0][None]
    model_part = rng.integers(0, 5, V).astype(np.int32)
    visible = (rng.random((1, V)) > 0.5).astype(np.uint8)
    visible[0, model_part == 4] = 0
    data = cloud[0][rng.integers(0, V, 300)] + rng.normal(0, 0.03, (300, 3))
    data_part = rng.integers(0, 5, 300).astype(np.int32)
    got = orc.icp_nearest(cloud, model_part, visible, [0, 300], data, data_part)
    for i in range(300):
        cand = np.nonzero((model_part == data_part[i]) & (visible[0] == 1))[0]
        if len(cand) == 0:
            assert got[i] == -1 and data_part[i] == 4
        else:
            assert got[i] == cand[np.argmin(((cloud[0][cand] - data[i]) ** 2).sum(axis=1))]


def test_icp_priors_oracle_matches_numpy(synth, orc, model_np):
    """The oracle's prior blocks against a numpy restatement built from GaussianMixture::residual and prec_cho."""
    from conftest import PRIOR_TXT
    prior = synth.load_pose_prior(PRIOR_TXT)
    og = orc.GMM(prior=prior)
    x = synth.make_icp_inputs(model_np, 2, 3, ndata=50)
    totals = np.array([400.0, 36.0])
    bp, bs = 0.2, 1.5
    cost, grad = orc.icp_priors(og, bp, bs, totals, x["w"], x["q"], np.zeros(2), np.zeros((2, 85)))
    arr = og.arrays()
    for b in range(2):
        v, wq = x["q"][b, 1:, :3], x["q"][b, 1:, 3]
        n = np.linalg.norm(v, axis=1)
        aa = (v / n[:, None] * np.sign(wq)[:, None] * (2 * np.arctan2(n, np.abs(wq)))[:, None]).reshape(-1)
        out, comp = og.residual(aa)
        sbp, sbs = bp * np.sqrt(totals[b]) / 15, bs * np.sqrt(totals[b]) / 15
        r = np.asarray(out) * sbp
        c = 0.5 * np.sum(r ** 2) + 0.5 * np.sum((x["w"][b] * sbs) ** 2)
        L = np.asarray(arr["prec_cho"])[comp].reshape(69, 69)
        g = np.zeros(85)
        g[6:75] = L @ r[:69] * 0.707106781186548 * sbp
        g[75:] = sbs * sbs * x["w"][b]
        assert abs(cost[b] - c) < 1e-10 * c
        assert np.max(np.abs(grad[b] - g)) < 1e-10 * np.max(np.abs(g))


def test_icp_minimiser_oracle_properties(synth, orc, model_np):
    """The oracle's restatement of the Ceres 1.14 minimiser AvatarOptimizer::optimize configures (LINE_SEARCH, BFGS, WOLFE, CUBIC;
    AvatarOptimizer.cpp:1313-1339).  Ceres is absent and the reference never tests this loop, so it is pinned by what the published
    algorithm guarantees: every accepted step satisfies the strong Wolfe conditions' sufficient decrease (cost strictly decreasing,
    cost_change equal to the difference of consecutive costs), the first record is the starting point, the termination test is the
    function tolerance |cost_change| <= 1e-4 * previous cost, the result re-evaluates to the reported final cost, and
    shapeEnabled = false leaves the shape keys untouched."""
    from conftest import PRIOR_TXT
    prior = synth.load_pose_prior(PRIOR_TXT)
    og = orc.GMM(prior=prior)
    om = orc.Model(model_np)
    x = synth.make_icp_inputs(model_np, 3, 77, ndata=600)
    args = (x["cptr"], x["mpt"], x["dpt"], x["dptr"], x["data"])
    r = om.icp_minimize(og, 0.05, 0.12, x["w"], x["p"], x["q"], *args, max_iterations=200, function_tolerance=1e-4, nthreads=3)
    totals = np.diff(x["cptr"]).astype(np.float64)
    for b in range(3):
        n = int(r["nrec"][b])
        t = r["trace"][b, :n]
        assert n == r["iterations"][b] + 1
        assert t[0, 1] == 0.0 and t[0, 4] == 0.0
        assert np.all(np.diff(t[:, 0]) < 0.0)
        assert np.allclose(t[1:, 1], t[:-1, 0] - t[1:, 0], rtol=0, atol=1e-9 * t[0, 0])
        assert np.all(t[1:, 4] > 0.0) and np.all(t[1:, 5] >= 1) and np.all(t[1:, 5] <= 20)
        if r["termination"][b] == 3:   # function tolerance: met at the last record, not before
            assert abs(t[-1, 1]) <= 1e-4 * t[-2, 0]
            assert np.all(np.abs(t[1:-1, 1]) > 1e-4 * t[:-2, 0])
        assert r["termination"][b] in (0, 3)
    assert np.all(np.abs(np.linalg.norm(r["q"], axis=-1) - 1.0) < 1e-12), "Plus keeps the quaternions on the unit sphere"
    e = om.icp_eval(r["w"], r["p"], r["q"], *args, nthreads=3)
    c, _ = orc.icp_priors(og, 0.05, 0.12, totals, r["w"], r["q"], e["cost"], e["grad"])
    assert np.max(np.abs(c - r["final_cost"]) / c) < 1e-12
    assert np.all(r["final_cost"] < 0.2 * r["initial_cost"])
    # pose only
    r2 = om.icp_minimize(og, 0.05, 0.12, x["w"], x["p"], x["q"], *args, shape=False, max_iterations=10, nthreads=3)
    assert np.array_equal(r2["w"], x["w"]) and np.all(r2["final_cost"] < r2["initial_cost"])
    # max_iterations = 0: nothing but the first evaluation
    r0 = om.icp_minimize(None, 0.0, 0.0, x["w"], x["p"], x["q"], *args, max_iterations=0)
    assert np.all(r0["nrec"] == 1) and np.all(r0["evaluations"] == 1) and np.array_equal(r0["q"], x["q"])
