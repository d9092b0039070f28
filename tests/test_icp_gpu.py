"""GPU parity: vertex-level ICP term (b3d_icp_create / b3d_icp_eval, csrc/icp.cu) against the oracle's literal restatement of
external/avatar/AvatarOptimizer.cpp:161-641 on the same seeded inputs.  Tolerances: float64 kernel 1e-10 relative (the two
sides sum in different orders and the GPU uses the closed form -2 [d]x G_parent of the reference's Jacobian blocks); float32
kernel 2e-4 of the gradient's largest entry, 1e-5 on the cost."""
import numpy as np
import pytest

from conftest import pkg

pytestmark = pytest.mark.gpu


def _run(synth, orc, model_np, gpu_model, b, seed, dtype, nd=None, shape=True):
    api = pkg("api")
    x = synth.make_icp_inputs(model_np, b, seed, ndata=nd)
    om = orc.Model(model_np)
    ref = om.icp_eval(x["w"], x["p"], x["q"], x["cptr"], x["mpt"], x["dpt"], x["dptr"], x["data"], shape=shape, nthreads=8)
    prob = api.IcpProblem(gpu_model, x["cptr"], x["mpt"], x["dpt"], x["dptr"], x["data"], dtype=dtype)
    got = prob.eval(x["w"], x["p"], x["q"], shape=shape)
    return ref, got, prob, x


@pytest.fixture(scope="module")
def gpu_model(model_np):
    return pkg("api").AvatarModel(model_np)


def test_icp_f64_matches_oracle(synth, orc, model_np, gpu_model):
    ref, got, _, _ = _run(synth, orc, model_np, gpu_model, 5, 31, np.float64, nd=3000)
    assert np.max(np.abs(got["cost"] - ref["cost"]) / ref["cost"]) < 1e-10
    scale = np.max(np.abs(ref["grad"]), axis=1, keepdims=True)
    assert np.max(np.abs(got["grad"] - ref["grad"]) / scale) < 1e-10


def test_icp_f32_matches_oracle(synth, orc, model_np, gpu_model):
    ref, got, _, _ = _run(synth, orc, model_np, gpu_model, 70, 32, np.float32)   # 3 body groups, the last one partial
    assert np.max(np.abs(got["cost"] - ref["cost"]) / ref["cost"]) < 1e-5
    scale = np.max(np.abs(ref["grad"]), axis=1, keepdims=True)
    assert np.max(np.abs(got["grad"] - ref["grad"]) / scale) < 2e-4


@pytest.mark.parametrize("dtype,ctol,gtol", [(np.float64, 1e-10, 1e-10), (np.float32, 1e-5, 2e-4)])
def test_icp_block_rigged_model(synth, orc, dtype, ctol, gtol):
    """The same parity on a model whose joints own contiguous vertex ranges (synth.make_model(rig="blocks")): a weight slot of the
    accumulation kernel then keeps its joint for hundreds of vertices (long per-joint partial sums, few flushes)."""
    m = synth.make_model(rig="blocks")
    gm = pkg("api").AvatarModel(m)
    ref, got, _, _ = _run(synth, orc, m, gm, 37, 35, dtype, nd=2500)
    assert np.max(np.abs(got["cost"] - ref["cost"]) / ref["cost"]) < ctol
    scale = np.max(np.abs(ref["grad"]), axis=1, keepdims=True)
    assert np.max(np.abs(got["grad"] - ref["grad"]) / scale) < gtol


def test_icp_pose_only_and_repeatable(synth, orc, model_np, gpu_model):
    ref, got, prob, x = _run(synth, orc, model_np, gpu_model, 3, 33, np.float64, nd=500, shape=False)
    assert np.all(got["grad"][:, 75:] == 0.0)
    scale = np.max(np.abs(ref["grad"]), axis=1, keepdims=True)
    assert np.max(np.abs(got["grad"] - ref["grad"]) / scale) < 1e-10
    again = prob.eval(x["w"], x["p"], x["q"], shape=False)   # the handle is reusable: same state, same answer to rounding of the atomics
    assert np.allclose(again["grad"], got["grad"], rtol=0, atol=1e-12 * scale.max())


def test_icp_empty_and_many_to_one_correspondences(synth, orc, model_np, gpu_model):
    """Body 0 has no correspondences at all, body 1 matches every data point to the same model point."""
    api = pkg("api")
    x = synth.make_icp_inputs(model_np, 2, 34, ndata=40)
    n1 = int(x["cptr"][2] - x["cptr"][1])
    mpt = np.full(n1, 1234, dtype=np.int32)
    cptr = np.array([0, 0, n1], dtype=np.int64)
    dpt = x["dpt"][x["cptr"][1]:]
    om = orc.Model(model_np)
    ref = om.icp_eval(x["w"], x["p"], x["q"], cptr, mpt, dpt, x["dptr"], x["data"])
    got = api.IcpProblem(gpu_model, cptr, mpt, dpt, x["dptr"], x["data"], dtype=np.float64).eval(x["w"], x["p"], x["q"])
    assert got["cost"][0] == 0.0 and np.all(got["grad"][0] == 0.0) and ref["cost"][0] == 0.0
    assert abs(got["cost"][1] - ref["cost"][1]) < 1e-10 * ref["cost"][1]
    assert np.max(np.abs(got["grad"][1] - ref["grad"][1])) < 1e-10 * np.max(np.abs(ref["grad"][1]))


def test_icp_descent_step_lowers_cost(synth, orc, model_np, gpu_model):
    """The gradient is a descent direction through the reference's Plus(): a small step against it lowers the cost."""
    ref, got, prob, x = _run(synth, orc, model_np, gpu_model, 4, 35, np.float32)
    g = got["grad"]
    step = 1e-3 / np.max(np.abs(g), axis=1, keepdims=True)
    d = -step * g
    q2 = np.stack([orc.quat_plus(x["q"][i], d[i, 3:75].reshape(24, 3)) for i in range(4)])
    new = prob.eval(x["w"] + d[:, 75:], x["p"] + d[:, :3], q2)
    pred = np.sum(g * d, axis=1)
    assert np.all(new["cost"] < got["cost"])
    assert np.allclose(new["cost"] - got["cost"], pred, rtol=0.2)


def _nn_case(synth, model_np, b, seed, nparts, nd):
    """Posed clouds, part labels from the heaviest joint (modelPartLabels, AvatarOptimizer.cpp:1296-1300, with partMap = joint mod
    nparts), a visibility mask, and data points labelled like the model point they were sampled near (a tenth mislabelled)."""
    rng = np.random.default_rng(seed)
    x = synth.make_icp_inputs(model_np, b, seed, ndata=nd)
    V = model_np["v_template"].shape[0]
    model_part = (np.argmax(model_np["weights"], axis=1) % nparts).astype(np.int32)
    fin = synth.make_forward_inputs(b, seed + 7)
    cloud = np.stack([synth.forward_numpy(model_np, fin["w"][i], x["p"][i], fin["r"][i])[0] for i in range(b)])
    visible = (rng.random((b, V)) > 0.3).astype(np.uint8)
    near = np.array([np.argmin(((cloud[i // nd][::7] - x["data"][i]) ** 2).sum(axis=1)) * 7 for i in range(b * nd)])
    data_part = model_part[near].copy()
    flip = rng.random(b * nd) < 0.1
    data_part[flip] = rng.integers(0, nparts, int(flip.sum()))
    return x, cloud, model_part, visible, data_part


def test_icp_nearest_f64_indices_bit_exact(synth, orc, model_np, gpu_model):
    api = pkg("api")
    x, cloud, model_part, visible, data_part = _nn_case(synth, model_np, 3, 41, 6, 1500)
    visible[1, model_part == 2] = 0                  # body 1: part 2 has no visible model point -> -1
    ref = orc.icp_nearest(cloud, model_part, visible, x["dptr"], x["data"], data_part, nthreads=8)
    got = api.icp_nearest(gpu_model, cloud, model_part, x["dptr"], x["data"], data_part, visible=visible)
    assert np.array_equal(got, ref)
    sel = slice(int(x["dptr"][1]), int(x["dptr"][2]))
    assert np.all(got[sel][data_part[sel] == 2] == -1) and np.any(got >= 0)
    # no visibility mask, one part: plain nearest neighbour over the whole cloud
    one = np.zeros_like(model_part)
    ref1 = orc.icp_nearest(cloud, one, None, x["dptr"], x["data"], np.zeros_like(data_part), nthreads=8)
    got1 = api.icp_nearest(gpu_model, cloud, one, x["dptr"], x["data"], np.zeros_like(data_part))
    assert np.array_equal(got1, ref1)
    brute = np.array([np.argmin(((cloud[0] - d) ** 2).sum(axis=1)) for d in x["data"][:200]])
    assert np.array_equal(got1[:200], brute)


def test_icp_nearest_f32_and_chain_into_problem(synth, orc, model_np, gpu_model):
    """float32 search: same indices as the oracle on the float-rounded inputs up to near-ties; its output feeds b3d_icp_create
    directly (dpt = NULL, unmatched points skipped) and the evaluation equals the oracle's on the explicit pair list."""
    api = pkg("api")
    x, cloud, model_part, visible, data_part = _nn_case(synth, model_np, 2, 42, 4, 2500)
    visible[0, model_part == 1] = 0                  # body 0: the data points of part 1 stay unmatched
    c32, d32 = cloud.astype(np.float32), x["data"].astype(np.float32)
    ref = orc.icp_nearest(c32.astype(np.float64), model_part, visible, x["dptr"], d32.astype(np.float64), data_part, nthreads=8)
    got = api.icp_nearest(gpu_model, c32, model_part, x["dptr"], d32, data_part, visible=visible, dtype=np.float32)
    diff = np.nonzero(got != ref)[0]
    assert len(diff) <= 2e-3 * len(ref)
    for i in diff:      # a different answer is a near-tie
        bdy = int(np.searchsorted(x["dptr"], i, side="right") - 1)
        da = np.sum((cloud[bdy][got[i]] - x["data"][i]) ** 2)
        db = np.sum((cloud[bdy][ref[i]] - x["data"][i]) ** 2)
        assert abs(da - db) < 1e-5 * max(da, db)
    keep = got >= 0
    assert keep.sum() < len(got)                      # some data points are unmatched (their part is invisible)
    prob = api.IcpProblem(gpu_model, x["dptr"], got, None, x["dptr"], x["data"], dtype=np.float64)
    out = prob.eval(x["w"], x["p"], x["q"])
    idx = np.arange(len(got))
    local = idx - x["dptr"][np.searchsorted(x["dptr"], idx, side="right") - 1]
    cptr = np.concatenate([[0], np.cumsum([keep[x["dptr"][i]:x["dptr"][i + 1]].sum() for i in range(2)])])
    om = orc.Model(model_np)
    ref_eval = om.icp_eval(x["w"], x["p"], x["q"], cptr, got[keep], local[keep], x["dptr"], x["data"])
    assert np.max(np.abs(out["cost"] - ref_eval["cost"]) / ref_eval["cost"]) < 1e-10
    assert np.max(np.abs(out["grad"] - ref_eval["grad"])) < 1e-10 * np.max(np.abs(ref_eval["grad"]))


def test_icp_visible_bit_exact_and_feeds_the_search(synth, orc):
    """pointVisible (AvatarOptimizer.cpp:1337-1352) on the smooth render model (it has faces): float64 mask bit-identical to the
    oracle's; the mask then restricts the correspondence search exactly as on the CPU."""
    api = pkg("api")
    rm = synth.make_render_model()
    gm = api.AvatarModel(rm)
    fin = synth.make_forward_inputs(4, 51)
    cloud, _, _ = api.smpl_forward(gm, fin["w"], fin["p"], fin["r"], dtype=np.float64)
    vis = api.icp_visible(gm, cloud)
    ref = orc.icp_visible(cloud, rm["f"])
    assert np.array_equal(vis, ref)
    assert vis.min() == 0 and vis.max() == 1   # both outcomes occur (the synthetic mesh is not consistently oriented: most vertices pass)
    rng = np.random.default_rng(3)
    V = cloud.shape[1]
    model_part = (np.argmax(rm["weights"], axis=1) % 5).astype(np.int32)
    nd = 800
    src = rng.integers(0, V, (4, nd))
    data = (cloud[np.arange(4)[:, None], src] + rng.normal(0, 0.02, (4, nd, 3))).reshape(-1, 3)
    data_part = model_part[src].reshape(-1)
    dptr = np.arange(5) * nd
    got = api.icp_nearest(gm, cloud, model_part, dptr, data, data_part, visible=vis)
    want = orc.icp_nearest(cloud, model_part, ref, dptr, data, data_part, nthreads=4)
    assert np.array_equal(got, want)
    assert np.all(vis.reshape(-1)[(np.repeat(np.arange(4), nd) * V + got)[got >= 0]] == 1)


def test_icp_large_batch_single_split(synth, orc, model_np, gpu_model):
    """4,736 bodies = 148 body groups: the launch takes one CTA per group over the whole vertex range (no vertex split), and the
    last tile of the sorted vertex order is partial.  The batch tiles 32 distinct bodies; every copy must reproduce the oracle."""
    api = pkg("api")
    base, reps = 32, 148
    x = synth.make_icp_inputs(model_np, base, 36, ndata=1500)
    om = orc.Model(model_np)
    ref = om.icp_eval(x["w"], x["p"], x["q"], x["cptr"], x["mpt"], x["dpt"], x["dptr"], x["data"], nthreads=8)
    npair, ndata = int(x["cptr"][-1]), int(x["dptr"][-1])
    cptr = np.concatenate([x["cptr"][:-1] + r * npair for r in range(reps)] + [[reps * npair]])
    dptr = np.concatenate([x["dptr"][:-1] + r * ndata for r in range(reps)] + [[reps * ndata]])
    prob = api.IcpProblem(gpu_model, cptr, np.tile(x["mpt"], reps), np.tile(x["dpt"], reps), dptr, np.tile(x["data"], (reps, 1)), dtype=np.float32)
    t = lambda a: np.tile(a, (reps,) + (1,) * (a.ndim - 1))
    got = prob.eval(t(x["w"]), t(x["p"]), t(x["q"]))
    cost = got["cost"].reshape(reps, base)
    grad = got["grad"].reshape(reps, base, 85)
    scale = np.max(np.abs(ref["grad"]), axis=1, keepdims=True)
    assert np.max(np.abs(cost - ref["cost"]) / ref["cost"]) < 1e-5
    assert np.max(np.abs(grad - ref["grad"]) / scale) < 2e-4


def test_icp_full_objective_with_priors(synth, orc, model_np, gpu_model):
    """ICP term + AvatarPosePriorCostFunctor + AvatarShapePriorCostFunctor (weights scaled by sqrt(#residual blocks) / 15) against
    the oracle; a quaternion with negative w and an identity quaternion exercise Eigen's AngleAxis conversion branches."""
    from conftest import PRIOR_TXT
    api = pkg("api")
    prior = synth.load_pose_prior(PRIOR_TXT)
    gg, og = api.GaussianMixture(prior=prior), orc.GMM(prior=prior)
    x = synth.make_icp_inputs(model_np, 6, 37, ndata=900)
    q = x["q"].copy()
    q[0, 5] *= -1.0
    q[1, 7] = [0.0, 0.0, 0.0, 1.0]
    om = orc.Model(model_np)
    base = om.icp_eval(x["w"], x["p"], q, x["cptr"], x["mpt"], x["dpt"], x["dptr"], x["data"], nthreads=4)
    totals = np.diff(x["cptr"]).astype(np.float64)
    prob = api.IcpProblem(gpu_model, x["cptr"], x["mpt"], x["dpt"], x["dptr"], x["data"], dtype=np.float64)
    for bp, bs, g in ((0.2, 1.0, gg), (0.0, 1.0, None), (0.2, 0.0, gg)):
        rc, rg = orc.icp_priors(og if bp > 0 else None, bp, bs, totals, x["w"], q, base["cost"], base["grad"])
        got = prob.eval_full(x["w"], x["p"], q, gmm=g, beta_pose=bp, beta_shape=bs)
        assert np.max(np.abs(got["cost"] - rc) / rc) < 1e-10
        assert np.max(np.abs(got["grad"] - rg) / np.max(np.abs(rg), axis=1, keepdims=True)) < 1e-10
        assert np.all(rc > base["cost"])


def test_icp_iteration_stays_on_the_device(synth, orc, model_np, gpu_model):
    """One ICP iteration with everything resident (B3D_ICP_DEVICE_INPUT): posed cloud from b3d_smpl_forward_dev, b3d_icp_visible,
    b3d_icp_nearest and b3d_icp_create on device pointers, then the evaluation — equal to the host-pointer chain."""
    import ctypes as C
    import torch
    api, capi = pkg("api"), pkg("_capi")
    lib = capi.lib()
    rm = synth.make_render_model()          # a model with faces (visibility needs them)
    gm = api.AvatarModel(rm)
    b, nd, nparts = 3, 1200, 4
    x = synth.make_icp_inputs(rm, b, 44, ndata=nd)
    r = synth.quat_to_r(x["q"])
    V = gm.numPoints
    model_part = (np.argmax(rm["weights"], axis=1) % nparts).astype(np.int32)
    rng = np.random.default_rng(5)
    data_part = rng.integers(0, nparts, b * nd).astype(np.int32)
    # host chain
    cloud, _, _ = api.smpl_forward(gm, x["w"], x["p"], r, dtype=np.float64)
    vis = api.icp_visible(gm, cloud)
    nn = api.icp_nearest(gm, cloud, model_part, x["dptr"], x["data"], data_part, visible=vis)
    want = api.IcpProblem(gm, x["dptr"], nn, None, x["dptr"], x["data"], dtype=np.float64).eval(x["w"], x["p"], x["q"])
    # device chain
    dev = torch.device("cuda:0")
    P = lambda t: C.c_void_p(t.data_ptr())
    tw, tp, tr, tq = [torch.from_numpy(np.ascontiguousarray(a, dtype=np.float64)).to(dev) for a in (x["w"], x["p"], r, x["q"])]
    tcloud = torch.empty(b, V, 3, dtype=torch.float64, device=dev)
    tjp, tjt = torch.empty(b, 24, 3, dtype=torch.float64, device=dev), torch.empty(b, 24, 12, dtype=torch.float64, device=dev)
    capi.check(lib.b3d_smpl_forward_dev(gm.h, b, capi.F64, P(tw), P(tp), P(tr), P(tcloud), P(tjp), P(tjt), 0, None), "forward")
    tvis = torch.empty(b, V, dtype=torch.uint8, device=dev)
    capi.check(lib.b3d_icp_visible(gm.h, b, capi.F64, P(tcloud), 1, P(tvis)), "visible")
    tdata = torch.from_numpy(np.ascontiguousarray(x["data"])).to(dev)
    tpart = torch.from_numpy(data_part).to(dev)
    tnn = torch.empty(b * nd, dtype=torch.int32, device=dev)
    dptr = np.ascontiguousarray(x["dptr"], dtype=np.int64)
    capi.check(lib.b3d_icp_nearest(gm.h, b, capi.F64, P(tcloud), nparts, capi.ptr(model_part), P(tvis), capi.ptr(dptr), P(tdata), P(tpart), 1,
                                   P(tnn)), "nearest")
    assert np.array_equal(tvis.cpu().numpy(), vis) and np.array_equal(tnn.cpu().numpy(), nn)
    h = C.c_void_p()
    capi.check(lib.b3d_icp_create(gm.h, b, capi.F64, capi.ptr(dptr), P(tnn), None, capi.ptr(dptr), P(tdata), 1, C.byref(h)), "create")
    tcost, tgrad = torch.empty(b, dtype=torch.float64, device=dev), torch.empty(b, 85, dtype=torch.float64, device=dev)
    capi.check(lib.b3d_icp_eval_dev(h, P(tw), P(tp), P(tq), 0, P(tcost), P(tgrad), None), "eval")
    torch.cuda.synchronize()
    lib.b3d_icp_destroy(h)
    assert np.allclose(tcost.cpu().numpy(), want["cost"], rtol=1e-12, atol=0)
    assert np.allclose(tgrad.cpu().numpy(), want["grad"], rtol=0, atol=1e-11 * np.max(np.abs(want["grad"])))


def _solve_inputs(synth, model_np, b, seed, ndata):
    x = synth.make_icp_inputs(model_np, b, seed, ndata=ndata)
    return x


@pytest.mark.parametrize("shape", [True, False])
def test_icp_solve_matches_oracle_minimiser(synth, orc, model_np, gpu_model, shape):
    """b3d_icp_solve (the ceres::Solve of AvatarOptimizer::optimize: LINE_SEARCH / BFGS / WOLFE / CUBIC, AvatarOptimizer.cpp:1313-1339,
    :1486) against the oracle's restatement of the same minimiser, float64 evaluation kernel, ICP term + both priors: per body the
    iteration records (cost, step size, line-search evaluations) and the minimiser.  Iteration, evaluation and line-search counts
    must be equal.  Tolerances: the two sides evaluate the same objective to 1e-10 (test_icp_full_objective_with_priors) and a
    quasi-Newton trajectory amplifies rounding differences from iteration to iteration (the inverse Hessian is built from gradient
    differences): cost trace 1e-10 relative over the first 12 iterations and 1e-8 over all 25, minimiser 1e-7 (measured on B200:
    5e-14, 1.6e-11 and 1e-11)."""
    from conftest import PRIOR_TXT
    api = pkg("api")
    prior = synth.load_pose_prior(PRIOR_TXT)
    gg, og = api.GaussianMixture(prior=prior), orc.GMM(prior=prior)
    B = 5
    x = _solve_inputs(synth, model_np, B, 52, 1500)
    om = orc.Model(model_np)
    kw = dict(max_iterations=25, function_tolerance=1e-4)
    ref = om.icp_minimize(og, 0.05, 0.12, x["w"], x["p"], x["q"], x["cptr"], x["mpt"], x["dpt"], x["dptr"], x["data"], shape=shape,
                          nthreads=5, **kw)
    prob = api.IcpProblem(gpu_model, x["cptr"], x["mpt"], x["dpt"], x["dptr"], x["data"], dtype=np.float64)
    got = prob.solve(x["w"], x["p"], x["q"], gmm=gg, beta_pose=0.05, beta_shape=0.12, shape=shape, **kw)
    worst = dict(cost_head=0.0, cost_tail=0.0, step=0.0, w=0.0, p=0.0, q=0.0)
    for b in range(B):
        assert got["nrec"][b] == ref["nrec"][b], (b, got["nrec"][b], ref["nrec"][b])
        assert got["iterations"][b] == ref["iterations"][b] and got["termination"][b] == ref["termination"][b]
        assert got["evaluations"][b] == ref["evaluations"][b]
        n = int(ref["nrec"][b])
        tg, tr = got["trace"][b, :n], ref["trace"][b, :n]
        assert np.array_equal(tg[:, 5:], tr[:, 5:]), "line-search evaluation / iteration counts per iteration"
        rel = np.abs(tg[:, 0] - tr[:, 0]) / tr[:, 0]
        worst["cost_head"] = max(worst["cost_head"], float(rel[:13].max()))
        worst["cost_tail"] = max(worst["cost_tail"], float(rel.max()))
        worst["step"] = max(worst["step"], float(np.max(np.abs(tg[1:, 4] - tr[1:, 4]) / tr[1:, 4])))
        assert np.all(np.diff(tg[:, 0]) < 0), "the Wolfe search only accepts decreasing steps"
    worst["w"] = float(np.max(np.abs(got["w"] - ref["w"])))
    worst["p"] = float(np.max(np.abs(got["p"] - ref["p"])))
    worst["q"] = float(np.max(np.abs(got["q"] - ref["q"])))
    print("\n[icp solve, shape=%s] iterations %s, evaluations %s, batched evaluations %d; worst deviations %s"
          % (shape, got["iterations"].tolist(), got["evaluations"].tolist(), got["batched_evaluations"], worst))
    assert worst["cost_head"] < 1e-10 and worst["cost_tail"] < 1e-8 and worst["step"] < 1e-6
    assert worst["w"] < 1e-7 and worst["p"] < 1e-7 and worst["q"] < 1e-7
    if not shape:
        assert np.array_equal(got["w"], x["w"]), "shapeEnabled = false leaves the shape keys alone"
    # the result is a better point of the objective the evaluator reports
    chk = prob.eval_full(got["w"], got["p"], got["q"], gmm=gg, beta_pose=0.05, beta_shape=0.12, shape=shape)
    assert np.max(np.abs(chk["cost"] - got["final_cost"]) / got["final_cost"]) < 1e-9
    assert np.all(got["final_cost"] < 0.5 * got["initial_cost"])


def test_icp_solve_float32_kernel_and_limits(synth, orc, model_np, gpu_model):
    """The float32 evaluation kernel under the same minimiser: it converges to the same basin (final cost within 2 % of the float64
    run).  max_iterations = 0 returns the starting point with one record; a batch whose bodies stop at different iterations keeps
    the finished bodies' results."""
    api = pkg("api")
    B = 4
    x = _solve_inputs(synth, model_np, B, 53, 1200)
    p64 = api.IcpProblem(gpu_model, x["cptr"], x["mpt"], x["dpt"], x["dptr"], x["data"], dtype=np.float64)
    p32 = api.IcpProblem(gpu_model, x["cptr"], x["mpt"], x["dpt"], x["dptr"], x["data"], dtype=np.float32)
    a = p64.solve(x["w"], x["p"], x["q"], beta_shape=0.1, max_iterations=30)
    b = p32.solve(x["w"], x["p"], x["q"], beta_shape=0.1, max_iterations=30)
    assert np.max(np.abs(a["final_cost"] - b["final_cost"]) / a["final_cost"]) < 2e-2
    z = p64.solve(x["w"], x["p"], x["q"], beta_shape=0.1, max_iterations=0)
    assert np.array_equal(z["w"], x["w"]) and np.array_equal(z["q"], x["q"]) and np.all(z["nrec"] == 1) and np.all(z["iterations"] == 0)
    assert np.allclose(z["initial_cost"], a["initial_cost"], rtol=1e-12)
    # different stopping points inside one batch: a loose tolerance stops some bodies early
    c = p64.solve(x["w"], x["p"], x["q"], beta_shape=0.1, max_iterations=30, function_tolerance=2e-2)
    assert len(set(c["iterations"].tolist())) > 1 or np.all(c["iterations"] < 30)
    for bb in range(B):
        n = int(c["nrec"][bb])
        assert np.allclose(c["trace"][bb, :n, 0], a["trace"][bb, :n, 0], rtol=1e-6)


def test_avatar_optimize_loop_matches_oracle_chain(synth, orc):
    """The whole AvatarOptimizer::optimize loop (AvatarOptimizer.cpp:1346-1516: occlusion mask -> findNN -> residual blocks -> BFGS
    solve -> rotations, icp_iters times) on the GPU against the same chain of oracle calls: identical correspondences in every ICP
    iteration, equal iteration / evaluation counts, final state to 1e-6."""
    from conftest import PRIOR_TXT
    api = pkg("api")
    rm = synth.make_render_model()          # a model with faces (the occlusion test needs them)
    gm, om = api.AvatarModel(rm), orc.Model(rm)
    prior = synth.load_pose_prior(PRIOR_TXT)
    gg, og = api.GaussianMixture(prior=prior), orc.GMM(prior=prior)
    B, nd, nparts = 3, 1200, 6
    rng = np.random.default_rng(11)
    V = rm["v_template"].shape[0]
    model_part = (np.argmax(rm["weights"], axis=1) % nparts).astype(np.int32)
    x = synth.make_icp_inputs(rm, B, 91, ndata=nd)
    # data: noisy samples of a differently posed copy of the body, labelled with the part of the model point they came from
    tgt = synth.make_forward_inputs(B, 93)
    tq = x["q"] + 0.05 * rng.normal(size=x["q"].shape)
    tq /= np.linalg.norm(tq, axis=-1, keepdims=True)
    tcloud, _, _ = om.forward(x["w"] + 0.2 * tgt["w"], x["p"] + 0.02, synth.quat_to_r(tq))
    src = rng.integers(0, V, (B, nd))
    data = (tcloud[np.arange(B)[:, None], src] + rng.normal(0, 0.005, (B, nd, 3))).reshape(-1, 3)
    data_part = model_part[src].reshape(-1)
    dptr = np.arange(B + 1, dtype=np.int64) * nd
    kw = dict(beta_pose=0.05, beta_shape=0.1)
    got = api.avatar_optimize(gm, x["w"], x["p"], x["q"], dptr, data, data_part, model_part, gmm=gg, icp_iters=3, max_iters_per_icp=12, **kw)
    # the same chain on the CPU
    w, p, q = x["w"].copy(), x["p"].copy(), x["q"].copy()
    for it in range(3):
        cloud, _, _ = om.forward(w, p, synth.quat_to_r(q))
        vis = orc.icp_visible(cloud, rm["f"])
        nn = orc.icp_nearest(cloud, model_part, vis, dptr, data, data_part, nthreads=3)
        h = got["history"][it]
        assert np.array_equal(nn, h["nn"]), f"ICP iteration {it}: correspondences differ"
        keep = nn >= 0                                                    # unmatched data points add no residual block
        local = (np.arange(B * nd) - np.repeat(dptr[:-1], nd)).astype(np.int32)
        cptr = np.concatenate([[0], np.cumsum([int(keep[dptr[i]:dptr[i + 1]].sum()) for i in range(B)])]).astype(np.int64)
        ref = om.icp_minimize(og, kw["beta_pose"], kw["beta_shape"], w, p, q, cptr, nn[keep], local[keep], dptr, data, max_iterations=12, nthreads=3)
        assert np.array_equal(ref["iterations"], h["iterations"]) and np.array_equal(ref["evaluations"], h["evaluations"])
        assert np.max(np.abs(ref["final_cost"] - h["final_cost"]) / ref["final_cost"]) < 1e-8
        w, p, q = ref["w"], ref["p"], ref["q"]
    assert np.max(np.abs(got["w"] - w)) < 1e-6 and np.max(np.abs(got["p"] - p)) < 1e-6 and np.max(np.abs(got["q"] - q)) < 1e-6
    assert np.all(got["history"][-1]["final_cost"] < got["history"][0]["initial_cost"])
