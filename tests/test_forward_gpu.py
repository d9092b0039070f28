"""Parity of the batched SMPL evaluator (b3d_smpl_forward, through the C-ABI) with the oracle's Avatar::update restatement.
Tolerance (BASELINE.json north_star): vertices 1e-5 relative; indexing outputs bit-exact."""
import numpy as np
import pytest

from conftest import pkg, rel_err

pytestmark = pytest.mark.gpu

TOL_F32 = 1e-5     # north_star: "SMPL vertices must match the reference's avatar/Eigen path to 1e-5 relative"
TOL_F64 = 1e-12


def test_parents_bit_exact(gmodel, omodel, synth):
    assert gmodel.parent.dtype == np.int32
    assert np.array_equal(gmodel.parent, omodel.constants()["parent"])
    assert np.array_equal(gmodel.parent, synth.PARENT)
    k = omodel.constants()
    assert rel_err(gmodel.initialJointPos, k["J0"]) < 1e-14
    assert rel_err(gmodel.jointShapeReg, k["JSR"]) < 1e-13
    assert np.max(np.abs(gmodel.offset - k["offset"])) < 1e-15


@pytest.mark.parametrize("B", [1, 7, 100, 1000])
@pytest.mark.parametrize("dtype,tol", [(np.float32, TOL_F32), (np.float64, TOL_F64)])
def test_forward_parity(api, gmodel, omodel, synth, B, dtype, tol):
    inp = synth.make_forward_inputs(B, 100 + B)
    c, jp, jt = api.smpl_forward(gmodel, inp["w"], inp["p"], inp["r"], dtype=dtype)
    # the oracle sees the same (possibly float32-rounded) inputs
    w, p, r = (np.asarray(inp[k], dtype=dtype).astype(np.float64) for k in ("w", "p", "r"))
    c0, jp0, jt0 = omodel.forward(w, p, r, nthreads=8)
    assert c.dtype == dtype and c.shape == (B, 6890, 3)
    for b in range(B):   # per-body relative error, as a per-frame user would measure it
        assert rel_err(c[b], c0[b]) < tol
    assert rel_err(jp, jp0) < tol and rel_err(jt, jt0) < tol


@pytest.mark.parametrize("lbs", ["tc", "sorted", "bodylane", "generic"])
@pytest.mark.parametrize("B", [1, 130, 1000])
def test_forward_parity_every_f32_kernel(api, gmodel, omodel, synth, lbs, B):
    """Each float32 skinning kernel (B3D_FWD_LBS_*), forced, against the oracle: the tensor-core shape blend (tcgen05 kind::tf32
    with a hi/lo split) has to meet the same 1e-5 as the CUDA-core kernels; partial body groups (B % 128 != 0) included."""
    inp = synth.make_forward_inputs(B, 300 + B)
    c, jp, jt = api.smpl_forward(gmodel, inp["w"], inp["p"], inp["r"], dtype=np.float32, lbs=lbs)
    w, p, r = (np.asarray(inp[k], dtype=np.float32).astype(np.float64) for k in ("w", "p", "r"))
    c0, jp0, jt0 = omodel.forward(w, p, r, nthreads=8)
    worst = max(rel_err(c[b], c0[b]) for b in range(B))
    assert worst < TOL_F32, (lbs, B, worst)


def test_forward_tc_matches_sorted_large_batch(api, gmodel, synth):
    """BASELINE-scale batch: the tensor-core kernel against the CUDA-core kernel on every (body, vertex) — catches a misplaced
    row, slot or body group that a small batch cannot reach."""
    B = 4096 + 37
    inp = synth.make_forward_inputs(B, 91)
    a, _, _ = api.smpl_forward(gmodel, inp["w"], inp["p"], inp["r"], dtype=np.float32, lbs="tc")
    b, _, _ = api.smpl_forward(gmodel, inp["w"], inp["p"], inp["r"], dtype=np.float32, lbs="sorted")
    err = np.max(np.abs(a.astype(np.float64) - b), axis=(1, 2)) / np.max(np.abs(b), axis=(1, 2))
    assert err.max() < 2e-6, err.max()


@pytest.mark.parametrize("splits", [1, 5, 24, 54, 108])
def test_forward_tc_every_split(api, gmodel, synth, splits):
    """The tensor-core kernel's CTAs own runs of sort tiles; every cut of the 108 tiles (down to one tile per CTA, where the
    copy / MMA / store pipelines barely fill) must give the same cloud as the CUDA-core kernel, with and without pose blendshapes."""
    B = 300
    inp = synth.make_forward_inputs(B, 17 + splits)
    api.lib().b3d_debug_tc_splits(splits)      # debug hook: force the vertex split (0 = automatic)
    try:
        for pd in (False, True):
            a, _, _ = api.smpl_forward(gmodel, inp["w"], inp["p"], inp["r"], dtype=np.float32, lbs="tc", posedirs=pd)
            b, _, _ = api.smpl_forward(gmodel, inp["w"], inp["p"], inp["r"], dtype=np.float32, lbs="generic" if pd else "sorted", posedirs=pd)
            err = np.max(np.abs(a.astype(np.float64) - b), axis=(1, 2)) / np.max(np.abs(b), axis=(1, 2))
            assert err.max() < (2e-5 if pd else 2e-6), (splits, pd, err.max())
    finally:
        api.lib().b3d_debug_tc_splits(0)


def test_forward_posedirs_extension(api, gmodel, omodel_pd, synth):
    inp = synth.make_forward_inputs(9, 5)
    c, _, _ = api.smpl_forward(gmodel, inp["w"], inp["p"], inp["r"], dtype=np.float64, posedirs=True)
    c0, _, _ = omodel_pd.forward(inp["w"], inp["p"], inp["r"])
    assert rel_err(c, c0) < 1e-12
    c32, _, _ = api.smpl_forward(gmodel, inp["w"], inp["p"], inp["r"], dtype=np.float32, posedirs=True)
    assert rel_err(c32, c0) < 2e-5


@pytest.mark.parametrize("B", [130, 700])
def test_forward_posedirs_tensor_core(api, gmodel, omodel_pd, synth, B):
    """Pose blendshapes fused into the tensor-core kernel (K = 208 more of TF32 on top of the shape blend, one pass): same
    2e-5 as the CUDA-core float32 posedirs path, against the oracle with posedirs on."""
    inp = synth.make_forward_inputs(B, 40 + B)
    c, _, _ = api.smpl_forward(gmodel, inp["w"], inp["p"], inp["r"], dtype=np.float32, posedirs=True, lbs="tc")
    w, p, r = (np.asarray(inp[k], dtype=np.float32).astype(np.float64) for k in ("w", "p", "r"))
    c0, _, _ = omodel_pd.forward(w, p, r, nthreads=8)
    worst = max(rel_err(c[b], c0[b]) for b in range(B))
    assert worst < 2e-5, worst
    # and it must differ from the no-posedirs cloud (the term is really there)
    c1, _, _ = api.smpl_forward(gmodel, inp["w"], inp["p"], inp["r"], dtype=np.float32, lbs="tc")
    assert rel_err(c1, c0) > 1e-4


def test_forward_edge_cases(api, gmodel, omodel, synth):
    # rest pose: w = 0, r = I, p = 0 -> cloud equals the template exactly in float64 up to the weight sum rounding
    w, p = np.zeros((1, 10)), np.zeros((1, 3))
    r = np.tile(np.eye(3).reshape(9), (1, 24, 1))
    c, jp, _ = api.smpl_forward(gmodel, w, p, r, dtype=np.float64)
    c0, jp0, _ = omodel.forward(w, p, r)
    assert rel_err(c, c0) < 1e-13 and np.max(np.abs(jp[0, 0])) == 0.0
    # B = 0 is a no-op; joints-only call (no cloud) works
    c, jp, jt = api.smpl_forward(gmodel, np.zeros((0, 10)), np.zeros((0, 3)), np.zeros((0, 24, 9)))
    assert c.shape[0] == 0
    inp = synth.make_forward_inputs(5, 1)
    c, jp, jt = api.smpl_forward(gmodel, inp["w"], inp["p"], inp["r"], want_cloud=False)
    assert c is None and rel_err(jp, omodel.forward(inp["w"], inp["p"], inp["r"], want_cloud=False)[1]) < 1e-12


def test_forward_linearity_large_batch(api, gmodel, synth):
    """Size-independent property at a BASELINE-scale batch: for fixed pose, the cloud is affine in the shape vector w, so
    f(w1) + f(w2) - f(0) == f(w1 + w2) — holds only if every (body, vertex) was written by the right thread."""
    B = 4096
    inp = synth.make_forward_inputs(B, 77)
    r, p = inp["r"], inp["p"]
    r[B // 2:] = r[:B // 2]; p[B // 2:] = p[:B // 2]
    w = inp["w"]
    w[B // 2:] = 0.0
    c, _, _ = api.smpl_forward(gmodel, w, p, r, dtype=np.float64)
    w2 = w.copy(); w2[:B // 2] *= 2.0
    c2, _, _ = api.smpl_forward(gmodel, w2, p, r, dtype=np.float64)
    # f(2w) - f(0) == 2 (f(w) - f(0))
    lhs = c2[:B // 2] - c[B // 2:]
    rhs = 2.0 * (c[:B // 2] - c[B // 2:])
    assert np.max(np.abs(lhs - rhs)) < 1e-12


def test_avatar_object_interface(api, gmodel, omodel, synth):
    av = api.Avatar(gmodel)
    av.p = np.array([0.0, 0.0, 3.0])
    av.r[0] = synth.initial_root_rotation().T.reshape(9)
    av.update()
    c0, jp0, jt0 = omodel.forward(av.w[None], av.p[None], av.r[None])
    assert rel_err(av.cloud, c0[0]) < 1e-13 and rel_err(av.jointPos, jp0[0]) < 1e-13 and rel_err(av.jointTrans, jt0[0]) < 1e-13


def test_forward_given_transforms_reskins_identically(synth, model_np):
    """B3D_FWD_GIVEN_TRANSFORMS: the skinning kernel alone, fed the transforms of an earlier call, writes the same cloud."""
    api, capi = pkg("api"), pkg("_capi")
    gm = api.AvatarModel(model_np)
    for b in (5, 700):     # tile-sorted kernel / tensor-core kernel
        fin = synth.make_forward_inputs(b, 77)
        cloud, jp, jt = api.smpl_forward(gm, fin["w"], fin["p"], fin["r"], dtype=np.float32)
        w = np.ascontiguousarray(fin["w"], dtype=np.float32)
        junk = np.zeros((b, 24, 9), dtype=np.float32)          # p / r are ignored
        again = np.zeros_like(cloud)
        capi.check(capi.lib().b3d_smpl_forward(gm.h, b, capi.F32, capi.ptr(w), capi.ptr(junk), capi.ptr(junk), capi.ptr(again), None,
                                               capi.ptr(jt), 0x100), "given transforms")
        assert np.array_equal(again, cloud)


def test_forward_at_roofline_batch_vs_oracle_sample(api, gmodel, omodel, synth):
    """The batch the roofline is quoted on (B = 65,536, float32, tensor-core kernel): 64 bodies spread over the batch (first and last
    body group, group edges, odd rows) against the oracle evaluated on the float32-rounded inputs; 1e-5 relative per body."""
    import ctypes as C
    import torch
    B = 65536
    capi, lib = api.capi, api.lib()
    inp = synth.make_forward_inputs(B, 5)
    dev = torch.device("cuda", 0)
    w, p, r = [torch.from_numpy(np.ascontiguousarray(inp[k], dtype=np.float32)).to(dev) for k in ("w", "p", "r")]
    cloud = torch.empty(B, gmodel.numPoints, 3, dtype=torch.float32, device=dev)
    jp, jt = torch.empty(B, 24, 3, dtype=torch.float32, device=dev), torch.empty(B, 24, 12, dtype=torch.float32, device=dev)
    P = lambda t: C.c_void_p(t.data_ptr())
    capi.check(lib.b3d_smpl_forward_dev(gmodel.h, B, capi.F32, P(w), P(p), P(r), P(cloud), P(jp), P(jt), 0,
                                        C.c_void_p(torch.cuda.current_stream().cuda_stream)), "forward 65536")
    torch.cuda.synchronize()
    rng = np.random.default_rng(1)
    pick = np.unique(np.concatenate([[0, 1, 127, 128, 129, 32767, 32768, B - 129, B - 128, B - 1], rng.integers(0, B, 54)]))
    f32 = lambda a: np.asarray(a, dtype=np.float32).astype(np.float64)
    c0, jp0, jt0 = omodel.forward(f32(inp["w"][pick]), f32(inp["p"][pick]), f32(inp["r"][pick]), nthreads=8)
    idx = torch.from_numpy(pick).to(dev)
    c = cloud.index_select(0, idx).cpu().numpy().astype(np.float64)
    err = np.max(np.abs(c - c0), axis=(1, 2)) / np.max(np.abs(c0), axis=(1, 2))
    assert err.max() < 1e-5, (err.max(), pick[int(np.argmax(err))])
    assert rel_err(jp.index_select(0, idx).cpu().numpy(), jp0) < 1e-5
    assert rel_err(jt.index_select(0, idx).cpu().numpy(), jt0) < 1e-5


@pytest.mark.parametrize("lbs", ["tc", "sorted"])
def test_forward_parity_block_rigged_model(api, orc, synth, lbs):
    """A model whose joints own contiguous vertex ranges (synth.make_model(rig="blocks")): the sorted kernels then keep their
    register-resident joint transforms over long runs of slots and across sort tiles - the opposite control flow of the default
    model (vertex v on joint v mod 24), where they change on every third slot.  Same 1e-5 against the oracle; partial body group
    and a batch big enough for several CTAs per body group."""
    m = synth.make_model(rig="blocks")
    gm, om = api.AvatarModel(m), orc.Model(m)
    B = 640 + 45
    inp = synth.make_forward_inputs(B, 77)
    c, jp, jt = api.smpl_forward(gm, inp["w"], inp["p"], inp["r"], dtype=np.float32, lbs=lbs)
    w, p, r = (np.asarray(inp[k], dtype=np.float32).astype(np.float64) for k in ("w", "p", "r"))
    c0, jp0, jt0 = om.forward(w, p, r, nthreads=8)
    worst = max(rel_err(c[b], c0[b]) for b in range(B))
    assert worst < TOL_F32, (lbs, worst)
    assert rel_err(jp, jp0) < TOL_F32 and rel_err(jt, jt0) < TOL_F32


@pytest.mark.parametrize("seed", [0, 1, 2, 3])
def test_forward_tc_random_rigs(api, synth, seed):
    """The host tables of the tensor-core kernel (slot order, register-set assignment) on rigs the two synthetic models do not
    produce: runs of 1 - 300 consecutive vertices share a random set of 1 - 4 joints (joint 0, which the unused register sets start
    on, over-represented), weights random.  Tensor-core kernel against the generic kernel on every (body, vertex)."""
    rng = np.random.default_rng(1000 + seed)
    m = dict(synth.make_model())
    V, J = m["weights"].shape
    wt = np.zeros((V, J), dtype=np.float32)
    v = 0
    while v < V:
        run = int(rng.integers(1, 301)) if seed % 2 == 0 else int(rng.integers(1, 12))
        nj = int(rng.integers(1, 5))
        js = rng.choice(J, size=nj, replace=False)
        if rng.random() < 0.3 and 0 not in js:
            js[0] = 0
        for u in range(v, min(V, v + run)):
            ww = rng.dirichlet(np.ones(nj)).astype(np.float32)
            ww = np.maximum(ww, np.float32(1e-3))
            wt[u, js] = ww / ww.sum(dtype=np.float32)
        v += run
    m["weights"] = wt
    gm = api.AvatarModel(m)
    B = 256 + 19
    inp = synth.make_forward_inputs(B, 500 + seed)
    b, _, _ = api.smpl_forward(gm, inp["w"], inp["p"], inp["r"], dtype=np.float32, lbs="generic")
    for lbs in ("tc", "sorted", "bodylane"):   # the two CUDA-core kernels with vertex orders of their own ride along
        a, _, _ = api.smpl_forward(gm, inp["w"], inp["p"], inp["r"], dtype=np.float32, lbs=lbs)
        err = np.max(np.abs(a.astype(np.float64) - b), axis=(1, 2)) / np.max(np.abs(b), axis=(1, 2))
        assert err.max() < 2e-6, (seed, lbs, err.max())


@pytest.mark.parametrize("maxnz", [6, 11])
def test_forward_more_than_four_weights_per_vertex(api, orc, synth, maxnz):
    """Rigs with up to 6 (weight table of 8) and up to 11 (table of 24) joints per vertex take the generic skinning kernel in both
    precisions (the sorted / tensor-core tables exist for 4-weight models only); same tolerances against the oracle."""
    rng = np.random.default_rng(40 + maxnz)
    m = dict(synth.make_model())
    V, J = m["weights"].shape
    wt = np.zeros((V, J), dtype=np.float32)
    for v in range(V):
        nj = int(rng.integers(1, maxnz + 1))
        js = rng.choice(J, size=nj, replace=False)
        ww = rng.dirichlet(np.ones(nj)).astype(np.float32) + np.float32(1e-3)
        wt[v, js] = ww / ww.sum(dtype=np.float32)
    m["weights"] = wt
    gm, om = api.AvatarModel(m), orc.Model(m)
    B = 37
    inp = synth.make_forward_inputs(B, 600 + maxnz)
    for dtype, tol in ((np.float32, TOL_F32), (np.float64, TOL_F64)):
        c, jp, jt = api.smpl_forward(gm, inp["w"], inp["p"], inp["r"], dtype=dtype)
        w, p, r = (np.asarray(inp[k], dtype=dtype).astype(np.float64) for k in ("w", "p", "r"))
        c0, jp0, jt0 = om.forward(w, p, r, nthreads=8)
        assert max(rel_err(c[b], c0[b]) for b in range(B)) < tol, (maxnz, dtype)
        assert rel_err(jp, jp0) < tol and rel_err(jt, jt0) < tol
