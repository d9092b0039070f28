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
