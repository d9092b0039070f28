"""b3d_model_load_npz: a model read by the library's own .npz reader is the same model as the one built from numpy's arrays."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("variant", ["plain", "compressed_f64_i64_fortran"])
def test_native_npz_model_equals_numpy_model(api, synth, model_np, tmp_path, variant):
    path = str(tmp_path / "model.npz")
    m = dict(model_np)
    if variant == "plain":
        np.savez(path, **m)
    else:   # what scripts/npz_fixer.py writes: int64 kintree with -1 at the root; plus float64 / Fortran-ordered 2-D members, deflated
        kt = m["kintree_table"].astype(np.int64)
        kt[kt == 0xFFFFFFFF] = -1
        m.update(kintree_table=kt, v_template=m["v_template"].astype(np.float64), weights=np.asfortranarray(m["weights"]),
                 J_regressor=np.asfortranarray(m["J_regressor"].astype(np.float64)), f=np.asfortranarray(m["f"].astype(np.int64)))
        np.savez_compressed(path, **m)
    a = api.AvatarModel(model_np)
    b = api.AvatarModel.from_npz_native(path, with_posedirs=True)
    assert b.numPoints == a.numPoints and b.numShapeKeys == a.numShapeKeys
    assert np.array_equal(a.parent, b.parent) and a.parent[0] == -1
    for k in ("initialJointPos", "jointShapeReg", "offset"):
        assert np.array_equal(getattr(a, k), getattr(b, k)), k
    fi = synth.make_forward_inputs(600, 4)
    for dt, pd in ((np.float32, False), (np.float64, False), (np.float32, True)):
        ca, ja, ta = api.smpl_forward(a, fi["w"], fi["p"], fi["r"], dtype=dt, posedirs=pd)
        cb, jb, tb = api.smpl_forward(b, fi["w"], fi["p"], fi["r"], dtype=dt, posedirs=pd)
        assert np.array_equal(ca, cb) and np.array_equal(ja, jb) and np.array_equal(ta, tb), (dt, pd)
    # faces travelled too: the overlay of one body is the same picture
    img = np.zeros((1, 90, 160, 3), dtype=np.uint8)
    cam = (144.0, 144.0, 80.0, 45.0)
    assert np.array_equal(api.render_overlay(a, ca[:1], cam, img.copy(), backface_cull=False), api.render_overlay(b, ca[:1], cam, img.copy(), backface_cull=False))


def test_native_npz_rejects_bad_models(api, model_np, tmp_path):
    capi = __import__("importlib").import_module("3dbodyanimation_b200._capi")
    m = dict(model_np)
    m.pop("posedirs")
    np.savez(tmp_path / "nopd.npz", **m)
    api.AvatarModel.from_npz_native(str(tmp_path / "nopd.npz"))
    with pytest.raises(capi.B3DError):
        api.AvatarModel.from_npz_native(str(tmp_path / "nopd.npz"), with_posedirs=True)
    bad = dict(m, weights=m["weights"][:, :23])
    np.savez(tmp_path / "shape.npz", **bad)
    with pytest.raises(capi.B3DError) as e:
        api.AvatarModel.from_npz_native(str(tmp_path / "shape.npz"))
    assert e.value.code == 6
    kt = m["kintree_table"].copy()
    kt[0, 0] = 0                                  # root whose parent is not -1: _ARK_ASSERT_EQ(parent[0], -1), AvatarModel.cpp:41
    np.savez(tmp_path / "root.npz", **dict(m, kintree_table=kt))
    with pytest.raises(capi.B3DError):
        api.AvatarModel.from_npz_native(str(tmp_path / "root.npz"))
