"""The library's own .npz reader (csrc/npz.cu: zip central directory, stored / deflated members, .npy headers, the reference's
loadFloatMatrix / loadUintMatrix views) against numpy on files numpy wrote.  Host code only: runs without a GPU.  Bit-exact."""
import os
import zipfile

import numpy as np
import pytest

from conftest import pkg


@pytest.fixture(scope="module")
def api_cpu():
    return pkg("api")          # importing the binding needs the library, not a device


def _arrays():
    rng = np.random.default_rng(2)
    return dict(a32=rng.normal(size=(7, 5)).astype(np.float32), a64=rng.normal(size=(6, 4)), f32F=np.asfortranarray(rng.normal(size=(5, 3)).astype(np.float32)),
                t3=rng.normal(size=(4, 3, 10)).astype(np.float32), t3F=np.asfortranarray(rng.normal(size=(4, 3, 10))),
                kt_u32=np.array([[0xFFFFFFFF, 0, 0, 1], [0, 1, 2, 3]], dtype=np.uint32), kt_i64=np.array([[-1, 0, 0, 1], [0, 1, 2, 3]], dtype=np.int64),
                f_i32F=np.asfortranarray(rng.integers(0, 100, size=(9, 3)).astype(np.int32)), big=rng.normal(size=(300, 700)))


@pytest.mark.parametrize("compressed", [False, True])
def test_members_match_numpy(api_cpu, tmp_path, compressed):
    arr = _arrays()
    path = str(tmp_path / "m.npz")
    (np.savez_compressed if compressed else np.savez)(path, **arr)
    with zipfile.ZipFile(path) as z:
        assert {i.compress_type for i in z.infolist()} == {zipfile.ZIP_DEFLATED if compressed else zipfile.ZIP_STORED}
    for k, a in arr.items():
        shape, word, fortran = api_cpu.npz_shape(path, k)
        assert shape == a.shape and word == a.dtype.itemsize and fortran == (a.flags.f_contiguous and not a.flags.c_contiguous)
    # 2-D float views: the logical matrix whatever the storage order / width
    for k in ("a32", "a64", "f32F", "big"):
        a = arr[k]
        assert np.array_equal(api_cpu.npz_float_matrix(path, k, *a.shape), a.astype(np.float64))
    # the reference reads v_template FLAT as 1 x 3V (AvatarModel.cpp:45-48): memory order, not logical order
    assert np.array_equal(api_cpu.npz_float_matrix(path, "f32F", 1, 15)[0], arr["f32F"].ravel(order="K").astype(np.float64))
    # shapedirs as a (3V x S) matrix: C order is the logical reshape; Fortran order is what Eigen's column-major map of the raw
    # buffer gives (Util.cpp:253-258), i.e. element (r, c) = buffer[r + 3V c]
    assert np.array_equal(api_cpu.npz_float_matrix(path, "t3", 12, 10), arr["t3"].reshape(12, 10).astype(np.float64))
    assert np.array_equal(api_cpu.npz_float_matrix(path, "t3F", 12, 10), arr["t3F"].ravel(order="K").reshape(10, 12).T)
    # integers: 32-bit as is, 64-bit cast (-1 -> 0xFFFFFFFF), Fortran order honoured
    assert np.array_equal(api_cpu.npz_uint_matrix(path, "kt_u32", 2, 4), arr["kt_u32"])
    assert np.array_equal(api_cpu.npz_uint_matrix(path, "kt_i64", 2, 4), arr["kt_u32"])
    assert np.array_equal(api_cpu.npz_uint_matrix(path, "f_i32F", 9, 3), arr["f_i32F"].astype(np.uint32))


def test_errors(api_cpu, tmp_path):
    capi = pkg("_capi")
    path = str(tmp_path / "m.npz")
    np.savez(path, a=np.zeros((2, 3), dtype=np.float32), i=np.zeros((2, 3), dtype=np.int32), be=np.zeros(3, dtype=">f4"), h=np.zeros(3, dtype=np.float16))
    with pytest.raises(capi.B3DError) as e:
        api_cpu.npz_shape(path, "missing")
    assert e.value.code == 4                                   # B3D_ENOFILE
    with pytest.raises(capi.B3DError) as e:
        api_cpu.npz_shape(str(tmp_path / "nope.npz"), "a")
    assert e.value.code == 4
    for name, fn, r, c in (("a", api_cpu.npz_float_matrix, 3, 3), ("i", api_cpu.npz_float_matrix, 2, 3), ("a", api_cpu.npz_uint_matrix, 2, 3),
                           ("be", api_cpu.npz_float_matrix, 1, 3), ("h", api_cpu.npz_float_matrix, 1, 3)):
        with pytest.raises(capi.B3DError) as e:
            fn(path, name, r, c)
        assert e.value.code == 6, name                         # B3D_EFORMAT
    open(tmp_path / "junk.npz", "wb").write(b"PK\x03\x04" + os.urandom(200))
    with pytest.raises(capi.B3DError) as e:
        api_cpu.npz_shape(str(tmp_path / "junk.npz"), "a")
    assert e.value.code == 6
    trunc = open(path, "rb").read()[:-30]
    open(tmp_path / "trunc.npz", "wb").write(trunc)
    with pytest.raises(capi.B3DError):
        api_cpu.npz_shape(str(tmp_path / "trunc.npz"), "a")
