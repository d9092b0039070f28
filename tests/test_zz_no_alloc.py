"""b3d_smpl_forward_noalloc_dev (include/b3d.h): the allocation-free variant of the batched SMPL evaluator, whose joint-transform
scratch is a file-scope `static __device__` array instead of stream-ordered memory.  It has no counterpart in the reference (added
for callers that cannot call cudaMalloc); the check is against the existing entry point b3d_smpl_forward_dev on the same inputs,
through the C-ABI, on two calls in a row (the second call finds the first call's contents in the shared scratch).

This file is named so that pytest collects it last: a failure here must not hide other tests.
It passes alone in a fresh process and after the rest of the suite."""
import ctypes as C

import numpy as np
import pytest

from conftest import rel_err

pytestmark = pytest.mark.gpu

TOL_F32 = 1e-5     # the tolerance of tests/test_forward_gpu.py (north_star: vertices to 1e-5 relative)
TOL_F64 = 1e-12
MAX_BODIES = 1024  # B3D_NOALLOC_MAX_BODIES
EINVAL = 1          # B3D_EINVAL


def _call(api, gmodel, inp, dtype, noalloc, flags=0):
    import torch
    capi, lib = api.capi, api.lib()
    dev = torch.device("cuda", 0)
    td = torch.float32 if dtype == np.float32 else torch.float64
    code = capi.F32 if dtype == np.float32 else capi.F64
    B = inp["w"].shape[0]
    w, p, r = [torch.from_numpy(np.ascontiguousarray(inp[k], dtype=dtype)).to(dev) for k in ("w", "p", "r")]
    cloud = torch.full((B, gmodel.numPoints, 3), float("nan"), dtype=td, device=dev)
    jp = torch.full((B, 24, 3), float("nan"), dtype=td, device=dev)
    P = lambda t: C.c_void_p(t.data_ptr())
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    if noalloc:
        rc = lib.b3d_smpl_forward_noalloc_dev(gmodel.h, B, code, P(w), P(p), P(r), P(cloud), P(jp), st)
    else:
        rc = lib.b3d_smpl_forward_dev(gmodel.h, B, code, P(w), P(p), P(r), P(cloud), P(jp), None, flags, st)
    capi.check(rc, "forward")
    torch.cuda.synchronize()
    return cloud.cpu().numpy(), jp.cpu().numpy()


@pytest.mark.parametrize("dtype,tol", [(np.float32, TOL_F32), (np.float64, TOL_F64)])
def test_noalloc_forward_equals_existing_entry_point_twice(api, gmodel, synth, dtype, tol):
    capi = api.capi
    sorted_flag = 0x20   # B3D_FWD_LBS_SORTED: the float32 kernel the allocation-free variant launches
    # two calls in a row with different batch sizes and inputs: the second finds the first one's transforms in the scratch
    for B, seed in ((MAX_BODIES, 31), (77, 32)):
        inp = synth.make_forward_inputs(B, seed)
        got_c, got_jp = _call(api, gmodel, inp, dtype, noalloc=True)
        ref_c, ref_jp = _call(api, gmodel, inp, dtype, noalloc=False)             # the library's default path (tcgen05 from 512 bodies)
        assert np.isfinite(got_c).all() and np.isfinite(got_jp).all()
        err = np.max(np.abs(got_c.astype(np.float64) - ref_c), axis=(1, 2)) / np.max(np.abs(ref_c), axis=(1, 2))
        assert err.max() < tol, (B, err.max())
        assert rel_err(got_jp, ref_jp) < tol
        same_c, same_jp = _call(api, gmodel, inp, dtype, noalloc=False, flags=sorted_flag if dtype == np.float32 else 0)
        assert np.array_equal(got_c, same_c) and np.array_equal(got_jp, same_jp), "same kernels, same inputs: bit-identical"


def test_noalloc_forward_rejects_batches_beyond_the_static_scratch(api, gmodel, synth):
    import torch
    capi, lib = api.capi, api.lib()
    B = MAX_BODIES + 1
    dev = torch.device("cuda", 0)
    w = torch.zeros(B, 10, device=dev); p = torch.zeros(B, 3, device=dev); r = torch.zeros(B, 24, 9, device=dev)
    P = lambda t: C.c_void_p(t.data_ptr())
    rc = lib.b3d_smpl_forward_noalloc_dev(gmodel.h, B, capi.F32, P(w), P(p), P(r), None, None, None)
    assert rc == EINVAL
    assert "B3D_NOALLOC_MAX_BODIES" in capi.last_error()
    # size 0 and null arguments behave like the existing entry point
    assert lib.b3d_smpl_forward_noalloc_dev(gmodel.h, 0, capi.F32, P(w), P(p), P(r), None, None, None) == 0
    assert lib.b3d_smpl_forward_noalloc_dev(None, 1, capi.F32, P(w), P(p), P(r), None, None, None) == EINVAL
