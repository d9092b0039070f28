"""ctypes binding of the CPU oracle (oracle/liboracle.so).  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this module;
the product package never does.  Parity unpinned by the reference (it ships no tests) — see oracle.hpp.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

NJ, NS, NPF, MAXKP, NFULL = 24, 10, 76, 17, 86
LS_DENSE_QR, LS_DENSE_CHOLESKY, LS_BLOCK_TRIDIAG = 0, 1, 2

KP_DTYPE = np.dtype([("jid", "<i4"), ("pad", "<i4"), ("u", "<f8"), ("v", "<f8")])
ITER_DTYPE = np.dtype([("cost", "<f8"), ("cost_change", "<f8"), ("gradient_max_norm", "<f8"), ("step_norm", "<f8"),
                       ("relative_decrease", "<f8"), ("radius", "<f8"), ("flags", "<i4"), ("aux", "<i4")])


class FitOptions(C.Structure):
    _fields_ = [("mode", C.c_int), ("max_iters", C.c_int), ("beta_pose", C.c_double), ("beta_shape", C.c_double),
                ("lambda_temp", C.c_double), ("fx", C.c_double), ("fy", C.c_double), ("cx", C.c_double), ("cy", C.c_double),
                ("huber", C.c_double), ("scale_lo", C.c_double), ("scale_hi", C.c_double), ("linear_solver", C.c_int),
                ("use_jets", C.c_int)]


class SeqOptions(C.Structure):
    _fields_ = [("skip", C.c_int32), ("wsize", C.c_int32), ("overlap", C.c_int32), ("it1", C.c_int32), ("it2", C.c_int32),
                ("num_shards", C.c_int32), ("nthreads", C.c_int32), ("pad", C.c_int32), ("beta_pose", C.c_double),
                ("beta_shape", C.c_double), ("lambda_temp", C.c_double), ("beta_shape_lock", C.c_double)]


def build(force: bool = False) -> str:
    so = os.path.join(_HERE, "liboracle.so")
    srcs = [os.path.join(_HERE, f) for f in os.listdir(_HERE) if f.endswith((".cpp", ".hpp"))]
    if force or not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
        subprocess.run(["make", "-C", _HERE, "liboracle.so"], check=True, capture_output=True)
    return so


def lib():
    global _LIB
    if _LIB is None:
        _LIB = C.CDLL(build())
        _LIB.orc_model_create.restype = C.c_void_p
        _LIB.orc_gmm_load.restype = C.c_void_p
        _LIB.orc_gmm_create.restype = C.c_void_p
        assert _LIB.orc_sizeof_iter_record() == ITER_DTYPE.itemsize
        assert _LIB.orc_sizeof_fit_options() == C.sizeof(FitOptions)
        assert _LIB.orc_sizeof_kp() == KP_DTYPE.itemsize
    return _LIB


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def fit_options(mode=0, max_iters=100, beta_pose=0.0, beta_shape=0.0, lambda_temp=0.0, cam=(432.0, 432.0, 240.0, 135.0),
                linear_solver=LS_DENSE_QR, use_jets=1) -> FitOptions:
    return FitOptions(mode, max_iters, beta_pose, beta_shape, lambda_temp, cam[0], cam[1], cam[2], cam[3], 3.0, 0.3, 3.0,
                      linear_solver, use_jets)


class Model:
    def __init__(self, model: dict, use_posedirs: bool = False):
        L = lib()
        self.V, self.J, self.S = model["v_template"].shape[0], model["kintree_table"].shape[1], model["shapedirs"].shape[2]
        pd = _f64(model["posedirs"].reshape(self.V * 3, -1)) if use_posedirs else None
        kt = np.ascontiguousarray(model["kintree_table"][0], dtype=np.uint32)
        self.h = C.c_void_p(L.orc_model_create(self.V, self.J, self.S, _p(_f64(model["v_template"])), _p(_f64(model["shapedirs"])),
                                               _p(_f64(model["J_regressor"])), _p(_f64(model["weights"])), _p(kt), _p(pd),
                                               0 if pd is None else pd.shape[1]))
        assert self.h.value

    def __del__(self):
        if getattr(self, "h", None) is not None and self.h.value:
            lib().orc_model_free(self.h)
            self.h = None

    def constants(self):
        parent = np.zeros(self.J, dtype=np.int32)
        j0, jsr, off = np.zeros((self.J, 3)), np.zeros((self.J, 3, self.S)), np.zeros((self.J, 3))
        lib().orc_model_get(self.h, _p(parent), _p(j0), _p(jsr), _p(off))
        return dict(parent=parent, J0=j0, JSR=jsr, offset=off)

    def forward(self, w, p, r, nthreads=1, want_cloud=True):
        w, p, r = _f64(w), _f64(p), _f64(r)
        b = w.shape[0]
        cloud = np.zeros((b, self.V, 3)) if want_cloud else None
        jp, jt = np.zeros((b, self.J, 3)), np.zeros((b, self.J, 12))
        lib().orc_forward(self.h, b, _p(w), _p(p), _p(r), _p(cloud), _p(jp), _p(jt), nthreads)
        return cloud, jp, jt

    def icp_eval(self, w, p, q, cptr, mpt, dpt, dptr, data, shape=True, want_jac=False, nthreads=1):
        """AvatarOptimizer's ICP term: cost (B,), grad (B, 85) [p | 24 x 3 rotation tangents | 10 shape]; with want_jac also
        resid (ncorr, 3) and the dense Jacobian (ncorr, 3, 85)."""
        w, p, q, data = _f64(w), _f64(p), _f64(q), _f64(data)
        b = w.shape[0]
        cptr, dptr = np.ascontiguousarray(cptr, dtype=np.int64), np.ascontiguousarray(dptr, dtype=np.int64)
        mpt, dpt = np.ascontiguousarray(mpt, dtype=np.int32), np.ascontiguousarray(dpt, dtype=np.int32)
        nc = 3 + 3 * self.J + self.S
        cost, grad = np.zeros(b), np.zeros((b, nc))
        resid = np.zeros((int(cptr[-1]), 3)) if want_jac else None
        jac = np.zeros((int(cptr[-1]), 3, nc)) if want_jac else None
        lib().orc_icp_eval(self.h, b, _p(w), _p(p), _p(q), _p(cptr), _p(mpt), _p(dpt), _p(dptr), _p(data), int(shape), _p(resid),
                           _p(jac), _p(cost), _p(grad), nthreads)
        return dict(cost=cost, grad=grad, resid=resid, jac=jac)

    def icp_minimize(self, gmm, beta_pose, beta_shape, w, p, q, cptr, mpt, dpt, dptr, data, shape=True, max_iterations=50,
                     function_tolerance=1e-4, gradient_tolerance=1e-10, parameter_tolerance=1e-8, trace_cap=128, nthreads=1):
        """The ceres::Solve of AvatarOptimizer::optimize (LINE_SEARCH, BFGS, WOLFE, CUBIC) on the ICP objective; returns the
        minimiser (w, p, q), per-iteration records and a summary per body."""
        w, p, q, data = _f64(w).copy(), _f64(p).copy(), _f64(q).copy(), _f64(data)
        b = w.shape[0]
        cptr, dptr = np.ascontiguousarray(cptr, dtype=np.int64), np.ascontiguousarray(dptr, dtype=np.int64)
        mpt, dpt = np.ascontiguousarray(mpt, dtype=np.int32), np.ascontiguousarray(dpt, dtype=np.int32)
        trace, nrec, summ = np.zeros((b, trace_cap, 7)), np.zeros(b, dtype=np.int32), np.zeros((b, 6))
        lib().orc_icp_minimize(self.h, gmm.h if gmm is not None else None, b, C.c_double(beta_pose), C.c_double(beta_shape), _p(cptr),
                               _p(mpt), _p(dpt), _p(dptr), _p(data), int(shape), int(max_iterations), C.c_double(function_tolerance),
                               C.c_double(gradient_tolerance), C.c_double(parameter_tolerance), _p(w), _p(p), _p(q), _p(trace),
                               int(trace_cap), _p(nrec), _p(summ), nthreads)
        return dict(w=w, p=p, q=q, trace=trace, nrec=nrec, initial_cost=summ[:, 0], final_cost=summ[:, 1],
                    iterations=summ[:, 2].astype(int), evaluations=summ[:, 3].astype(int), restarts=summ[:, 4].astype(int),
                    termination=summ[:, 5].astype(int))

    def reproj_check(self, with_shape, jid, u, v, cam, R0, x86):
        out = [np.zeros(2), np.zeros(2), np.zeros((2, NFULL)), np.zeros(2), np.zeros((2, NFULL))]
        lib().orc_reproj_check(self.h, int(with_shape), int(jid), C.c_double(u), C.c_double(v), _p(_f64(cam)), _p(_f64(R0)),
                               _p(_f64(x86)), *[_p(o) for o in out])
        return out

    def fit_single(self, gmm, opts: FitOptions, kps, nkp, R0, sim3, w, nthreads=1, want_trace=True):
        b = kps.shape[0]
        kps = np.ascontiguousarray(kps, dtype=KP_DTYPE)
        nkp = np.ascontiguousarray(nkp, dtype=np.int32)
        R0, sim3, w = _f64(R0), _f64(sim3).copy(), _f64(w).copy()
        jaa = np.zeros((b, self.J, 3))
        usable = np.zeros(b, dtype=np.uint8)
        stride = opts.max_iters + 2
        trace = np.zeros((b, stride), dtype=ITER_DTYPE) if want_trace else None
        niters, term = np.zeros(b, dtype=np.int32), np.zeros(b, dtype=np.int32)
        lib().orc_fit_single_batch(self.h, gmm.h if gmm is not None else None, C.byref(opts), b, _p(kps), _p(nkp), _p(R0), _p(sim3),
                                   _p(w), _p(jaa), _p(usable), _p(trace), stride, _p(niters), _p(term), nthreads)
        return dict(sim3=sim3, w=w, jointAA=jaa, usable=usable, trace=trace, niters=niters, term=term)

    def fit_window(self, opts: FitOptions, kps, nkp, R0, poses, w, want_trace=True):
        f = kps.shape[0]
        kps = np.ascontiguousarray(kps, dtype=KP_DTYPE)
        nkp = np.ascontiguousarray(nkp, dtype=np.int32)
        R0, poses, w = _f64(R0), _f64(poses).copy(), _f64(w).copy()
        cap = opts.max_iters + 2
        trace = np.zeros(cap, dtype=ITER_DTYPE) if want_trace else None
        niters, term = C.c_int32(0), C.c_int32(0)
        ok = lib().orc_fit_window(self.h, C.byref(opts), f, _p(kps), _p(nkp), _p(R0), _p(poses), _p(w), _p(trace), cap,
                                  C.byref(niters), C.byref(term))
        return dict(poses=poses, w=w, usable=bool(ok), trace=None if trace is None else trace[:niters.value],
                    niters=niters.value, term=term.value)

    def fit_sequence(self, base: FitOptions, so: SeqOptions, kps, nkp, poses, R0, W):
        n = kps.shape[0]
        kps = np.ascontiguousarray(kps, dtype=KP_DTYPE)
        nkp = np.ascontiguousarray(nkp, dtype=np.int32)
        poses, R0, W = _f64(poses).copy(), _f64(R0).copy(), _f64(W).copy()
        r, p = np.zeros((n, self.J, 9)), np.zeros((n, 3))
        nwin = len(range(0, n, so.wsize - so.overlap))
        cap = 1 + nwin + max(so.num_shards, 1)
        wi, wt = np.zeros(cap, dtype=np.int32), np.zeros(cap, dtype=np.int32)
        nfail = lib().orc_fit_sequence(self.h, C.byref(base), C.byref(so), n, _p(kps), _p(nkp), _p(poses), _p(R0), _p(W), _p(r),
                                       _p(p), _p(wi), _p(wt))
        return dict(poses=poses, R0=R0, W=W, r=r, p=p, nfail=nfail, win_iters=wi, win_term=wt)


class GMM:
    def __init__(self, path: str | None = None, prior: dict | None = None):
        st = C.c_int(0)
        if path is not None:
            self.h = C.c_void_p(lib().orc_gmm_load(path.encode(), C.byref(st)))
        else:
            self.h = C.c_void_p(lib().orc_gmm_create(prior["nComps"], prior["nDims"], _p(_f64(prior["weights"])),
                                                     _p(_f64(prior["means"])), _p(_f64(prior["covs"])), C.byref(st)))
        self.status = st.value
        nc, nd = C.c_int(0), C.c_int(0)
        lib().orc_gmm_dims(self.h, C.byref(nc), C.byref(nd))
        self.nComps, self.nDims = nc.value, nd.value

    def __del__(self):
        if getattr(self, "h", None) is not None and self.h.value:
            lib().orc_gmm_free(self.h)
            self.h = None

    def arrays(self):
        nc, nd = self.nComps, self.nDims
        out = dict(weights=np.zeros(nc), means=np.zeros((nc, nd)), covs=np.zeros((nc, nd, nd)), prec_cho=np.zeros((nc, nd, nd)),
                   consts_log=np.zeros(nc))
        lib().orc_gmm_get(self.h, _p(out["weights"]), _p(out["means"]), _p(out["covs"]), _p(out["prec_cho"]), _p(out["consts_log"]))
        return out

    def residual(self, x):
        out = np.zeros(self.nDims + 1)
        c = lib().orc_gmm_residual(self.h, _p(_f64(x)), _p(out))
        return out, c


def icp_priors(gmm, beta_pose, beta_shape, totals, w, q, cost, grad):
    """Adds AvatarOptimizer's pose-prior and shape-prior blocks (weights scaled by sqrt(#ICP residual blocks) / 15) to cost (B,) and
    grad (B, 85); returns the new arrays."""
    w, q, totals = _f64(w), _f64(q), _f64(totals)
    cost, grad = _f64(cost).copy(), _f64(grad).copy()
    lib().orc_icp_priors(gmm.h if gmm is not None else None, w.shape[0], C.c_double(beta_pose), C.c_double(beta_shape), _p(totals), _p(w),
                         _p(q), _p(cost), _p(grad))
    return cost, grad


def icp_visible(cloud, faces):
    """pointVisible of AvatarOptimizer's ICP loop for B posed clouds (B, V, 3); faces (F, 3)."""
    cloud = _f64(cloud)
    faces = np.ascontiguousarray(faces, dtype=np.uint32)
    vis = np.zeros(cloud.shape[:2], dtype=np.uint8)
    lib().orc_icp_visible(cloud.shape[0], cloud.shape[1], faces.shape[0], _p(faces), _p(cloud), _p(vis))
    return vis


def icp_nearest(cloud, model_part, visible, dptr, data, data_part, nthreads=1):
    """findNN (invert = true) for B bodies: cloud (B, V, 3), model_part (V,), visible (B, V) uint8 or None, data points of body b at
    dptr[b]:dptr[b+1].  Returns the matched model point per data point (-1: none)."""
    cloud, data = _f64(cloud), _f64(data)
    b, v = cloud.shape[0], cloud.shape[1]
    model_part = np.ascontiguousarray(model_part, dtype=np.int32)
    data_part = np.ascontiguousarray(data_part, dtype=np.int32)
    dptr = np.ascontiguousarray(dptr, dtype=np.int64)
    vis = None if visible is None else np.ascontiguousarray(visible, dtype=np.uint8)
    out = np.full(data.shape[0], -2, dtype=np.int32)
    lib().orc_icp_nearest(b, v, _p(cloud), _p(model_part), _p(vis), _p(dptr), _p(data), _p(data_part), _p(out), nthreads)
    return out


def quat_plus(x, d):
    """FakeQuaternionParameterization::Plus on n quaternions (x, y, z, w)."""
    x, d = _f64(x).reshape(-1, 4), _f64(d).reshape(-1, 3)
    out = np.zeros_like(x)
    lib().orc_quat_plus(x.shape[0], _p(x), _p(d), _p(out))
    return out


def aa_to_matrix(aa):
    R = np.zeros(9)
    lib().orc_aa_to_matrix(_p(_f64(aa)), _p(R))
    return R


def writeback(poses, R0):
    poses, R0 = _f64(poses), _f64(R0).copy()
    f = poses.shape[0]
    r, p = np.zeros((f, NJ, 9)), np.zeros((f, 3))
    lib().orc_writeback(f, _p(poses), _p(R0), _p(r), _p(p))
    return R0, r, p


def map_mediapipe(landmarks, width, height, variant="utils"):
    """MediaPipe-33 -> PixelKP list, restating load_mp_json (include/Utils.h:61-99; variant 'main_single' follows the
    divergent copy in src/main_single_frame.cpp:69-102).  ``landmarks``: parsed JSON (list of dicts).  Pure-Python, small."""
    MP_MAP = [-1, 23, 24, -1, 25, 26, -1, 27, 28, -1, 31, 32, -1, -1, -1, 0, 11, 12, 13, 14, 15, 16, -1, -1]
    USE_SMPL = [1, 2, 4, 5, 7, 8, 10, 11, 15, 16, 17, 18, 19, 20, 21, 0, 0]

    def num(j, i, key):
        if not isinstance(j, list) or i >= len(j) or not isinstance(j[i], dict):
            return None
        v = j[i].get(key)
        return float(v) if isinstance(v, (int, float)) and not isinstance(v, bool) else None

    j = landmarks
    if variant == "main_single" and (not isinstance(j, list) or len(j) < 33):
        raise ValueError("reference loader would throw on a null landmark (src/main_single_frame.cpp:74-80)")
    midvis_default = 1.0 if variant == "utils" else 0.0

    def mid(a, b):
        xa, ya, xb, yb = num(j, a, "x"), num(j, a, "y"), num(j, b, "x"), num(j, b, "y")
        if None in (xa, ya, xb, yb):
            return None
        va, vb = num(j, a, "visibility"), num(j, b, "visibility")
        va = midvis_default if va is None else va
        vb = midvis_default if vb is None else vb
        return 0.5 * (xa + xb), 0.5 * (ya + yb), min(va, vb)

    pel, che = mid(23, 24), mid(11, 12)
    out = []
    for sid in USE_SMPL:
        if sid == 0:
            if pel is None:
                continue
            x, y, vis = pel
        elif sid == 6:
            if che is None:
                continue
            x, y, vis = che
        else:
            mp = MP_MAP[sid]
            if mp < 0:
                continue
            x, y = num(j, mp, "x"), num(j, mp, "y")
            if x is None or y is None:
                continue
            vis = num(j, mp, "visibility")
            vis = 1.0 if vis is None else vis
        if vis < 0.5:
            continue
        out.append((sid, x * width, y * height))
    return out


def fill_triangle(img, pts, color):
    """cv2.fillConvexPoly(img, pts, (color,)*ch, LINE_8) restated (oracle/render.cpp); img H x W x ch uint8, in place."""
    h, w, ch = img.shape
    assert img.dtype == np.uint8 and img.flags.c_contiguous
    xy = np.ascontiguousarray(pts, dtype=np.int32).reshape(6)
    lib().orc_fill_triangle(_p(img), w, h, ch, _p(xy), int(color))
    return img


def render_overlay(cloud, faces, cam, images, backface_cull=True, want_order=False):
    """renderSMPLMesh (include/RenderSMPLMesh.h:16-110) for B bodies: cloud (B, V, 3) any float dtype (converted to double exactly),
    faces (F, 3), images (B, H, W, 3) uint8 modified in place.  Returns (images, drawn (B,), order (B, F) or None)."""
    cloud = np.ascontiguousarray(cloud, dtype=np.float64)
    b, v, _ = cloud.shape
    faces = np.ascontiguousarray(faces, dtype=np.int32)
    f = faces.shape[0]
    assert images.dtype == np.uint8 and images.flags.c_contiguous and images.shape[0] == b and images.shape[3] == 3
    h, w = images.shape[1:3]
    drawn = np.zeros(b, dtype=np.int32)
    order = np.zeros((b, f), dtype=np.int32) if want_order else None
    lib().orc_render_overlay(b, v, f, _p(cloud), _p(faces), C.c_double(cam[0]), C.c_double(cam[1]), C.c_double(cam[2]), C.c_double(cam[3]),
                             w, h, 1 if backface_cull else 0, _p(images), _p(drawn), _p(order))
    return images, drawn, order
