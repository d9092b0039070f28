"""Host-side mirror of the reference's interface for the hot path, on top of the C-ABI (include/b3d.h).

Names, argument meaning and error behaviour follow the reference (jonH34400/3DBodyAnimation):
  AvatarModel / Avatar / Avatar.update        external/avatar/include/Avatar.h:64-220, Avatar.cpp:22-75
  GaussianMixture                             external/avatar/include/GaussianMixture.h:9-49
  PixelKP, Sim3Params, FramePoseParams        include/Sim3BA.h:9-19, include/MultiFrameBA.h:9-14
  OptimizePoseReprojection                    include/Sim3BA.h:515-683
  OptimizePoseShapeReprojection               include/Sim3BA.h:348-511
  mean_pixel_error                            include/Utils.h:102-115
plus the batched calls the GPU wants (smpl_forward, fit_single_batch).  Every computation goes through
lib3dba_b200.so; there is no numpy/CPU fallback.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass, field

import numpy as np

from . import _capi as capi
from ._capi import F32, F64, FIT_POSE, FIT_POSE_SHAPE, ITER_DTYPE, KP_DTYPE, MAXKP, NJ, NPF, NS, check, lib, ptr

# MediaPipe-33 -> SMPL-24 tables (include/Utils.h:18-23): bit-exact indexing outputs
MP_MAP = np.array([-1, 23, 24, -1, 25, 26, -1, 27, 28, -1, 31, 32, -1, -1, -1, 0, 11, 12, 13, 14, 15, 16, -1, -1], dtype=np.int32)
USE_SMPL = np.array([1, 2, 4, 5, 7, 8, 10, 11, 15, 16, 17, 18, 19, 20, 21, 0, 0], dtype=np.int32)


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def fit_opts(mode=FIT_POSE, max_iters=100, beta_pose=0.0, beta_shape=0.0, lambda_temp=0.0, cam=(432.0, 432.0, 240.0, 135.0),
             huber=3.0, scale_bounds=(0.3, 3.0)) -> capi.FitOpts:
    return capi.FitOpts(mode, max_iters, beta_pose, beta_shape, lambda_temp, cam[0], cam[1], cam[2], cam[3], huber,
                        scale_bounds[0], scale_bounds[1])


class AvatarModel:
    """ark::AvatarModel: built from the arrays of SMPL's model.npz (a dict or an .npz path)."""

    def __init__(self, model):
        if isinstance(model, str):
            with np.load(model) as z:
                model = {k: z[k] for k in z.files}
        self._keep = {}
        dt = np.float64 if model["v_template"].dtype == np.float64 else np.float32
        for k in ("v_template", "shapedirs", "J_regressor", "weights"):
            self._keep[k] = np.ascontiguousarray(model[k], dtype=dt)
        kt = np.ascontiguousarray(np.asarray(model["kintree_table"])[0].astype(np.uint32))
        faces = np.ascontiguousarray(model["f"], dtype=np.uint32) if "f" in model else None
        pd = np.ascontiguousarray(model["posedirs"], dtype=dt) if "posedirs" in model else None
        self._keep.update(kt=kt, faces=faces, pd=pd)
        v, j, s = self._keep["v_template"].shape[0], kt.shape[0], self._keep["shapedirs"].shape[2]
        desc = capi.ModelDesc(v, j, s, 0 if faces is None else faces.shape[0], F64 if dt == np.float64 else F32, 0,
                              ptr(self._keep["v_template"]), ptr(self._keep["shapedirs"]), ptr(self._keep["J_regressor"]),
                              ptr(self._keep["weights"]), ptr(kt), ptr(faces), ptr(pd))
        h = C.c_void_p()
        check(lib().b3d_model_create(C.byref(desc), C.byref(h)), "b3d_model_create")
        self.h = h
        self.numPoints, self.numJoints, self.numShapeKeys = v, j, s
        self.parent = np.zeros(NJ, dtype=np.int32)
        check(lib().b3d_model_get_parents(self.h, ptr(self.parent)), "b3d_model_get_parents")
        self.initialJointPos, self.jointShapeReg, self.offset = np.zeros((NJ, 3)), np.zeros((NJ, 3, NS)), np.zeros((NJ, 3))
        check(lib().b3d_model_get_constants(self.h, ptr(self.initialJointPos), ptr(self.jointShapeReg), ptr(self.offset)),
              "b3d_model_get_constants")
        self.useJointShapeRegressor = True
        self.posePrior = None      # GaussianMixture or None (AvatarModel.cpp:296 loads pose_prior.txt when present)

    @classmethod
    def from_npz_native(cls, path: str, with_posedirs: bool = False):
        """The same model through the library's own .npz reader (b3d_model_load_npz: zip + inflate + .npy parsing in C++ with
        the array views of external/avatar/AvatarModel.cpp:23-102) instead of numpy."""
        self = cls.__new__(cls)
        self._keep = {}
        h = C.c_void_p()
        check(lib().b3d_model_load_npz(os.fsencode(path), 1 if with_posedirs else 0, C.byref(h)), "b3d_model_load_npz")
        self.h = h
        shp, nd = (C.c_int64 * 4)(), C.c_int32(0)
        check(lib().b3d_npz_shape(os.fsencode(path), b"shapedirs", C.byref(nd), shp, None, None), "b3d_npz_shape")
        self.numPoints, self.numJoints, self.numShapeKeys = int(lib().b3d_model_num_verts(h)), NJ, int(shp[2])
        self.parent = np.zeros(NJ, dtype=np.int32)
        check(lib().b3d_model_get_parents(self.h, ptr(self.parent)), "b3d_model_get_parents")
        self.initialJointPos, self.jointShapeReg, self.offset = np.zeros((NJ, 3)), np.zeros((NJ, 3, NS)), np.zeros((NJ, 3))
        check(lib().b3d_model_get_constants(self.h, ptr(self.initialJointPos), ptr(self.jointShapeReg), ptr(self.offset)),
              "b3d_model_get_constants")
        self.useJointShapeRegressor = True
        self.posePrior = None
        return self

    def __del__(self):
        h = getattr(self, "h", None)
        if h is not None and h.value and lib is not None:      # (lib is None while the interpreter shuts down)
            lib().b3d_model_destroy(h)
            self.h = None


def npz_shape(path: str, name: str):
    """(shape, element bytes, fortran_order) of one member of an .npz, read by the library (no GPU needed)."""
    shp, nd, ws, fo = (C.c_int64 * 4)(), C.c_int32(0), C.c_int32(0), C.c_int32(0)
    check(lib().b3d_npz_shape(os.fsencode(path), name.encode(), C.byref(nd), shp, C.byref(ws), C.byref(fo)), "b3d_npz_shape")
    return tuple(int(shp[i]) for i in range(nd.value)), ws.value, bool(fo.value)


def npz_float_matrix(path: str, name: str, rows: int, cols: int) -> np.ndarray:
    """util::loadFloatMatrix(npz[name], rows, cols) (external/avatar/Util.cpp:249-271), row-major float64."""
    out = np.zeros((rows, cols))
    check(lib().b3d_npz_float_matrix(os.fsencode(path), name.encode(), C.c_int64(rows), C.c_int64(cols), ptr(out)), "b3d_npz_float_matrix")
    return out


def npz_uint_matrix(path: str, name: str, rows: int, cols: int) -> np.ndarray:
    """util::loadUintMatrix(npz[name], rows, cols) (external/avatar/Util.cpp:272-298), row-major uint32."""
    out = np.zeros((rows, cols), dtype=np.uint32)
    check(lib().b3d_npz_uint_matrix(os.fsencode(path), name.encode(), C.c_int64(rows), C.c_int64(cols), ptr(out)), "b3d_npz_uint_matrix")
    return out


class GaussianMixture:
    """ark::GaussianMixture.  nComps == -1 when the prior file is missing (GaussianMixture.cpp:15-19)."""

    def __init__(self, path: str | None = None, prior: dict | None = None):
        self.h = None
        self.nComps = -1
        self.nDims = 0
        h = C.c_void_p()
        if path is not None:
            rc = lib().b3d_gmm_load(path.encode(), C.byref(h))
            if rc == 4:      # B3D_ENOFILE: warn-and-continue, like the reference
                return
            check(rc, "b3d_gmm_load")
        elif prior is not None:
            check(lib().b3d_gmm_create(int(prior["nComps"]), int(prior["nDims"]), ptr(_f64(prior["weights"])),
                                       ptr(_f64(prior["means"])), ptr(_f64(prior["covs"])), C.byref(h)), "b3d_gmm_create")
        else:
            return
        self.h = h
        nc, nd = C.c_int32(0), C.c_int32(0)
        check(lib().b3d_gmm_dims(self.h, C.byref(nc), C.byref(nd)), "b3d_gmm_dims")
        self.nComps, self.nDims = nc.value, nd.value

    def arrays(self):
        out = dict(prec_cho=np.zeros((self.nComps, self.nDims, self.nDims)), consts_log=np.zeros(self.nComps))
        check(lib().b3d_gmm_get(self.h, ptr(out["prec_cho"]), ptr(out["consts_log"])), "b3d_gmm_get")
        return out

    def residual(self, x):
        """GaussianMixture::residual for one (69,) or a batch (B, 69) of poses; returns (out (B,70), comp (B,))."""
        x = np.atleast_2d(_f64(x))
        out, comp = np.zeros((x.shape[0], self.nDims + 1)), np.zeros(x.shape[0], dtype=np.int32)
        check(lib().b3d_gmm_residual(self.h, x.shape[0], ptr(x), ptr(out), ptr(comp)), "b3d_gmm_residual")
        return out, comp

    def __del__(self):
        h = getattr(self, "h", None)
        if h is not None and h.value:
            lib().b3d_gmm_destroy(h)
            self.h = None


def _gmm_handle(gmm):
    return gmm.h if (gmm is not None and gmm.h is not None and gmm.nComps > 0) else None


# ---------------------------------------------------------------------------------------------------------
# batched calls
# ---------------------------------------------------------------------------------------------------------
LBS_FLAGS = {None: 0, "tc": 0x10, "sorted": 0x20, "bodylane": 0x40, "generic": 0x80}   # B3D_FWD_LBS_* (include/b3d.h)


def smpl_forward(model: AvatarModel, w, p, r, dtype=np.float64, posedirs=False, want_cloud=True, lbs=None):
    """Batched Avatar::update.  w (B,10), p (B,3), r (B,24,9 col-major).  Returns cloud (B,V,3), jointPos (B,24,3),
    jointTrans (B,24,12) in ``dtype`` (float32 or float64)."""
    dt = np.dtype(dtype)
    code = F64 if dt == np.float64 else F32
    w = np.ascontiguousarray(w, dtype=dt).reshape(-1, NS)
    b = w.shape[0]
    p = np.ascontiguousarray(p, dtype=dt).reshape(b, 3)
    r = np.ascontiguousarray(r, dtype=dt).reshape(b, NJ, 9)
    cloud = np.empty((b, model.numPoints, 3), dtype=dt) if want_cloud else None
    jp, jt = np.empty((b, NJ, 3), dtype=dt), np.empty((b, NJ, 12), dtype=dt)
    check(lib().b3d_smpl_forward(model.h, b, code, ptr(w), ptr(p), ptr(r), ptr(cloud), ptr(jp), ptr(jt),
                                 (1 if posedirs else 0) | LBS_FLAGS[lbs]),
          "b3d_smpl_forward")
    return cloud, jp, jt


def fit_single_batch(model: AvatarModel, gmm, opts: capi.FitOpts, kps, nkp, R0, sim3, w, want_trace=True):
    """B independent single-frame fits (frame loop of src/main_single_frame.cpp:192-249)."""
    kps = np.ascontiguousarray(kps, dtype=KP_DTYPE).reshape(-1, MAXKP)
    b = kps.shape[0]
    nkp = np.ascontiguousarray(nkp, dtype=np.int32).reshape(b)
    R0, sim3, w = _f64(R0).reshape(b, 9), _f64(sim3).reshape(b, 7).copy(), _f64(w).reshape(b, NS).copy()
    jaa = np.zeros((b, NJ, 3))
    usable = np.zeros(b, dtype=np.uint8)
    stride = opts.max_iters + 2
    trace = np.zeros((b, stride), dtype=ITER_DTYPE) if want_trace else None
    niters, term = np.zeros(b, dtype=np.int32), np.zeros(b, dtype=np.int32)
    check(lib().b3d_fit_single(model.h, _gmm_handle(gmm), C.byref(opts), b, ptr(kps), ptr(nkp), ptr(R0), ptr(sim3), ptr(w),
                               ptr(jaa), ptr(usable), ptr(trace), stride, ptr(niters), ptr(term)), "b3d_fit_single")
    return dict(sim3=sim3, w=w, jointAA=jaa, usable=usable, trace=trace, niters=niters, term=term)


class Workspace:
    """Reusable device buffers + one CUDA stream for fit_frames (b3d_workspace)."""

    def __init__(self):
        h = C.c_void_p()
        check(lib().b3d_workspace_create(C.byref(h)), "b3d_workspace_create")
        self.h = h

    def __del__(self):
        h = getattr(self, "h", None)
        if h is not None and h.value:
            lib().b3d_workspace_destroy(h)
            self.h = None


def fit_frames(ws: Workspace, model: AvatarModel, gmm, opts: capi.FitOpts, kps, nkp, R0, sim3, w, out=None, cloud_dtype=np.float32,
               want_cloud=False):
    """The per-frame body of src/main_single_frame.cpp:205-255 for B frames in one call: fit -> write-back -> update ->
    mean pixel error.  Host arrays in, host arrays out; sim3 and w are updated in place when they are float64 C arrays.
    ``out`` may carry preallocated (e.g. pinned) output arrays keyed jointAA/usable/niters/term/r/p/jointPos/px_err/cloud."""
    kps = np.ascontiguousarray(kps, dtype=KP_DTYPE).reshape(-1, MAXKP)
    b = kps.shape[0]
    nkp = np.ascontiguousarray(nkp, dtype=np.int32).reshape(b)
    R0 = _f64(R0).reshape(b, 9)
    sim3 = sim3 if (isinstance(sim3, np.ndarray) and sim3.dtype == np.float64 and sim3.flags.c_contiguous) else _f64(sim3).copy()
    w = w if (isinstance(w, np.ndarray) and w.dtype == np.float64 and w.flags.c_contiguous) else _f64(w).copy()
    out = {} if out is None else out
    cdt = np.dtype(cloud_dtype)

    def buf(name, shape, dt):
        a = out.get(name)
        if a is None:
            a = out[name] = np.zeros(shape, dtype=dt)
        return a

    jaa, us = buf("jointAA", (b, NJ, 3), np.float64), buf("usable", (b,), np.uint8)
    ni, te = buf("niters", (b,), np.int32), buf("term", (b,), np.int32)
    r, p, jp = buf("r", (b, NJ, 9), np.float64), buf("p", (b, 3), np.float64), buf("jointPos", (b, NJ, 3), np.float64)
    pe = buf("px_err", (b,), np.float64)
    cloud = buf("cloud", (b, model.numPoints, 3), cdt) if want_cloud else None
    check(lib().b3d_fit_frames(ws.h, model.h, _gmm_handle(gmm), C.byref(opts), b, ptr(kps), ptr(nkp), ptr(R0), ptr(sim3), ptr(w),
                               ptr(jaa), ptr(us), ptr(ni), ptr(te), ptr(r), ptr(p), ptr(jp), ptr(pe), ptr(cloud),
                               F64 if cdt == np.float64 else F32), "b3d_fit_frames")
    out.update(sim3=sim3, w=w)
    return out


def fit_window(model: AvatarModel, opts: capi.FitOpts, kps, nkp, R0, poses, w, want_trace=True):
    """One OptimizeMultiFrame solve (include/MultiFrameBA.h:33-177) over F frames; w is the shared shape block."""
    kps = np.ascontiguousarray(kps, dtype=KP_DTYPE).reshape(-1, MAXKP)
    f = kps.shape[0]
    nkp = np.ascontiguousarray(nkp, dtype=np.int32).reshape(f)
    R0, poses, w = _f64(R0).reshape(f, 9), _f64(poses).reshape(f, NPF).copy(), _f64(w).reshape(NS).copy()
    cap = opts.max_iters + 2
    trace = np.zeros(cap, dtype=ITER_DTYPE) if want_trace else None
    usable, niters, term = np.zeros(1, dtype=np.uint8), np.zeros(1, dtype=np.int32), np.zeros(1, dtype=np.int32)
    check(lib().b3d_fit_window(model.h, C.byref(opts), f, ptr(kps), ptr(nkp), ptr(R0), ptr(poses), ptr(w), ptr(usable), ptr(trace), cap,
                               ptr(niters), ptr(term)), "b3d_fit_window")
    n = int(niters[0])
    return dict(poses=poses, w=w, usable=bool(usable[0]), trace=None if trace is None else trace[:n], niters=n, term=int(term[0]))


def seq_opts(skip=10, wsize=20, overlap=5, it1=1000, it2=60, num_shards=1, beta_pose=5.0, beta_shape=25.0, lambda_temp=3.0,
             beta_shape_lock=1e5) -> capi.SeqOpts:
    """Defaults are the code defaults of src/main_multi_frame.cpp:29-36 (it2 = 60 and beta_shape_lock = 1e5 are hard-coded there)."""
    return capi.SeqOpts(skip, wsize, overlap, it1, it2, num_shards, beta_pose, beta_shape, lambda_temp, beta_shape_lock)


def fit_sequences(model: AvatarModel, base: capi.FitOpts, so: capi.SeqOpts, kps, nkp, poses, R0, W, seq_first=None, seq_len=None):
    """3dba_multi's anchor + sliding-window schedule for one or several back-to-back sequences (b3d_fit_sequences)."""
    kps = np.ascontiguousarray(kps, dtype=KP_DTYPE).reshape(-1, MAXKP)
    n = kps.shape[0]
    nkp = np.ascontiguousarray(nkp, dtype=np.int32).reshape(n)
    poses, R0, W = _f64(poses).reshape(n, NPF).copy(), _f64(R0).reshape(n, 9).copy(), _f64(W).reshape(n, NS).copy()
    sf = np.ascontiguousarray([0] if seq_first is None else seq_first, dtype=np.int32)
    sl = np.ascontiguousarray([n] if seq_len is None else seq_len, dtype=np.int32)
    r, p = np.zeros((n, NJ, 9)), np.zeros((n, 3))
    stride = so.wsize - so.overlap
    cap = int(sum(1 + len(range(0, int(l), stride)) + max(so.num_shards, 1) for l in sl))
    it, te = np.zeros(cap, dtype=np.int32), np.zeros(cap, dtype=np.int32)
    nfail = C.c_int32(0)
    check(lib().b3d_fit_sequences(model.h, C.byref(base), C.byref(so), len(sf), ptr(sf), ptr(sl), n, ptr(kps), ptr(nkp), ptr(poses), ptr(R0),
                                  ptr(W), ptr(r), ptr(p), ptr(it), ptr(te), C.byref(nfail)), "b3d_fit_sequences")
    return dict(poses=poses, R0=R0, W=W, r=r, p=p, nfail=nfail.value, solve_iters=it, solve_term=te)


def writeback(poses, R0):
    """a15: R0 <- R(rootAA) R0; r[0] = R0, r[j] = R(jointAA_j); p = rootT.  poses (F,76), R0 (F,9)."""
    poses = _f64(poses).reshape(-1, NPF)
    f = poses.shape[0]
    R0 = _f64(R0).reshape(f, 9).copy()
    r, p = np.zeros((f, NJ, 9)), np.zeros((f, 3))
    check(lib().b3d_writeback(f, ptr(poses), ptr(R0), ptr(r), ptr(p)), "b3d_writeback")
    return R0, r, p


def mean_pixel_error_batch(kps, nkp, jointPos, cam):
    kps = np.ascontiguousarray(kps, dtype=KP_DTYPE).reshape(-1, MAXKP)
    f = kps.shape[0]
    nkp = np.ascontiguousarray(nkp, dtype=np.int32).reshape(f)
    jp = _f64(jointPos).reshape(f, NJ, 3)
    err = np.zeros(f)
    check(lib().b3d_mean_pixel_error(f, ptr(kps), ptr(nkp), ptr(jp), C.c_double(cam[0]), C.c_double(cam[1]), C.c_double(cam[2]),
                                     C.c_double(cam[3]), ptr(err)), "b3d_mean_pixel_error")
    return err


RENDER_NO_CULL = 1   # B3D_RENDER_NO_CULL


def render_overlay(model: AvatarModel, cloud, cam, images, backface_cull=True):
    """smpl::render::renderSMPLMesh (include/RenderSMPLMesh.h:16-110) for B bodies: cloud (B, V, 3) float32 / float64 in camera
    coordinates, images (B, H, W, 3) uint8 painted in place (and returned).  See b3d_render_overlay for the fill rule."""
    cloud = np.asarray(cloud)
    dt = np.float32 if cloud.dtype == np.float32 else np.float64
    cloud = np.ascontiguousarray(cloud, dtype=dt).reshape(-1, model.numPoints, 3)
    b = cloud.shape[0]
    if not (isinstance(images, np.ndarray) and images.dtype == np.uint8 and images.flags.c_contiguous and images.ndim == 4
            and images.shape[0] == b and images.shape[3] == 3):
        raise ValueError("images must be a C-contiguous (B, H, W, 3) uint8 array")
    h, w = images.shape[1:3]
    check(lib().b3d_render_overlay(model.h, b, F32 if dt == np.float32 else F64, ptr(cloud), C.c_double(cam[0]), C.c_double(cam[1]),
                                   C.c_double(cam[2]), C.c_double(cam[3]), w, h, 0 if backface_cull else RENDER_NO_CULL, ptr(images)),
          "b3d_render_overlay")
    return images


def renderSMPLMesh(model: AvatarModel, cloud, img, fx, fy, cx, cy, fill=True, backface_cull=True, wireframe=False):
    """The reference's call (RenderSMPLMesh.h:16-24) for one body; ``img`` (H, W, 3) uint8 is painted in place.  The face list is
    the model's.  wireframe = True is not available (both call sites pass false)."""
    if wireframe:
        raise NotImplementedError("wireframe overlay (cv::polylines LINE_AA) is not built")
    if fill:
        render_overlay(model, np.asarray(cloud)[None], (fx, fy, cx, cy), img[None], backface_cull)
    return img


# ---------------------------------------------------------------------------------------------------------
# the reference's per-object interface
# ---------------------------------------------------------------------------------------------------------
@dataclass
class PixelKP:           # include/Sim3BA.h:9
    jid: int
    u: float
    v: float


@dataclass
class Sim3Params:        # include/Sim3BA.h:11-19: data = [s, aa(3), t(3)]
    data: np.ndarray = field(default_factory=lambda: np.array([1.0, 0, 0, 0, 0, 0, 0], dtype=np.float64))


@dataclass
class FramePoseParams:   # include/MultiFrameBA.h:9-14
    scale: float = 1.0
    rootAA: np.ndarray = field(default_factory=lambda: np.zeros(3))
    rootT: np.ndarray = field(default_factory=lambda: np.zeros(3))
    jointAA: np.ndarray = field(default_factory=lambda: np.zeros((NJ, 3)))

    def pack(self) -> np.ndarray:
        return np.concatenate([[self.scale], self.rootAA, self.rootT, np.asarray(self.jointAA)[1:].ravel()])

    def unpack(self, v):
        self.scale, self.rootAA, self.rootT = float(v[0]), np.array(v[1:4]), np.array(v[4:7])
        self.jointAA = np.vstack([np.zeros((1, 3)), np.asarray(v[7:76]).reshape(NJ - 1, 3)])


class Avatar:
    """ark::Avatar: public w, p, r in; cloud, jointPos, jointTrans out of update() (Avatar.h:191-219)."""

    def __init__(self, model: AvatarModel):
        self.model = model
        self.w = np.zeros(NS)
        self.p = np.zeros(3)
        self.r = np.tile(np.eye(3).reshape(9), (NJ, 1))     # col-major 3x3 per joint (identity either way)
        self.cloud = np.zeros((model.numPoints, 3))
        self.jointPos = np.zeros((NJ, 3))
        self.jointTrans = np.zeros((NJ, 12))

    def update(self):
        c, jp, jt = smpl_forward(self.model, self.w[None], self.p[None], self.r[None], dtype=np.float64)
        self.cloud, self.jointPos, self.jointTrans = c[0], jp[0], jt[0]


def _kp_array(kps):
    arr = np.zeros((1, MAXKP), dtype=KP_DTYPE)
    if len(kps) > MAXKP:
        raise ValueError(f"at most {MAXKP} keypoints per frame")
    for i, k in enumerate(kps):
        arr[0, i] = (k.jid, 0, k.u, k.v)
    return arr, np.array([len(kps)], dtype=np.int32)


def _optimize_single(mode, model, avatar, kps, fx, fy, cx, cy, valid_joint_ids, initSim3, max_iters, betaPose, betaShape, gmm):
    valid = set(int(j) for j in valid_joint_ids)
    if valid:                                            # Sim3BA.h:410-413: a non-empty list filters, an empty one does not
        kps = [k for k in kps if k.jid in valid]
    arr, nkp = _kp_array(kps)
    o = fit_opts(mode, max_iters, betaPose, betaShape, 0.0, (fx, fy, cx, cy))
    res = fit_single_batch(model, gmm, o, arr, nkp, avatar.r[0][None], initSim3.data[None], avatar.w[None])
    # write-back (Sim3BA.h:481-507 / 649-679)
    pose = np.concatenate([res["sim3"][0], res["jointAA"][0, 1:].ravel()])
    R0, r, p = writeback(pose[None], avatar.r[0][None])
    avatar.r = r[0].copy()
    avatar.p = p[0].copy()
    avatar.w = res["w"][0].copy()
    initSim3.data = res["sim3"][0].copy()
    report = f"iterations {int(res['niters'][0])}, termination {int(res['term'][0])}, final cost {res['trace'][0]['cost'][max(int(res['niters'][0]) - 1, 0)]:.6e}"
    return bool(res["usable"][0]), report


def OptimizePoseReprojection(model, avatar, kps, fx, fy, cx, cy, valid_joint_ids, initSim3, max_iters=100, betaPose=0.0,
                             betaShape=0.0, gmm=None):
    """include/Sim3BA.h:515-525.  Returns (IsSolutionUsable, report)."""
    return _optimize_single(FIT_POSE, model, avatar, kps, fx, fy, cx, cy, valid_joint_ids, initSim3, max_iters, betaPose, betaShape, gmm)


def OptimizePoseShapeReprojection(model, avatar, kps, fx, fy, cx, cy, valid_joint_ids, initSim3, max_iters=100, betaPose=0.0,
                                  betaShape=0.0, gmm=None):
    """include/Sim3BA.h:348-358.  Returns (IsSolutionUsable, report)."""
    return _optimize_single(FIT_POSE_SHAPE, model, avatar, kps, fx, fy, cx, cy, valid_joint_ids, initSim3, max_iters, betaPose, betaShape, gmm)


def mean_pixel_error(kps, avatar: Avatar, fx, fy, cx, cy) -> float:
    """include/Utils.h:102-115 (uses the avatar's posed joints: scale is not applied)."""
    arr, nkp = _kp_array(kps)
    return float(mean_pixel_error_batch(arr, nkp, avatar.jointPos[None], (fx, fy, cx, cy))[0])


class IcpProblem:
    """The ICP data term of ark::AvatarOptimizer::optimize for a batch of bodies with fixed correspondences (one ICP iteration,
    external/avatar/AvatarOptimizer.cpp:1386-1412): model point mpt[e] <-> point dpt[e] of body b's data cloud for e in
    [cptr[b], cptr[b+1]), the cloud of body b starting at point dptr[b] of ``data``.  ``eval`` returns what Ceres' line-search
    minimiser asks the problem for at a state (w, p, q): the cost and its gradient in the tangent space of the parameter blocks
    (p | 24 quaternions under FakeQuaternionParameterization | w)."""

    def __init__(self, model: AvatarModel, cptr, mpt, dpt, dptr, data, dtype=np.float32):
        dt = np.dtype(dtype)
        self.model = model
        self.cptr = np.ascontiguousarray(cptr, dtype=np.int64)
        self.dptr = np.ascontiguousarray(dptr, dtype=np.int64)
        self.B = self.cptr.shape[0] - 1
        mpt = np.ascontiguousarray(mpt, dtype=np.int32)
        dpt = None if dpt is None else np.ascontiguousarray(dpt, dtype=np.int32)   # None: pair e is data point e (icp_nearest output)
        data = np.ascontiguousarray(data, dtype=dt).reshape(-1, 3)
        if mpt.shape[0] != self.cptr[-1] or (dpt is not None and dpt.shape[0] != self.cptr[-1]) or data.shape[0] != self.dptr[-1]:
            raise ValueError("IcpProblem: offsets do not match the pair / data arrays")
        if mpt.size and mpt.max() >= model.numPoints:
            raise ValueError("IcpProblem: model point out of range")
        if dpt is None and not np.array_equal(self.cptr, self.dptr):
            raise ValueError("IcpProblem: dpt=None needs one pair per data point (cptr == dptr)")
        h = C.c_void_p()
        check(lib().b3d_icp_create(model.h, self.B, F64 if dt == np.float64 else F32, ptr(self.cptr), ptr(mpt), ptr(dpt),
                                   ptr(self.dptr), ptr(data), 0, C.byref(h)), "b3d_icp_create")
        self.h = h

    def __del__(self):
        h = getattr(self, "h", None)
        if h is not None and h.value and lib is not None:
            lib().b3d_icp_destroy(h)
            self.h = None

    def eval(self, w, p, q, shape=True):
        w, p, q = _f64(w).reshape(self.B, NS), _f64(p).reshape(self.B, 3), _f64(q).reshape(self.B, NJ, 4)
        cost, grad = np.empty(self.B), np.empty((self.B, 85))
        check(lib().b3d_icp_eval(self.h, ptr(w), ptr(p), ptr(q), 0 if shape else 2, ptr(cost), ptr(grad)), "b3d_icp_eval")
        return dict(cost=cost, grad=grad)

    def eval_full(self, w, p, q, gmm=None, beta_pose=0.0, beta_shape=0.0, shape=True):
        """The whole objective of one ICP iteration of AvatarOptimizer::optimize (:1386-1428): ICP term + pose prior (betaPose,
        GMM) + shape prior (betaShape), the prior weights scaled by sqrt(#residual blocks) / 15."""
        w, p, q = _f64(w).reshape(self.B, NS), _f64(p).reshape(self.B, 3), _f64(q).reshape(self.B, NJ, 4)
        cost, grad = np.empty(self.B), np.empty((self.B, 85))
        check(lib().b3d_icp_eval_full(self.h, _gmm_handle(gmm), C.c_double(beta_pose), C.c_double(beta_shape), ptr(w), ptr(p), ptr(q),
                                      0 if shape else 2, ptr(cost), ptr(grad)), "b3d_icp_eval_full")
        return dict(cost=cost, grad=grad)


    def solve(self, w, p, q, gmm=None, beta_pose=0.0, beta_shape=0.0, shape=True, max_iterations=50, function_tolerance=1e-4,
              gradient_tolerance=1e-10, parameter_tolerance=1e-8, trace_cap=128):
        """The ceres::Solve of one ICP iteration (external/avatar/AvatarOptimizer.cpp:1486; LINE_SEARCH / BFGS / WOLFE / CUBIC,
        options :1313-1339) for every body of the batch, on the GPU: returns the minimiser (w, p, q), the per-iteration records
        (cost, cost_change, gradient_max_norm, step_norm, step_size, line-search evaluations, line-search iterations) and a
        summary per body."""
        w, p, q = _f64(w).reshape(self.B, NS).copy(), _f64(p).reshape(self.B, 3).copy(), _f64(q).reshape(self.B, NJ, 4).copy()
        o = capi.IcpSolveOpts(int(max_iterations), 0 if shape else 2, function_tolerance, gradient_tolerance, parameter_tolerance,
                              beta_pose, beta_shape)
        trace, nrec, summ = np.zeros((self.B, trace_cap, 7)), np.zeros(self.B, dtype=np.int32), np.zeros((self.B, 6))
        evals = C.c_int32(0)
        check(lib().b3d_icp_solve(self.h, _gmm_handle(gmm), C.byref(o), ptr(w), ptr(p), ptr(q), ptr(trace), trace_cap, ptr(nrec),
                                  ptr(summ), C.byref(evals)), "b3d_icp_solve")
        return dict(w=w, p=p, q=q, trace=trace, nrec=nrec, initial_cost=summ[:, 0], final_cost=summ[:, 1],
                    iterations=summ[:, 2].astype(int), evaluations=summ[:, 3].astype(int), restarts=summ[:, 4].astype(int),
                    termination=summ[:, 5].astype(int), batched_evaluations=int(evals.value))


def icp_nearest(model: AvatarModel, cloud, model_part, dptr, data, data_part, visible=None, dtype=np.float64):
    """findNN with invert = true (external/avatar/AvatarOptimizer.cpp:841-907): for every data point the nearest visible model point
    with the same part label (-1: none).  cloud (B, V, 3), model_part (V,), visible (B, V) or None, body b's data points at
    dptr[b]:dptr[b+1].  The result pairs with ``IcpProblem(model, dptr, nn, None, dptr, data)``."""
    dt = np.dtype(dtype)
    cloud = np.ascontiguousarray(cloud, dtype=dt).reshape(-1, model.numPoints, 3)
    b = cloud.shape[0]
    model_part = np.ascontiguousarray(model_part, dtype=np.int32).reshape(model.numPoints)
    dptr = np.ascontiguousarray(dptr, dtype=np.int64).reshape(b + 1)
    data = np.ascontiguousarray(data, dtype=dt).reshape(-1, 3)
    data_part = np.ascontiguousarray(data_part, dtype=np.int32).reshape(data.shape[0])
    if data.shape[0] != dptr[-1]:
        raise ValueError("icp_nearest: dptr does not match the data array")
    vis = None if visible is None else np.ascontiguousarray(visible, dtype=np.uint8).reshape(b, model.numPoints)
    out = np.full(data.shape[0], -2, dtype=np.int32)
    nparts = int(max(model_part.max(), data_part.max() if data_part.size else 0)) + 1
    check(lib().b3d_icp_nearest(model.h, b, F64 if dt == np.float64 else F32, ptr(cloud), nparts, ptr(model_part), ptr(vis), ptr(dptr),
                                ptr(data), ptr(data_part), 0, ptr(out)), "b3d_icp_nearest")
    return out


def icp_visible(model: AvatarModel, cloud, dtype=np.float64):
    """pointVisible of AvatarOptimizer's ICP loop (external/avatar/AvatarOptimizer.cpp:1337-1352): (B, V) bytes, 1 where the vertex
    belongs to a face turned towards the camera in the posed cloud (B, V, 3).  The model needs its faces."""
    dt = np.dtype(dtype)
    cloud = np.ascontiguousarray(cloud, dtype=dt).reshape(-1, model.numPoints, 3)
    vis = np.empty(cloud.shape[:2], dtype=np.uint8)
    check(lib().b3d_icp_visible(model.h, cloud.shape[0], F64 if dt == np.float64 else F32, ptr(cloud), 0, ptr(vis)), "b3d_icp_visible")
    return vis


def avatar_optimize(model: AvatarModel, w, p, q, dptr, data, data_part, model_part, *, gmm=None, beta_pose=0.0, beta_shape=0.0,
                    icp_iters=3, max_iters_per_icp=20, enable_occlusion=True, shape=True, dtype=np.float64, function_tolerance=1e-4):
    """ark::AvatarOptimizer::optimize (external/avatar/AvatarOptimizer.cpp:1246-1520) for a batch of bodies, every stage on the GPU:
    per ICP iteration the posed cloud (Avatar::update), the occlusion mask (:1347-1365), the correspondences (findNN, invert = true,
    :1391-1394), the residual blocks of the matched pairs (:1396-1474) and the BFGS line-search solve (:1486); the quaternions the
    solve returns are the next iteration's rotations (:1491-1494).  Body b's data points are data[dptr[b]:dptr[b+1]] with part labels
    data_part; model_part labels the model points (modelPartLabels, :1306-1310).  Returns the final state and one summary per
    ICP iteration."""
    from . import synth
    w, p, q = _f64(w).copy(), _f64(p).copy(), _f64(q).copy()
    b = w.shape[0]
    dptr = np.ascontiguousarray(dptr, dtype=np.int64)
    history = []
    for _ in range(int(icp_iters)):
        r = synth.quat_to_r(q)
        cloud, _, _ = smpl_forward(model, w, p, r, dtype=np.float64)
        vis = icp_visible(model, cloud) if enable_occlusion else None
        nn = icp_nearest(model, cloud, model_part, dptr, data, data_part, visible=vis, dtype=np.float64)
        prob = IcpProblem(model, dptr, nn, None, dptr, data, dtype=dtype)
        res = prob.solve(w, p, q, gmm=gmm, beta_pose=beta_pose, beta_shape=beta_shape, shape=shape, max_iterations=max_iters_per_icp,
                         function_tolerance=function_tolerance)
        w, p, q = res["w"], res["p"], res["q"]
        history.append(dict(matched=np.array([int((nn[dptr[i]:dptr[i + 1]] >= 0).sum()) for i in range(b)]), initial_cost=res["initial_cost"],
                            final_cost=res["final_cost"], iterations=res["iterations"], evaluations=res["evaluations"],
                            termination=res["termination"], nn=nn))
        del prob
    return dict(w=w, p=p, q=q, history=history)
