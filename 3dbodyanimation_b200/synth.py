"""Deterministic synthetic SMPL-shaped model and keypoint workloads.

The licensed SMPL ``model.npz`` is not available offline (reference ``.MISSING_LARGE_BLOBS``), so every
config of BASELINE.json runs on a random-init model with SMPL's shapes (6890 v / 13776 f / 24 j / 10 betas)
whose array names, shapes and dtypes are the ones ``ark::AvatarModel`` reads
(reference external/avatar/AvatarModel.cpp:26-102).  Keypoints are produced by the reference's own forward
model (include/Sim3BA.h:34-88 and :126-227 restated in numpy) from random ground-truth poses.

numpy only; shared by tests/, bench.py and __graft_entry__.smoke() so that the oracle and the CUDA path
see bit-identical inputs.
"""
from __future__ import annotations

import numpy as np

SEED = 20240607
V, J, S, NF, P = 6890, 24, 10, 13776, 207
NPF = 76          # per-frame parameters: s, rootAA(3), rootT(3), jointAA[1..23](69)
MAXKP = 17

PARENT = np.array([-1, 0, 0, 0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 9, 9, 12, 13, 14, 16, 17, 18, 19, 20, 21], dtype=np.int32)

REST_JOINTS = np.array([
    (0.00, 0.00, 0.00), (0.07, -0.09, 0.00), (-0.07, -0.09, 0.00), (0.00, 0.11, -0.03),
    (0.10, -0.47, 0.00), (-0.10, -0.47, 0.00), (0.00, 0.25, -0.02), (0.09, -0.87, -0.03),
    (-0.09, -0.87, -0.03), (0.00, 0.30, 0.00), (0.11, -0.93, 0.09), (-0.11, -0.93, 0.09),
    (0.00, 0.51, -0.03), (0.08, 0.42, -0.02), (-0.08, 0.42, -0.02), (0.00, 0.58, 0.02),
    (0.17, 0.45, -0.02), (-0.17, 0.45, -0.02), (0.43, 0.44, -0.04), (-0.43, 0.44, -0.04),
    (0.68, 0.44, -0.04), (-0.68, 0.44, -0.04), (0.77, 0.43, -0.05), (-0.77, 0.43, -0.05)], dtype=np.float64)

# MediaPipe-33 -> SMPL-24 tables (reference include/Utils.h:18-23).  USE_SMPL is std::array<int,17> with 15
# initialisers: the two trailing zeros make the pelvis (midpoint of MP 23/24) appear twice.
MP_MAP = np.array([-1, 23, 24, -1, 25, 26, -1, 27, 28, -1, 31, 32, -1, -1, -1, 0, 11, 12, 13, 14, 15, 16, -1, -1], dtype=np.int32)
USE_SMPL = np.array([1, 2, 4, 5, 7, 8, 10, 11, 15, 16, 17, 18, 19, 20, 21, 0, 0], dtype=np.int32)

KP_DTYPE = np.dtype([("jid", "<i4"), ("pad", "<i4"), ("u", "<f8"), ("v", "<f8")])   # PixelKP, 24 bytes

# camera of the 480x270 fixture images: f = max(H, W); fx = fy = 0.9 f (src/main_single_frame.cpp:171-176)
CAM_W, CAM_H = 480, 270
FX = FY = 0.9 * max(CAM_W, CAM_H)
CX, CY = 0.5 * CAM_W, 0.5 * CAM_H


def initial_root_rotation() -> np.ndarray:
    """r[0] = AngleAxis(pi, Y).toRotationMatrix() * diag(1,-1,1) (src/main_single_frame.cpp:208-212), row-major 3x3."""
    c, s = np.cos(np.pi), np.sin(np.pi)
    ry = np.array([[c, 0.0, s], [0.0, 1.0, 0.0], [-s, 0.0, c]])
    return ry @ np.diag([1.0, -1.0, 1.0])


def make_model(seed: int = SEED, with_posedirs: bool = False, rig: str = "cyclic") -> dict:
    """Random-init model with SMPL's key names / shapes / dtypes (float32 arrays, uint32 kintree and faces).
    rig = "cyclic" (default, what every test, fixture and headline number uses): vertex v hangs on joint v mod 24, so any run of
    consecutive vertices mixes all joints; rig = "blocks": every joint owns a contiguous range of vertex numbers, the way a mesh
    that was modelled body part by body part is numbered (bench.py reports the skinning kernel on it as a second data point)."""
    rng = np.random.default_rng(seed)
    if rig not in ("cyclic", "blocks"):
        raise ValueError("rig must be 'cyclic' or 'blocks'")
    cluster = np.arange(V) % J if rig == "cyclic" else (np.arange(V) * J) // V
    v_template = (REST_JOINTS[cluster] + rng.normal(0.0, 0.06, size=(V, 3))).astype(np.float32)
    j_reg = np.zeros((J, V), dtype=np.float32)
    for j in range(J):
        idx = np.nonzero(cluster == j)[0]
        j_reg[j, idx] = np.float32(1.0 / len(idx))
    first_child = np.full(J, -1, dtype=np.int64)
    for j in range(1, J):
        if first_child[PARENT[j]] < 0:
            first_child[PARENT[j]] = j
    weights = np.zeros((V, J), dtype=np.float64)
    dir_w = rng.dirichlet([4.0, 2.0, 1.0, 1.0], size=V)
    for v in range(V):
        j = int(cluster[v])
        pj = int(PARENT[j]) if PARENT[j] >= 0 else j
        cj = int(first_child[j]) if first_child[j] >= 0 else j
        gj = int(PARENT[pj]) if PARENT[pj] >= 0 else j
        for k, jj in enumerate((j, pj, cj, gj)):
            weights[v, jj] += dir_w[v, k]
    weights = weights.astype(np.float32)
    weights /= weights.sum(axis=1, keepdims=True, dtype=np.float32)
    shapedirs = rng.normal(0.0, 0.01, size=(V, 3, S)).astype(np.float32)
    faces = np.empty((NF, 3), dtype=np.uint32)
    a = rng.integers(0, V, size=NF)
    b = (a + rng.integers(1, V, size=NF)) % V
    c = rng.integers(0, V, size=NF)
    clash = (c == a) | (c == b)
    while clash.any():
        c[clash] = rng.integers(0, V, size=int(clash.sum()))
        clash = (c == a) | (c == b)
    faces[:, 0], faces[:, 1], faces[:, 2] = a, b, c
    kintree = np.zeros((2, J), dtype=np.uint32)
    kintree[0] = PARENT.astype(np.int64).astype(np.uint32)       # root -> 0xFFFFFFFF (scripts/npz_fixer.py convention)
    kintree[1] = np.arange(J, dtype=np.uint32)
    model = dict(v_template=v_template, shapedirs=shapedirs, J_regressor=j_reg, weights=weights,
                 kintree_table=kintree, f=faces)
    if with_posedirs:
        model["posedirs"] = rng.normal(0.0, 0.002, size=(V, 3, P)).astype(np.float32)
    return model


def make_mesh_faces(model: dict) -> np.ndarray:
    """A mesh-like face list for the overlay renderer (make_model's ``f`` joins random vertices: fine for the loader, useless
    as a surface).  The template is 24 blobs of ~287 vertices around the rest joints; each blob is triangulated as a closed
    star-shaped surface (convex hull of the vertex directions seen from the blob centre = spherical Delaunay), normals
    outward: 2 n - 4 small triangles per blob, ~13.7 k faces like SMPL's 13,776.  uint32 (F, 3)."""
    from scipy.spatial import ConvexHull
    vt = np.asarray(model["v_template"], dtype=np.float64)
    cluster = np.arange(V) % J
    out = []
    for j in range(J):
        idx = np.nonzero(cluster == j)[0]
        d = vt[idx] - vt[idx].mean(axis=0)
        u = d / np.linalg.norm(d, axis=1, keepdims=True)
        tri = ConvexHull(u).simplices
        a, b, c = u[tri[:, 0]], u[tri[:, 1]], u[tri[:, 2]]
        flip = np.einsum("ij,ij->i", np.cross(b - a, c - a), a + b + c) < 0
        tri[flip] = tri[flip][:, ::-1]
        out.append(idx[tri])
    return np.ascontiguousarray(np.concatenate(out), dtype=np.uint32)


def make_render_model(radius: float = 0.09, seed: int = SEED) -> dict:
    """make_model() with a smooth skin for the overlay renderer: every blob's vertices are moved onto a sphere of ``radius``
    metres around the blob centre (directions kept), and ``f`` is make_mesh_faces of that template.  Faces then project to a
    couple of pixels each at 480 x 270 like SMPL's; the raw random template gives spiky blobs with ~10 x the overdraw."""
    model = make_model(seed)
    vt = model["v_template"].astype(np.float64)
    cluster = np.arange(V) % J
    for j in range(J):
        idx = np.nonzero(cluster == j)[0]
        c = vt[idx].mean(axis=0)
        d = vt[idx] - c
        vt[idx] = c + radius * d / np.linalg.norm(d, axis=1, keepdims=True)
    model = dict(model, v_template=vt.astype(np.float32))
    model["f"] = make_mesh_faces(model)
    return model


def load_pose_prior(path: str) -> dict:
    """pose_prior.txt: 'nComps nDims', weights, means (row-major), covariances (scripts/convert_gmm_to_avatar.py:17-29)."""
    with open(path) as fh:
        tok = fh.read().split()
    nc, nd = int(tok[0]), int(tok[1])
    vals = np.array(tok[2:], dtype=np.float64)
    w = vals[:nc]
    means = vals[nc:nc + nc * nd].reshape(nc, nd)
    covs = vals[nc + nc * nd:nc + nc * nd + nc * nd * nd].reshape(nc, nd, nd)
    return dict(nComps=nc, nDims=nd, weights=w.copy(), means=means.copy(), covs=covs.copy())


# ---------------------------------------------------------------------------------------------------------
# forward model used only to synthesise keypoints
# ---------------------------------------------------------------------------------------------------------
def _rodrigues_rotate(aa: np.ndarray, p: np.ndarray) -> np.ndarray:
    th2 = float(aa @ aa)
    if th2 > np.finfo(np.float64).eps:
        th = np.sqrt(th2)
        w = aa / th
        return p * np.cos(th) + np.cross(w, p) * np.sin(th) + w * (w @ p) * (1.0 - np.cos(th))
    return p + np.cross(aa, p)


def model_constants(model: dict) -> dict:
    """initialJointPos (24x3), jointShapeReg (24x3x10) and rest offsets in float64 (AvatarModel.cpp:111-127)."""
    vt = model["v_template"].astype(np.float64)
    sd = model["shapedirs"].astype(np.float64)
    jr = model["J_regressor"].astype(np.float64)
    j0 = jr @ vt
    jsr = np.einsum("jv,vcs->jcs", jr, sd)
    rel = j0 - j0[0]
    off = np.zeros((J, 3))
    off[1:] = rel[1:] - rel[PARENT[1:]]
    return dict(J0=j0, JSR=jsr, offset=off)


def project_joint(consts: dict, jid: int, s: float, raa, t, jaa, w, R0, with_shape: bool) -> np.ndarray:
    """ReprojCost / ReprojCostShape forward (Sim3BA.h:34-88, 126-227): pixel (u, v) of SMPL joint ``jid``."""
    off, jsr = consts["offset"], consts["JSR"]

    def bone(j):
        b = off[j].copy()
        if with_shape:
            b += jsr[j] @ w - (jsr[PARENT[j]] @ w if PARENT[j] >= 0 else 0.0)
        return b

    pos = bone(jid)
    cur = jid
    while PARENT[cur] != -1 and PARENT[cur] != 0:
        pa = int(PARENT[cur])
        pos = _rodrigues_rotate(jaa[pa], pos) + bone(pa)
        cur = pa
    pc = _rodrigues_rotate(np.asarray(raa, dtype=np.float64), R0 @ pos)
    pc = s * pc + t
    return np.array([FX * pc[0] / pc[2] + CX, FY * pc[1] / pc[2] + CY])


def _gmm_sample(prior: dict, chol: np.ndarray, rng) -> np.ndarray:
    c = int(rng.choice(prior["nComps"], p=prior["weights"] / prior["weights"].sum()))
    return prior["means"][c] + 0.5 * (chol[c] @ rng.normal(size=prior["nDims"]))


def make_frames(model: dict, n: int, seed: int, *, shape: bool, prior: dict | None = None, drop_prob: float = 0.0,
                noise_px: float = 1.0) -> dict:
    """Independent frames for C1 (shape=False, prior=None) / C2 (shape=True, prior=GMM).

    Returns kps (n x 17 KP_DTYPE), nkp (n,), R0 (n x 9 col-major), sim3 init (n x 7), w init (n x 10) and the
    ground truth.  Initialisation is exactly the mains': s=1, aa=0, t=(0,0,3), jointAA=0, w=0.
    """
    rng = np.random.default_rng(seed)
    consts = model_constants(model)
    R0 = initial_root_rotation()
    chol = np.linalg.cholesky(prior["covs"]) if prior is not None else None
    kps = np.zeros((n, MAXKP), dtype=KP_DTYPE)
    nkp = np.zeros(n, dtype=np.int32)
    gt = np.zeros((n, NPF + S))
    for f in range(n):
        jaa = np.zeros((J, 3))
        if prior is not None:
            jaa[1:] = _gmm_sample(prior, chol, rng).reshape(J - 1, 3)
        else:
            jaa[1:] = rng.normal(0.0, 0.25, size=(J - 1, 3))
        raa = rng.normal(0.0, 0.15, size=3)
        t = np.array([rng.uniform(-0.4, 0.4), rng.uniform(-0.3, 0.3), rng.uniform(2.5, 3.5)])
        s = rng.uniform(0.85, 1.15)
        w = rng.normal(size=S) if shape else np.zeros(S)
        gt[f, 0], gt[f, 1:4], gt[f, 4:7], gt[f, 7:76], gt[f, 76:] = s, raa, t, jaa[1:].ravel(), w
        k = 0
        for sid in USE_SMPL:
            uv = project_joint(consts, int(sid), s, raa, t, jaa, w, R0, shape) + rng.normal(0.0, noise_px, size=2)
            if drop_prob > 0.0 and rng.uniform() < drop_prob:
                continue
            kps[f, k] = (int(sid), 0, uv[0], uv[1])
            k += 1
        nkp[f] = k
    sim3 = np.tile(np.array([1.0, 0, 0, 0, 0, 0, 3.0]), (n, 1))
    return dict(kps=kps, nkp=nkp, R0=np.tile(R0.T.reshape(9), (n, 1)).copy(), sim3=sim3, w=np.zeros((n, S)), gt=gt)


def make_sequence(model: dict, n: int, seed: int, *, drop_prob: float = 0.0, noise_px: float = 1.0) -> dict:
    """Smooth trajectory for C3/C5: every parameter is a per-sequence mean plus three sinusoids (periods 40-400 frames)."""
    rng = np.random.default_rng(seed)
    consts = model_constants(model)
    R0 = initial_root_rotation()
    tt = np.arange(n)[:, None]

    def traj(mean, amp):
        mean = np.atleast_1d(np.asarray(mean, dtype=np.float64))
        out = np.tile(mean, (n, 1))
        for _ in range(3):
            period = rng.uniform(40.0, 400.0, size=mean.shape)
            phase = rng.uniform(0.0, 2 * np.pi, size=mean.shape)
            out = out + (amp / 3.0) * np.sin(2 * np.pi * tt / period + phase)
        return out

    jaa_all = traj(rng.normal(0.0, 0.15, size=3 * (J - 1)), 0.25)
    raa_all = traj(rng.normal(0.0, 0.10, size=3), 0.15)
    t_all = traj([rng.uniform(-0.2, 0.2), rng.uniform(-0.15, 0.15), rng.uniform(2.8, 3.2)], 0.2)
    s_all = traj([rng.uniform(0.95, 1.05)], 0.05)[:, 0]
    w = rng.normal(size=S)
    kps = np.zeros((n, MAXKP), dtype=KP_DTYPE)
    nkp = np.zeros(n, dtype=np.int32)
    gt = np.zeros((n, NPF + S))
    for f in range(n):
        jaa = np.zeros((J, 3))
        jaa[1:] = jaa_all[f].reshape(J - 1, 3)
        gt[f, 0], gt[f, 1:4], gt[f, 4:7], gt[f, 7:76], gt[f, 76:] = s_all[f], raa_all[f], t_all[f], jaa_all[f], w
        k = 0
        for sid in USE_SMPL:
            uv = project_joint(consts, int(sid), s_all[f], raa_all[f], t_all[f], jaa, w, R0, True) + rng.normal(0.0, noise_px, size=2)
            if drop_prob > 0.0 and rng.uniform() < drop_prob:
                continue
            kps[f, k] = (int(sid), 0, uv[0], uv[1])
            k += 1
        nkp[f] = k
    poses = np.zeros((n, NPF))
    poses[:, 0] = 1.0
    poses[:, 6] = 3.0                       # FramePoseParams init (src/main_multi_frame.cpp:96-101)
    return dict(kps=kps, nkp=nkp, R0=np.tile(R0.T.reshape(9), (n, 1)).copy(), poses=poses, W=np.zeros((n, S)), gt=gt)


def _rodrigues_batch(aa: np.ndarray, p: np.ndarray) -> np.ndarray:
    """ceres::AngleAxisRotatePoint for arrays of axis-angles (n, 3) and points (n, 3) (the Rodrigues branch; the synthetic
    trajectories never come near theta = 0)."""
    th = np.linalg.norm(aa, axis=1, keepdims=True)
    w = aa / np.maximum(th, 1e-300)
    c, sn = np.cos(th), np.sin(th)
    return p * c + np.cross(w, p) * sn + w * np.sum(w * p, axis=1, keepdims=True) * (1.0 - c)


def make_sequence_fast(model: dict, n: int, seed: int, *, noise_px: float = 1.0) -> dict:
    """make_sequence for long sequences (C5: 100,000 frames), vectorised over the frames: the same kind of trajectory (per-sequence
    mean plus three sinusoids per parameter) and the same forward model, all 17 keypoint slots per frame.  Not bit-identical to
    make_sequence (the random draws come in a different order)."""
    rng = np.random.default_rng(seed)
    consts = model_constants(model)
    off, jsr = consts["offset"], consts["JSR"]
    R0 = initial_root_rotation()
    tt = np.arange(n)[:, None]

    def traj(mean, amp):
        mean = np.atleast_1d(np.asarray(mean, dtype=np.float64))
        out = np.tile(mean, (n, 1))
        for _ in range(3):
            period = rng.uniform(40.0, 400.0, size=mean.shape)
            phase = rng.uniform(0.0, 2 * np.pi, size=mean.shape)
            out = out + (amp / 3.0) * np.sin(2 * np.pi * tt / period + phase)
        return out

    jaa_all = traj(rng.normal(0.0, 0.15, size=3 * (J - 1)), 0.25)
    raa_all = traj(rng.normal(0.0, 0.10, size=3), 0.15)
    t_all = traj([rng.uniform(-0.2, 0.2), rng.uniform(-0.15, 0.15), rng.uniform(2.8, 3.2)], 0.2)
    s_all = traj([rng.uniform(0.95, 1.05)], 0.05)[:, 0]
    w = rng.normal(size=S)
    bone = off + jsr @ w
    bone[1:] -= (jsr @ w)[PARENT[1:]]                 # bone(j) = offset_j + (JSR_j - JSR_parent(j)) w ; the root has no parent term
    jaa = np.concatenate([np.zeros((n, 1, 3)), jaa_all.reshape(n, J - 1, 3)], axis=1)
    kps = np.zeros((n, MAXKP), dtype=KP_DTYPE)
    for k, sid in enumerate(USE_SMPL):
        pos = np.tile(bone[sid], (n, 1))
        cur = int(sid)
        while PARENT[cur] != -1 and PARENT[cur] != 0:
            pa = int(PARENT[cur])
            pos = _rodrigues_batch(jaa[:, pa], pos) + bone[pa]
            cur = pa
        pc = _rodrigues_batch(raa_all, pos @ R0.T)
        pc = s_all[:, None] * pc + t_all
        uv = np.stack([FX * pc[:, 0] / pc[:, 2] + CX, FY * pc[:, 1] / pc[:, 2] + CY], axis=1) + rng.normal(0.0, noise_px, size=(n, 2))
        kps["jid"][:, k] = int(sid)
        kps["u"][:, k], kps["v"][:, k] = uv[:, 0], uv[:, 1]
    gt = np.zeros((n, NPF + S))
    gt[:, 0], gt[:, 1:4], gt[:, 4:7], gt[:, 7:76], gt[:, 76:] = s_all, raa_all, t_all, jaa_all, w
    poses = np.zeros((n, NPF))
    poses[:, 0] = 1.0
    poses[:, 6] = 3.0
    return dict(kps=kps, nkp=np.full(n, MAXKP, dtype=np.int32), R0=np.tile(R0.T.reshape(9), (n, 1)).copy(), poses=poses, W=np.zeros((n, S)), gt=gt)


def make_forward_inputs(b: int, seed: int) -> dict:
    """C4 inputs: w ~ N(0,1), r from AA ~ N(0, 0.3^2) (col-major 3x3 per joint), p as the frame generator."""
    rng = np.random.default_rng(seed)
    w = rng.normal(size=(b, S))
    aa = rng.normal(0.0, 0.3, size=(b, J, 3))
    th = np.linalg.norm(aa, axis=-1, keepdims=True)
    ax = aa / np.maximum(th, 1e-300)
    c, s = np.cos(th)[..., None], np.sin(th)[..., None]
    K = np.zeros((b, J, 3, 3))
    K[..., 0, 1], K[..., 0, 2] = -ax[..., 2], ax[..., 1]
    K[..., 1, 0], K[..., 1, 2] = ax[..., 2], -ax[..., 0]
    K[..., 2, 0], K[..., 2, 1] = -ax[..., 1], ax[..., 0]
    R = c * np.eye(3) + (1 - c) * ax[..., :, None] * ax[..., None, :] + s * K       # row-major R[a, b]
    r = np.ascontiguousarray(np.swapaxes(R, -1, -2)).reshape(b, J, 9)              # col-major storage
    p = np.stack([rng.uniform(-0.4, 0.4, b), rng.uniform(-0.3, 0.3, b), rng.uniform(2.5, 3.5, b)], axis=1)
    return dict(w=w, r=r, p=p)


def forward_numpy(model: dict, w: np.ndarray, p: np.ndarray, r: np.ndarray, use_posedirs: bool = False):
    """Independent vectorised numpy restatement of Avatar::update (external/avatar/Avatar.cpp:22-75) for one body.

    r: (24, 9) col-major.  Returns cloud (V,3), jointPos (24,3), jointTrans (24,12 col-major 3x4).
    Used by the tests to cross-check the C++ oracle.
    """
    vt = model["v_template"].astype(np.float64)
    sd = model["shapedirs"].astype(np.float64)
    wt = model["weights"].astype(np.float64)
    c = model_constants(model)
    R = r.reshape(J, 3, 3).transpose(0, 2, 1)
    shaped = vt + sd @ w
    if use_posedirs:
        pf = (R[1:] - np.eye(3)).reshape(-1)
        shaped = shaped + model["posedirs"].astype(np.float64) @ pf
    jp = c["J0"] + c["JSR"] @ w
    G = np.zeros((J, 3, 4))
    G[0, :, :3], G[0, :, 3] = R[0], p
    for i in range(1, J):
        pa = PARENT[i]
        G[i, :, :3] = G[pa, :, :3] @ R[i]
        G[i, :, 3] = G[pa, :, 3] + G[pa, :, :3] @ (jp[i] - jp[pa])
    posed = G[:, :, 3].copy()
    G[:, :, 3] -= np.einsum("jab,jb->ja", G[:, :, :3], jp)
    T = np.einsum("vj,jab->vab", wt, G)
    cloud = np.einsum("vab,vb->va", T[:, :, :3], shaped) + T[:, :, 3]
    return cloud, posed, G.transpose(0, 2, 1).reshape(J, 12)


def make_icp_inputs(model: dict, b: int, seed: int, *, ndata: int | None = None, noise: float = 0.02, miss: float = 0.2) -> dict:
    """Inputs of AvatarOptimizer's ICP term (external/avatar/AvatarOptimizer.cpp:1246-1460) for b bodies: state (w, p, unit
    quaternions q in Eigen's (x, y, z, w) order), a data cloud per body (a perturbed, differently posed copy of the model cloud,
    shuffled) and correspondences in the reference's shape: per model point a list of data points (here 0, 1 or 2 of them; a
    fraction `miss` of the model points has none), flattened to pairs (mpt, dpt) with per-body offsets cptr / dptr."""
    rng = np.random.default_rng(seed)
    V = model["v_template"].shape[0]
    nd = V if ndata is None else ndata
    fin = make_forward_inputs(b, seed + 1)
    aa = rng.normal(0.0, 0.3, size=(b, J, 3))
    th = np.linalg.norm(aa, axis=-1, keepdims=True)
    q = np.concatenate([np.sin(th / 2) * aa / np.maximum(th, 1e-300), np.cos(th / 2)], axis=-1)
    tgt = make_forward_inputs(b, seed + 2)
    cloud = [forward_numpy(model, fin["w"][i] + 0.3 * tgt["w"][i], fin["p"][i], tgt["r"][i])[0] for i in range(b)]
    data, mpt, dpt, cptr, dptr = [], [], [], [0], [0]
    for i in range(b):
        src = rng.integers(0, V, nd)                       # data point k is a noisy sample of model point src[k]
        data.append(cloud[i][src] + rng.normal(0.0, noise, (nd, 3)))
        keep = rng.random(nd) > miss
        m = src[keep]
        d = np.nonzero(keep)[0]
        order = np.argsort(m, kind="stable")               # the reference walks model points in order (:1396-1405)
        mpt.append(m[order]); dpt.append(d[order])
        cptr.append(cptr[-1] + len(m)); dptr.append(dptr[-1] + nd)
    return dict(w=fin["w"], p=fin["p"], q=q, data=np.concatenate(data), mpt=np.concatenate(mpt).astype(np.int32),
                dpt=np.concatenate(dpt).astype(np.int32), cptr=np.array(cptr, dtype=np.int64), dptr=np.array(dptr, dtype=np.int64))


def quat_to_r(q: np.ndarray) -> np.ndarray:
    """Unit quaternions (..., 4) in (x, y, z, w) order -> col-major 3x3 (..., 9) (Eigen toRotationMatrix)."""
    x, y, z, w = q[..., 0], q[..., 1], q[..., 2], q[..., 3]
    R = np.stack([1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w),
                  2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w),
                  2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)], axis=-1).reshape(q.shape[:-1] + (3, 3))
    return np.ascontiguousarray(np.swapaxes(R, -1, -2)).reshape(q.shape[:-1] + (9,))
