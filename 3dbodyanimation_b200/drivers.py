"""Drop-in replacements for the reference's two executables on top of the C-ABI.

  run_single / main_single   3dba_single   src/main_single_frame.cpp:105-283
  run_multi  / main_multi    3dba_multi    src/main_multi_frame.cpp:15-236
  load_mp_json               MediaPipe-33 JSON -> PixelKP list, include/Utils.h:61-99 (``variant="utils"``, what both
                             executables call) and the divergent private copy src/main_single_frame.cpp:69-102
                             (``variant="main_single"``)
  list_sorted                include/Utils.h:33-41
  image_size                 stands in for cv::imread(images.front()).cols/rows (only the header is read)

Same positional arguments, defaults, console lines and ``log.csv`` format (``frame,mean_pixel_error_px,time_ms``) as the
reference.  What differs, on purpose:
  * 3dba_single fits ALL frames of the folder in ONE batched call (the frames are independent: each one starts from a fresh
    avatar, src/main_single_frame.cpp:205-221), so ``time_ms`` is the batch time divided by the number of fitted frames —
    the same convention the reference's multi-frame driver uses for its anchors and windows (main_multi_frame.cpp:146,190).
  * the overlays (``frame_<i>_render.png`` / ``frame_<i>_multi.png``) come from the CUDA rasteriser (b3d_render_overlay): same
    triangles, colours and order as renderSMPLMesh, OpenCV's LINE_8 coverage instead of LINE_AA; frames are
    decoded here (PNG) or by Pillow (JPEG / BMP) instead of cv::imread, RGB instead of BGR (the overlay is grey either way).
    ``--no-render`` skips them; ``--save-params`` also writes the fitted parameters (and for ``single`` the float32 clouds) to
    ``out_dir/params.npz``.  Models without a face list (``f``) cannot be rendered: use ``--no-render``.
Every computation goes through lib3dba_b200.so (api.py); there is no CPU path here.
"""
from __future__ import annotations

import json
import os
import struct
import sys
import time

import numpy as np

from . import api
from ._capi import FIT_POSE, FIT_POSE_SHAPE, KP_DTYPE, MAXKP, NJ, NPF, NS

MP_MAP = [int(x) for x in api.MP_MAP]
USE_SMPL = [int(x) for x in api.USE_SMPL]          # 17 entries: the two trailing zeros emit the pelvis twice
IMAGE_EXTS = (".png", ".jpg", ".jpeg", ".bmp")


# ---------------------------------------------------------------------------------------------------------
# files
# ---------------------------------------------------------------------------------------------------------
def list_sorted(folder: str, exts) -> list[str]:
    """Regular files of ``folder`` whose extension (lower-cased) is in ``exts``, sorted by path (Utils.h:24-41)."""
    exts = tuple(e.lower() for e in exts)
    out = [os.path.join(folder, n) for n in os.listdir(folder)
           if os.path.isfile(os.path.join(folder, n)) and os.path.splitext(n)[1].lower() in exts]
    return sorted(out)


def image_size(path: str) -> tuple[int, int]:
    """(width, height) from the header of a PNG / JPEG / BMP file."""
    with open(path, "rb") as f:
        head = f.read(32)
        if head[:8] == b"\x89PNG\r\n\x1a\n" and head[12:16] == b"IHDR":
            w, h = struct.unpack(">II", head[16:24])
            return int(w), int(h)
        if head[:2] == b"BM":
            w, h = struct.unpack("<ii", head[18:26])
            return int(w), abs(int(h))
        if head[:2] == b"\xff\xd8":
            f.seek(2)
            while True:
                b = f.read(1)
                if not b:
                    break
                if b != b"\xff":
                    continue
                m = f.read(1)
                while m == b"\xff":
                    m = f.read(1)
                if not m:
                    break
                code = m[0]
                if code in (0x01,) or 0xD0 <= code <= 0xD9:
                    continue
                seg = f.read(2)
                if len(seg) < 2:
                    break
                n = struct.unpack(">H", seg)[0]
                if 0xC0 <= code <= 0xCF and code not in (0xC4, 0xC8, 0xCC):
                    data = f.read(5)
                    h, w = struct.unpack(">HH", data[1:5])
                    return int(w), int(h)
                f.seek(n - 2, 1)
    raise ValueError(f"Failed to read {path}")


def write_png(path: str, img: np.ndarray) -> None:
    """(H, W, 3) uint8 RGB -> 8-bit truecolour PNG (zlib only; stands in for cv::imwrite)."""
    import zlib
    img = np.ascontiguousarray(img, dtype=np.uint8)
    h, w, _ = img.shape

    def chunk(tag, data):
        return struct.pack(">I", len(data)) + tag + data + struct.pack(">I", zlib.crc32(tag + data))

    raw = np.concatenate([np.zeros((h, 1), dtype=np.uint8), img.reshape(h, w * 3)], axis=1).tobytes()
    with open(path, "wb") as f:
        f.write(b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, 2, 0, 0, 0)) +
                chunk(b"IDAT", zlib.compress(raw, 6)) + chunk(b"IEND", b""))


def _read_png(path: str):
    """Non-interlaced 8-bit grey / RGB / RGBA PNG -> (H, W, 3) uint8, or None when the file is another flavour."""
    import zlib
    data = open(path, "rb").read()
    if data[:8] != b"\x89PNG\r\n\x1a\n":
        return None
    pos, idat, hdr = 8, [], None
    while pos + 8 <= len(data):
        n, tag = struct.unpack(">I", data[pos:pos + 4])[0], data[pos + 4:pos + 8]
        body = data[pos + 8:pos + 8 + n]
        if tag == b"IHDR":
            hdr = struct.unpack(">IIBBBBB", body)
        elif tag == b"IDAT":
            idat.append(body)
        elif tag == b"IEND":
            break
        pos += 12 + n
    if hdr is None or hdr[2] != 8 or hdr[3] not in (0, 2, 6) or hdr[6] != 0:
        return None
    w, h, ch = hdr[0], hdr[1], {0: 1, 2: 3, 6: 4}[hdr[3]]
    raw = np.frombuffer(zlib.decompress(b"".join(idat)), dtype=np.uint8).reshape(h, 1 + w * ch)
    out = np.zeros((h, w * ch), dtype=np.uint8)
    prev = np.zeros(w * ch, dtype=np.int32)
    for y in range(h):
        ft, line = int(raw[y, 0]), raw[y, 1:].astype(np.int32)
        if ft == 0:
            cur = line
        elif ft == 2:
            cur = (line + prev) & 255
        else:                                  # Sub / Average / Paeth depend on the pixel to the left: byte by byte
            cur = np.zeros(w * ch, dtype=np.int32)
            for i in range(w * ch):
                a = cur[i - ch] if i >= ch else 0
                b, c = prev[i], (prev[i - ch] if i >= ch else 0)
                if ft == 1:
                    pr = a
                elif ft == 3:
                    pr = (a + b) >> 1
                else:
                    pa, pb, pc = abs(b - c), abs(a - c), abs(a + b - 2 * c)
                    pr = a if (pa <= pb and pa <= pc) else (b if pb <= pc else c)
                cur[i] = (line[i] + pr) & 255
        out[y] = cur
        prev = cur
    out = out.reshape(h, w, ch)
    return np.repeat(out, 3, axis=2) if ch == 1 else np.ascontiguousarray(out[:, :, :3])


def read_image(path: str) -> np.ndarray:
    """cv::imread stand-in: (H, W, 3) uint8 RGB.  PNG is decoded here; JPEG / BMP (and exotic PNGs) through Pillow when it is
    installed.  Raises ValueError otherwise (the caller then paints over a black frame)."""
    if os.path.splitext(path)[1].lower() == ".png":
        img = _read_png(path)
        if img is not None:
            return img
    try:
        from PIL import Image
    except ImportError as e:
        raise ValueError(f"Failed to read {path}: no decoder for this format") from e
    with Image.open(path) as im:
        return np.ascontiguousarray(np.asarray(im.convert("RGB"), dtype=np.uint8))


def _background(images, i, W, H):
    if images is not None and i < len(images):
        try:
            img = read_image(images[i])
            if img.shape[:2] == (H, W):
                return img
            print(f"[WARN] {images[i]} is {img.shape[1]}x{img.shape[0]}, expected {W}x{H}; painting over black", file=sys.stderr)
        except ValueError as e:
            print(f"[WARN] {e}", file=sys.stderr)
    return np.zeros((H, W, 3), dtype=np.uint8)


def render_frames(model, clouds, frame_ids, images, W, H, out_dir, suffix, chunk=64):
    """renderSMPLMesh(cloud, faces, frame.clone(), fx, fy, cx, cy, fill, cull, no wireframe) + imwrite for each fitted frame
    (main_single_frame.cpp:271-277 -> frame_<i>_render.png, main_multi_frame.cpp:206-229 -> frame_<i>_multi.png)."""
    cam = intrinsics(W, H)
    for s in range(0, len(frame_ids), chunk):
        ids = frame_ids[s:s + chunk]
        imgs = np.stack([_background(images, int(i), W, H) for i in ids])
        api.render_overlay(model, clouds[s:s + chunk], cam, imgs)
        for k, i in enumerate(ids):
            write_png(os.path.join(out_dir, f"frame_{int(i)}_{suffix}.png"), imgs[k])


# ---------------------------------------------------------------------------------------------------------
# keypoints
# ---------------------------------------------------------------------------------------------------------
def _number(j, idx, key):
    """json[idx][key] when it is a JSON number, else None (safe_coord / safe_number, Utils.h:44-58)."""
    if not isinstance(j, list) or idx >= len(j) or not isinstance(j[idx], dict):
        return None
    v = j[idx].get(key)
    if isinstance(v, bool) or not isinstance(v, (int, float)):
        return None
    return float(v)


def map_landmarks(j, W: int, H: int, variant: str = "utils") -> list[api.PixelKP]:
    """The body of load_mp_json on parsed JSON.  Keypoints come out in USE_SMPL order; a landmark is kept iff both
    coordinates are numbers and visibility >= 0.5 (missing visibility counts as 1.0; for the two midpoints the
    ``main_single`` copy defaults it to 0.0, main_single_frame.cpp:84-85)."""
    if variant not in ("utils", "main_single"):
        raise ValueError(f"unknown loader variant {variant!r}")
    if variant == "main_single" and (not isinstance(j, list) or len(j) < 33):
        # j[23]["x"] on a null / short array throws inside nlohmann::json (main_single_frame.cpp:74-80)
        raise ValueError("landmark array has fewer than 33 entries")
    mid_default = 1.0 if variant == "utils" else 0.0

    def midpoint(a, b):
        c = [_number(j, a, "x"), _number(j, a, "y"), _number(j, b, "x"), _number(j, b, "y")]
        if any(x is None for x in c):
            return None
        va, vb = _number(j, a, "visibility"), _number(j, b, "visibility")
        va = mid_default if va is None else va
        vb = mid_default if vb is None else vb
        return 0.5 * (c[0] + c[2]), 0.5 * (c[1] + c[3]), min(va, vb)

    pelvis, chest = midpoint(23, 24), midpoint(11, 12)
    out = []
    for sid in USE_SMPL:
        if sid == 0:
            got = pelvis
        elif sid == 6:
            got = chest
        else:
            mp = MP_MAP[sid]
            got = None
            if mp >= 0:
                x, y = _number(j, mp, "x"), _number(j, mp, "y")
                if x is not None and y is not None:
                    vis = _number(j, mp, "visibility")
                    got = (x, y, 1.0 if vis is None else vis)
        if got is None or got[2] < 0.5:
            continue
        out.append(api.PixelKP(sid, got[0] * W, got[1] * H))
    return out


def load_mp_json(path: str, W: int, H: int, variant: str = "utils") -> list[api.PixelKP]:
    with open(path) as f:
        return map_landmarks(json.load(f), W, H, variant)


def pack_keypoints(frames: list[list[api.PixelKP]]):
    """List of per-frame PixelKP lists -> (kps (F, 17) KP_DTYPE, nkp (F,)) as the C-ABI wants them."""
    kps = np.zeros((len(frames), MAXKP), dtype=KP_DTYPE)
    nkp = np.zeros(len(frames), dtype=np.int32)
    for f, fr in enumerate(frames):
        if len(fr) > MAXKP:
            raise ValueError(f"frame {f}: at most {MAXKP} keypoints per frame")
        nkp[f] = len(fr)
        for i, k in enumerate(fr):
            kps[f, i] = (k.jid, 0, k.u, k.v)
    return kps, nkp


# ---------------------------------------------------------------------------------------------------------
# shared set-up
# ---------------------------------------------------------------------------------------------------------
def intrinsics(W: int, H: int):
    """fx = fy = 0.9 max(H, W), principal point in the image centre (main_single_frame.cpp:171-176)."""
    f = float(max(H, W))
    return 0.9 * f, 0.9 * f, 0.5 * W, 0.5 * H


def initial_root_rotation() -> np.ndarray:
    """r[0] = AngleAxis(pi, Y) * diag(1, -1, 1), column-major 9 vector (main_single_frame.cpp:208-212)."""
    c, s = np.cos(np.pi), np.sin(np.pi)
    ry = np.array([[c, 0.0, s], [0.0, 1.0, 0.0], [-s, 0.0, c]])
    return (ry @ np.diag([1.0, -1.0, 1.0])).T.reshape(9).copy()


def load_model(smpl_path: str):
    """ark::AvatarModel(model_dir): model.npz + pose_prior.txt of one directory (AvatarModel.cpp:17-23, :296).  A path to
    the .npz itself is accepted too (the prior is then looked up next to it)."""
    if os.path.isdir(smpl_path):
        npz, prior = os.path.join(smpl_path, "model.npz"), os.path.join(smpl_path, "pose_prior.txt")
    else:
        npz, prior = smpl_path, os.path.join(os.path.dirname(smpl_path) or ".", "pose_prior.txt")
    model = api.AvatarModel(npz)
    model.posePrior = api.GaussianMixture(prior)     # nComps == -1 when the file is missing, like the reference
    return model


def _g(x: float) -> str:
    return "%g" % x          # std::ostream << double: 6 significant digits


class _Log:
    """log.csv opened in append mode; header only when the file is new (main_multi_frame.cpp:39-42)."""

    def __init__(self, out_dir: str):
        path = os.path.join(out_dir, "log.csv")
        new = not os.path.exists(path)
        self.f = open(path, "a")
        if new:
            self.f.write("frame,mean_pixel_error_px,time_ms\n")

    def row(self, frame: int, px: float, ms: float):
        self.f.write(f"{frame},{_g(px)},{_g(ms)}\n")

    def flush(self):
        self.f.flush()

    def close(self):
        self.f.close()


def _frame_size(img_folder, size):
    if size is not None:
        return int(size[0]), int(size[1]), None
    images = list_sorted(img_folder, IMAGE_EXTS)
    if not images:
        return None, None, images
    W, H = image_size(images[0])
    return W, H, images


# ---------------------------------------------------------------------------------------------------------
# 3dba_single
# ---------------------------------------------------------------------------------------------------------
def run_single(model, keypoints, W: int, H: int, max_iters=100, beta_pose=20.0, beta_shape=30.0, opt_shape=False, use_gmm=False,
               workspace=None, want_cloud=False):
    """Frame loop of src/main_single_frame.cpp:192-255 for a list of per-frame PixelKP lists, batched.

    Frames without keypoints are skipped (:201-204).  Every other frame starts from a fresh avatar (w = 0, p = (0,0,3),
    r[0] = Ry(pi) flipY, identity joints; sim3 = (1, 0, t = p)), is fitted with OptimizePose[Shape]Reprojection using its
    own keypoints' joint ids as valid_joint_ids, updated, and scored with mean_pixel_error.
    Returns dict(frames=indices fitted, px_err, usable, niters, term, sim3, w, jointAA, r, p, jointPos[, cloud], ms)."""
    fx, fy, cx, cy = intrinsics(W, H)
    idx = [i for i, k in enumerate(keypoints) if len(k) > 0]
    kps, nkp = pack_keypoints([keypoints[i] for i in idx])
    b = len(idx)
    out = dict(frames=np.array(idx, dtype=np.int64), ms=0.0)
    if b == 0:
        return out
    R0 = np.tile(initial_root_rotation(), (b, 1))
    sim3 = np.tile(np.array([1.0, 0, 0, 0, 0, 0, 3.0]), (b, 1))
    w = np.zeros((b, NS))
    opts = api.fit_opts(FIT_POSE_SHAPE if opt_shape else FIT_POSE, max_iters, beta_pose, beta_shape, 0.0, (fx, fy, cx, cy))
    ws = workspace or api.Workspace()
    t0 = time.perf_counter()
    res = api.fit_frames(ws, model, model.posePrior if use_gmm else None, opts, kps, nkp, R0, sim3, w, want_cloud=want_cloud)
    out["ms"] = (time.perf_counter() - t0) * 1e3
    out.update(res)
    return out


def main_single(argv=None) -> int:
    argv = list(sys.argv if argv is None else argv)
    size, save, render = None, False, True
    rest = []
    it = iter(argv[1:])
    for a in it:                                   # three extensions, all optional and absent from the reference's CLI
        if a == "--size":
            wh = next(it, "").lower().split("x")
            size = (int(wh[0]), int(wh[1]))
        elif a == "--save-params":
            save = True
        elif a == "--no-render":
            render = False
        else:
            rest.append(a)
    if len(rest) < 4:
        print("usage: 3dba_single <SMPL.npz> <kps_folder> <images_folder> <out_dir>\n"
              "                   [max_iters=100] [beta_pose=20] [beta_shape=30]\n"
              "                   [--opt-shape] [--use-gmm]")
        return 0
    smpl_path, kps_folder, img_folder, out_dir = rest[:4]
    max_iters, beta_pose, beta_shape, opt_shape, use_gmm = 100, 20.0, 30.0, False, False
    seen = 0
    for a in rest[4:]:
        if a == "--opt-shape":
            opt_shape = True
            continue
        if a == "--use-gmm":
            use_gmm = True
            continue
        try:
            val = float(a)
        except ValueError:
            print(f"[WARN] Unknown arg ignored: {a}", file=sys.stderr)
            continue
        if seen == 0:
            max_iters = max(1, int(val))
        elif seen == 1:
            beta_pose = val
        elif seen == 2:
            beta_shape = val
        seen += 1
    print(f"[ARGS] max_iters={max_iters}  beta_pose={_g(beta_pose)}  beta_shape={_g(beta_shape)}"
          f"  opt_shape={'true' if opt_shape else 'false'}  use_gmm={'true' if use_gmm else 'false'}", flush=True)
    os.makedirs(out_dir, exist_ok=True)

    W, H, images = _frame_size(img_folder, size)
    if W is None:
        print(f"No images in {img_folder}", file=sys.stderr)
        return 1
    jsons = list_sorted(kps_folder, (".json",))
    if not jsons:
        print(f"No JSON files in {kps_folder}", file=sys.stderr)
        return 1
    model = load_model(smpl_path)

    frames = []
    for i, jp in enumerate(jsons):
        if images is not None and i >= len(images):
            print(f"No image for frame {i}", file=sys.stderr)
            break
        k = load_mp_json(jp, W, H)
        if not k:
            print(f"Frame {i} has no valid keypoints; skipping.", file=sys.stderr)
        frames.append(k)
    print(f"Pose prior components: {model.posePrior.nComps}  (GMM {'ON' if use_gmm else 'OFF'})")
    res = run_single(model, frames, W, H, max_iters, beta_pose, beta_shape, opt_shape, use_gmm, want_cloud=save or render)
    n = len(res["frames"])
    log = _Log(out_dir)
    for k, i in enumerate(res["frames"]):
        log.row(int(i), float(res["px_err"][k]), res["ms"] / n)
    log.close()
    if save and n:
        np.savez(os.path.join(out_dir, "params.npz"), **{k: res[k] for k in
                 ("frames", "sim3", "w", "jointAA", "r", "p", "jointPos", "px_err", "usable", "niters", "term", "cloud") if k in res})
    if render and n:
        render_frames(model, res["cloud"], res["frames"], images, W, H, out_dir, "render")
    print("Done.")
    return 0


# ---------------------------------------------------------------------------------------------------------
# 3dba_multi
# ---------------------------------------------------------------------------------------------------------
def run_multi(model, keypoints, W: int, H: int, it1=1000, skip=10, wsize=20, overlap=5, beta_pose=5.0, beta_shape=25.0,
              lambda_temp=3.0, it2=60, beta_shape_lock=1e5, log=None, echo=None, on_solve=None):
    """Anchor stage + sliding windows of src/main_multi_frame.cpp:80-201 for one sequence.

    Each OptimizeMultiFrame is one b3d_fit_window call, in the reference's order (the windows overlap, so they are a serial
    chain).  State kept as the reference keeps it: ``poses`` are the FramePoseParams (anchors work on a COPY, :117, so the
    stage-1 result only reaches the avatars), ``R0`` is avatars[i]->r[0] and accumulates R(rootAA) at every write-back
    (MultiFrameBA.h:160-163), ``Wsh`` is avatars[i]->w (shared after stage 1, :154; a window writes its front avatar's).
    ``log.row(frame, px, ms)`` gets the anchors' rows, then every window's rows, like log.csv.  ``on_solve(ids, opts, inputs,
    result)`` (tests) sees every solve's inputs (R0, poses, w as they went in) and its result with the iteration trace."""
    n = len(keypoints)
    fx, fy, cx, cy = cam = intrinsics(W, H)
    kps, nkp = pack_keypoints(keypoints)
    poses = np.zeros((n, NPF))
    poses[:, 0], poses[:, 6] = 1.0, 3.0
    R0 = np.tile(initial_root_rotation(), (n, 1))
    Wsh = np.zeros((n, NS))
    r = np.tile(np.eye(3).reshape(9), (n, NJ, 1))
    r[:, 0] = R0
    p = np.tile(np.array([0.0, 0.0, 3.0]), (n, 1))
    say = echo or (lambda s: None)

    def solve(ids, P, beta_s, iters):
        """OptimizeMultiFrame over frames ``ids`` with pose blocks P; returns usable, updated P.  Writes the avatars back."""
        o = api.fit_opts(FIT_POSE_SHAPE, iters, beta_pose, beta_s, lambda_temp, cam)
        before = dict(R0=R0[ids].copy(), poses=np.array(P, copy=True), w=Wsh[ids[0]].copy()) if on_solve else None
        t0 = time.perf_counter()
        res = api.fit_window(model, o, kps[ids], nkp[ids], R0[ids], P, Wsh[ids[0]], want_trace=on_solve is not None)
        if on_solve:
            on_solve(ids, o, before, res)
        R0n, rn, pn = api.writeback(res["poses"], R0[ids])
        ms = (time.perf_counter() - t0) * 1e3
        R0[ids], r[ids], p[ids] = R0n, rn, pn
        Wsh[ids[0]] = res["w"]
        return res, ms

    def score(ids):
        _, jp, _ = api.smpl_forward(model, Wsh[ids], p[ids], r[ids], dtype=np.float64, want_cloud=False)
        return api.mean_pixel_error_batch(kps[ids], nkp[ids], jp, cam)

    rows, fails = [], 0
    anchors = np.arange(0, n, skip)
    say(f"[INFO] stage-1  anchor frames = {len(anchors)}")
    res, ms_anchor = solve(anchors, poses[anchors].copy(), beta_shape, it1)
    fails += 0 if res["usable"] else 1
    say(f"[INFO] stage-1 done  ({'success' if res['usable'] else 'fail'})  in {_g(ms_anchor)} ms\n"
        f"iterations {res['niters']}, termination {res['term']}")
    for fid, px in zip(anchors, score(anchors)):
        rows.append((int(fid), float(px), ms_anchor / len(anchors)))
    Wsh[:] = Wsh[0]
    if log:
        for row in rows:
            log.row(*row)
        log.flush()

    stride = wsize - overlap
    for s in range(0, n, stride):
        e = min(s + wsize, n)
        ids = np.arange(s, e)
        res, ms_win = solve(ids, poses[ids], beta_shape_lock, it2)
        fails += 0 if res["usable"] else 1
        say(f"[INFO] window [{s},{e})  solving …  -> {'OK' if res['usable'] else 'FAIL'}  ({_g(ms_win)} ms)")
        poses[ids] = res["poses"]
        new = [(int(i), float(px), ms_win / (e - s)) for i, px in zip(ids, score(ids))]
        rows += new
        if log:
            for row in new:
                log.row(*row)
            log.flush()
    return dict(poses=poses, R0=R0, W=Wsh, r=r, p=p, rows=rows, nfail=fails)


def main_multi(argv=None) -> int:
    argv = list(sys.argv if argv is None else argv)
    size, save, render = None, False, True
    rest = []
    it = iter(argv[1:])
    for a in it:
        if a == "--size":
            wh = next(it, "").lower().split("x")
            size = (int(wh[0]), int(wh[1]))
        elif a == "--save-params":
            save = True
        elif a == "--no-render":
            render = False
        else:
            rest.append(a)
    if len(rest) < 4:
        print("usage: 3dba <SMPL.npz> <kps_folder> <image_folder> <out_dir>\n"
              "             [max_iters=120] [max_iters_stage2=120] [anchor_skip=15] [window=30] [overlap=10]\n"
              "             [beta_pose=5.0] [beta_shape=25.0]")
        return 0
    smpl_path, kps_folder, img_folder, out_dir = rest[:4]
    opt = rest[4:]

    def arg(i, conv, default):
        return conv(opt[i]) if len(opt) > i else default

    def atoi(s):                                   # std::atoi: 0 on garbage
        try:
            return int(float(s))
        except ValueError:
            return 0

    def atof(s):
        try:
            return float(s)
        except ValueError:
            return 0.0

    it1 = arg(0, atoi, 1000)
    _it2 = arg(1, atoi, 500)                       # parsed and never used: the window solves hard-code 60 (:181)
    skip, wsize, overlap = arg(2, atoi, 10), arg(3, atoi, 20), arg(4, atoi, 5)
    beta_pose, beta_shape, lam = arg(5, atof, 5.0), arg(6, atof, 25.0), arg(7, atof, 3.0)
    os.makedirs(out_dir, exist_ok=True)
    log = _Log(out_dir)

    jsons = list_sorted(kps_folder, (".json",))
    W, H, images = _frame_size(img_folder, size)
    if W is None or not jsons or (images is not None and len(images) != len(jsons)):
        print("image / json count mismatch", file=sys.stderr)
        return 1
    print(f"[INFO] frames: \"{img_folder}\"  = {len(jsons):4d}\n[INFO] anchor skip     : {skip}\n"
          f"[INFO] window / overlap: {wsize} / {overlap}\n"
          f"[INFO] β_pose={_g(beta_pose)}  β_shape={_g(beta_shape)}  λ_temp={_g(lam)}", flush=True)
    if skip <= 0 or wsize - overlap <= 0:
        raise ValueError("anchor_skip and window - overlap must be positive")
    model = load_model(smpl_path)
    frames = [load_mp_json(jp, W, H) for jp in jsons]
    res = run_multi(model, frames, W, H, it1, skip, wsize, overlap, beta_pose, beta_shape, lam, log=log, echo=print)
    log.close()
    if save:
        np.savez(os.path.join(out_dir, "params.npz"), **{k: res[k] for k in ("poses", "R0", "W", "r", "p")})
    if render:
        clouds, _, _ = api.smpl_forward(model, res["W"], res["p"], res["r"], dtype=np.float32)
        render_frames(model, clouds, np.arange(len(frames)), images, W, H, out_dir, "multi")
        print(f"[INFO] rendering finished, saved to  \"{out_dir}\"")
    print("done.")
    return 0


if __name__ == "__main__":
    prog = sys.argv[1] if len(sys.argv) > 1 else ""
    if prog in ("single", "3dba_single"):
        sys.exit(main_single(["3dba_single"] + sys.argv[2:]))
    if prog in ("multi", "3dba_multi", "3dba"):
        sys.exit(main_multi(["3dba"] + sys.argv[2:]))
    print("usage: python -m 3dbodyanimation_b200.drivers {single|multi} <args of 3dba_single / 3dba_multi>")
    sys.exit(0)
