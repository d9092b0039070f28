"""Generates tests/golden/render_cv2.npz with the REAL OpenCV (cv2 in this image): the painter's loop of the reference's
renderSMPLMesh (include/RenderSMPLMesh.h:16-110) written with numpy + cv2.fillConvexPoly, on posed bodies of the synthetic model.
Stored: the inputs' recipe (seeds), the LINE_8 images the oracle / CUDA rasteriser must reproduce bit for bit, and the LINE_AA
images of the reference's actual call for the record (how far the missing edge blend is from it).
    python tools/make_render_golden.py"""
import importlib
import os
import sys

import cv2
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
synth = importlib.import_module("3dbodyanimation_b200.synth")


def render_cv2(cloud, faces, img, fx, fy, cx, cy, line_type, cull=True):
    cloud = np.asarray(cloud, dtype=np.float64)
    z = cloud[:, 2]
    valid = z > 1e-6
    with np.errstate(divide="ignore", invalid="ignore"):
        u = (fx * cloud[:, 0] / z + cx).astype(np.float32)
        v = (fy * cloud[:, 1] / z + cy).astype(np.float32)
    rnd = lambda a: np.where(a >= 0, np.floor(a + np.float32(0.5)), -np.floor(-a + np.float32(0.5)))    # std::round on float
    items = []
    for f, (i0, i1, i2) in enumerate(faces):
        if not (valid[i0] and valid[i1] and valid[i2]):
            continue
        v0, v1, v2 = cloud[i0], cloud[i1], cloud[i2]
        e1, e2 = v1 - v0, v2 - v0
        n = np.array([e1[1] * e2[2] - e1[2] * e2[1], e1[2] * e2[0] - e1[0] * e2[2], e1[0] * e2[1] - e1[1] * e2[0]])
        if cull and n[2] >= 0.0:
            continue
        c = (v0 + v1 + v2) / 3.0
        cn = np.sqrt(c[0] * c[0] + c[1] * c[1] + c[2] * c[2])
        nn = np.sqrt(n[0] * n[0] + n[1] * n[1] + n[2] * n[2])
        view, nh = -c / (cn if cn > 0 else 1.0), n / (nn if nn > 0 else 1.0)
        shade = min(max(nh[0] * view[0] + nh[1] * view[1] + nh[2] * view[2], 0.0), 1.0)
        col = int(np.floor(220 * shade + 0.5))
        pts = np.array([[rnd(u[i]), rnd(v[i])] for i in (i0, i1, i2)], dtype=np.int32)
        items.append(((v0[2] + v1[2] + v2[2]) / 3.0, f, pts, col))
    items.sort(key=lambda t: -t[0])          # stable: ties keep face order
    for _, _, pts, col in items:
        cv2.fillConvexPoly(img, pts.reshape(-1, 1, 2), (col, col, col), line_type)
    return len(items)


def bodies(model, n, seed):
    fi = synth.make_forward_inputs(n, seed)
    rng = np.random.default_rng(seed)
    p = np.stack([rng.uniform(-0.8, 0.8, n), rng.uniform(-0.5, 0.5, n), rng.uniform(2.0, 4.5, n)], axis=1)
    p[0] = (0.0, 0.0, 3.0)
    p[-1] = (1.1, -0.9, 1.2)                  # close and half out of the frame: exercises cv::clipLine and the border clamps
    cloud = np.stack([synth.forward_numpy(model, fi["w"][b], p[b], fi["r"][b])[0] for b in range(n)])
    return np.ascontiguousarray(cloud, dtype=np.float32)


if __name__ == "__main__":
    model = synth.make_model()
    W, H = synth.CAM_W, synth.CAM_H
    cam = (synth.FX, synth.FY, synth.CX, synth.CY)
    cloud = bodies(model, 4, 77)
    faces = synth.make_mesh_faces(model).astype(np.int64)
    yy, xx = np.mgrid[0:H, 0:W]
    bg = np.stack([(xx // 2) % 256, (yy * 255) // (H - 1), ((xx + yy) // 3) % 256], axis=2).astype(np.uint8)   # smooth: compresses
    out8, outaa, drawn = [], [], []
    for b in range(cloud.shape[0]):
        a, c = bg.copy(), bg.copy()
        drawn.append(render_cv2(cloud[b], faces, a, *cam, cv2.LINE_8))
        render_cv2(cloud[b], faces, c, *cam, cv2.LINE_AA)
        out8.append(a)
        outaa.append(c)
    out8, outaa = np.stack(out8), np.stack(outaa)
    diff = (out8 != outaa).any(axis=3)
    body = (out8 != bg[None]).any(axis=3)
    print("mean |LINE_8 - LINE_AA| over body pixels (grey levels):",
          [float(np.abs(out8[b].astype(float) - outaa[b])[body[b]].mean()) for b in range(len(drawn))], "body px", body.reshape(len(drawn), -1).sum(axis=1))
    print("cv2", cv2.__version__, "drawn", drawn, "pixels LINE_8 != LINE_AA per body:", diff.reshape(len(drawn), -1).sum(axis=1),
          "max |diff|", int(np.abs(out8.astype(int) - outaa.astype(int)).max()))
    import hashlib
    sha = np.array([hashlib.sha256(out8[b].tobytes()).hexdigest() for b in range(len(drawn))])
    # all four images by hash, the border-clipped one in full (kept small on purpose)
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "render_cv2.npz"), sha256=sha, line8_last=out8[-1], drawn=np.array(drawn),
                        line_aa_absdiff_body_mean=np.array([np.abs(out8[b].astype(float) - outaa[b])[body[b]].mean() for b in range(len(drawn))]),
                        seed=np.array([77]), cv2_version=np.array(cv2.__version__))
