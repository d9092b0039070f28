"""Overlay renderer oracle (oracle/render.cpp: renderSMPLMesh restated, OpenCV's LINE_8 polygon fill restated) against
  * the committed golden images that the REAL OpenCV produced (tests/golden/render_cv2.npz, tools/make_render_golden.py);
  * cv2 itself when it is importable on this machine (random triangles inside and across the borders, whole painter loops).
Bit-exact: pixel / index work."""
import hashlib
import importlib.util
import os
import sys

import numpy as np
import pytest

from conftest import GOLDEN, ROOT


def _golden_tool():
    spec = importlib.util.spec_from_file_location("make_render_golden", os.path.join(ROOT, "tools", "make_render_golden.py"))
    mod = importlib.util.module_from_spec(spec)
    sys.modules["make_render_golden"] = mod
    return spec, mod


def _background(H, W):
    yy, xx = np.mgrid[0:H, 0:W]
    return np.stack([(xx // 2) % 256, (yy * 255) // (H - 1), ((xx + yy) // 3) % 256], axis=2).astype(np.uint8)


def golden_scene(synth):
    """Inputs of tools/make_render_golden.py rebuilt without cv2 (the recipe is seeds only)."""
    model = synth.make_model()
    n, seed = 4, 77
    fi = synth.make_forward_inputs(n, seed)
    rng = np.random.default_rng(seed)
    p = np.stack([rng.uniform(-0.8, 0.8, n), rng.uniform(-0.5, 0.5, n), rng.uniform(2.0, 4.5, n)], axis=1)
    p[0] = (0.0, 0.0, 3.0)
    p[-1] = (1.1, -0.9, 1.2)
    cloud = np.stack([synth.forward_numpy(model, fi["w"][b], p[b], fi["r"][b])[0] for b in range(n)]).astype(np.float32)
    faces = synth.make_mesh_faces(model)
    bg = _background(synth.CAM_H, synth.CAM_W)
    return model, np.ascontiguousarray(cloud), faces, np.ascontiguousarray(np.tile(bg[None], (n, 1, 1, 1)))


def test_mesh_faces_are_a_surface(synth):
    model = synth.make_model()
    f = synth.make_mesh_faces(model)
    assert f.dtype == np.uint32 and f.shape[1] == 3 and 13000 < f.shape[0] < 14000 and f.max() < synth.V
    # closed surfaces: every undirected edge is shared by exactly two faces, once in each direction
    e = np.concatenate([f[:, [0, 1]], f[:, [1, 2]], f[:, [2, 0]]]).astype(np.int64)
    key = e[:, 0] * synth.V + e[:, 1]
    rev = e[:, 1] * synth.V + e[:, 0]
    assert len(np.unique(key)) == len(key) and np.array_equal(np.sort(key), np.sort(rev))


def test_oracle_matches_opencv_golden(orc, synth):
    g = np.load(os.path.join(GOLDEN, "render_cv2.npz"))
    model, cloud, faces, img = golden_scene(synth)
    cam = (synth.FX, synth.FY, synth.CX, synth.CY)
    out, drawn, order = orc.render_overlay(cloud, faces, cam, img, want_order=True)
    assert np.array_equal(drawn, g["drawn"])
    assert [hashlib.sha256(out[b].tobytes()).hexdigest() for b in range(4)] == list(g["sha256"])
    assert np.array_equal(out[-1], g["line8_last"])
    # draw list: far to near, stable
    z = cloud.astype(np.float64)[0][:, 2]
    d = (z[faces[order[0, :drawn[0]], 0]] + z[faces[order[0, :drawn[0]], 1]] + z[faces[order[0, :drawn[0]], 2]]) / 3.0
    assert np.all(np.diff(d) <= 0) and np.all(order[0, drawn[0]:] == -1)
    # the missing LINE_AA blend, for the record: ~10 grey levels mean over the body pixels
    assert 5.0 < float(g["line_aa_absdiff_body_mean"].mean()) < 20.0


def test_oracle_cull_flag_and_invalid_vertices(orc, synth):
    model, cloud, faces, img = golden_scene(synth)
    cam = (synth.FX, synth.FY, synth.CX, synth.CY)
    a, da, _ = orc.render_overlay(cloud[:1], faces, cam, img[:1].copy(), backface_cull=True)
    b, db, _ = orc.render_overlay(cloud[:1], faces, cam, img[:1].copy(), backface_cull=False)
    assert db[0] == faces.shape[0] and da[0] < db[0]
    # closed surfaces: front faces mostly hide the back faces (painter's order by mean depth is not exact visibility where the
    # skinned blobs interpenetrate), so the two pictures differ in few pixels
    assert np.mean(np.any(a != b, axis=3)) < 0.15
    behind = cloud[:1].copy()
    behind[0, :, 2] -= 10.0                      # everything behind the camera: nothing valid, image untouched
    c, dc, _ = orc.render_overlay(behind, faces, cam, img[:1].copy())
    assert dc[0] == 0 and np.array_equal(c, img[:1])
    near = cloud[:1].copy()
    near[0, ::7, 2] = 2e-6                       # valid but absurdly close: pixel coordinates beyond 2^20, faces dropped
    d, dd, _ = orc.render_overlay(near, faces, cam, img[:1].copy())
    assert 0 < dd[0] < da[0]


cv2 = pytest.importorskip("cv2") if importlib.util.find_spec("cv2") else None


@pytest.mark.skipif(cv2 is None, reason="OpenCV not importable here; the committed golden images cover it")
def test_triangle_fill_matches_cv2(orc):
    rng = np.random.default_rng(3)
    W, H = 64, 48
    for t in range(6000):
        kind = t % 5
        if kind == 0:
            pts = rng.integers(0, [W, H], size=(3, 2))
        elif kind == 1:
            pts = rng.integers(-40, [W + 40, H + 40], size=(3, 2))
        elif kind == 2:
            pts = rng.integers(-5, [W + 5, H + 5], size=(1, 2)) + rng.integers(-4, 5, size=(3, 2))
        elif kind == 3:
            pts = rng.integers(-3000, [W + 3000, H + 3000], size=(3, 2))
        else:
            pts = rng.integers(0, [W, H], size=(1, 2)) + rng.integers(-1, 2, size=(3, 2))       # degenerate / repeated points
        pts = pts.astype(np.int32)
        a, b = np.zeros((H, W, 3), np.uint8), np.zeros((H, W, 3), np.uint8)
        cv2.fillConvexPoly(a, pts.reshape(-1, 1, 2), (200, 200, 200), cv2.LINE_8)
        orc.fill_triangle(b, pts, 200)
        assert np.array_equal(a, b), pts.tolist()


@pytest.mark.skipif(cv2 is None, reason="OpenCV not importable here; the committed golden images cover it")
def test_painter_loop_matches_cv2(orc, synth):
    spec, tool = _golden_tool()
    spec.loader.exec_module(tool)
    model, cloud, faces, img = golden_scene(synth)
    cam = (synth.FX, synth.FY, synth.CX, synth.CY)
    for cull in (True, False):
        ref = img[1].copy()
        n = tool.render_cv2(cloud[1], faces.astype(np.int64), ref, *cam, cv2.LINE_8, cull=cull)
        out, drawn, _ = orc.render_overlay(cloud[1:2], faces, cam, img[1:2].copy(), backface_cull=cull)
        assert drawn[0] == n and np.array_equal(out[0], ref)
    # random-vertex faces of make_model: huge overlapping triangles, heavy clipping, small image
    rf = np.asarray(model["f"][:600], dtype=np.int64)
    small = np.ascontiguousarray(_background(90, 160))
    camS = (144.0, 144.0, 80.0, 45.0)
    ref = small.copy()
    tool.render_cv2(cloud[3], rf, ref, *camS, cv2.LINE_8, cull=False)
    out, _, _ = orc.render_overlay(cloud[3:4], rf, camS, small[None].copy(), backface_cull=False)
    assert np.array_equal(out[0], ref)
