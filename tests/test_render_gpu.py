"""CUDA overlay rasteriser (b3d_render_overlay, csrc/render.cu) against the oracle's renderSMPLMesh restatement and the committed
OpenCV golden images.  Bit-exact: every pixel of every image."""
import hashlib
import os

import numpy as np
import pytest

from conftest import GOLDEN
from test_render_cpu import _background, golden_scene

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def scene(api, synth):
    model, cloud, faces, img = golden_scene(synth)
    gm = api.AvatarModel(dict(model, f=faces))
    return gm, model, cloud, faces, img


def test_render_matches_opencv_golden_and_oracle(api, orc, synth, scene):
    gm, model, cloud, faces, img = scene
    cam = (synth.FX, synth.FY, synth.CX, synth.CY)
    g = np.load(os.path.join(GOLDEN, "render_cv2.npz"))
    out = api.render_overlay(gm, cloud, cam, img.copy())
    assert [hashlib.sha256(out[b].tobytes()).hexdigest() for b in range(4)] == list(g["sha256"])
    assert np.array_equal(out[-1], g["line8_last"])
    for cull in (True, False):
        ref, _, _ = orc.render_overlay(cloud, faces, cam, img.copy(), backface_cull=cull)
        got = api.render_overlay(gm, cloud, cam, img.copy(), backface_cull=cull)
        assert np.array_equal(got, ref), cull
    # float64 clouds take the other instantiation; the float32 values converted exactly give the same picture
    got64 = api.render_overlay(gm, cloud.astype(np.float64), cam, img.copy())
    assert np.array_equal(got64, out)
    # the reference's per-body call
    one = img[0].copy()
    api.renderSMPLMesh(gm, cloud[0], one, *cam, True, True, False)
    assert np.array_equal(one, out[0])


def test_render_batch_after_forward(api, orc, omodel, synth, scene):
    """300 bodies (more than one per SM: the scratch planes are reused) posed by the CUDA evaluator, other image size."""
    gm, model, _, faces, _ = scene
    n = 300
    fi = synth.make_forward_inputs(n, 5)
    rng = np.random.default_rng(9)
    p = np.stack([rng.uniform(-1.5, 1.5, n), rng.uniform(-1.0, 1.0, n), rng.uniform(1.0, 6.0, n)], axis=1)
    cloud, _, _ = api.smpl_forward(gm, fi["w"], p, fi["r"], dtype=np.float32)
    W, H = 320, 200
    cam = (0.9 * W, 0.9 * W, 0.5 * W, 0.5 * H)
    img = np.ascontiguousarray(np.tile(_background(H, W)[None], (n, 1, 1, 1)))
    got = api.render_overlay(gm, cloud, cam, img.copy())
    ref, drawn, _ = orc.render_overlay(cloud, faces, cam, img.copy())
    bad = [b for b in range(n) if not np.array_equal(got[b], ref[b])]
    assert not bad, bad[:10]
    assert drawn.min() > 3000


def test_render_degenerate_inputs(api, orc, synth, scene):
    gm, model, cloud, faces, img = scene
    cam = (synth.FX, synth.FY, synth.CX, synth.CY)
    c = np.repeat(cloud[:1], 5, axis=0).copy()
    c[0, :, 2] -= 10.0                       # behind the camera: untouched image
    c[1, ::7, 2] = 2e-6                      # projected beyond +-2^20 px: those faces are dropped
    c[2, :, 2] = 3.0                         # a flat sheet: every face has the same depth -> ties go to the higher face index
    c[3, :, 0] += 40.0                       # far off to the right: every edge is clipped away
    c[4, :, 2] = np.where(np.arange(c.shape[1]) % 2 == 0, 0.02, c[4, :, 2])   # huge triangles across the whole frame
    ref, drawn, _ = orc.render_overlay(c, faces, cam, img[:1].repeat(5, axis=0).copy(), backface_cull=False)
    got = api.render_overlay(gm, c, cam, img[:1].repeat(5, axis=0).copy(), backface_cull=False)
    for b in range(5):
        assert np.array_equal(got[b], ref[b]), b
    assert np.array_equal(got[0], img[0]) and np.array_equal(got[3], img[0]) and drawn[0] == 0


def test_render_errors(api, synth, model_np):
    nofaces = api.AvatarModel({k: v for k, v in model_np.items() if k != "f"})
    with pytest.raises(Exception):
        api.render_overlay(nofaces, np.zeros((1, synth.V, 3), np.float32), (1.0, 1.0, 0.0, 0.0), np.zeros((1, 8, 8, 3), np.uint8))
    with pytest.raises(ValueError):
        api.render_overlay(nofaces, np.zeros((1, synth.V, 3), np.float32), (1.0, 1.0, 0.0, 0.0), np.zeros((1, 8, 8), np.uint8))
