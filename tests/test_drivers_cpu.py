"""Host logic of the executables' replacements (3dbodyanimation_b200/drivers.py) that needs no GPU: the MediaPipe loader against
the oracle's restatement on every shipped fixture frame and on hand-made edge cases (bit-exact: index / byte work), file
listing, image headers, the log.csv format."""
import glob
import json
import os
import struct
import zlib

import numpy as np
import pytest

from conftest import GOLDEN, pkg


@pytest.fixture(scope="module")
def drv():
    return pkg("drivers")


def _same(a, b):
    return [(k.jid, k.u, k.v) for k in a] == [tuple(k) for k in b]


def test_loader_matches_oracle_on_fixture(drv, orc):
    files = sorted(glob.glob(os.path.join(GOLDEN, "keypoints_video1", "*.json")))
    assert len(files) == 38
    empty = 0
    for f in files:
        lm = json.load(open(f))
        got = drv.load_mp_json(f, 480, 270)
        assert _same(got, orc.map_mediapipe(lm, 480, 270, "utils"))           # exact: same doubles, same order
        if not lm:
            empty += 1
            assert got == []
            with pytest.raises(ValueError):
                drv.map_landmarks(lm, 480, 270, "main_single")
        else:
            assert _same(drv.map_landmarks(lm, 480, 270, "main_single"), orc.map_mediapipe(lm, 480, 270, "main_single"))
            assert [k.jid for k in got].count(0) in (0, 2)                      # USE_SMPL's trailing zeros: pelvis twice
    assert empty == 5


def test_loader_edge_cases(drv, orc):
    base = [{"x": 0.01 * i, "y": 0.02 * i, "z": 0.0, "visibility": 0.9} for i in range(33)]
    cases = []
    a = json.loads(json.dumps(base)); del a[23]["visibility"]; cases.append(a)          # midpoint default differs per variant
    a = json.loads(json.dumps(base)); a[25]["x"] = "0.3"; cases.append(a)               # not a number -> dropped
    a = json.loads(json.dumps(base)); a[26]["y"] = None; cases.append(a)
    a = json.loads(json.dumps(base)); a[27]["visibility"] = 0.49; a[28]["visibility"] = 0.5; cases.append(a)
    a = json.loads(json.dumps(base)); a[11]["x"] = True; cases.append(a)                # JSON boolean is not a number
    a = json.loads(json.dumps(base)); del a[13]["visibility"]; cases.append(a)          # plain landmark: default 1.0
    a = json.loads(json.dumps(base)); a[24] = {}; cases.append(a)                       # no pelvis, no left hip -> 14 keypoints
    for lm in cases:
        for variant in ("utils", "main_single"):
            assert _same(drv.map_landmarks(lm, 640, 360, variant), orc.map_mediapipe(lm, 640, 360, variant)), variant
    assert len(drv.map_landmarks(cases[0], 640, 360, "utils")) == 17 and len(drv.map_landmarks(cases[0], 640, 360, "main_single")) == 15
    assert len(drv.map_landmarks(cases[6], 640, 360, "utils")) == 14
    assert drv.map_landmarks(None, 640, 360) == [] and drv.map_landmarks([], 640, 360) == []
    assert drv.map_landmarks(base[:20], 640, 360) != []                                 # short arrays: what exists is used
    with pytest.raises(ValueError):
        drv.map_landmarks(base[:20], 640, 360, "main_single")
    kps, nkp = drv.pack_keypoints([drv.map_landmarks(base, 640, 360), []])
    assert kps.shape == (2, 17) and list(nkp) == [17, 0] and list(kps[0]["jid"]) == drv.USE_SMPL
    assert kps[0]["u"][0] == 0.23 * 640 and kps[0]["v"][0] == 0.46 * 360


def _png(path, w, h):
    def chunk(tag, data):
        return struct.pack(">I", len(data)) + tag + data + struct.pack(">I", zlib.crc32(tag + data))
    raw = b"".join(b"\x00" + b"\x00" * (3 * w) for _ in range(h))
    open(path, "wb").write(b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, 2, 0, 0, 0)) +
                           chunk(b"IDAT", zlib.compress(raw)) + chunk(b"IEND", b""))


def test_files_and_image_headers(drv, tmp_path):
    _png(tmp_path / "b.png", 480, 270)
    open(tmp_path / "a.bmp", "wb").write(b"BM" + b"\x00" * 16 + struct.pack("<ii", 123, -45) + b"\x00" * 28)
    jpg = b"\xff\xd8" + b"\xff\xe0" + struct.pack(">H", 16) + b"JFIF\x00" + b"\x00" * 9 + b"\xff\xdb" + struct.pack(">H", 67) + b"\x00" * 65 + \
          b"\xff\xc0" + struct.pack(">HBHHB", 17, 8, 720, 1280, 3) + b"\x00" * 9 + b"\xff\xd9"
    open(tmp_path / "c.JPG", "wb").write(jpg)
    open(tmp_path / "notes.txt", "w").write("x")
    os.mkdir(tmp_path / "d.png")
    listed = drv.list_sorted(str(tmp_path), drv.IMAGE_EXTS)
    assert [os.path.basename(p) for p in listed] == ["a.bmp", "b.png", "c.JPG"]
    assert drv.image_size(listed[0]) == (123, 45) and drv.image_size(listed[1]) == (480, 270) and drv.image_size(listed[2]) == (1280, 720)
    with pytest.raises(ValueError):
        drv.image_size(str(tmp_path / "notes.txt"))
    assert drv.intrinsics(480, 270) == (432.0, 432.0, 240.0, 135.0)
    assert drv.intrinsics(270, 480) == (432.0, 432.0, 135.0, 240.0)


def test_log_format(drv, tmp_path):
    log = drv._Log(str(tmp_path))
    log.row(0, 12.3456789, 1.0)
    log.row(10, 1e-7, 123456789.0)
    log.close()
    log = drv._Log(str(tmp_path))        # second run appends without a second header
    log.row(3, 2.5, 0.125)
    log.close()
    assert open(tmp_path / "log.csv").read() == "frame,mean_pixel_error_px,time_ms\n0,12.3457,1\n10,1e-07,1.23457e+08\n3,2.5,0.125\n"


def test_usage_and_missing_inputs(drv, tmp_path, capsys):
    assert drv.main_single(["3dba_single"]) == 0 and "usage: 3dba_single" in capsys.readouterr().out
    assert drv.main_multi(["3dba", "a", "b"]) == 0 and "usage: 3dba " in capsys.readouterr().out
    os.mkdir(tmp_path / "img"); os.mkdir(tmp_path / "kps")
    assert drv.main_single(["3dba_single", "m.npz", str(tmp_path / "kps"), str(tmp_path / "img"), str(tmp_path / "out")]) == 1
    assert "No images in" in capsys.readouterr().err
    assert drv.main_single(["3dba_single", "m.npz", str(tmp_path / "kps"), str(tmp_path / "img"), str(tmp_path / "out"), "--size", "480x270"]) == 1
    assert "No JSON files in" in capsys.readouterr().err
    assert drv.main_multi(["3dba", "m.npz", str(tmp_path / "kps"), str(tmp_path / "img"), str(tmp_path / "out")]) == 1
    assert "image / json count mismatch" in capsys.readouterr().err


def test_png_round_trip_and_foreign_pngs(drv, tmp_path):
    rng = np.random.default_rng(0)
    img = rng.integers(0, 256, size=(37, 53, 3), dtype=np.uint8)
    drv.write_png(str(tmp_path / "a.png"), img)
    assert drv.image_size(str(tmp_path / "a.png")) == (53, 37)
    assert np.array_equal(drv.read_image(str(tmp_path / "a.png")), img)
    PIL = pytest.importorskip("PIL.Image")
    assert np.array_equal(np.asarray(PIL.open(tmp_path / "a.png").convert("RGB")), img)       # a standard decoder reads what we write
    smooth = np.stack([np.add.outer(np.arange(40), np.arange(64)) % 256] * 3, axis=2).astype(np.uint8)
    smooth[..., 1] = smooth[..., 1] // 2
    PIL.fromarray(smooth).save(tmp_path / "b.png", optimize=True)                             # adaptive filters (Sub / Up / Avg / Paeth)
    assert np.array_equal(drv.read_image(str(tmp_path / "b.png")), smooth)
    PIL.fromarray(smooth[..., 0]).save(tmp_path / "g.png")                                    # greyscale
    assert np.array_equal(drv.read_image(str(tmp_path / "g.png")), np.repeat(smooth[..., :1], 3, axis=2))
    PIL.fromarray(smooth).save(tmp_path / "c.jpg", quality=95)                                # JPEG goes through Pillow
    assert drv.read_image(str(tmp_path / "c.jpg")).shape == (40, 64, 3) and drv.image_size(str(tmp_path / "c.jpg")) == (64, 40)
