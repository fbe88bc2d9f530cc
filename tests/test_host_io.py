"""C++ host I/O (acmmp_b200/host/io.hpp) on the reference's file formats: cam text, pair.txt,
.dmb, PGM/PPM, binary PLY - through build/io_selftest.  CPU only."""
import os
import subprocess

import numpy as np
import pytest

import refbind
from acmmp_b200 import lib as L

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "build", "io_selftest")

pytestmark = pytest.mark.skipif(not os.path.exists(EXE), reason="build/io_selftest not built (make cli)")


def test_formats_round_trip(tmp_path):
    d = str(tmp_path)
    rng = np.random.default_rng(0)
    with open(os.path.join(d, "pair.txt"), "w") as f:          # select_dtu_cams.py:28-35 layout
        f.write("3\n0\n3 1 2.5 2 1.0 7 0.0\n1\n2 0 1.0 2 1.0\n2\n1 0 -1\n")
    os.makedirs(os.path.join(d, "cams"))
    K = [2892.33, 0, 823.205, 0, 2883.18, 619.07, 0, 0, 1]
    R = np.linalg.qr(rng.normal(size=(3, 3)))[0].astype(np.float32).reshape(9)
    cam = L.Camera.make(K, R, [1.5, -2.25, 600.125], 1600, 1200, 425.0, 935.0)
    refbind._write_cam(os.path.join(d, "cams", "00000000_cam.txt"), cam)
    m = rng.normal(size=(7, 5, 3)).astype(np.float32)
    refbind._write_dmb(os.path.join(d, "in.dmb"), m)
    img = rng.integers(0, 256, size=(6, 9, 3), dtype=np.uint8)   # RGB in the PPM
    with open(os.path.join(d, "img.ppm"), "wb") as f:
        f.write(b"P6\n9 6\n255\n" + img.tobytes())
    pts = rng.normal(size=(11, 9)).astype(np.float32)
    pts[:, 6:] = rng.integers(0, 256, size=(11, 3))
    pts[3, 0] = np.inf
    pts.tofile(os.path.join(d, "points.bin"))
    out = subprocess.run([EXE, d], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    kv = dict(l.split("=", 1) for l in out.stdout.splitlines() if "=" in l and not l.startswith("ref="))
    refs = [l for l in out.stdout.splitlines() if l.startswith("ref=")]
    assert kv["problems"] == "3"
    assert refs == ["ref=0 src=1,2,", "ref=1 src=0,2,", "ref=2 src="]           # scores <= 0 dropped
    assert np.array_equal(np.float32([float(x) for x in kv["K"].strip(",").split(",")]), np.float32(K))
    assert np.array_equal(np.float32([float(x) for x in kv["R"].strip(",").split(",")]), R)
    assert kv["t"] == "1.5,-2.25,600.125" and kv["depth"] == "425,935"
    assert kv["dmb"] == "7,5,3"
    assert open(os.path.join(d, "out.dmb"), "rb").read() == open(os.path.join(d, "in.dmb"), "rb").read()
    assert kv["image"] == "6,9,3"
    gray = np.fromfile(os.path.join(d, "gray.dmb"), np.float32, offset=16).reshape(6, 9)
    b, g, r = img[..., 2].astype(np.int64), img[..., 1].astype(np.int64), img[..., 0].astype(np.int64)
    assert np.array_equal(gray, ((b * 1868 + g * 9617 + r * 4899 + 8192) >> 14).astype(np.float32))
    try:
        import cv2
        assert np.abs(gray - cv2.cvtColor(img, cv2.COLOR_RGB2GRAY).astype(np.float32)).max() <= 1   # OpenCV's own rounding
    except ImportError:
        pass
    blob = open(os.path.join(d, "out.ply"), "rb").read()
    end = blob.index(b"end_header\n") + 11
    assert b"element vertex 11\n" in blob[:end] and b"format binary_little_endian 1.0" in blob[:end]
    rec = np.frombuffer(blob[end:], dtype=np.dtype([("f", "<f4", 6), ("c", "u1", 3)]))
    assert len(rec) == 11
    exp = pts.copy()
    exp[3, :3] = 0                                                     # non-finite coordinates are zeroed (ACMMP.cpp:414-418)
    assert np.array_equal(rec["f"], exp[:, :6])
    assert np.array_equal(rec["c"], exp[:, [8, 7, 6]].astype(np.uint8))  # red, green, blue = c2, c1, c0


def test_png_reader_against_cv2(tmp_path):
    """ReadPng16 (acmmp_b200/host/io.hpp; zlib) returns what cv2.imread(path, -1) returns: 8/16-bit,
    grey / RGB / RGBA, every filter type the encoder picks - the 16-bit prior images of the seeded mode."""
    cv2 = pytest.importorskip("cv2")
    exe = os.path.join(ROOT, "build", "io_selftest")
    if not os.path.exists(exe):
        pytest.skip("build/io_selftest not built (make cli)")
    rng = np.random.default_rng(3)
    ramp = (np.add.outer(np.arange(37), np.arange(53)) * 400).astype(np.uint16)           # smooth: sub / up / paeth filters
    cases = {
        "g16": rng.integers(0, 65536, (37, 53)).astype(np.uint16),
        "ramp16": ramp,
        "c16": np.stack([ramp, ramp[::-1], rng.integers(0, 65536, (37, 53)).astype(np.uint16)], -1),
        "g8": rng.integers(0, 256, (20, 31)).astype(np.uint8),
        "c8": (np.stack([ramp, ramp.T[:37, :53] if False else ramp // 3, ramp // 7], -1) >> 8).astype(np.uint8),
        "a16": rng.integers(0, 65536, (9, 11, 4)).astype(np.uint16),
    }
    for name, img in cases.items():
        path = str(tmp_path / (name + ".png"))
        assert cv2.imwrite(path, img)
        back = cv2.imread(path, -1)
        assert (back == img).all()
        out = str(tmp_path / (name + ".bin"))
        r = subprocess.run([exe, "--png", path, out], capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
        rows, cols, cn, depth = [int(x) for x in r.stdout.strip().split("=")[1].split(",")]
        assert (rows, cols) == img.shape[:2] and cn == (1 if img.ndim == 2 else img.shape[2]) and depth == 8 * img.dtype.itemsize
        mine = np.fromfile(out, np.uint16).reshape(img.shape)
        assert (mine == img).all(), name
    bad = str(tmp_path / "bad.png")
    open(bad, "wb").write(b"not a png")
    assert subprocess.run([exe, "--png", bad, str(tmp_path / "x.bin")], capture_output=True).returncode == 12


@pytest.mark.parametrize("shape,dst", [((48, 64, 1), (32, 24)), ((48, 64, 3), (32, 24)), ((37, 53, 3), (80, 61)), ((37, 53, 1), (19, 11)),
                                       ((30, 40, 1), (40, 30)), ((96, 128, 3), (100, 75))])
def test_resize_linear_8bit_equals_cv2(tmp_path, shape, dst):
    """ResizeLinear8 (io.hpp) = cv::resize(INTER_LINEAR) on 8-bit images, incl. the 2 x 2 INTER_AREA fast path:
    what RunFusion applies to colour images and masks whose size differs from the depth map's
    (ACMMP.cpp:179-203, acmmp_definitions.cpp:897)."""
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(sum(shape) + dst[0])
    img = rng.integers(0, 256, size=shape, dtype=np.uint8)
    path = str(tmp_path / ("in.ppm" if shape[2] == 3 else "in.pgm"))
    with open(path, "wb") as f:
        f.write((b"P6" if shape[2] == 3 else b"P5") + b"\n%d %d\n255\n" % (shape[1], shape[0]) + img.tobytes())
    out = str(tmp_path / "out.bin")
    r = subprocess.run([EXE, "--resize8", path, out, str(dst[0]), str(dst[1])], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    got = np.fromfile(out, np.uint8).reshape(dst[1], dst[0], shape[2])
    src = img[..., ::-1] if shape[2] == 3 else img[..., 0]          # the reader keeps BGR in memory
    want = cv2.resize(np.ascontiguousarray(src), dst, interpolation=cv2.INTER_LINEAR).reshape(dst[1], dst[0], shape[2])
    assert np.array_equal(got, want), int(np.abs(got.astype(int) - want.astype(int)).max())


def test_dmb_reader_rejects_bad_headers(tmp_path):
    def write(name, hdr, payload):
        p = str(tmp_path / name)
        with open(p, "wb") as f:
            f.write(np.asarray(hdr, np.int32).tobytes() + payload)
        return p
    ok1 = write("d.dmb", [1, 3, 4, 1], np.zeros(12, np.float32).tobytes())
    ok3 = write("n.dmb", [1, 3, 4, 3], np.zeros(36, np.float32).tobytes())
    run = lambda p, kind: subprocess.run([EXE, "--dmb", p, kind], capture_output=True, text=True).returncode  # noqa: E731
    assert run(ok1, "depth") == 0 and run(ok3, "normal") == 0 and run(ok1, "any") == 0
    assert run(ok3, "depth") != 0 and run(ok1, "normal") != 0            # channel count is part of the contract
    assert run(write("neg.dmb", [1, -3, 4, 1], b""), "any") != 0
    assert run(write("huge.dmb", [1, 2 ** 30, 2 ** 30, 1], b""), "any") != 0
    assert run(write("ch.dmb", [1, 3, 4, 2], np.zeros(24, np.float32).tobytes()), "any") != 0
    assert run(write("type.dmb", [2, 3, 4, 1], np.zeros(12, np.float32).tobytes()), "any") != 0
    assert run(write("short.dmb", [1, 3, 4, 1], np.zeros(5, np.float32).tobytes()), "any") != 0


def test_decoded_image_content_cache(tmp_path):
    """ImageCache (host/io.hpp): one decode per file, validated by size + mtime, LRU under a byte budget"""
    a, b = tmp_path / "a.pgm", tmp_path / "b.pgm"
    rng = np.random.default_rng(0)
    for path, (h, w) in ((a, (30, 40)), (b, (20, 50))):
        with open(path, "wb") as f:
            f.write(b"P5\n%d %d\n255\n" % (w, h) + rng.integers(0, 256, (h, w), dtype=np.uint8).tobytes())
    out = subprocess.run([EXE, "--cache", str(a), str(b)], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    lines = dict(l.split("=", 1) for l in out.stdout.strip().splitlines())
    assert lines["after_three_reads"] == "1,2,%d same=1" % (30 * 40 + 20 * 50)       # second read of a: a hit
    assert lines["after_rewrite"] == "1,3,%d first=1,2,7" % (2 + 20 * 50)             # changed file: decoded again
    assert lines["after_budget"] == "1,2"                                             # the least recently used entry (b) was dropped
    assert lines["final"] == "1,4"                                                    # ... so b is decoded again
