"""Loads tests/golden/ref_golden.npz (outputs of the reference's own CUDA kernels,
see tests/golden/make_ref_golden.py)."""
import os

import numpy as np

from acmmp_b200 import lib as L

PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_golden.npz")


def load():
    g = np.load(PATH)
    cams = [L.Camera.make(r[0:9], r[9:18], r[18:21], int(r[22]), int(r[21]), r[23], r[24]) for r in g["cams"]]
    images = [x.astype(np.float32) for x in g["images_u8"]]
    return g, images, cams


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def depth_normal_ok(pa, pb):
    """fractions of pixels with depth within 0.1 % and normal within 0.5 degrees."""
    da, db = pa[..., 3].astype(np.float64), pb[..., 3].astype(np.float64)
    rel = np.abs(da - db) / np.maximum(np.abs(db), 1e-12)
    na, nb = pa[..., :3].astype(np.float64), pb[..., :3].astype(np.float64)
    na = na / np.maximum(np.linalg.norm(na, axis=-1, keepdims=True), 1e-12)
    nb = nb / np.maximum(np.linalg.norm(nb, axis=-1, keepdims=True), 1e-12)
    ang = np.degrees(np.arccos(np.clip((na * nb).sum(-1), -1, 1)))
    return rel <= 1e-3, ang <= 0.5
