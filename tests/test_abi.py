"""The C-ABI library loads on a CPU-only machine and exports every symbol that
include/acmmp_b200.h declares; context-less host entry points work without a GPU;
compute entry points fail loudly (no CPU fallback)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from acmmp_b200 import host, lib as L

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "acmmp_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(acmmp_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    handle = L.load()
    names = declared_symbols()
    assert len(names) >= 50
    missing = [n for n in names if not hasattr(handle, n)]
    assert not missing, missing
    assert sorted(set(L.EXPORTS)) == names, set(names) ^ set(L.EXPORTS)


def test_struct_layouts():
    assert C.sizeof(L.Camera) == 100          # reference struct Camera, acmmp_definitions.h:47-55
    p = L.Params()
    L.load().acmmp_default_params(C.byref(p))
    assert (p.max_iterations, p.patch_size, p.radius_increment, p.top_k, p.max_image_size) == (2, 11, 2, 4, 3200)
    assert (p.sigma_spatial, p.sigma_color) == (5.0, 3.0)
    assert abs(p.baseline - 0.54) < 1e-7
    assert L.load().acmmp_version().decode().endswith("sm_100a")


def test_no_device_means_loud_failure_not_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(L.AcmmpError):
        host.ACMMP(0)
    img = np.zeros((8, 8), np.float32)
    with pytest.raises(L.AcmmpError):
        host.resize_linear(img, 4, 4)


def test_delaunay_is_host_only_and_correct():
    from scipy.spatial import Delaunay
    rs = np.random.default_rng(3)
    # general position: compare triangle sets with qhull
    pts = np.unique(rs.integers(0, 4000, (600, 2)), axis=0)
    tri = host.DelaunayTriangulation(pts)
    ref = Delaunay(pts)
    a = {tuple(sorted(map(tuple, t))) for t in tri.tolist()}
    b = {tuple(sorted(map(tuple, t))) for t in pts[ref.simplices].tolist()}
    assert len(a) == len(b)
    assert len(a & b) >= 0.99 * len(b)
    # degenerate input (quasi-grid with co-circular points): same count, same area, empty circumcircles
    gx, gy = np.meshgrid(np.arange(0, 200, 5), np.arange(0, 150, 5), indexing="ij")
    pts = np.stack([gx.ravel() + rs.integers(0, 2, gx.size), gy.ravel()], 1)
    tri = host.DelaunayTriangulation(pts)
    ref = Delaunay(pts)
    assert len(tri) == len(ref.simplices)

    def area(t):
        t = np.asarray(t, np.float64)
        return 0.5 * np.abs((t[:, 1, 0] - t[:, 0, 0]) * (t[:, 2, 1] - t[:, 0, 1]) - (t[:, 1, 1] - t[:, 0, 1]) * (t[:, 2, 0] - t[:, 0, 0]))
    assert area(tri).sum() == area(pts[ref.simplices]).sum()
    assert (area(tri) > 0).all()
    # Delaunay property with exact integers: no point strictly inside a circumcircle
    P = pts.astype(object)
    for t in tri[:: max(1, len(tri) // 60)]:
        (ax, ay), (bx, by), (cx, cy) = [(int(x), int(y)) for x, y in t]
        orient = (bx - ax) * (cy - ay) - (by - ay) * (cx - ax)
        for (dx, dy) in pts[:: 7]:
            dx, dy = int(dx), int(dy)
            a0, a1 = ax - dx, ay - dy
            b0, b1 = bx - dx, by - dy
            c0, c1 = cx - dx, cy - dy
            det = (a0 * a0 + a1 * a1) * (b0 * c1 - b1 * c0) - (b0 * b0 + b1 * b1) * (a0 * c1 - a1 * c0) + (c0 * c0 + c1 * c1) * (a0 * b1 - a1 * b0)
            assert det * (1 if orient > 0 else -1) <= 0


def test_delaunay_edge_cases():
    assert len(host.DelaunayTriangulation(np.zeros((0, 2), np.int32))) == 0
    assert len(host.DelaunayTriangulation([[0, 0], [5, 5]])) == 0
    assert len(host.DelaunayTriangulation([[0, 0], [5, 5], [10, 10], [15, 15]])) == 0     # collinear
    assert len(host.DelaunayTriangulation([[0, 0], [5, 0], [0, 5], [0, 0]])) == 1          # duplicate dropped
    assert len(host.DelaunayTriangulation([[0, 0], [5, 0], [0, 5], [5, 5]])) == 2          # co-circular square


def test_delaunay_does_not_depend_on_the_thread_count():
    """the top levels of the divide-and-conquer recursion run on separate threads, each half in its own edge
    arena; appended arenas reproduce the sequential edge numbering, so triangles AND their order (which decides
    the raster overwrite, acmmp_definitions.cpp:346-352) are identical for any thread count."""
    lib = L.load()
    rng = np.random.default_rng(3)
    cases = []
    # support points as the planar prior makes them: one per 5 x 5 cell, ~15 % of the cells empty
    for w, h in ((1600, 1200), (640, 480)):
        cy, cx = np.mgrid[0:h // 5, 0:w // 5]
        keep = rng.random(cx.shape) > 0.15
        pts = np.stack([cx[keep] * 5 + rng.integers(0, 5, keep.sum()), cy[keep] * 5 + rng.integers(0, 5, keep.sum())], -1)
        cases.append(pts[rng.permutation(len(pts))])
    cases.append(np.stack(np.meshgrid(np.arange(0, 400, 5), np.arange(0, 300, 5)), -1).reshape(-1, 2))   # exact grid: all co-circular
    cases.append(rng.integers(0, 3000, (20000, 2)))
    try:
        for pts in cases:
            assert lib.acmmp_set_delaunay_threads(1) == 0
            a = host.DelaunayTriangulation(pts)
            for t in (2, 4, 8, 64):
                assert lib.acmmp_set_delaunay_threads(t) == 0
                b = host.DelaunayTriangulation(pts)
                assert a.shape == b.shape and (a == b).all(), (len(pts), t)
    finally:
        lib.acmmp_set_delaunay_threads(0)
    assert lib.acmmp_set_delaunay_threads(-1) != 0
