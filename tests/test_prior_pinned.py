"""oracle/prior.c pinned on the REFERENCE'S OWN planar-prior functions - ACMMP::GetSupportPoints,
GetPriorPlaneParams, GetDepthFromPlaneParam (ACMMP.cpp:868-958), compiled from where they lie into
oracle/_ref/libacmmp_ref_full.so (oracle/ref_full_harness.cu) - on seeded inputs.  Host code only: no GPU needed.
(The raster + range check of ProcessProblem needs a run of the reference program: tests/test_gpu_prior_pinned.py.)"""
import numpy as np
import pytest

import oraclebind as ob
import refbind
from acmmp_b200 import host, lib as L

pytestmark = pytest.mark.skipif(not refbind.full_available(), reason="oracle/_ref/libacmmp_ref_full.so not built (needs /root/reference)")


def _camera(w, h):
    K = [[2892.0 * w / 1600.0, 0, w / 2.0 - 3.25], [0, 2880.0 * w / 1600.0, h / 2.0 + 1.5], [0, 0, 1]]
    return L.Camera.make(K, np.eye(3), [0, 0, 0], w, h, 425.0, 935.0)


@pytest.mark.parametrize("size", [(96, 72), (323, 241), (640, 480)])      # 323 x 241: ragged last cells
def test_support_points_equal_the_reference(size):
    w, h = size
    rng = np.random.default_rng(w)
    cost = rng.uniform(0.0, 0.4, (h, w)).astype(np.float32)
    cost[rng.random((h, w)) < 0.05] = 2.0                               # failed pixels
    cost[: h // 4] += 0.2                                               # cells without a pixel under 0.1
    cost[5:10, 5:10] = 0.05                                             # ties inside a cell: first in scan order wins
    mine = ob.support_points(cost)
    ref = refbind.full_support_points(_camera(w, h), cost)
    assert len(ref) > 50 and mine.shape == ref.shape
    assert (mine == ref).all()


def test_plane_fit_and_depth_equal_the_reference():
    w, h = 320, 240
    cam = _camera(w, h)
    rng = np.random.default_rng(7)
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float32)
    depth = (600.0 + 0.3 * xx - 0.2 * yy + 5.0 * np.sin(xx / 17.0) + rng.normal(0, 0.5, (h, w))).astype(np.float32)
    pts = np.unique(np.stack([rng.integers(0, w, 700), rng.integers(0, h, 700)], -1), axis=0).astype(np.int32)
    tris = host.DelaunayTriangulation(pts).reshape(-1, 6)
    assert len(tris) > 1000
    ref_planes = refbind.full_prior_planes(cam, depth, tris)
    _, _, mine = ob.prior_from_triangles(cam, depth, tris, 100.0, 2000.0)
    # same plane: unit normal and offset (the reference's cv::SVD::solveZ is the stand-in's null vector here; both
    # normalise by |n| afterwards, so only float rounding of the last steps may differ)
    assert np.abs(mine - ref_planes).max() <= 2e-4 * np.abs(ref_planes).max()
    assert (np.abs(mine[:, :3] - ref_planes[:, :3]).max(1) <= 2e-6).mean() >= 0.99
    assert (ref_planes[:, 3] > 0).all() and (mine[:, 3] > 0).all()       # sign rule, ACMMP.cpp:943-950
    # GetDepthFromPlaneParam at the triangle's own vertices gives back the depth map
    xy = tris[:, :2]
    d_ref = refbind.full_depths_from_planes(cam, ref_planes, xy)
    K = np.array(cam.K, np.float32)
    pl = ref_planes
    d_mine = (-pl[:, 3] * K[0] / ((xy[:, 0] - K[2]) * pl[:, 0] + (K[0] / K[4]) * (xy[:, 1] - K[5]) * pl[:, 1] + K[0] * pl[:, 2])).astype(np.float32)
    assert np.abs(d_mine - d_ref).max() <= 1e-3
    assert np.abs(d_ref - depth[xy[:, 1], xy[:, 0]]).max() <= 0.05
