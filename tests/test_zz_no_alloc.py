"""acmmp_prior_raster_noalloc (include/acmmp_b200.h; no counterpart in the reference):
the planar-prior raster with its scratch in a static __device__ array, for callers that cannot
cudaMalloc.  Called through the C ABI on device pointers and checked bit for bit against the existing entry point
(acmmp_build_planar_prior + acmmp_get_planar_prior), on two calls in a row whose sizes differ, so that the second
call meets the first call's leftovers in the shared scratch.  Collected last on purpose (the file name).
pytest -m gpu."""
import ctypes as C

import numpy as np
import pytest

import oraclebind as ob

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
if not torch.cuda.is_available():
    pytest.skip("no CUDA device", allow_module_level=True)

from acmmp_b200 import host, lib as L  # noqa: E402
from acmmp_b200.scene import make_scene  # noqa: E402


def _existing_entry_point(w, h, seed):
    """acmmp_build_planar_prior on a finished run: its prior planes / labels, and the inputs the raster saw"""
    sc = make_scene(5, w, h, 4, seed=seed)
    ids = [0] + sc["pairs"][0]
    m = host.ACMMP(0)
    m.InputInitialization([sc["images"][i] for i in ids], [sc["cams"][i] for i in ids])
    m.RunPatchMatch(11 + seed)
    planes, cost, _, _ = m.get_state(with_rng=False)
    ns, nt = m.BuildPlanarPrior()
    prior, mask = m.GetPlanarPrior()
    p = m.params()
    m.close()
    pts = ob.support_points(cost)                         # GetSupportPoints (ACMMP.cpp:868-894), column-major cell order
    tris = host.DelaunayTriangulation(pts).reshape(-1, 6)
    assert ns == len(pts) and nt == len(tris) and nt > 100
    return sc["cams"][0], p.depth_min, p.depth_max, planes, tris, prior, mask


def _no_alloc(cam, dmin, dmax, planes, tris, stream=None):
    lib = L.load()
    dev = torch.device("cuda", 0)
    d_planes = torch.from_numpy(np.ascontiguousarray(planes, np.float32)).to(dev)
    d_tris = torch.from_numpy(np.ascontiguousarray(tris, np.int32)).to(dev)
    h, w = planes.shape[:2]
    d_prior = torch.full((h, w, 4), float("nan"), dtype=torch.float32, device=dev)
    d_mask = torch.full((h, w), 0x7FFFFFFF, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    err = C.create_string_buffer(256)
    c = cam.copy()
    rc = lib.acmmp_prior_raster_noalloc(0, C.byref(c), dmin, dmax, C.c_void_p(d_planes.data_ptr()), C.c_void_p(d_tris.data_ptr()), len(tris),
                                        C.c_void_p(d_prior.data_ptr()), C.c_void_p(d_mask.data_ptr()), C.c_void_p(stream), err, len(err))
    assert rc == 0, err.value.decode()
    torch.cuda.synchronize()
    return d_prior.cpu().numpy(), d_mask.cpu().numpy().view(np.uint32)


def test_no_alloc_prior_raster_equals_build_planar_prior_twice_in_a_row():
    big = _existing_entry_point(320, 240, seed=2)
    small = _existing_entry_point(160, 120, seed=3)
    for cam, dmin, dmax, planes, tris, prior, mask in (big, small, big):   # the smaller problem meets the larger one's leftovers
        got_prior, got_mask = _no_alloc(cam, dmin, dmax, planes, tris)
        assert (got_mask == mask).all(), float((got_mask == mask).mean())
        assert (got_prior.view(np.uint32) == prior.view(np.uint32)).all()
        assert 0.3 < (mask > 0).mean() <= 1.0


def test_no_alloc_prior_raster_rejects_what_does_not_fit():
    lib = L.load()
    cam = make_scene(2, 64, 48, 1, seed=0)["cams"][0].copy()
    err = C.create_string_buffer(256)
    dummy = torch.zeros(16, device="cuda")
    ptr = C.c_void_p(dummy.data_ptr())
    cam.width, cam.height = 3200, 2400                    # 7.68 M pixels > ACMMP_NOALLOC_MAX_PIXELS
    assert lib.acmmp_prior_raster_noalloc(0, C.byref(cam), 1.0, 2.0, ptr, ptr, 10, ptr, ptr, None, err, len(err)) == -1
    assert b"ACMMP_NOALLOC_MAX_PIXELS" in err.value
    cam.width, cam.height = 64, 48
    assert lib.acmmp_prior_raster_noalloc(0, C.byref(cam), 1.0, 2.0, ptr, ptr, 153601, ptr, ptr, None, err, len(err)) == -1
    assert b"ACMMP_NOALLOC_MAX_TRIANGLES" in err.value
    assert lib.acmmp_prior_raster_noalloc(0, C.byref(cam), 1.0, 2.0, None, ptr, 1, ptr, ptr, None, None, 0) == -1
