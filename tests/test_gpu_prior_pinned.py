"""The planar prior that the REFERENCE PROGRAM builds - triangle raster, plane fit, range check inside ProcessProblem
(acmmp_definitions.cpp:306-378), captured at its call of CudaPlanarPriorInitialization by oracle/ref_full_harness.cu -
against (a) the CPU restatement oracle/prior.c and (b) the device construction, fed with the depth map and triangles
the reference used.  pytest -m gpu."""
import ctypes as C

import numpy as np
import pytest

import oraclebind as ob

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
if not torch.cuda.is_available():
    pytest.skip("no CUDA device", allow_module_level=True)

import refbind  # noqa: E402
from acmmp_b200 import lib as L  # noqa: E402
from acmmp_b200.scene import make_scene  # noqa: E402

needs_full = pytest.mark.skipif(not refbind.full_available(), reason="oracle/_ref/libacmmp_ref_full.so not built")


def _device_raster(cam, dmin, dmax, depth, tris):
    lib = L.load()
    h, w = depth.shape
    planes = np.zeros((h, w, 4), np.float32)
    planes[..., 3] = depth
    d_planes = torch.from_numpy(planes).cuda()
    d_tris = torch.from_numpy(np.ascontiguousarray(tris, np.int32)).cuda()
    d_prior = torch.zeros((h, w, 4), dtype=torch.float32, device="cuda")
    d_mask = torch.zeros((h, w), dtype=torch.int32, device="cuda")
    err = C.create_string_buffer(256)
    c = cam.copy()
    rc = lib.acmmp_prior_raster_noalloc(0, C.byref(c), dmin, dmax, C.c_void_p(d_planes.data_ptr()), C.c_void_p(d_tris.data_ptr()), len(tris),
                                        C.c_void_p(d_prior.data_ptr()), C.c_void_p(d_mask.data_ptr()), None, err, len(err))
    assert rc == 0, err.value.decode()
    torch.cuda.synchronize()
    return d_prior.cpu().numpy(), d_mask.cpu().numpy().view(np.uint32)


@needs_full
@pytest.mark.parametrize("size", [(320, 240), (640, 480)])
def test_prior_of_the_reference_program_equals_oracle_and_device(tmp_path, size):
    w, h = size
    sc = make_scene(5, w, h, 4, seed=8)
    dense = str(tmp_path / "dense")
    refbind.write_dense_folder(dense, sc["cams"], sc["images"], sc["pairs"])
    cap = refbind.full_capture_prior(dense, idx=0, seed=99)
    depth, tris, ref_planes, ref_mask = cap["depth"], cap["tris"], cap["planes"], cap["mask"].astype(np.uint32)
    assert depth.shape == (h, w) and len(tris) > 500 and len(ref_planes) == len(tris)   # every triangle lies inside the image
    assert (ref_mask > 0).mean() > 0.5
    cam = sc["cams"][0]
    dmin, dmax = float(np.float32(cam.depth_min) * np.float32(0.6)), float(np.float32(cam.depth_max) * np.float32(1.2))
    covered = ref_mask > 0
    ref_prior = np.zeros((h, w, 4), np.float32)
    ref_prior[covered] = ref_planes[ref_mask[covered] - 1]               # CudaPlanarPriorInitialization's table, ACMMP.cpp:819-827
    # (a) the CPU restatement: labels identical, planes to float rounding (solveZ stand-in vs cross product)
    orc_prior, orc_mask, orc_planes = ob.prior_from_triangles(cam, depth, tris, dmin, dmax)
    assert (orc_mask == ref_mask).mean() >= 0.9999, (orc_mask == ref_mask).mean()
    scale = np.abs(ref_planes).max()
    assert np.abs(orc_planes - ref_planes).max() <= 2e-4 * scale
    both = covered & (orc_mask == ref_mask)
    assert np.abs(orc_prior[both] - ref_prior[both]).max() <= 2e-4 * scale
    # (b) the device construction on the same inputs
    dev_prior, dev_mask = _device_raster(cam, dmin, dmax, depth, tris)
    assert (dev_mask == ref_mask).mean() >= 0.9999, (dev_mask == ref_mask).mean()
    both = covered & (dev_mask == ref_mask)
    assert np.abs(dev_prior[both] - ref_prior[both]).max() <= 2e-4 * scale
    assert (np.abs(dev_prior[both][:, :3] - ref_prior[both][:, :3]).max(1) <= 2e-6).mean() >= 0.99
    # uncovered pixels carry a zero plane and label 0 everywhere
    assert (dev_prior[dev_mask == 0] == 0).all()
    print("labels equal: oracle %.6f device %.6f; covered %.3f; triangles %d" %
          ((orc_mask == ref_mask).mean(), (dev_mask == ref_mask).mean(), covered.mean(), len(tris)))
