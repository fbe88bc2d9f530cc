"""Box staging of the sweep kernel (north_star: source patches staged in shared memory via TMA tiles): on R8 problems a
64 x 32-byte box per (tile, source view) is loaded by cp.async.bulk.tensor and half of the covered taps are filtered from
it with the texture unit's weights.  The results must not depend on the setting - bit for bit - in every mode
(photometric, planar prior, geometric consistency, multi-scale scan).  pytest -m gpu."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
if not torch.cuda.is_available():
    pytest.skip("no CUDA device", allow_module_level=True)

from acmmp_b200 import host  # noqa: E402
from acmmp_b200 import lib as L  # noqa: E402
from acmmp_b200.scene import make_scene  # noqa: E402


def same(a, b):
    return a.shape == b.shape and (np.ascontiguousarray(a).view(np.uint32) == np.ascontiguousarray(b).view(np.uint32)).all()


def one_view(sc, staging, geom=False, prior=False, seed=21):
    n = len(sc["cams"])
    order = [0] + list(sc["pairs"][0])
    imgs = [np.clip(np.rint(sc["images"][i]), 0, 255).astype(np.float32) for i in order]
    cams = [sc["cams"][i] for i in order]
    m = host.ACMMP()
    m.SetTextureFormat(L.TEX_AUTO)
    m.SetBoxStaging(staging)
    if geom:
        m.SetGeomConsistencyParams(False)
    m.InputInitialization(imgs, cams, depths=[sc["depths"][i] for i in order] if geom else None)
    assert m.TextureIsU8()
    if geom:
        h, w = imgs[0].shape
        nrm = np.zeros((h, w, 3), np.float32)
        nrm[..., 2] = -1.0
        m.CudaSpaceInitialization(warm_start=(sc["depths"][0] * 1.01, nrm, np.full((h, w), 0.3, np.float32)))
    m.RunPatchMatch(seed)
    if prior:
        m.SetPlanarPriorParams()
        m.BuildPlanarPrior()
        m.RunPatchMatch(seed + 1)
    out = (m.GetPlaneHypotheses().copy(), m.GetCosts().copy())
    m.close()
    return out


@pytest.mark.parametrize("mode", ["photometric", "prior", "geom"])
def test_single_view_results_do_not_depend_on_box_staging(mode):
    sc = make_scene(5, 640, 480, 4, seed=2)
    kw = dict(geom=mode == "geom", prior=mode == "prior")
    off = one_view(sc, False, **kw)
    on = one_view(sc, True, **kw)
    assert same(off[0], on[0]) and same(off[1], on[1])
    assert (off[1] < 0.3).mean() > 0.5                      # a converged result, not a trivial one


def test_odd_sizes_fall_back_to_the_texture_unit():
    # 650 is no multiple of 16: no tensor map over the layer, the kernel without staging runs; a 64-row image: every box is clamped to the border
    for (w, h) in ((650, 480), (640, 64)):
        sc = make_scene(4, w, h, 3, seed=5)
        off = one_view(sc, False)
        on = one_view(sc, True)
        assert same(off[0], on[0]) and same(off[1], on[1])


def test_scan_results_do_not_depend_on_box_staging():
    sc = make_scene(5, 1024, 768, 4, seed=3, device="cuda", span_deg=(10.0, 8.0))
    outs = []
    for staging in (False, True):
        scan = host.Scan(sc["cams"], sc["pairs"])
        scan.set_box_staging(staging)
        for v, im in enumerate(sc["images"]):
            scan.set_image_u8(v, np.clip(np.rint(im), 0, 255).astype(np.uint8))
        for level in range(scan.num_levels()):
            scan.begin_level(level)
            for stage in range(3):
                scan.run_stage(stage, 4)
        assert scan.level_is_u8()
        outs.append([scan.get_view(v) for v in range(5)])
        scan.close()
    for v in range(5):
        for k in range(3):
            assert same(outs[0][v][k], outs[1][v][k]), (v, k)
