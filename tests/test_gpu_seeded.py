"""Seeded mode on the device (the reference's `-p`): the conversion of the 16-bit
prior images into plane hypotheses (acmmp_prior_planes_from_maps, csrc/seeded.cu) against the oracle -
itself bit-exact against the reference's own pSampler (tests/test_seeded.py) - and the scan schedule
started from them (acmmp_scan_set_prior_maps) against the same steps made one by one through the
operator API.  pytest -m gpu."""
import numpy as np
import pytest

import oraclebind as ob

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
if not torch.cuda.is_available():
    pytest.skip("no CUDA device", allow_module_level=True)

from acmmp_b200 import host  # noqa: E402
from acmmp_b200.scene import make_scene  # noqa: E402
from seeded_inputs import encode_priors  # noqa: E402
from test_seeded import scene_priors  # noqa: E402

MASK64 = (1 << 64) - 1


@pytest.mark.parametrize("scale", [1, 2, 3])
def test_device_planes_equal_oracle(scale):
    sc = make_scene(3, 192, 144, 2, seed=5)
    for v in range(3):
        cam = sc["cams"][v].copy()
        rows, cols = 144 // scale, 192 // scale
        if scale > 1:
            cam.K[0] /= scale; cam.K[2] /= scale; cam.K[4] /= scale; cam.K[5] /= scale
            cam.width, cam.height = cols, rows
        d16, n16 = scene_priors(sc, v)
        mine = host.prior_planes_from_maps(cam, d16, n16, rows, cols)
        ref = ob.prior_planes_from_maps(cam, d16, n16, rows, cols)
        assert (mine.view(np.uint32) == ref.view(np.uint32)).all()
    with pytest.raises(Exception):
        host.prior_planes_from_maps(cam, d16[:10], n16[:10], rows, cols)      # prior images smaller than the level


def camera_frame_normals(sc, v):
    """the scene's planes are fronto-parallel enough: normals towards the camera"""
    h, w = sc["depths"][v].shape
    n = np.zeros((h, w, 3))
    n[..., 2] = -1.0
    return n


def test_scan_seeded_stage_equals_operator_api_steps():
    """level 0, photometric stage with priors: scan result == InputInitialization -> SetPlanarPrior(oracle
    planes) -> RunPatchMatch -> planar prior -> RunPatchMatch on one ACMMP object, same seeds - bit for bit."""
    sc = make_scene(5, 320, 240, 4, seed=6)
    seed = 41
    priors = [encode_priors(sc["cams"][v], sc["depths"][v], camera_frame_normals(sc, v), noise=0.01, seed=v) for v in range(5)]
    s = host.Scan(sc["cams"], sc["pairs"])
    for v, im in enumerate(sc["images"]):
        s.set_image(v, im)
    for v in (0, 2, 3):                                   # views 1 and 4 stay unseeded
        s.set_prior_maps(v, *priors[v])
    assert s.num_levels() == 1
    s.begin_level(0)
    s.run_stage(host.STAGE_PHOTO, seed)
    s.sync()
    got = {v: s.get_view(v) for v in range(5)}
    s.close()
    for v in (0, 1, 3):
        ids = [v] + list(sc["pairs"][v])
        m = host.ACMMP()
        m.InputInitialization([sc["images"][i] for i in ids], [sc["cams"][i] for i in ids])
        if v != 1:
            m.SetPlanarPrior(ob.prior_planes_from_maps(sc["cams"][v], priors[v][0], priors[v][1], 240, 320))
        sv = (seed * 1000003 + v * 8191) & MASK64
        m.RunPatchMatch(sv)
        m.SetPlanarPriorParams()
        m.BuildPlanarPrior()
        m.RunPatchMatch(sv + 1)
        depth, normal, cost = m.GetMaps()
        m.close()
        assert (depth.view(np.uint32) == got[v][0].view(np.uint32)).all(), v
        assert (normal.view(np.uint32) == got[v][1].view(np.uint32)).all(), v
        assert (cost.view(np.uint32) == got[v][2].view(np.uint32)).all(), v


def test_seeds_steer_the_result():
    """good priors give a better photometric stage than random initialisation; removing them restores it"""
    sc = make_scene(5, 320, 240, 4, seed=6)

    def stage(prior_views, wrong=False):
        s = host.Scan(sc["cams"], sc["pairs"])
        for v, im in enumerate(sc["images"]):
            s.set_image(v, im)
        for v in prior_views:
            d = sc["depths"][v] * (1.3 if wrong else 1.0)
            s.set_prior_maps(v, *encode_priors(sc["cams"][v], d, camera_frame_normals(sc, v)))
        if wrong:
            for v in prior_views:
                s.set_prior_maps(v)                      # taken back: the unseeded schedule again
        s.begin_level(0)
        s.run_stage(host.STAGE_PHOTO, 7)
        s.sync()
        out = [s.get_view(v)[0] for v in range(5)]
        s.close()
        return out

    plain = stage([])
    seeded = stage(range(5))
    removed = stage(range(5), wrong=True)
    err = lambda maps: np.mean([np.mean(np.abs(m - g)[8:-8, 8:-8] / g[8:-8, 8:-8] < 0.002) for m, g in zip(maps, sc["depths"])])  # noqa: E731
    assert err(seeded) >= err(plain) and err(seeded) > 0.9
    assert all((a.view(np.uint32) == b.view(np.uint32)).all() for a, b in zip(plain, removed))
