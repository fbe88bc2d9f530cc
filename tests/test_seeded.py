"""Seeded mode (-p), CPU side: the restatement of pSampler::GetPriorPlaneEstimate
(oracle/upsample_fusion.c: orc_prior_planes_from_maps) against the reference's OWN function compiled
from where it lies (oracle/_ref/libacmmp_ref_host_ieee.so)."""
import numpy as np
import pytest

import oraclebind as ob
import refbind
from acmmp_b200.scene import make_scene
from seeded_inputs import encode_priors

needs_ref_host = pytest.mark.skipif(not refbind.host_available(), reason="oracle/_ref/libacmmp_ref_host*.so not built")


def scene_priors(sc, v, noise=0.01):
    cam = sc["cams"][v]
    h, w = sc["depths"][v].shape
    rng = np.random.default_rng(v)
    n = rng.normal(size=(h, w, 3)) * 0.2 + np.array([0.0, 0.0, -1.0])
    n[::7, ::5] *= -1.0                           # some normals face away: the sampler flips them
    return encode_priors(cam, sc["depths"][v], n, noise=noise, seed=v)


@needs_ref_host
@pytest.mark.parametrize("scale", [1, 2])
def test_prior_planes_restatement_is_bit_exact_against_reference(scale, tmp_path):
    sc = make_scene(3, 160, 120, 2, seed=5)
    dense = str(tmp_path / "dense")
    for v in range(3):
        d16, n16 = scene_priors(sc, v)
        refbind.write_prior_images(dense, v, d16, n16)
    for v in range(3):
        cam = sc["cams"][v].copy()
        rows, cols = 120 // scale, 160 // scale
        if scale > 1:                            # the camera of the coarser level
            cam.K[0] /= scale; cam.K[2] /= scale; cam.K[4] /= scale; cam.K[5] /= scale
            cam.width, cam.height = cols, rows
        d16, n16 = scene_priors(sc, v)
        ref = refbind.host_prior_planes(dense, 3, v, cam, rows, cols)
        mine = ob.prior_planes_from_maps(cam, d16, n16, rows, cols)
        assert (ref.view(np.uint32) == mine.view(np.uint32)).all()
        # planes face the camera and reproduce the coded depth
        assert (mine[..., 2] < 0).mean() > 0.99
        ys, xs = np.mgrid[0:rows, 0:cols]
        K = np.array(cam.K[:], np.float64)
        depth = -mine[..., 3] / (mine[..., 0] * (xs - K[2]) / K[0] + mine[..., 1] * (ys - K[5]) / K[4] + mine[..., 2])
        coded = d16[::scale, ::scale][:rows, :cols] * ((cam.depth_max - cam.depth_min) / 65535.0) + cam.depth_min
        assert np.abs(depth - coded).max() <= 1e-4 * coded.max()
