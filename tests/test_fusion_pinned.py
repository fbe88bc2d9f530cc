"""Pins the fusion restatement (oracle/upsample_fusion.c: orc_fusion) against the reference's
OWN RunFusion (acmmp_definitions.cpp:828-1043), compiled from where it lies by `make ref`
(oracle/ref_host_harness.cpp, oracle/_ref/libacmmp_ref_host*.so) and run on the reference's
own on-disk layout.  CPU only."""
import tempfile

import numpy as np
import pytest

import oraclebind as ob
import refbind
from acmmp_b200.scene import make_scene
from fusion_inputs import noisy_maps

needs_ref_host = pytest.mark.skipif(not refbind.host_available(), reason="oracle/_ref/libacmmp_ref_host*.so not built")


def both(sc, depths, normals, colors, scalar, thresh, build):
    with tempfile.TemporaryDirectory() as d:
        ref_pts = refbind.host_fusion(d, sc["cams"], depths, normals, colors, sc["pairs"], scalar, thresh, build=build)
    cnt, pts = ob.fusion(sc["cams"], depths, normals, sc["pairs"], scalar, thresh, colors=colors, want_points=True)
    return ref_pts, cnt, pts


@needs_ref_host
@pytest.mark.parametrize("thresh,scalar", [(1, 0.3), (2, 0.3), (3, 0.1)])
def test_restatement_is_bit_exact_against_reference_runfusion_ieee_build(thresh, scalar):
    sc = make_scene(6, 160, 120, 4, seed=2)
    depths, normals, colors = noisy_maps(sc, 11)
    ref_pts, cnt, pts = both(sc, depths, normals, colors, scalar, thresh, "ieee")
    assert cnt > 1000 and len(ref_pts) == cnt
    # same points in the same order, same normals; the PLY stores the colour bytes reversed (ACMMP.cpp:411-430)
    assert (ref_pts[:, :6].view(np.uint32) == pts[:, :6].view(np.uint32)).all()
    assert (ref_pts[:, 6:] == pts[:, [8, 7, 6]]).all()


@needs_ref_host
def test_restatement_against_reference_runfusion_with_reference_flags():
    """-O3 -ffast-math (src/CMakeLists.txt:28) reassociates the float arithmetic: the accepted
    set stays the same here, coordinates move by a few ulp."""
    sc = make_scene(6, 160, 120, 4, seed=2)
    depths, normals, colors = noisy_maps(sc, 11)
    ref_pts, cnt, pts = both(sc, depths, normals, colors, 0.3, 1, "reference_flags")
    assert abs(len(ref_pts) - cnt) <= max(1, int(1e-3 * cnt))          # north_star: fused count within 0.1 %
    if len(ref_pts) == cnt:
        scale = np.abs(pts[:, :3]).max()
        assert np.abs(ref_pts[:, :3] - pts[:, :3]).max() <= 1e-6 * scale
        assert (ref_pts[:, 3:6] == pts[:, 3:6]).all()


@needs_ref_host
def test_used_list_quirk_is_exercised():
    """the per-view used_list (acmmp_definitions.cpp:936) matters on these inputs: resetting it
    per pixel (the upstream behaviour) gives a different count, so the pin covers the quirk"""
    sc = make_scene(6, 160, 120, 4, seed=2)
    depths, normals, colors = noisy_maps(sc, 11, holes=0.2)
    ref_pts, cnt, _ = both(sc, depths, normals, colors, 0.3, 2, "ieee")
    assert len(ref_pts) == cnt
    # without stale coordinates every accepted pixel masks only its own consistent sources
    from fusion_inputs import fusion_no_stale
    assert fusion_no_stale(sc["cams"], depths, normals, sc["pairs"], 0.3, 2) != cnt


# ---- prior-aware fusion ---------------------------------------------------

def second_maps(sc, seed):
    """the second (depth, normal) set of every view: another noisy run with its own holes"""
    d2, n2, _ = noisy_maps(sc, seed, noise=0.003, holes=0.15, outliers=0.05)
    return d2, n2


@needs_ref_host
@pytest.mark.parametrize("thresh,scalar,penalty", [(1, 0.3, 1), (2, 0.3, 1), (1, 0.1, 0), (2, 0.2, 2)])
def test_prior_fusion_restatement_is_bit_exact_against_reference(thresh, scalar, penalty):
    sc = make_scene(6, 160, 120, 4, seed=2)
    depths, normals, colors = noisy_maps(sc, 11, holes=0.1)
    depths2, normals2 = second_maps(sc, 12)
    with tempfile.TemporaryDirectory() as d:
        ref_pts = refbind.host_prior_fusion(d, sc["cams"], depths, normals, depths2, normals2, colors, sc["pairs"], scalar, thresh,
                                            penalty, build="ieee")
    cnt, pts = ob.prior_fusion(sc["cams"], depths, normals, depths2, normals2, sc["pairs"], scalar, thresh, penalty, colors=colors)
    assert cnt > 1000 and len(ref_pts) == cnt
    assert (ref_pts[:, :6].view(np.uint32) == pts[:, :6].view(np.uint32)).all()
    assert (ref_pts[:, 6:] == pts[:, [8, 7, 6]]).all()
    # both hypotheses win somewhere: points carry normals of set 0 and of set 1
    key = lambda a: set(map(bytes, np.ascontiguousarray(a).view(np.uint8).reshape(len(a), -1)))  # noqa: E731
    n0 = key(np.concatenate([x.reshape(-1, 3) for x in normals]))
    n1 = key(np.concatenate([x.reshape(-1, 3) for x in normals2]))
    got = key(pts[:, 3:6])
    assert len(got & n0) > 100 and len(got & n1) > 100


@needs_ref_host
def test_prior_fusion_differs_from_plain_fusion():
    sc = make_scene(5, 128, 96, 3, seed=4)
    depths, normals, colors = noisy_maps(sc, 3, holes=0.1)
    depths2, normals2 = second_maps(sc, 5)
    cnt, _ = ob.prior_fusion(sc["cams"], depths, normals, depths2, normals2, sc["pairs"], 0.3, 1, 1, colors=colors)
    plain = ob.fusion(sc["cams"], depths, normals, sc["pairs"], 0.3, 1)
    assert cnt != plain and cnt > 0
