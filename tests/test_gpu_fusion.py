"""GPU depth-map fusion (acmmp_fusion_*) against the CPU restatement of the reference's
RunFusion (oracle/upsample_fusion.c: orc_fusion, acmmp_definitions.cpp:828-1043).
pytest -m gpu."""
import os
import sys

import numpy as np
import pytest

import oraclebind as ob

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
if not torch.cuda.is_available():
    pytest.skip("no CUDA device", allow_module_level=True)

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

from acmmp_b200 import host  # noqa: E402
from acmmp_b200.scene import make_scene  # noqa: E402
from fusion_inputs import noisy_maps  # noqa: E402


def point_keys(pts):
    return set(map(bytes, np.ascontiguousarray(pts[:, :3]).view(np.uint8).reshape(len(pts), 12)))


@pytest.mark.parametrize("thresh,scalar", [(1, 0.3), (2, 0.3), (3, 0.1)])
def test_fusion_matches_cpu_restatement(thresh, scalar):
    sc = make_scene(6, 160, 120, 4, seed=2)
    depths, normals, colors = noisy_maps(sc, 11)
    cnt, ref_pts = ob.fusion(sc["cams"], depths, normals, sc["pairs"], scalar, thresh, colors=colors, want_points=True)
    pts, per_view = host.RunFusion(sc["cams"], sc["pairs"], depths, normals, colors=colors, consistency_scalar=scalar,
                                   con_num_thresh=thresh)
    assert cnt > 1000 and sum(per_view) == len(pts)
    # north_star: fused point count within 0.1 %
    assert abs(len(pts) - cnt) <= max(1, int(1e-3 * cnt)), (len(pts), cnt)
    a, b = point_keys(pts), point_keys(ref_pts)
    assert len(a & b) >= 0.999 * max(len(a), len(b)), (len(a), len(b), len(a & b))
    if len(pts) == cnt:
        same = (pts.view(np.uint32) == ref_pts.view(np.uint32)).all(1).mean()
        assert same >= 0.999, same      # same points, same order, same normals / colours


def test_fusion_respects_masks_and_is_repeatable():
    sc = make_scene(5, 128, 96, 3, seed=4)
    depths, normals, _ = noisy_maps(sc, 3)
    f = host.Fusion(sc["cams"], sc["pairs"])
    masks = [np.zeros(d.shape, np.uint8) for d in depths]
    masks[0][:, :64] = 1
    for v in range(5):
        f.set_view(v, depths[v], normals[v], mask=masks[v])
    n1 = f.run(0.3, 1)
    p1, per1 = f.points()
    n2 = f.run(0.3, 1)               # masks restart from the ones handed over
    p2, per2 = f.points()
    assert n1 == n2 and per1 == per2 and (p1.view(np.uint32) == p2.view(np.uint32)).all()
    st = f.stats()
    assert all(r >= 1 for r in st["rounds"])
    # no point of view 0 comes from the masked half
    cam = sc["cams"][0]
    first = p1[:per1[0], :3]
    R = np.array(cam.R[:], np.float64).reshape(3, 3)
    t = np.array(cam.t[:], np.float64)
    K = np.array(cam.K[:], np.float64).reshape(3, 3)
    uvw = (K @ (R @ first.T.astype(np.float64) + t[:, None]))
    u = uvw[0] / uvw[2]
    assert per1[0] > 0 and u.min() > 63.0
    # the working mask of a source view marks the pixels used by earlier views
    assert f.mask(sc["pairs"][0][0]).sum() > 0
    f.close()


# ---- prior-aware fusion (acmmp_fusion_run_prior_aware; RunPriorAwareFusion, acmmp_definitions.cpp:443-826) ----

@pytest.mark.parametrize("thresh,scalar,penalty", [(1, 0.3, 1), (2, 0.3, 1), (1, 0.1, 0), (2, 0.2, 2)])
def test_prior_aware_fusion_matches_cpu_restatement(thresh, scalar, penalty):
    """the restatement is itself bit-exact against the reference's own function (tests/test_fusion_pinned.py)"""
    sc = make_scene(6, 160, 120, 4, seed=2)
    depths, normals, colors = noisy_maps(sc, 11, holes=0.1)
    depths2, normals2, _ = noisy_maps(sc, 12, noise=0.003, holes=0.15, outliers=0.05)
    cnt, ref_pts = ob.prior_fusion(sc["cams"], depths, normals, depths2, normals2, sc["pairs"], scalar, thresh, penalty, colors=colors)
    pts, per_view = host.RunPriorAwareFusion(sc["cams"], sc["pairs"], depths, normals, depths2, normals2, colors=colors,
                                             consistency_scalar=scalar, num_consistent_thresh=thresh, single_match_penalty=penalty)
    assert cnt > 1000 and sum(per_view) == len(pts)
    assert abs(len(pts) - cnt) <= max(1, int(1e-3 * cnt)), (len(pts), cnt)      # north_star: fused count within 0.1 %
    a, b = point_keys(pts), point_keys(ref_pts)
    assert len(a & b) >= 0.999 * max(len(a), len(b)), (len(a), len(b), len(a & b))
    if len(pts) == cnt:
        same = (pts.view(np.uint32) == ref_pts.view(np.uint32)).all(1).mean()
        assert same >= 0.999, same


def test_prior_aware_fusion_masks_and_state():
    sc = make_scene(5, 128, 96, 3, seed=4)
    depths, normals, _ = noisy_maps(sc, 3, holes=0.1)
    depths2, normals2, _ = noisy_maps(sc, 5, holes=0.2)
    f = host.Fusion(sc["cams"], sc["pairs"])
    masks = [np.zeros(d.shape, np.uint8) for d in depths]
    masks[1][:48] = 1
    for v in range(5):
        f.set_view(v, depths[v], normals[v], mask=masks[v])
    with pytest.raises(Exception):
        f.run_prior_aware(0.3, 1, 1)              # second maps missing
    for v in range(5):
        f.set_view_second(v, depths2[v], normals2[v])
    n1 = f.run_prior_aware(0.3, 1, 1)
    p1, per1 = f.points()
    cnt, ref = ob.prior_fusion(sc["cams"], depths, normals, depths2, normals2, sc["pairs"], 0.3, 1, 1, masks=masks)
    assert abs(n1 - cnt) <= max(1, int(1e-3 * cnt))
    n2 = f.run_prior_aware(0.3, 1, 1)             # repeatable
    p2, per2 = f.points()
    assert n1 == n2 and per1 == per2 and (p1.view(np.uint32) == p2.view(np.uint32)).all()
    n3 = f.run(0.3, 1)                            # the plain fusion still runs on the same object
    assert n3 > 0
    f.close()
