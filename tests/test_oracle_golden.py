"""Pins the CPU oracle (oracle/*.c) against golden vectors produced by the
REFERENCE's own CUDA kernels on a B200 (tests/golden/ref_golden.npz).

Tolerances: the reference build evaluates 1/x, sqrt, exp with MUFU approximations
(--use_fast_math) and samples source images through the texture unit; the oracle uses
exact libm and an emulation of that filter with the hardware's own 2-D weights (multiples
of 1/256, rule measured on B200 - oracle/cost.c: orc_tex_bilinear).  A sample whose source
coordinate lands within ~1e-5 px of a 1/256 weight step can still take the neighbouring
weight, which moves the cost by up to ~2e-3.  Hence: mean |d| <= 1e-5, >= 99 % of samples
within 1e-4 (north_star's NCC tolerance), max <= 5e-3.  The product CUDA path itself is held
to bit-exactness against the same vectors in test_gpu_parity.py."""
import numpy as np

import golden_util as gu
import oraclebind as ob


def _problem(with_depths=False):
    g, images, cams = gu.load()
    depths = [d for d in g["gt_depths"]] if with_depths else None
    return g, ob.Problem(images, cams, depths=depths), cams


def test_rng_state_equals_curand_init_of_reference_kernel():
    g, pr, _ = _problem()
    mine = ob.rng_init_image(int(g["seed"][0]), pr.w, pr.h)
    assert (mine == g["rng_after_seeded_init"]).all()


def test_homography_against_reference_kernel():
    g, pr, cams = _problem()
    for tag in ("p0", "p1", "p2"):
        pl = g["planes_" + tag].reshape(-1, 4)[::97]
        for v in range(1, pr.n):
            ref = g["homography_" + tag][v - 1]
            mine = np.stack([ob.homography(cams[0], cams[v], p) for p in pl])
            scale = np.abs(ref).max(axis=1, keepdims=True)
            assert np.abs(mine - ref).max() <= 2e-5 * scale.max()
            assert (np.abs(mine - ref) / scale).max() <= 2e-5


def test_ncc_against_reference_kernel():
    g, pr, _ = _problem()
    for tag in ("p0", "p1", "p2"):
        for v in range(1, pr.n):
            mine = pr.ncc_map(g["planes_" + tag], v)
            ref = g["ncc_" + tag][v - 1]
            d = np.abs(mine - ref)
            assert d.mean() <= 1e-5, (tag, v, d.mean())
            assert (d <= 1e-4).mean() >= 0.99, (tag, v, (d <= 1e-4).mean())
            assert d.max() <= 5e-3, (tag, v, d.max())
            # the out-of-image / flat-patch exits agree exactly
            assert ((mine == 2.0) == (ref == 2.0)).mean() >= 0.999


def test_initial_cost_and_view_masks_against_reference_kernel():
    g, pr, _ = _problem()
    for tag in ("p0", "p1", "p2"):
        pr.set_seeded(g["planes_" + tag])
        pr.init(ob.INIT_SEEDED)
        ref_c, ref_s = g["seeded_cost_" + tag], g["seeded_sel_" + tag]
        d = np.abs(pr.costs - ref_c)
        assert d.mean() <= 1e-5 and d.max() <= 5e-3
        # view-selection masks: equal wherever the k-th cost is not a near tie
        ncc = np.stack([g["ncc_" + tag][v] for v in range(pr.n - 1)], -1)
        srt = np.sort(ncc, -1)
        kth = srt[..., 3]
        near_tie = (np.abs(ncc - kth[..., None]) < 2e-3).sum(-1) > 1
        assert (pr.sel == ref_s)[~near_tie].mean() >= 0.999
        assert (pr.sel == ref_s).mean() >= 0.97


def test_geom_consistency_cost_against_reference_kernel():
    g, pr, _ = _problem(with_depths=True)
    pl = g["planes_p1"]
    for v in range(1, pr.n):
        ref = g["geom_cost_p1"][v - 1]
        mine = np.zeros_like(ref)
        for y in range(0, pr.h, 3):
            for x in range(0, pr.w, 3):
                mine[y, x] = pr.geom_cost(x, y, pl[y, x], v)
        m = np.zeros_like(ref, bool)
        m[::3, ::3] = True
        d = np.abs(mine - ref)[m]
        # the integer truncation of the projected pixel flips for coordinates within float noise of an integer
        assert np.median(d) <= 1e-3 and (d <= 2e-2).mean() >= 0.97


def test_random_init_planes_follow_the_reference_stream():
    g, pr, _ = _problem()
    pr.rng_init(int(g["seed"][0]))
    pr.init(ob.INIT_RANDOM)
    ref = g["rinit_planes"]
    # same draws -> same depth and normal up to approximate sqrt / division
    assert np.abs(pr.planes[..., :3] - ref[..., :3]).max() <= 1e-4
    assert (np.abs(pr.planes[..., 3] - ref[..., 3]) <= 1e-3 * np.abs(ref[..., 3]) + 1e-2).mean() >= 0.999
    assert (pr.rng == g["rinit_rng"]).all(-1).mean() >= 0.999   # rejection loop length can differ on a boundary draw
    d = np.abs(pr.costs - g["rinit_costs"])
    assert np.nanmean(d) <= 2e-4


def test_jbu_against_reference_kernel():
    g, images, _ = gu.load()
    mine = ob.jbu_depth(images[0], g["jbu_in_depth"])
    ref = g["jbu_out"]
    assert np.abs(mine - ref).max() <= 2e-4 * np.abs(ref).max()


def test_full_run_statistics_against_reference_kernels():
    """PatchMatch amplifies 1-ulp differences (arg-min / acceptance decisions), so a
    full run is compared statistically, next to the reference's own run-to-run noise
    (it has a same-colour read race)."""
    g, pr, _ = _problem()
    pr.run(ob.INIT_RANDOM, int(g["seed"][0]))
    ref, rerun = g["final_planes"], g["final_planes_rerun"]
    inner = (slice(8, -8), slice(8, -8))
    gt = g["gt_depths"][0]
    err_o = np.abs(pr.planes[..., 3] - gt) / gt
    err_r = np.abs(ref[..., 3] - gt) / gt
    acc_o, acc_r = (err_o[inner] < 0.01).mean(), (err_r[inner] < 0.01).mean()
    assert abs(acc_o - acc_r) <= 0.03, (acc_o, acc_r)
    dk, nk = gu.depth_normal_ok(pr.planes, ref)
    dk2, nk2 = gu.depth_normal_ok(rerun, ref)
    assert dk[inner].mean() >= 0.85, dk[inner].mean()
    # informational: the reference against itself
    assert dk2[inner].mean() >= 0.9
