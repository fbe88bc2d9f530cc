#!/usr/bin/env python
"""Stage-by-stage parity of the product CUDA path against the reference's own
kernels (oracle/_ref) on a B200: RNG stream, fixed-hypothesis init (NCC + view
masks), random init, single half-sweeps from identical state (step-exact), tail,
geometric consistency, planar prior, hierarchy init, JBU, full runs, timing and
the reference's run-to-run self-noise.  Writes gpurun_out/parity_<tag>.json."""
import argparse
import json
import os
import sys
import time
import traceback

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from acmmp_b200 import host, lib as L  # noqa: E402
from acmmp_b200.scene import make_scene  # noqa: E402
import refbind  # noqa: E402


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def cmp_f32(a, b, mask=None):
    a = np.asarray(a, np.float32)
    b = np.asarray(b, np.float32)
    if mask is not None:
        a = a[mask]
        b = b[mask]
    same = (bits(a) == bits(b)) | (np.isnan(a) & np.isnan(b))
    d = np.abs(a.astype(np.float64) - b.astype(np.float64))
    d = d[np.isfinite(d)]
    return {"n": int(a.size), "bit_equal_frac": float(same.mean()) if a.size else 1.0,
            "max_abs": float(d.max()) if d.size else 0.0, "mean_abs": float(d.mean()) if d.size else 0.0}


def plane_stats(pa, pb, mask=None):
    """planes (H,W,4): fraction of pixels whose 4 floats are all bit-identical."""
    same = (bits(pa) == bits(pb)).all(-1)
    if mask is not None:
        same = same[mask]
    return float(same.mean())


def depth_normal_tol(pa, pb, mask=None):
    """fraction of pixels with depth within 0.1 % and normal within 0.5 deg
    (planes after the tail: world normal + depth)."""
    da, db = pa[..., 3].astype(np.float64), pb[..., 3].astype(np.float64)
    rel = np.abs(da - db) / np.maximum(np.abs(db), 1e-12)
    na, nb = pa[..., :3].astype(np.float64), pb[..., :3].astype(np.float64)
    na /= np.maximum(np.linalg.norm(na, axis=-1, keepdims=True), 1e-12)
    nb /= np.maximum(np.linalg.norm(nb, axis=-1, keepdims=True), 1e-12)
    ang = np.degrees(np.arccos(np.clip((na * nb).sum(-1), -1, 1)))
    ok_d, ok_n = rel <= 1e-3, ang <= 0.5
    if mask is not None:
        ok_d, ok_n = ok_d[mask], ok_n[mask]
    return {"depth_ok": float(ok_d.mean()), "normal_ok": float(ok_n.mean()), "both_ok": float((ok_d & ok_n).mean())}


def color_mask(h, w, color):
    yy, xx = np.mgrid[0:h, 0:w]
    return ((xx + yy) & 1) == color


def interior(h, w, b):
    m = np.zeros((h, w), bool)
    m[b:h - b, b:w - b] = True
    return m


def gt_planes(scene, idx, rng, noise=0.0):
    """per-pixel plane (camera frame normal + offset) from the ground-truth depth,
    optionally perturbed: fixed hypotheses for the seeded branch."""
    cam = scene["cams"][idx]
    h, w = scene["height"], scene["width"]
    depth = scene["depths"][idx].astype(np.float32).copy()
    if noise > 0:
        depth *= (1 + noise * (rng.random((h, w), dtype=np.float32) - 0.5)).astype(np.float32)
    n = np.zeros((h, w, 3), np.float32)
    n[..., 2] = -1.0
    if noise > 0:
        n[..., 0] = noise * 4 * (rng.random((h, w), dtype=np.float32) - 0.5)
        n[..., 1] = noise * 4 * (rng.random((h, w), dtype=np.float32) - 0.5)
        n /= np.linalg.norm(n, axis=-1, keepdims=True)
    ys, xs = np.mgrid[0:h, 0:w].astype(np.float32)
    K = np.array(cam.K, np.float32)
    X = depth * (xs - K[2]) / K[0]
    Y = depth * (ys - K[5]) / K[4]
    d = -(n[..., 0] * X + n[..., 1] * Y + n[..., 2] * depth)
    return np.concatenate([n, d[..., None]], -1).astype(np.float32)


def problem(scene, idx):
    ids = [idx] + scene["pairs"][idx]
    return [scene["images"][i] for i in ids], [scene["cams"][i] for i in ids], ids


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--width", type=int, default=640)
    ap.add_argument("--height", type=int, default=480)
    ap.add_argument("--views", type=int, default=5)
    ap.add_argument("--nsrc", type=int, default=4)
    ap.add_argument("--seed", type=int, default=1234)
    ap.add_argument("--tag", default="r1")
    ap.add_argument("--big", action="store_true", help="also time a 1600x1200 x 10-source view")
    args = ap.parse_args()
    out = {"config": vars(args)}
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    rep_path = os.path.join(ROOT, "gpurun_out", f"parity_{args.tag}.json")

    def save():
        with open(rep_path, "w") as f:
            json.dump(out, f, indent=1)

    def section(name, fn):
        t = time.time()
        try:
            out[name] = fn()
        except Exception as e:  # keep going: every section is independent evidence
            out[name] = {"error": repr(e), "trace": traceback.format_exc()[-1500:]}
        out[name + "_s"] = round(time.time() - t, 2)
        print(f"== {name}: {json.dumps(out[name])[:3000]}", flush=True)
        save()

    rs = np.random.default_rng(0)
    scene = make_scene(args.views, args.width, args.height, args.nsrc, seed=0)
    h, w = args.height, args.width
    images, cams, ids = problem(scene, 0)
    mine = host.ACMMP(0)
    ref = refbind.RefACMMP(0)
    ref.set_problem(images, cams)
    mine.InputInitialization(images, cams)

    # ---- 1. RNG stream: curand_init(seed, y, x) ------------------------------------
    def s_rng():
        r_mine = host.rng_init_states(args.seed, w, h)
        ref.set_flags(seeded=1)
        ref.set_planes(2, gt_planes(scene, 0, rs))
        ref.run_init(args.seed)
        r_ref = ref.get_state()[3]
        return {"bit_equal_frac": float((r_mine == r_ref).all(-1).mean())}
    section("rng_init", s_rng)

    # ---- 2. fixed-hypothesis init: NCC cost + view-selection masks ------------------
    def s_seeded():
        res = {}
        for noise in (0.0, 0.02, 0.3):
            pl = gt_planes(scene, 0, rs, noise)
            ref.set_flags(seeded=1)
            ref.set_planes(2, pl)
            ref.run_init(args.seed)
            rp, rc_, rsel, _, _ = ref.get_state()
            mine.InputInitialization(images, cams)
            mine.SetPlanarPrior(pl)
            mine.run_init(args.seed)
            mp, mc, msel, _ = mine.get_state()
            r = {"cost": cmp_f32(mc, rc_), "mask_equal_frac": float((msel == rsel).mean()),
                 "plane_bit_equal": plane_stats(mp, rp), "cost_mean": float(rc_.mean())}
            # per-view NCC through the reference's ComputeBilateralNCC probe
            res[f"noise_{noise}"] = r
        return res
    section("seeded_init", s_seeded)

    # ---- 3. random init --------------------------------------------------------------
    def s_random():
        ref.set_flags(seeded=0, geom=0, planar_prior=0, hierarchy=0, upsample=0)
        ref.run_init(args.seed)
        nonlocal mine
        rp, rc_, rsel, rrng, _ = ref.get_state()
        mine.close()
        mine = host.ACMMP(0)
        mine.InputInitialization(images, cams)
        mine.run_init(args.seed)
        mp, mc, msel, mrng = mine.get_state()
        return {"plane_bit_equal": plane_stats(mp, rp), "cost": cmp_f32(mc, rc_),
                "mask_equal_frac": float((msel == rsel).mean()), "rng_equal_frac": float((mrng == rrng).all(-1).mean()),
                "plane_w": cmp_f32(mp[..., 3], rp[..., 3]), "plane_x": cmp_f32(mp[..., 0], rp[..., 0])}
    section("random_init", s_random)

    # ---- 4. step-exact half sweeps from identical state --------------------------------
    def sweep_compare(flags_desc, masks=None):
        res = {}
        for it in (0, 1):
            for color in (0, 1):
                rp0, rc0, rsel0, rrng0, _ = ref.get_state()
                mine.set_state(rp0, rc0, rsel0, rrng0)
                ref.run_sweep(it, color)
                mine.run_sweep(it, color)
                rp, rc_, rsel, rrng, _ = ref.get_state()
                mp, mc, msel, mrng = mine.get_state()
                # the reference once more from the same state: its own determinism
                ref.set_planes(0, rp0); ref.set_scalars(0, rc0); ref.set_scalars(2, rsel0); ref.set_rng(rrng0)
                ref.run_sweep(it, color)
                rp2 = ref.get_state()[0]
                cm = color_mask(h, w, color)
                inner = cm & interior(h, w, 4)
                same = (bits(mp) == bits(rp)).all(-1)
                bad = inner & ~same
                ys, xs = np.nonzero(bad)
                r = {
                    "plane_bit_equal_all": plane_stats(mp, rp, cm),
                    "plane_bit_equal_interior": plane_stats(mp, rp, inner),
                    "ref_rerun_bit_equal_interior": plane_stats(rp2, rp, inner),
                    "cost": cmp_f32(mc, rc_, inner), "mask_equal": float((msel == rsel)[inner].mean()),
                    "rng_equal": float((mrng == rrng).all(-1)[inner].mean()),
                    "other_colour_untouched": plane_stats(mp, rp0, ~cm),
                    "n_bad": int(bad.sum()),
                    "bad_y_mod32": np.bincount(ys % 32, minlength=32).tolist(),
                    "bad_x_mod32": np.bincount(xs % 32, minlength=32).tolist(),
                    "bad_cost_equal_frac": float((bits(mc) == bits(rc_))[bad].mean()) if bad.any() else 1.0,
                    "bad_rng_equal_frac": float((mrng == rrng).all(-1)[bad].mean()) if bad.any() else 1.0,
                }
                if masks is not None:
                    r["equal_mask_pos"] = plane_stats(mp, rp, inner & (masks > 0))
                    r["equal_mask_zero"] = plane_stats(mp, rp, inner & (masks == 0))
                samples = []
                for k in range(min(6, len(ys))):
                    y, x = int(ys[k * max(1, len(ys) // 6)]), int(xs[k * max(1, len(ys) // 6)])
                    samples.append({"xy": [x, y], "mine": mp[y, x].tolist() + [float(mc[y, x])],
                                    "ref": rp[y, x].tolist() + [float(rc_[y, x])],
                                    "before": rp0[y, x].tolist() + [float(rc0[y, x])]})
                r["samples"] = samples
                res[f"it{it}_c{color}"] = r
        return res

    def s_sweeps():
        ref.set_flags(seeded=0, geom=0, planar_prior=0, hierarchy=0, upsample=0)
        ref.run_init(args.seed)
        return sweep_compare("photometric")
    section("sweeps_photometric", s_sweeps)

    # ---- 5. tail ---------------------------------------------------------------------------
    def s_tail():
        rp0, rc0, rsel0, rrng0, _ = ref.get_state()
        mine.set_state(rp0, rc0, rsel0, rrng0)
        ref.run_tail()
        mine.run_tail()
        rp = ref.get_state()[0]
        mp = mine.get_state()[0]
        return {"plane_bit_equal": plane_stats(mp, rp), "depth": cmp_f32(mp[..., 3], rp[..., 3])}
    section("tail", s_tail)

    # ---- 6. full photometric run: mine vs ref, ref vs ref (self-noise), timing ----------------
    def s_full():
        ref.set_flags(seeded=0, geom=0, planar_prior=0, hierarchy=0, upsample=0)
        ms_ref = ref.run(args.seed)
        ra = ref.get_state()
        ms_ref2 = ref.run(args.seed)
        rb = ref.get_state()
        mine.InputInitialization(images, cams)
        mine.RunPatchMatch(args.seed)
        tm = mine.last_timing()
        mp, mc, _, _ = mine.get_state()
        gt = scene["depths"][0]
        inner = interior(h, w, 8)
        err_m = np.abs(mp[..., 3] - gt) / gt
        err_r = np.abs(ra[0][..., 3] - gt) / gt
        return {"ref_ms": [ms_ref, ms_ref2], "mine": tm,
                "ref_self_noise": {"plane_bit_equal": plane_stats(ra[0], rb[0]), **depth_normal_tol(ra[0], rb[0], inner)},
                "mine_vs_ref": {"plane_bit_equal": plane_stats(mp, ra[0]), **depth_normal_tol(mp, ra[0], inner)},
                "gt_err_lt_1pct": {"mine": float((err_m[inner] < 0.01).mean()), "ref": float((err_r[inner] < 0.01).mean())},
                "gt_err_lt_0.2pct": {"mine": float((err_m[inner] < 0.002).mean()), "ref": float((err_r[inner] < 0.002).mean())}}
    section("full_photometric", s_full)

    # ---- 7. geometric consistency: step-exact sweeps with depth maps ----------------------------
    def s_geom():
        # warm start = the reference's photometric result (world normal + depth)
        rp, rc_, rsel, rrng, _ = ref.get_state()
        depth, normal, cost = rp[..., 3].copy(), rp[..., :3].copy(), rc_.copy()
        dmaps = [depth] + [scene["depths"][i] for i in ids[1:]]
        ref.set_flags(geom=1, seeded=0, planar_prior=0, hierarchy=0, upsample=0)
        for i, d in enumerate(dmaps):
            ref.set_depth_map(i, d)
        ref.set_planes(0, rp)
        ref.set_scalars(0, cost)
        m = host.ACMMP(0)
        m.SetGeomConsistencyParams(False)
        m.InputInitialization(images, cams, depths=dmaps)
        m.CudaSpaceInitialization(warm_start=(depth, normal, cost))
        ref.run_init(args.seed)
        m.run_init(args.seed)
        r1 = ref.get_state()
        m1 = m.get_state()
        res = {"init": {"plane_bit_equal": plane_stats(m1[0], r1[0]), "cost": cmp_f32(m1[1], r1[1]),
                        "mask_equal": float((m1[2] == r1[2]).mean())}}
        nonlocal mine
        mine = m
        return res
    section("geom_init", s_geom)

    def s_geom_sweeps():
        return sweep_compare("geom")
    section("sweeps_geom", s_geom_sweeps)

    # ---- 8. planar prior: step-exact sweeps with a supplied prior ---------------------------------
    def s_prior():
        nonlocal mine
        pl = gt_planes(scene, 0, rs, 0.01)
        masks = np.zeros((h, w), np.uint32)
        masks[h // 6: 5 * h // 6, w // 6: 5 * w // 6] = 1
        masks[(rs.random((h, w)) < 0.1)] = 0
        ref.set_flags(geom=0, seeded=0, planar_prior=1, hierarchy=0, upsample=0)
        ref.set_planes(1, pl)
        ref.set_scalars(3, masks)
        ref.run_init(args.seed)   # random init branch (the reference re-randomises, ACMMP.cu:626)
        m = host.ACMMP(0)
        m.SetPlanarPriorParams()
        m.InputInitialization(images, cams)
        m.CudaPlanarPriorInitialization(pl, masks)
        m.run_init(args.seed)
        r1, m1 = ref.get_state(), m.get_state()
        mine = m
        res = {"init_plane_bit_equal": plane_stats(m1[0], r1[0])}
        res.update(sweep_compare("prior", masks))
        return res
    section("sweeps_prior", s_prior)

    # ---- 9. JBU + hierarchy (upsample) init ------------------------------------------------------------
    def s_jbu():
        small = make_scene(args.views, w // 2, h // 2, args.nsrc, seed=0)
        d_small = small["depths"][0]
        a = host.RunJBU(images[0], d_small)
        b = refbind.ref_jbu(images[0], d_small)
        return {"depth": cmp_f32(a, b)}
    section("jbu", s_jbu)

    def s_hier():
        small = make_scene(args.views, w // 2, h // 2, args.nsrc, seed=0)
        cn = np.zeros((h // 2, w // 2, 3), np.float32)
        Rm = np.array(cams[0].R, np.float32).reshape(3, 3)
        ncam = np.array([0.05, -0.03, -1.0], np.float32)
        ncam /= np.linalg.norm(ncam)
        cn[:] = Rm.T @ ncam                      # world-frame normal
        cn += 0.02 * (rs.random(cn.shape, dtype=np.float32) - 0.5)
        cc = (0.05 + 0.3 * rs.random((h // 2, w // 2), dtype=np.float32)).astype(np.float32)
        fine_depth = host.RunJBU(images[0], small["depths"][0])
        ref.set_flags(geom=0, seeded=0, planar_prior=0, hierarchy=1, upsample=1, scaled_cols=float(w // 2), scaled_rows=float(h // 2))
        ref.set_planes(3, np.concatenate([cn, cc[..., None]], -1))
        fp = np.zeros((h, w, 4), np.float32)
        fp[..., 3] = fine_depth
        ref.set_planes(0, fp)
        ref.run_init(args.seed)
        r1 = ref.get_state()
        m = host.ACMMP(0)
        m.SetHierarchyParams()
        m.InputInitialization(images, cams)
        m.CudaSpaceInitialization(hierarchy=(cn, cc, fine_depth))
        m.run_init(args.seed)
        m1 = m.get_state()
        return {"plane_bit_equal": plane_stats(m1[0], r1[0]), "cost": cmp_f32(m1[1], r1[1]),
                "mask_equal": float((m1[2] == r1[2]).mean()), "normal_x": cmp_f32(m1[0][..., 0], r1[0][..., 0])}
    section("hierarchy_init", s_hier)

    # ---- 10. DTU-shape single view timing (1600x1200, 10 sources) ----------------------------------------
    def s_big():
        sc = make_scene(11, 1600, 1200, 10, seed=1, device="cuda", span_deg=(10.0, 8.0))
        im, cm_, _ = problem(sc, 5)
        r = refbind.RefACMMP(0)
        r.set_problem(im, cm_)
        t_ref = [r.run(args.seed) for _ in range(2)]
        ra = r.get_state()
        m = host.ACMMP(0)
        m.InputInitialization(im, cm_)
        tms = []
        for _ in range(3):
            m.RunPatchMatch(args.seed)
            tms.append(m.last_timing())
        mp = m.get_state()[0]
        inner = interior(1200, 1600, 8)
        fp32, mufu = host.measure_fp32_peak(0)
        return {"ref_ms": t_ref, "mine": tms, "speedup": t_ref[-1] / tms[-1]["total_ms"],
                "mine_vs_ref": depth_normal_tol(mp, ra[0], inner), "fp32_tflops": fp32, "mufu_gops": mufu}
    if args.big:
        section("dtu_view_timing", s_big)
    save()
    print("report:", rep_path)


if __name__ == "__main__":
    main()
