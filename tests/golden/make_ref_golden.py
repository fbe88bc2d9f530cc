#!/usr/bin/env python
"""Generates tests/golden/ref_golden.npz: outputs of the REFERENCE's own CUDA
kernels (oracle/_ref/libacmmp_ref.so = /root/reference/src/ACMMP.cu compiled for
sm_100 with the reference's flags, driven by oracle/ref_harness.cu) on a small
seeded scene, stored together with the inputs.  Run on a B200 box:

    python tests/golden/make_ref_golden.py            # writes gpurun_out/ref_golden.npz

and copy the file to tests/golden/.  The CPU oracle (oracle/*.c) and the product
CUDA path are both tested against these vectors (tests/test_oracle_golden.py,
tests/test_gpu_parity.py).  The reference ships no golden vectors of its own
(SURVEY.md section 4), so these are the pins."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import refbind  # noqa: E402
from acmmp_b200.scene import make_scene  # noqa: E402
from acmmp_b200 import lib as L  # noqa: E402

W, H, NV, NSRC, SEED = 96, 72, 5, 4, 1234


def cam_to_array(c):
    return np.concatenate([np.array(c.K, np.float32), np.array(c.R, np.float32), np.array(c.t, np.float32),
                           np.array([c.height, c.width, c.depth_min, c.depth_max], np.float32)])


def gt_planes(depth, cam, rs, noise):
    h, w = depth.shape
    d = depth.astype(np.float32).copy()
    n = np.zeros((h, w, 3), np.float32)
    n[..., 2] = -1.0
    if noise > 0:
        d *= (1 + noise * (rs.random((h, w), dtype=np.float32) - 0.5)).astype(np.float32)
        n[..., 0] = noise * 4 * (rs.random((h, w), dtype=np.float32) - 0.5)
        n[..., 1] = noise * 4 * (rs.random((h, w), dtype=np.float32) - 0.5)
        n /= np.linalg.norm(n, axis=-1, keepdims=True)
    ys, xs = np.mgrid[0:h, 0:w].astype(np.float32)
    K = np.array(cam.K, np.float32)
    X = d * (xs - K[2]) / K[0]
    Y = d * (ys - K[5]) / K[4]
    off = -(n[..., 0] * X + n[..., 1] * Y + n[..., 2] * d)
    return np.concatenate([n, off[..., None]], -1).astype(np.float32)


def main():
    rs = np.random.default_rng(7)
    sc = make_scene(NV, W, H, NSRC, seed=0)
    ids = [0] + sc["pairs"][0]
    images = [sc["images"][i] for i in ids]
    cams = [sc["cams"][i] for i in ids]
    depths = [sc["depths"][i] for i in ids]
    out = {"images_u8": np.stack(images).astype(np.uint8), "cams": np.stack([cam_to_array(c) for c in cams]),
           "gt_depths": np.stack(depths).astype(np.float32), "seed": np.array([SEED], np.uint64)}
    ref = refbind.RefACMMP(0)
    ref.set_problem(images, cams)

    # fixed hypotheses: exact, mildly and strongly perturbed
    for tag, noise in (("p0", 0.0), ("p1", 0.02), ("p2", 0.3)):
        pl = gt_planes(depths[0], cams[0], rs, noise)
        out[f"planes_{tag}"] = pl
        out[f"ncc_{tag}"] = np.stack([ref.ncc(pl, v) for v in range(1, len(ids))])
        out[f"homography_{tag}"] = np.stack([ref.homography(pl.reshape(-1, 4)[::97], v) for v in range(1, len(ids))])
        ref.set_flags(seeded=1, geom=0, planar_prior=0, hierarchy=0, upsample=0)
        ref.set_planes(2, pl)
        ref.run_init(SEED)
        p, c, s, r, _ = ref.get_state()
        out[f"seeded_cost_{tag}"] = c
        out[f"seeded_sel_{tag}"] = s
        if tag == "p0":
            out["rng_after_seeded_init"] = r   # = curand_init(seed, y, x), no draws in this branch

    # geometric consistency cost against the ground-truth source depth maps
    for i, d in enumerate(depths):
        ref.set_depth_map(i, d)
    out["geom_cost_p1"] = np.stack([ref.geom_cost(out["planes_p1"], v) for v in range(1, len(ids))])

    # random initialisation (RNG stream + random plane + cost + mask)
    ref.set_flags(seeded=0, geom=0, planar_prior=0, hierarchy=0, upsample=0)
    ref.run_init(SEED)
    p, c, s, r, _ = ref.get_state()
    out.update(rinit_planes=p, rinit_costs=c, rinit_sel=s, rinit_rng=r)
    # one black half-sweep from that state, then the rest of the run
    ref.run_sweep(0, 0)
    p, c, s, r, _ = ref.get_state()
    out.update(sweep00_planes=p, sweep00_costs=c, sweep00_sel=s, sweep00_rng=r)
    ref.run_sweep(0, 1)
    ref.run_sweep(1, 0)
    ref.run_sweep(1, 1)
    p, c, s, r, _ = ref.get_state()
    out.update(sweeps_planes=p, sweeps_costs=c)
    ref.run_tail()
    p, c, s, r, _ = ref.get_state()
    out.update(final_planes=p, final_costs=c)
    # the same full run once more: the reference's own run-to-run noise
    ref.run(SEED)
    p2 = ref.get_state()[0]
    out["final_planes_rerun"] = p2

    # JBU depth upsample (2x) of the half-resolution ground-truth depth
    small = make_scene(NV, W // 2, H // 2, NSRC, seed=0)
    out["jbu_in_depth"] = small["depths"][0].astype(np.float32)
    out["jbu_out"] = refbind.ref_jbu(images[0], small["depths"][0])

    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    path = os.path.join(ROOT, "gpurun_out", "ref_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
