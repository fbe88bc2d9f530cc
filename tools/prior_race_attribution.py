#!/usr/bin/env python
"""Why do planar-prior half-sweeps of the reference build differ from the product's (and from themselves)?

Input: gpurun_out/parity_dump_<tag>.npz written by tools/parity_study.py --dump 320x240x5x4 - the state S before
a planar-prior half-sweep, the reference kernel's output R (and its re-run R2) and the product's output M.  At
320 x 240 the reference launch is ONE wave of blocks (80 <= 148): photometric and geometric half-sweeps are
bit-identical to the product's there, so whatever differs under a prior is not the inter-block race.

Finding (profiles/r2_prior_race_attribution.txt): the reference adopts a neighbour's plane in the MIDDLE of the
kernel - `plane_hypotheses[center] = plane_hypotheses[positions[max_cost_idx]]` (ACMMP.cu:1129-1131, :1141-1143),
re-loading the neighbour from global memory (SASS: LDG.E.128 R8 at 0x19ee0 / 0x1a150, STG.E.128 at 0x1a040 /
0x1a230) after its cost was computed from the value read at the start.  Same-colour near candidates belong to other
warps of the same block, which reach that point earlier or later: the re-load sees the neighbour's OWN adoption of
the same launch.  The pixel then carries a plane that is two propagation hops away from it, with the cost of the
one-hop plane.  Every plane of S is unique (random initialisation), so the hop count of an output plane can be read
off exactly."""
import collections
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def arm_positions(Sc):
    """arg-min candidate of the 8 arms (ACMMP.cu:812-988), as (8, H, W, 2) [y, x]"""
    H, W = Sc.shape
    yy, xx = np.mgrid[0:H, 0:W]
    P = np.pad(Sc, 24, constant_values=np.nan)
    C = lambda dy, dx: P[24 + dy:24 + dy + H, 24 + dx:24 + dx + W]  # noqa: E731
    pos = np.zeros((8, H, W, 2), int)
    dirs = [(-1, 0), (-1, 0), (1, 0), (1, 0), (0, -1), (0, -1), (0, 1), (0, 1)]
    for a, (dy, dx) in enumerate(dirs):
        if a & 1:
            best, bp = C(3 * dy, 3 * dx).copy(), np.stack([yy + 3 * dy, xx + 3 * dx], -1)
            for i in range(1, 11):
                k = 3 + 2 * i
                c = C(k * dy, k * dx)
                take = (best < c) if a == 7 else (c < best)      # right_far keeps the largest (ACMMP.cu:879)
                best, bp = np.where(take, c, best), np.where(take[..., None], np.stack([yy + k * dy, xx + k * dx], -1), bp)
        else:
            best, bp = C(dy, dx).copy(), np.stack([yy + dy, xx + dx], -1)
            lat = (0, 1) if dx == 0 else (1, 0)
            for i in range(3):
                for sg in (-1, 1):
                    oy, ox = (2 + i) * dy + sg * i * lat[0], (2 + i) * dx + sg * i * lat[1]
                    c = C(oy, ox)
                    take = c < best
                    best, bp = np.where(take, c, best), np.where(take[..., None], np.stack([yy + oy, xx + ox], -1), bp)
        pos[a] = bp
    return pos


def candidate_pattern():
    pat = set()
    for dy, dx in [(-1, 0), (1, 0), (0, -1), (0, 1)]:
        lat = (0, 1) if dx == 0 else (1, 0)
        pat.add((dy, dx))
        for i in range(3):
            for sg in (-1, 1):
                pat.add(((2 + i) * dy + sg * i * lat[0], (2 + i) * dx + sg * i * lat[1]))
        for i in range(11):
            pat.add(((3 + 2 * i) * dy, (3 + 2 * i) * dx))
    return pat


def analyse(d, key, out):
    color = int(key[-1])
    S, Sc, R, Rc = d[key + "_S_planes"], d[key + "_S_costs"], d[key + "_R_planes"], d[key + "_R_costs"]
    M, Mc, R2 = d[key + "_Mo_planes"], d[key + "_Mo_costs"], d[key + "_R2_planes"]
    masks = d["prior_masks"]
    h, w = Sc.shape
    yy, xx = np.mgrid[0:h, 0:w]
    inner = (((xx + yy) & 1) == color) & (yy >= 24) & (yy < h - 24) & (xx >= 24) & (xx < w - 24)
    eq = (bits(M) == bits(R)).all(-1)
    stable = (bits(R2) == bits(R)).all(-1)
    pos = arm_positions(Sc)
    pat = candidate_pattern()
    flat = bits(S.reshape(-1, 4))
    index = {}
    for i in range(flat.shape[0]):
        index[flat[i].tobytes()] = i
    out.append(f"== half-sweep {key}: {int(inner.sum())} pixels at least 24 px from the border; product = reference on {eq[inner].mean():.4f}, "
               f"reference = its own re-run on {stable[inner].mean():.4f}; planes of S that are unique: {len(index)} of {h * w}")

    def classify(sel, A):
        c = collections.Counter()
        rel = collections.Counter()
        ys, xs = np.nonzero(sel)
        for y, x in zip(ys, xs):
            i = index.get(bits(A[y, x]).tobytes())
            if i is None:
                c["not a plane of S (a refinement candidate won)"] += 1
                continue
            sy, sx = divmod(i, w)
            off = (sy - y, sx - x)
            if off == (0, 0):
                c["the pixel's own plane"] += 1
            elif off in pat:
                c["one hop: a candidate of the pixel"] += 1
            else:
                hit = None
                for a in range(8):
                    qy, qx = pos[a, y, x]
                    if ((qy + qx) & 1) == color and (sy - qy, sx - qx) in pat:
                        hit = (qy, qx)
                        break
                if hit is None:
                    c["elsewhere"] += 1
                else:
                    c["TWO hops: a candidate of the same-colour pixel chosen for one of the arms"] += 1
                    same_block = (y // 32 == hit[0] // 32) and (x // 32 == hit[1] // 32)
                    if same_block:
                        dw = (hit[0] % 32) // 2 - (y % 32) // 2
                        rel["same block, same warp" if dw == 0 else ("same block, earlier warp" if dw < 0 else "same block, later warp")] += 1
                    else:
                        rel["another block"] += 1
        return c, rel, len(ys)

    for name, mk in (("covered by the prior", masks > 0), ("not covered", masks == 0)):
        for label, sel, A in (("reference output where it DIFFERS from the product (and repeats itself)", inner & ~eq & stable & mk, R),
                              ("product output at the same pixels", inner & ~eq & stable & mk, M),
                              ("reference output where it EQUALS the product", inner & eq & mk, R)):
            c, rel, n = classify(sel, A)
            out.append(f"  {name}: {label}: {n} pixels")
            for k, v in c.most_common():
                out.append(f"      {v / max(n, 1):7.4f}  {k}")
            if rel:
                tot = sum(rel.values())
                out.append("      the neighbour the second hop went through sits in: " + ", ".join(f"{k} {v / tot:.3f}" for k, v in rel.most_common()))
        bad = inner & ~eq & stable & mk
        if bad.any():
            out.append(f"  {name}: cost stored by the reference = cost stored by the product on {(bits(Rc) == bits(Mc))[bad].mean():.4f} of the differing pixels "
                       "(the cost is the one-hop plane's; only the plane moved on)")


def main():
    path = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "parity_dump_r2a.npz")
    d = np.load(path)
    out = [__doc__.split("Finding")[0].strip(), ""]
    for key in ("prior_c0", "prior_c1"):
        analyse(d, key, out)
    for key in ("photo_c0", "geom_c0"):
        R, M = d[key + "_R_planes"], d[key + "_Mo_planes"]
        h, w = R.shape[:2]
        yy, xx = np.mgrid[0:h, 0:w]
        inner = (((xx + yy) & 1) == 0) & (yy >= 24) & (yy < h - 24) & (xx >= 24) & (xx < w - 24)
        out.append(f"== control, half-sweep {key} without a prior: product = reference on {(bits(M) == bits(R)).all(-1)[inner].mean():.5f} of the same pixels")
    print("\n".join(out))


if __name__ == "__main__":
    main()
