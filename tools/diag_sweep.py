#!/usr/bin/env python
"""Diagnoses the first half-sweep against the reference golden: where and how the
product differs."""
import os, sys, json
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import golden_util as gu
from acmmp_b200 import host
g, images, cams = gu.load()
h, w = images[0].shape
m = host.ACMMP(0)
m.InputInitialization(images, cams)
m.run_init(int(g["seed"][0]))
p0, c0, s0, r0 = m.get_state()
m.run_sweep(0, 0)
p1, c1, s1, r1 = m.get_state()
rp, rc = g["sweep00_planes"], g["sweep00_costs"]
yy, xx = np.mgrid[0:h, 0:w]
black = ((xx + yy) & 1) == 0
same = (gu.bits(p1) == gu.bits(rp)).all(-1)
bad = black & ~same
print("black px", black.sum(), "bad", bad.sum(), "frac", bad.sum() / black.sum())
ys, xs = np.nonzero(bad)
print("bad by border distance<4:", ((ys < 4) | (ys >= h - 4) | (xs < 4) | (xs >= w - 4)).sum())
print("bad y hist (rows of 8):", np.bincount(ys // 8, minlength=h // 8).tolist())
print("bad x hist (cols of 8):", np.bincount(xs // 8, minlength=w // 8).tolist())
print("rng equal among bad:", (r1 == g["sweep00_rng"]).all(-1)[bad].mean())
print("cost equal among bad:", (gu.bits(c1) == gu.bits(rc))[bad].mean())
print("ref unchanged among bad:", (gu.bits(rp) == gu.bits(p0)).all(-1)[bad].mean(), " mine unchanged:", (gu.bits(p1) == gu.bits(p0)).all(-1)[bad].mean())
# does the reference's plane equal some neighbour's initial plane (pure propagation)?
def is_copy(P, y, x):
    for dy in range(-25, 26):
        for dx in range(-25, 26):
            yy_, xx_ = y + dy, x + dx
            if 0 <= yy_ < h and 0 <= xx_ < w and (gu.bits(P[y, x]) == gu.bits(p0[yy_, xx_])).all():
                return (dx, dy)
    return None
n_ref_copy = n_mine_copy = 0
samples = []
for y, x in list(zip(ys, xs))[:400]:
    a, b = is_copy(rp, y, x), is_copy(p1, y, x)
    n_ref_copy += a is not None; n_mine_copy += b is not None
    if len(samples) < 12:
        samples.append({"xy": [int(x), int(y)], "ref_copy_of": a, "mine_copy_of": b, "ref": rp[y, x].tolist() + [float(rc[y, x])],
                        "mine": p1[y, x].tolist() + [float(c1[y, x])], "before": p0[y, x].tolist() + [float(c0[y, x])]})
print("of first", min(400, len(ys)), "bad: ref is copy of an initial plane:", n_ref_copy, "mine:", n_mine_copy)
for s in samples:
    print(json.dumps(s))
