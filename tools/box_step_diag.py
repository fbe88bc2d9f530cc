"""Diagnostics: planar-prior run with box staging off / on, stepped half-sweep by half-sweep; prints the first state
(planes / costs / selected views / RNG) that differs and where.  Found the contraction-order divergence of prior_dot
(sweep_kernel.cu.inc) in round 2."""
import os, sys, numpy as np
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
from acmmp_b200 import host, lib as L
from acmmp_b200.scene import make_scene
sc = make_scene(5, 640, 480, 4, seed=2)
order = [0] + list(sc["pairs"][0])
imgs = [np.clip(np.rint(sc["images"][i]), 0, 255).astype(np.float32) for i in order]
cams = [sc["cams"][i] for i in order]
ms = []
for staging in (False, True):
    m = host.ACMMP(); m.SetTextureFormat(L.TEX_AUTO); m.SetBoxStaging(staging)
    m.InputInitialization(imgs, cams)
    m.RunPatchMatch(21)
    m.SetPlanarPriorParams(); m.BuildPlanarPrior()
    ms.append(m)
pp = [m.GetPlanarPrior() for m in ms]
print("prior planes equal", (pp[0][0].view(np.uint32) == pp[1][0].view(np.uint32)).all(), "masks equal", (pp[0][1] == pp[1][1]).all())
for m in ms: m.run_init(22)
names = ["planes", "costs", "sel", "rng"]
def cmp(tag):
    st = [m.get_state() for m in ms]
    bad = False
    for k in range(4):
        a, b = st[0][k], st[1][k]
        d = a.view(np.uint32) != b.view(np.uint32)
        if d.ndim == 3: d = d.any(-1)
        if d.any():
            bad = True
            ys, xs = np.nonzero(d)
            print(tag, names[k], "differs on", int(d.sum()), "px; first", [(int(y), int(x)) for y, x in zip(ys[:6], xs[:6])])
            y, x = int(ys[0]), int(xs[0])
            print("    off", st[0][k][y, x], "on", st[1][k][y, x], "mask", pp[0][1][y, x], "color", (y + x) & 1)
    return bad
cmp("init")
for it in range(3):
    for color in (0, 1):
        for m in ms: m.run_sweep(it, color)
        if cmp("it %d color %d" % (it, color)): sys.exit(0)
print("all equal")
