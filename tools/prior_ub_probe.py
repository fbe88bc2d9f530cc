#!/usr/bin/env python
"""What do the uninitialised `plane_hypotheses_now` registers of the reference hold in planar-prior sweeps
(ACMMP.cu:1149)?  Runs single half-sweeps from identical state against the reference's own
kernels and prints the fraction of bit-equal interior pixels, split by prior coverage, for the two compat
modes (0 = ACMMP_COMPAT_REF_SM100, the pinning derived with this tool; 1 = ACMMP_COMPAT_UPSTREAM).
The pinning was found by trying, behind temporary switches, which plane the refinement starts from
(live plane / best arm's plane / centre plane / last arm's plane) with and without the shadowed
`depth_now` of ACMMP.cu:1120: profiles/r1q_prior_ub_probe.txt."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from acmmp_b200 import host  # noqa: E402
from acmmp_b200.scene import make_scene  # noqa: E402
import refbind  # noqa: E402
from gpu_parity import gt_planes, problem, bits, interior, color_mask  # noqa: E402

exps = [int(x, 0) for x in (sys.argv[1:] or ["0", "1"])]
w, h = 640, 480
scene = make_scene(5, w, h, 4, seed=0)
images, cams, ids = problem(scene, 0)
rs = np.random.default_rng(0)
pl = gt_planes(scene, 0, rs, 0.01)
masks = np.zeros((h, w), np.uint32)
masks[h // 6: 5 * h // 6, w // 6: 5 * w // 6] = 1
masks[(rs.random((h, w)) < 0.1)] = 0
ref = refbind.RefACMMP(0)
ref.set_problem(images, cams)
ref.set_flags(geom=0, seeded=0, planar_prior=1, hierarchy=0, upsample=0)
ref.set_planes(1, pl)
ref.set_scalars(3, masks)
ref.run_init(1234)
m = host.ACMMP(0)
m.SetPlanarPriorParams()
m.InputInitialization(images, cams)
m.CudaPlanarPriorInitialization(pl, masks)
m.run_init(1234)
out = {}
for it in (0, 1):
    for color in (0, 1):
        rp0, rc0, rsel0, rrng0, _ = ref.get_state()
        ref.run_sweep(it, color)
        rp, rc_, rsel, rrng, _ = ref.get_state()
        ref.set_planes(0, rp0); ref.set_scalars(0, rc0); ref.set_scalars(2, rsel0); ref.set_rng(rrng0)
        ref.run_sweep(it, color)
        rp2 = ref.get_state()[0]
        inner = color_mask(h, w, color) & interior(h, w, 4)
        line = {"ref_rerun": float((bits(rp2) == bits(rp)).all(-1)[inner].mean())}
        for e in exps:
            m.SetCompatMode(e)
            m.set_state(rp0, rc0, rsel0, rrng0)
            m.run_sweep(it, color)
            mp, mc, msel, mrng = m.get_state()
            same = (bits(mp) == bits(rp)).all(-1)
            csame = bits(mc) == bits(rc_)
            line[f"compat{e}"] = {"all": round(float(same[inner].mean()), 4), "mask>0": round(float(same[inner & (masks > 0)].mean()), 4),
                               "mask=0": round(float(same[inner & (masks == 0)].mean()), 4), "cost": round(float(csame[inner].mean()), 4)}
        out[f"it{it}_c{color}"] = line
        print(f"it{it}_c{color}", json.dumps(line), flush=True)
        # continue from the reference's state (first run)
        ref.set_planes(0, rp); ref.set_scalars(0, rc_); ref.set_scalars(2, rsel); ref.set_rng(rrng)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "prior_ub_probe.json"), "w"), indent=1)
