#!/usr/bin/env python
"""Full-ACMMP schedule on a synthetic scan, then depth-map fusion of its final maps on the
device; optionally checks the fused point count against the CPU restatement (oracle)."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from acmmp_b200 import host  # noqa: E402
from acmmp_b200.scene import make_scene  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--views", type=int, default=12)
ap.add_argument("--width", type=int, default=1600)
ap.add_argument("--height", type=int, default=1200)
ap.add_argument("--nsrc", type=int, default=10)
ap.add_argument("--oracle", action="store_true", help="also run the CPU restatement (slow at full size)")
ap.add_argument("--repeat", type=int, default=2)
args = ap.parse_args()

sc = make_scene(args.views, args.width, args.height, args.nsrc, seed=1, device="cuda",
                span_deg=(10.0, 8.0) if args.views <= 11 else None)
scan = host.Scan(sc["cams"], sc["pairs"])
for v, im in enumerate(sc["images"]):
    scan.set_image(v, im)
t0 = time.time()
scan.run_all(1234)
scan.sync()
print("schedule wall s", round(time.time() - t0, 2), flush=True)
fus = host.Fusion(sc["cams"], sc["pairs"])
for v in range(args.views):
    d, n, _ = scan.view_device(v)
    fus.set_view(v, d, n, on_device=True)
for r in range(args.repeat):
    t0 = time.time()
    npts = fus.run(0.3, 1)
    wall = time.time() - t0
    st = fus.stats()
    print(json.dumps({"points": npts, "wall_s": round(wall, 3), "score_ms": round(st["score_ms"], 1), "greedy_ms": round(st["greedy_ms"], 1),
                      "emit_ms": round(st["emit_ms"], 1), "rounds_max": max(st["rounds"]), "rounds_mean": round(float(np.mean(st["rounds"])), 1),
                      "points_per_s": round(npts / wall)}), flush=True)
pts, per = fus.points()
print("per view", per[:8], "...")
if args.oracle:
    import oraclebind as ob
    maps = [scan.get_view(v) for v in range(args.views)]
    t0 = time.time()
    cnt = ob.fusion(sc["cams"], [m[0] for m in maps], [m[1] for m in maps], sc["pairs"], 0.3, 1)
    print(json.dumps({"oracle_points": cnt, "oracle_wall_s": round(time.time() - t0, 1), "rel_diff": abs(cnt - npts) / max(cnt, 1)}), flush=True)
