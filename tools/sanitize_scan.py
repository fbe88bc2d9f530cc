#!/usr/bin/env python
"""Workload for compute-sanitizer: the whole schedule - photometric + planar prior,
geom x2 - over BASELINE configs[0]'s shape (5 views, 640 x 480) plus fusion, through the C ABI.

    compute-sanitizer --tool racecheck  python tools/sanitize_scan.py     (shared-memory hazards)
    compute-sanitizer --tool initcheck  python tools/sanitize_scan.py     (reads of uninitialised global memory)
    compute-sanitizer --tool memcheck   python tools/sanitize_scan.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from acmmp_b200 import host  # noqa: E402
from acmmp_b200.scene import make_scene  # noqa: E402

w, h = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (640, 480)
sc = make_scene(5, w, h, 4, seed=0)
scan = host.Scan(sc["cams"], sc["pairs"])
for v, im in enumerate(sc["images"]):
    scan.set_image_u8(v, np.clip(np.rint(im), 0, 255).astype(np.uint8))
scan.run_all(seed=7)
scan.sync()
maps = [scan.get_view(v) for v in range(5)]
pts, per = host.RunFusion(sc["cams"], sc["pairs"], [m[0] for m in maps], [m[1] for m in maps])
gt = sc["depths"][0]
ok = float((np.abs(maps[0][0] - gt) / gt < 0.01)[8:-8, 8:-8].mean())
print(f"scan of 5 views {w}x{h} done: view 0 depth within 1 % on {ok:.3f}, fused points {len(pts)}")
scan.close()
