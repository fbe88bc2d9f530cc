#!/usr/bin/env python
"""One DTU-shape view (1600x1200, 10 sources) through RunPatchMatch: the short
command that ncu wraps (profiles/), also prints per-phase device times."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from acmmp_b200 import host  # noqa: E402
from acmmp_b200.scene import make_scene  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--width", type=int, default=1600)
ap.add_argument("--height", type=int, default=1200)
ap.add_argument("--nsrc", type=int, default=10)
ap.add_argument("--runs", type=int, default=2)
ap.add_argument("--device", default="cuda")
ap.add_argument("--tex", default="auto", choices=["auto", "r32f"], help="texel format of the views (auto: R8 for 8-bit images)")
args = ap.parse_args()
sc = make_scene(args.nsrc + 1, args.width, args.height, args.nsrc, seed=1, device=args.device, span_deg=(10.0, 8.0))
ref = (args.nsrc + 1) // 2
ids = [ref] + sc["pairs"][ref]
m = host.ACMMP(0)
from acmmp_b200 import lib as L  # noqa: E402
m.SetTextureFormat(L.TEX_AUTO if args.tex == "auto" else L.TEX_R32F)
m.InputInitialization([sc["images"][i] for i in ids], [sc["cams"][i] for i in ids])
for r in range(args.runs):
    m.RunPatchMatch(1234)
    t = m.last_timing()
    planes, cost = m.GetPlaneHypotheses(), m.GetCosts()
    import zlib
    t["crc"] = zlib.crc32(planes.tobytes()) ^ zlib.crc32(cost.tobytes())
    print(json.dumps(t))
print("launches", m.launch_count())
