"""Pair scoring of the input converter at c5 scale (300 views, 20 k keypoints each): device time of
acmmp_pair_scores next to the oracle restatement of the reference's calc_score on a sample of pairs.
    python tools/pairs_bench.py [n_images] [n_points] [sample_pairs]"""
import json
import os
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import colmap_model as cm  # noqa: E402
import convert_util as cu  # noqa: E402
from acmmp_b200 import convert  # noqa: E402


def main():
    V = int(sys.argv[1]) if len(sys.argv) > 1 else 300
    P = int(sys.argv[2]) if len(sys.argv) > 2 else 200000
    S = int(sys.argv[3]) if len(sys.argv) > 3 else 24
    m = cm.make_model(seed=1, n_images=V, n_points=P, visibility=0.1, keypoints_without_point=2000, duplicates=10)
    with tempfile.TemporaryDirectory() as d:
        cm.write_binary(m, d)
        t0 = time.time()
        model = convert.read_model(d, ".bin")
        t_read = time.time() - t0
    E = convert.extrinsic_matrices(model)
    convert.pair_scores(model, E)                       # warm-up (context, sort temp)
    t0 = time.time()
    score, ms = convert.pair_scores(model, E, return_ms=True)
    wall = time.time() - t0
    rng = np.random.default_rng(0)
    pairs = [tuple(sorted(int(x) for x in rng.choice(V, 2, replace=False))) for _ in range(S)]
    t0 = time.time()
    ref = cu.oracle_scores(model, E, pairs=pairs)
    t_cpu = time.time() - t0
    same = all(score[i, j] == ref[i, j] for i, j in pairs)
    npairs = V * (V - 1) // 2
    print(json.dumps({"images": V, "points": P, "keypoints": int(model.obs_offsets[-1]), "pairs": npairs,
                      "device_ms": round(ms, 3), "call_wall_ms": round(wall * 1e3, 1), "read_model_s": round(t_read, 2),
                      "oracle_s_per_pair": round(t_cpu / S, 4), "oracle_all_pairs_s_extrapolated": round(t_cpu / S * npairs, 1),
                      "sample_pairs_equal": bool(same), "pairs_per_s_device": round(npairs / (ms * 1e-3), 1)}))


if __name__ == "__main__":
    main()
