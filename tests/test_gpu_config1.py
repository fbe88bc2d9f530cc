"""BASELINE.json configs[0] end to end at its stated size: a synthetic 5-view 640 x 480 scene as a COLMAP sparse model
-> the converter (python -m acmmp_b200.convert = colmap2mvsnet_acm.py) -> build/acmmp_main (= main_ACMMP.cpp) -> the
reference's OWN RunFusion reading the output folder; beside it the reference PROGRAM on the same converted folder
(its depth maps are the oracle).  pytest -m gpu."""
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
if not torch.cuda.is_available():
    pytest.skip("no CUDA device", allow_module_level=True)

import colmap_model as cm  # noqa: E402
import refbind  # noqa: E402
from acmmp_b200.scene import make_scene  # noqa: E402

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "build", "acmmp_main")


def colmap_model_of(sc, n_points=4000, seed=0):
    """sparse model of a make_scene scene: points on the true surface, observed wherever they are visible"""
    rng = np.random.default_rng(seed)
    w, h, n = sc["width"], sc["height"], len(sc["cams"])
    rig = sc["rig"]
    K = rig[0][0]
    pts = []
    for v in range(n):                                        # back-project random pixels of every view
        _, R, t = rig[v]
        xs, ys = rng.integers(8, w - 8, n_points // n), rng.integers(8, h - 8, n_points // n)
        d = sc["depths"][v][ys, xs].astype(np.float64)
        Xc = np.stack([d * (xs - K[0, 2]) / K[0, 0], d * (ys - K[1, 2]) / K[1, 1], d], 1)
        pts.append((R.T @ (Xc - t).T).T)
    pts = np.concatenate(pts)
    images = []
    for v in range(n):
        _, R, t = rig[v]
        Xc = (R @ pts.T).T + t
        x, y = K[0, 0] * Xc[:, 0] / Xc[:, 2] + K[0, 2], K[1, 1] * Xc[:, 1] / Xc[:, 2] + K[1, 2]
        xi, yi = np.rint(x).astype(int), np.rint(y).astype(int)
        ok = (xi >= 0) & (xi < w) & (yi >= 0) & (yi < h) & (Xc[:, 2] > 0)
        vis = np.zeros(len(pts), bool)
        vis[ok] = np.abs(sc["depths"][v][yi[ok], xi[ok]] - Xc[ok, 2]) < 0.01 * Xc[ok, 2]      # not occluded
        idx = np.flatnonzero(vis)
        images.append(dict(id=v + 1, qvec=cm._quat_from_R(R), tvec=t, camera_id=1, name="view_%02d.jpg" % v,
                           xys=np.stack([x[idx], y[idx]], 1), point3D_ids=(idx + 1).astype(np.int64)))
    return dict(cameras={1: ("PINHOLE", w, h, [float(K[0, 0]), float(K[1, 1]), float(K[0, 2]), float(K[1, 2])])}, images=images,
                point_ids=np.arange(1, len(pts) + 1, dtype=np.int64), xyz=pts, rgb=rng.integers(0, 256, (len(pts), 3)), error=rng.random(len(pts)))


@pytest.mark.skipif(not (refbind.host_available() and refbind.full_available()), reason="oracle/_ref not built")
def test_config1_converter_cli_and_reference_fusion_640x480(tmp_path):
    if not os.path.exists(EXE):
        pytest.skip("build/acmmp_main not built (make cli)")
    n, w, h = 5, 640, 480
    sc = make_scene(n, w, h, 4, seed=0)
    colmap = str(tmp_path / "colmap")
    os.makedirs(os.path.join(colmap, "images"))
    model = colmap_model_of(sc)
    cm.write_text(model, os.path.join(colmap, "sparse"))
    for v in range(n):                                        # the raw container under the .jpg name (no codec in the stand-in OpenCV)
        img = np.clip(np.rint(sc["images"][v]), 0, 255).astype(np.uint8)
        with open(os.path.join(colmap, "images", "view_%02d.jpg" % v), "wb") as f:
            f.write(b"ACMRAW1 %d %d 1\n" % img.shape + img.tobytes())
    dense = str(tmp_path / "dense")
    out = subprocess.run([sys.executable, "-m", "acmmp_b200.convert", "--dense_folder", colmap, "--save_folder", dense], capture_output=True, text=True, cwd=ROOT)
    assert out.returncode == 0, out.stderr
    pair = open(os.path.join(dense, "pair.txt")).read().split()
    assert int(pair[0]) == n and len(os.listdir(os.path.join(dense, "cams"))) == n
    out = subprocess.run([EXE, dense, "--seed", "5"], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr + out.stdout
    ours_xyz, _ = refbind.read_ply(os.path.join(dense, "ACMMP", "ACMMP_model.ply"))
    ours_depth = [refbind.read_dmb(os.path.join(dense, "ACMMP", "2333_%08d" % i, "depths_geom.dmb")) for i in range(n)]
    # 1. the reference's own RunFusion on the files the CLI wrote: the same cloud
    import ctypes as C
    lib = C.CDLL(refbind.REF_HOST_PATHS["ieee"])
    lib.ref_host_run_fusion.argtypes = [C.c_char_p, C.c_char_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int),
                                        C.c_int, C.c_float, C.c_int]
    tok = [float(x) for x in pair[1:]]
    ids, offs, flat, k = [], [0], [], 0
    for _ in range(n):
        ids.append(int(tok[k])); m = int(tok[k + 1]); k += 2
        for j in range(m):
            if tok[k + 2 * j + 1] > 0:
                flat.append(int(tok[k + 2 * j]))
        k += 2 * m
        offs.append(len(flat))
    assert ids == list(range(n)) and min(b - a for a, b in zip(offs, offs[1:])) >= 2
    outdir = os.path.join(dense, "ACMMP")
    rc = lib.ref_host_run_fusion(dense.encode(), outdir.encode(), n, (C.c_int * n)(*ids), (C.c_int * (n + 1))(*offs), (C.c_int * len(flat))(*flat), 1, 0.3, 1)
    assert rc == 0
    ref_fused, _ = refbind.read_ply(os.path.join(outdir, "ACMMP_model.ply"))
    assert len(ours_xyz) > 300000
    assert abs(len(ref_fused) - len(ours_xyz)) <= 1e-3 * len(ref_fused), (len(ref_fused), len(ours_xyz))     # north_star: 0.1 %
    # 2. the reference PROGRAM on the same converted folder: its depth maps are the oracle
    refbind.full_run(dense, seed=99, output_dir="/REF")
    ref_xyz, _ = refbind.read_ply(os.path.join(dense, "REF", "ACMMP_model.ply"))
    inner = (slice(8, -8), slice(8, -8))
    for i in range(n):
        ref = refbind.read_dmb(os.path.join(dense, "REF", "2333_%08d" % i, "depths_geom.dmb"))
        gt = sc["depths"][i]
        acc_ref = (np.abs(ref - gt)[inner] / gt[inner] < 0.01).mean()
        acc_our = (np.abs(ours_depth[i] - gt)[inner] / gt[inner] < 0.01).mean()
        agree = (np.abs(ref - ours_depth[i])[inner] / gt[inner] < 0.001).mean()
        print("view %d: depth<1%% ref %.4f ours %.4f; within 0.1%% of each other %.4f" % (i, acc_ref, acc_our, agree))
        assert acc_ref >= 0.95 and abs(acc_our - acc_ref) <= 0.005
        assert agree >= 0.97
    print("fused points: reference program %d, product %d (%+.3f %%)" % (len(ref_xyz), len(ours_xyz), 100.0 * (len(ours_xyz) - len(ref_xyz)) / len(ref_xyz)))
    assert abs(len(ours_xyz) - len(ref_xyz)) <= 0.005 * len(ref_xyz), (len(ours_xyz), len(ref_xyz))
