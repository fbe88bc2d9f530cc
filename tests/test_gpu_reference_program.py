"""The reference PROGRAM (its kernels, host class, ProcessProblem, JBU and RunFusion compiled from
where they lie, oracle/ref_full_harness.cu) against the product on the same on-disk scene:
depth maps and fused point clouds.  pytest -m gpu."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
if not torch.cuda.is_available():
    pytest.skip("no CUDA device", allow_module_level=True)

import refbind  # noqa: E402
from acmmp_b200 import host  # noqa: E402
from acmmp_b200.scene import make_scene  # noqa: E402

needs_full = pytest.mark.skipif(not refbind.full_available(), reason="oracle/_ref/libacmmp_ref_full.so not built")


def ours(sc, seed):
    scan = host.Scan(sc["cams"], sc["pairs"])
    for v, im in enumerate(sc["images"]):
        scan.set_image(v, np.clip(np.rint(im), 0, 255).astype(np.float32))
    scan.run_all(seed)
    scan.sync()
    maps = [scan.get_view(v) for v in range(len(sc["cams"]))]
    pts, per = host.RunFusion(sc["cams"], sc["pairs"], [m[0] for m in maps], [m[1] for m in maps])
    scan.close()
    return maps, pts, per


@needs_full
@pytest.mark.parametrize("size", [(320, 240), (1280, 960)])     # one scale / two scales (JBU + hierarchy)
def test_reference_program_and_product_agree(tmp_path, size):
    w, h = size
    sc = make_scene(5, w, h, 4, seed=8)
    dense = str(tmp_path / "dense")
    refbind.write_dense_folder(dense, sc["cams"], sc["images"], sc["pairs"])
    t_pm, t_fu = refbind.full_run(dense, seed=99)
    ref_maps = [refbind.read_dmb(os.path.join(dense, "ACMMP", "2333_%08d" % i, "depths_geom.dmb")) for i in range(5)]
    ref_xyz, _ = refbind.read_ply(os.path.join(dense, "ACMMP", "ACMMP_model.ply"))
    maps, pts, per = ours(sc, 99)
    print("reference program: patchmatch %.2f s, fusion %.2f s; points ref %d ours %d" % (t_pm, t_fu, len(ref_xyz), len(pts)))
    for i in range(5):
        gt = sc["depths"][i]
        inner = (slice(8, -8), slice(8, -8))
        acc_ref = (np.abs(ref_maps[i] - gt)[inner] / gt[inner] < 0.01).mean()
        acc_our = (np.abs(maps[i][0] - gt)[inner] / gt[inner] < 0.01).mean()
        agree = (np.abs(ref_maps[i] - maps[i][0])[inner] / gt[inner] < 0.001).mean()
        print("view %d: acc<1%% ref %.4f ours %.4f, depth within 0.1%% of each other %.4f" % (i, acc_ref, acc_our, agree))
        assert acc_ref >= 0.9 and abs(acc_our - acc_ref) <= 0.01
        assert agree >= (0.93 if w < 1000 else 0.97)
    # fused clouds: different random streams and the reference's races move individual pixels; the count is a
    # scene-level quantity.  Yardstick: the reference program against ITSELF with another seed, on the same folder.
    dense2 = str(tmp_path / "dense2")
    refbind.write_dense_folder(dense2, sc["cams"], sc["images"], sc["pairs"])
    refbind.full_run(dense2, seed=100)
    ref2_xyz, _ = refbind.read_ply(os.path.join(dense2, "ACMMP", "ACMMP_model.ply"))
    d_our, d_ref = abs(len(pts) - len(ref_xyz)) / len(ref_xyz), abs(len(ref2_xyz) - len(ref_xyz)) / len(ref_xyz)
    print("fused points %dx%d: reference %d, reference (other seed) %d (%.3f %%), product %d (%.3f %%)"
          % (w, h, len(ref_xyz), len(ref2_xyz), 100 * d_ref, len(pts), 100 * d_our))
    # north_star: +-0.1 % at the two-scale size (measured on B200: 0.08 %), never further away than the reference is
    # from itself plus that; the tiny one-scale case is noisier (0.40 %)
    assert d_our <= (0.001 if w >= 1000 else 0.006) + d_ref, (len(pts), len(ref_xyz), len(ref2_xyz))
