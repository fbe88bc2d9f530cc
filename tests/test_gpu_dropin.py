"""The drop-in proof: the reference's UNCHANGED ACMMP.h and acmmp_definitions.cpp (ProcessProblem,
JointBilateralUpsampling, RunFusion, compiled from where they lie) linked against oracle/dropin/ACMMP_dropin.cpp -
class ACMMP / RunJBU / the .dmb and camera helpers re-implemented over the C ABI of libacmmp_b200.so - and driven by
the reference's main() loop.  Its output folder is compared with the product's own resident schedule (host.Scan) and
with the reference program itself.  pytest -m gpu."""
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

needs_dropin = pytest.mark.skipif(not refbind.dropin_available(), reason="oracle/_ref/libacmmp_dropin.so not built (make dropin)")


@needs_dropin
@pytest.mark.parametrize("size", [(320, 240), (1280, 960)])     # one scale / two scales (JBU + hierarchy through the drop-in)
def test_reference_driver_over_the_c_abi(tmp_path, size):
    w, h = size
    n = 5
    sc = make_scene(n, w, h, 4, seed=8)
    dense = str(tmp_path / "dense")
    refbind.write_dense_folder(dense, sc["cams"], sc["images"], sc["pairs"])
    refbind.dropin_run(dense, seed=99, fuse=True)
    out = os.path.join(dense, "ACMMP")
    maps = []
    for i in range(n):
        folder = os.path.join(out, "2333_%08d" % i)
        depth = refbind.read_dmb(os.path.join(folder, "depths_geom.dmb"))
        normal = refbind.read_dmb(os.path.join(folder, "normals.dmb"))
        cost = refbind.read_dmb(os.path.join(folder, "costs.dmb"))
        photo = refbind.read_dmb(os.path.join(folder, "depths.dmb"))
        assert depth.shape == (h, w) and normal.shape == (h, w, 3) and cost.shape == (h, w) and photo.shape == (h, w)
        maps.append((depth, normal))
    xyz, _ = refbind.read_ply(os.path.join(out, "ACMMP_model.ply"))      # written by the reference's own RunFusion
    # the product's own schedule on the same images (other random streams: the drop-in numbers its runs in call order)
    scan = host.Scan(sc["cams"], sc["pairs"])
    for v, im in enumerate(sc["images"]):
        scan.set_image(v, np.clip(np.rint(im), 0, 255).astype(np.float32))
    scan.run_all(99)
    scan.sync()
    ours = [scan.get_view(v) for v in range(n)]
    pts, _ = host.RunFusion(sc["cams"], sc["pairs"], [m[0] for m in ours], [m[1] for m in ours])
    scan.close()
    inner = (slice(8, -8), slice(8, -8))
    for i in range(n):
        gt = sc["depths"][i]
        acc_drop = (np.abs(maps[i][0] - gt)[inner] / gt[inner] < 0.01).mean()
        acc_ours = (np.abs(ours[i][0] - gt)[inner] / gt[inner] < 0.01).mean()
        agree = (np.abs(maps[i][0] - ours[i][0])[inner] / gt[inner] < 0.001).mean()
        print("view %d: acc<1%% drop-in %.4f scan %.4f, within 0.1%% of each other %.4f" % (i, acc_drop, acc_ours, agree))
        assert acc_drop >= 0.9 and abs(acc_drop - acc_ours) <= 0.01
        assert agree >= (0.93 if w < 1000 else 0.97)
        assert np.abs(np.linalg.norm(maps[i][1], axis=-1) - 1)[inner].max() < 1e-3
    print("fused points: drop-in %d, scan %d" % (len(xyz), len(pts)))
    assert abs(len(xyz) - len(pts)) <= 0.01 * len(pts), (len(xyz), len(pts))
