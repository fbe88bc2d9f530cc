"""Two real GPUs (pytest -m gpu on a box with >= 2 devices; skipped otherwise): the sharded scan with the library's own
NCCL exchange - `acmmp_main --gpus 2` (one thread + one resident scan per device, ncclCommInitAll, ncclAllGather on the
scans' streams) - must write exactly the maps of the same scan run as two logical shards on one GPU."""
import os
import subprocess

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
if not torch.cuda.is_available() or torch.cuda.device_count() < 2:
    pytest.skip("needs two CUDA devices", allow_module_level=True)

import refbind  # noqa: E402
from acmmp_b200 import host  # noqa: E402
from acmmp_b200.scene import make_scene  # noqa: E402

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "build", "acmmp_main")


class DevPtr:
    def __init__(self, ptr, n):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": "<f4", "data": (int(ptr), False), "version": 2}


def logical_shards(sc, world, seed):
    scans = [host.Scan(sc["cams"], sc["pairs"], rank=r, world=world) for r in range(world)]
    for s in scans:
        for v, im in enumerate(sc["images"]):
            s.set_image_u8(v, np.clip(np.rint(im), 0, 255).astype(np.uint8))
    for level in range(scans[0].num_levels()):
        for s in scans:
            s.begin_level(level)
        for stage in range(3):
            for s in scans:
                s.run_stage(stage, seed)
                s.sync()
            infos = [s.exchange_info() for s in scans]
            seg = infos[0][1] * infos[0][2]
            bufs = [torch.as_tensor(DevPtr(i[0], seg * world), device="cuda") for i in infos]
            for r in range(world):
                for q in range(world):
                    if q != r:
                        bufs[q][r * seg:(r + 1) * seg].copy_(bufs[r][r * seg:(r + 1) * seg])
            torch.cuda.synchronize()
    out = {}
    for s in scans:
        for v in s.owned():
            out[v] = s.get_view(v)
        s.close()
    return out


def test_cli_two_gpus_equals_two_logical_shards(tmp_path):
    if not os.path.exists(EXE):
        pytest.skip("build/acmmp_main not built (make cli)")
    assert host.nccl_version() > 0
    n, w, h = 5, 1280, 960                                    # two scales: JBU + hierarchy across the exchange
    sc = make_scene(n, w, h, 4, seed=8, device="cuda", span_deg=(10.0, 8.0))
    dense = str(tmp_path / "dense")
    refbind.write_dense_folder(dense, sc["cams"], sc["images"], sc["pairs"])
    out = subprocess.run([EXE, dense, "--seed", "7", "--gpus", "2"], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr + out.stdout
    ref = logical_shards(sc, 2, 7)
    for v in range(n):
        folder = os.path.join(dense, "ACMMP", "2333_%08d" % v)
        depth = refbind.read_dmb(os.path.join(folder, "depths_geom.dmb"))
        normal = refbind.read_dmb(os.path.join(folder, "normals.dmb"))
        cost = refbind.read_dmb(os.path.join(folder, "costs.dmb"))
        assert (depth.view(np.uint32) == ref[v][0].view(np.uint32)).all(), v
        assert (normal.view(np.uint32) == ref[v][1].view(np.uint32)).all(), v
        assert (cost.view(np.uint32) == ref[v][2].view(np.uint32)).all(), v
        gt = sc["depths"][v]
        assert (np.abs(depth - gt)[8:-8, 8:-8] / gt[8:-8, 8:-8] < 0.01).mean() >= 0.93
    xyz, _ = refbind.read_ply(os.path.join(dense, "ACMMP", "ACMMP_model.ply"))
    assert len(xyz) > 500000
