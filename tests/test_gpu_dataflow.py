"""The dataflow schedule (acmmp_scan_run_dataflow: one call, (stage, view) tasks in rounds, records instead of exchange
buffers) must give the staged schedule's maps bit for bit - same seeds, Jacobi geom-1 order - on one GPU; on a box with
two GPUs, two real ranks exchanging the records with grouped ncclBroadcast must give them too.  pytest -m gpu."""
import threading

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
if not torch.cuda.is_available():
    pytest.skip("no CUDA device", allow_module_level=True)

from acmmp_b200 import host  # noqa: E402
from acmmp_b200.scene import make_scene  # noqa: E402


def load(scan, sc):
    for v, im in enumerate(sc["images"]):
        scan.set_image_u8(v, np.clip(np.rint(im), 0, 255).astype(np.uint8))


def staged(sc, seed, **kw):
    scan = host.Scan(sc["cams"], sc["pairs"], sequential_geom=False)
    if kw:
        scan.set_schedule(**kw)
    load(scan, sc)
    photo = None
    for level in range(scan.num_levels()):
        scan.begin_level(level)
        for stage in range(1 if kw.get("photometric_only") else 3):
            scan.run_stage(stage, seed)
            if level == scan.num_levels() - 1 and stage == 0:
                photo = [scan.get_view(v)[0] for v in range(len(sc["cams"]))]
    out = [scan.get_view(v) for v in range(len(sc["cams"]))]
    scan.close()
    return out, photo


def same(a, b):
    return a.shape == b.shape and (a.view(np.uint32) == b.view(np.uint32)).all()


@pytest.mark.parametrize("shape", [(6, 1024, 768, 3), (5, 320, 240, 4)])
def test_dataflow_equals_staged_jacobi_bit_for_bit(shape):
    n, w, h, k = shape
    sc = make_scene(n, w, h, k, seed=3, device="cuda", span_deg=(10.0, 8.0))
    ref, ref_photo = staged(sc, 11)
    scan = host.Scan(sc["cams"], sc["pairs"])
    load(scan, sc)
    scan.keep_photometric(True)
    for rep in range(2):                                      # a second pass on the same scan: records and level store are reused
        scan.run_dataflow(11)
        st = scan.dataflow_stats()
        assert st["rounds"] == scan.num_levels() * 3 * n and st["lag"] == 1
        for v in range(n):
            d, nr, c = scan.get_view(v)
            assert same(d, ref[v][0]) and same(nr, ref[v][1]) and same(c, ref[v][2]), (rep, v)
            assert same(scan.photometric_depth(v), ref_photo[v]), (rep, v)
        scan.reset()
    # the staged calls still work on the same scan afterwards
    scan.set_sequential_geom(False)
    for level in range(scan.num_levels()):
        scan.begin_level(level)
        for stage in range(3):
            scan.run_stage(stage, 11)
    for v in range(n):
        d, nr, c = scan.get_view(v)
        assert same(d, ref[v][0]) and same(nr, ref[v][1]) and same(c, ref[v][2]), v
    scan.close()


def test_dataflow_photometric_only_schedule():
    sc = make_scene(5, 320, 240, 4, seed=4, device="cuda")
    ref, _ = staged(sc, 5, single_scale=True, photometric_only=True)
    scan = host.Scan(sc["cams"], sc["pairs"])
    scan.set_schedule(single_scale=True, photometric_only=True)
    load(scan, sc)
    scan.run_dataflow(5)
    for v in range(5):
        d, nr, c = scan.get_view(v)
        assert same(d, ref[v][0]) and same(nr, ref[v][1]) and same(c, ref[v][2]), v
    scan.close()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two CUDA devices")
def test_dataflow_two_ranks_over_nccl_equals_staged():
    n, w, h, k = 7, 1024, 768, 4
    sc = make_scene(n, w, h, k, seed=6, device="cuda", span_deg=(10.0, 8.0))
    ref, _ = staged(sc, 9)
    world = 2
    scans = [host.Scan(sc["cams"], sc["pairs"], device=r, rank=r, world=world) for r in range(world)]
    for s in scans:
        load(s, sc)
    host.Scan.comm_init_all(scans)
    errors = [None] * world

    def drive(r):
        try:
            scans[r].run_dataflow(9)
            scans[r].sync()
        except Exception as e:                                # noqa: BLE001
            errors[r] = e

    threads = [threading.Thread(target=drive, args=(r,)) for r in range(world)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert errors == [None] * world, errors
    assert scans[0].dataflow_stats()["rounds"] <= -(-6 * n // world) + 2
    for r in range(world):                                     # EVERY rank holds every view's final maps
        for v in range(n):
            d, nr, c = scans[r].get_view(v)
            assert same(d, ref[v][0]) and same(nr, ref[v][1]) and same(c, ref[v][2]), (r, v)
    for s in scans:
        s.close()
