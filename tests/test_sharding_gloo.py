"""world_size-2 test of the N>1 host logic on CPU (gloo): ownership, slot layout and
the stage-boundary all-gather deliver every view's depth map to every rank, byte
for byte, including the uneven 49-views-over-2-ranks case."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from acmmp_b200 import sharding


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n_views, stride, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    slots = sharding.num_slots(n_views, world)
    buf = torch.full((world * slots * stride,), float("nan"))
    for v in sharding.owned_views(n_views, rank, world):
        s = sharding.slot_index(v, world, slots)
        buf[s * stride:(s + 1) * stride] = torch.arange(stride, dtype=torch.float32) + 1000.0 * v
    sharding.all_gather_segments(buf, rank, world, dist)
    ok = True
    for v in range(n_views):
        s = sharding.slot_index(v, world, slots)
        exp = torch.arange(stride, dtype=torch.float32) + 1000.0 * v
        ok = ok and bool((buf[s * stride:(s + 1) * stride] == exp).all())
    np.save(os.path.join(out_dir, f"ok_{rank}.npy"), np.array([ok]))
    dist.barrier()
    dist.destroy_process_group()


def test_ownership_and_slots_partition_the_views():
    for n, world in [(49, 1), (49, 2), (49, 8), (5, 3), (300, 8)]:
        slots = sharding.num_slots(n, world)
        seen = {}
        for r in range(world):
            for v in sharding.owned_views(n, r, world):
                assert v not in seen
                seen[v] = sharding.slot_index(v, world, slots)
                assert r * slots <= seen[v] < (r + 1) * slots       # inside the owner's segment
        assert sorted(seen) == list(range(n))
        assert len(set(seen.values())) == n


def test_all_gather_of_depth_segments_world2(tmp_path):
    world, port = 2, _free_port()
    mp.spawn(_worker, args=(world, port, 49, 37, str(tmp_path)), nprocs=world, join=True)
    for r in range(world):
        assert np.load(tmp_path / f"ok_{r}.npy")[0]
