"""View sharding arithmetic and the stage-boundary depth-map exchange.

Host-side plumbing only.  Mirrors scan_owns / scan_slot_index of
csrc/capi_scan.cu.inc: rank r of `world` owns the views v with v % world == r; the
exchange buffer is [world segments][slots][stride] floats and view v lives in
segment v % world, slot v // world.  After every stage each rank contributes its
own segment and receives the others - only depth maps cross views in ACMMP
(ACMMP.cpp:608-635).

On the GPU that exchange is the LIBRARY's: acmmp_scan_all_gather (ncclAllGather in
place on the scan's stream, csrc/capi_nccl.cu.inc) or, without the stage barrier,
acmmp_scan_run_dataflow.  all_gather_segments below is the same exchange over any
torch.distributed backend: what tests/test_sharding_gloo.py runs with gloo on CPU
to cover the rank / slot arithmetic, and a transport for callers without NCCL
(acmmp_scan_exchange hands out the buffer)."""


def owned_views(n_views, rank, world):
    return [v for v in range(n_views) if v % world == rank]


def num_slots(n_views, world):
    return (n_views + world - 1) // world


def slot_index(view, world, slots):
    return (view % world) * slots + view // world


def all_gather_segments(buf, rank, world, dist):
    """buf: 1-D tensor of world * seg elements whose [rank*seg, (rank+1)*seg) part is
    valid; on return every part is.  `dist` = torch.distributed (initialised)."""
    if world == 1:
        return buf
    seg = buf.numel() // world
    mine = buf[rank * seg:(rank + 1) * seg].clone()
    dist.all_gather_into_tensor(buf, mine)
    return buf
