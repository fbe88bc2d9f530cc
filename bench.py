#!/usr/bin/env python
"""bench.py - views/s of full ACMMP (multi-scale + planar prior + geometric
consistency) on a synthetic DTU-shaped scan, 1..8 B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

One "step" = the whole schedule of the reference's main() (main_ACMMP.cpp:89-176)
over the 49-view scan, sharded by view across the N ranks (strong scaling: the
scan is fixed, depth maps are all-gathered over NCCL after every stage).
Prints ONE JSON line (rank 0).  See DESIGN.md "Measurement" for the definitions.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

FLOP_PER_NCC = 1538.0          # 36 x 38 + 170
# dram__bytes_read.sum + dram__bytes_write.sum of one sweep_kernel launch at 1600x1200, 10 sources
# (ncu --set full, profiles/r1g_sweep_init_full_summary.txt: 311 MB read + 121 MB written, average of 8
# launches); ~1 % of what HBM could move in that time
SWEEP_DRAM_BYTES_PER_LAUNCH = 432e6
# the same for sweep_kernel<R8> (finest level, R8 unorm texels)
SWEEP_DRAM_BYTES_PER_LAUNCH_U8 = 250e6   # profiles/r2ac_sweep_full_summary.txt (4 CTAs/SM): 160 MB read + 90 MB written (r1o, 3 CTAs: 148 + 67)
HYP_PER_RUN = 29               # 1 (init) + 2 x 14; 30 for the hierarchy (upsample-init) run


# BASELINE.json `configs`: the workload each --config runs (views, width, height, sources, camera layout)
CONFIGS = {
    # configs[2], the one `metric` is quoted on (default)
    "full": dict(views=49, width=1600, height=1200, nsrc=10, layout="grid",
                 name="configs[2]: DTU-shape {W}x{H} x {n} views, {ns} sources, full ACMMP ({nl} scales x (photometric+planar prior, geom x2))",
                 metric="views/s, DTU-shape 1600x1200 49-view full ACMMP (multi-scale + planar prior + geometric consistency)"),
    # configs[1]: isolates RunPatchMatch (ACMMP.cu:1378-1456) from the schedule - the kernel-vs-kernel line
    "photometric": dict(views=49, width=1600, height=1200, nsrc=10, layout="grid",
                        name="configs[1]: DTU-shape {W}x{H} x {n} views, {ns} sources, single scale, photometric only (one RunPatchMatch per view)",
                        metric="views/s, DTU-shape 1600x1200 49-view single-scale photometric PatchMatch"),
    # configs[3]: 6048x4032 is capped to 3200 (ACMMP.h:36) -> 3 scales, 2 JBU steps
    "eth3d": dict(views=11, width=6048, height=4032, nsrc=10, layout="grid",
                  name="configs[3]: ETH3D-high-res-shape {W}x{H} x {n} views, {ns} sources, full ACMMP ({nl} scales, JBU between them)",
                  metric="views/s, ETH3D-shape 6048x4032 11-view full ACMMP (3 scales + JBU)"),
    # configs[4]
    "tnt": dict(views=300, width=1920, height=1080, nsrc=20, layout="ring",
                name="configs[4]: Tanks&Temples-shape {W}x{H} x {n} views, {ns} sources, full ACMMP ({nl} scales) + fusion",
                metric="views/s, Tanks&Temples-shape 1920x1080 300-view full ACMMP + fusion"),
}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="full", choices=sorted(CONFIGS))
    ap.add_argument("--views", type=int, default=None)
    ap.add_argument("--width", type=int, default=None)
    ap.add_argument("--height", type=int, default=None)
    ap.add_argument("--nsrc", type=int, default=None)
    ap.add_argument("--seed", type=int, default=1)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--staged", action="store_true", help="N > 1: the stage-by-stage loop (view %% N, all-gather after every stage) instead of the dataflow schedule")
    ap.add_argument("--tex-format", default="auto", choices=["auto", "r32f"],
                    help="auto: R8 unorm texels on levels that are not resized (bit-identical samples); r32f: float texels everywhere")
    ap.add_argument("--ref-views", type=int, default=4, help="views of the reference arm's kernel-only pass (bounded sample)")
    ap.add_argument("--ref-full-scan", action="store_true", help="reference arm: every view of the scan per step instead of the bounded sub-scan")
    ap.add_argument("--ref-kernel-only-child", action="store_true", help=argparse.SUPPRESS)
    ap.add_argument("--no-ref-p2", action="store_true", help="reference arm: skip the extra step with patch P2 (#define DEBUG removed)")
    args = ap.parse_args()
    preset = CONFIGS[args.config]
    for k in ("views", "width", "height", "nsrc"):
        if getattr(args, k) is None:
            setattr(args, k, preset[k])
    args.layout = preset["layout"]
    return args


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.stop, self.th = index, [], threading.Event(), None

    def _loop(self):
        while not self.stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i", str(self.index)],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            self.stop.wait(0.2)

    def __enter__(self):
        self.th = threading.Thread(target=self._loop, daemon=True)
        self.th.start()
        return self

    def __exit__(self, *a):
        self.stop.set()
        self.th.join(timeout=6)

    def summary(self):
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 2 + i and r[2 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def dist_setup(args):
    import torch
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29512")
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", local))
    return rank, local, world, dist


def make_scan_inputs(args, device):
    """Synthetic DTU-shaped scan: 49 cameras on a 7x7 cap, f = 2892 px,
    depth range as in DTU cam files, 10 nearest views as sources; rendered on the
    device, integer-valued grey images like decoded JPEGs."""
    from acmmp_b200.scene import make_scene
    sc = make_scene(args.views, args.width, args.height, args.nsrc, seed=args.seed, device=device, as_numpy=False, layout=args.layout)
    return sc


class DevPtr:
    """lets torch alias a raw device pointer (CUDA array interface)."""

    def __init__(self, ptr, n):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": "<f4", "data": (int(ptr), False), "version": 2}


def hyp_evals_per_view(scan_sizes, runs_per_level=4):
    """(w, h) per level -> hypothesis evaluations of one view: `runs_per_level` RunPatchMatch per level (4 in full
    ACMMP, 1 in the photometric-only schedule), 29 per pixel per run, 30 for the hierarchy run."""
    total = 0
    for lvl, (w, h) in enumerate(scan_sizes):
        total += w * h * (runs_per_level * HYP_PER_RUN + (1 if lvl > 0 else 0))
    return total


def run_ours(args):
    import torch
    from acmmp_b200 import host
    rank, local, world, dist = dist_setup(args)
    dev = torch.device("cuda", local)
    sc = make_scan_inputs(args, dev)
    n = args.views
    imgs_dev = [im.contiguous().float() for im in sc["images"]]
    # e2e input = what the reference reads: 8-bit grey images (cv::imread GRAYSCALE), in pinned host memory;
    # the float conversion of ACMMP.cpp:541 runs on the device (acmmp_scan_set_image_u8)
    imgs_host = [torch.empty(im.shape, dtype=torch.uint8).pin_memory() for im in imgs_dev]
    for h_, d_ in zip(imgs_host, imgs_dev):
        assert bool((d_ == d_.round()).all()) and float(d_.min()) >= 0 and float(d_.max()) <= 255
        h_.copy_(d_.to(torch.uint8))
    torch.cuda.synchronize()
    imgs_np = [t.numpy() for t in imgs_host]
    from acmmp_b200 import lib as L
    scan = host.Scan(sc["cams"], sc["pairs"], device=local, rank=rank, world=world,
                     tex_format=L.TEX_R32F if args.tex_format == "r32f" else L.TEX_AUTO)
    photo_only = args.config == "photometric"
    if photo_only:
        scan.set_schedule(single_scale=True, photometric_only=True)
    stages = (host.STAGE_PHOTO,) if photo_only else (host.STAGE_PHOTO, host.STAGE_GEOM0, host.STAGE_GEOM1)
    runs_per_level = 1 if photo_only else 4            # RunPatchMatch per view per level
    owned = scan.owned()
    # N > 1: the whole scan as a dataflow of (stage, view) tasks (acmmp_scan_run_dataflow) - no stage barrier
    dataflow = world > 1 and not args.staged
    n_levels = scan.num_levels()
    if dataflow:
        _, task_rank, n_rounds, df_lag = host.plan_dataflow(sc["pairs"], n_levels, world, photo_only)
        per = 1 if photo_only else 3
        runs_rank = [0] * n_levels               # RunPatchMatch calls of this rank per level
        for t, r in enumerate(task_rank.tolist()):
            g = t // n
            if r == rank:
                runs_rank[g // per] += (1 if photo_only else 2) if g % per == 0 else 1
    else:
        runs_rank = [(1 if photo_only else 4) * len(owned)] * n_levels
    # size of the maps a step hands back = the finest level (the 3200 cap of ACMMP.h:36 applies to eth3d)
    cap = 3200
    fmax = max(args.width, args.height)
    if fmax > cap:
        f = min(np.float32(cap) / np.float32(args.width), np.float32(cap) / np.float32(args.height))
        W, H = int(round(float(np.float32(args.width) * f))), int(round(float(np.float32(args.height) * f)))
    else:
        W, H = args.width, args.height
    outs = {v: (torch.empty((H, W), dtype=torch.float32).pin_memory().numpy(),
                torch.empty((H, W, 3), dtype=torch.float32).pin_memory().numpy(),
                torch.empty((H, W), dtype=torch.float32).pin_memory().numpy()) for v in owned}
    n_exchanges = [0]

    # multi-GPU data plane = the library's own NCCL calls on the scan's stream (acmmp_scan_all_gather: no host sync, no
    # staging copy); torch.distributed only carries the 128-byte NCCL unique id, the barrier and the max over ranks
    if world > 1:
        uid = torch.zeros(128, dtype=torch.uint8, device=dev)
        if rank == 0:
            uid.copy_(torch.frombuffer(bytearray(host.nccl_unique_id()), dtype=torch.uint8))
        dist.broadcast(uid, src=0)
        scan.comm_init(bytes(uid.cpu().numpy().tobytes()))

    def exchange(ptr, stride, slots):
        if photo_only:          # no stage reads another view's depth map in the photometric-only schedule
            return
        gather(ptr, stride, slots)

    def gather(ptr, stride, slots):
        if world == 1:
            return
        scan.all_gather()
        n_exchanges[0] += 1

    # kernel time per pyramid level (CUDA events on the launching stream, collected after each level;
    # the collection waits for the level's last kernel, which the next level's set-up does anyway)
    level_tm = {}
    level_u8 = {}

    def level_done(level):
        t = scan.timing()
        acc = level_tm.setdefault(level, {"init_ms": 0.0, "sweeps_ms": 0.0, "tail_ms": 0.0, "total_ms": 0.0})
        for k in acc:
            acc[k] += t[k]
        level_u8[level] = scan.level_is_u8()

    def step(e2e):
        scan.reset()
        for v in range(n):
            if e2e:
                scan.set_image_u8(v, imgs_np[v], sync=False)
            else:
                scan.set_image(v, imgs_dev[v].data_ptr(), on_device=True)
        if dataflow:
            scan.run_dataflow(1234)
        else:
            scan.run_all(seed=1234, exchange=exchange, level_done=level_done, stages=stages)
        if e2e:
            for v in owned:
                scan.get_view(v, out=outs[v])

    def barrier():
        scan.sync()
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()

    def timed(k, e2e):
        barrier()
        scan.mark(0)
        for _ in range(k):
            step(e2e)
        scan.mark(1)
        barrier()
        ms = torch.tensor([scan.elapsed_ms(0, 1)], device=dev, dtype=torch.float64)
        if dist is not None:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    for _ in range(args.warmup):
        step(False)
    barrier()
    scan.timing()                      # clear the kernel-time log
    scan.host_stats()
    l0 = scan.timing()["launches"]
    level_tm.clear()
    with ClockSampler(local) as clocks:
        ms_dev = timed(args.steps, False)
    if dataflow:
        for lv in range(n_levels):
            t = scan.level_stats(lv)
            level_tm[lv] = {k: t[k] for k in ("init_ms", "sweeps_ms", "tail_ms", "total_ms")}
            level_u8[lv] = t["u8"]
    trace = None
    if dataflow:                        # when each rank finished each round of the last timed step
        tr = torch.tensor(scan.dataflow_trace(), device=dev, dtype=torch.float32)
        allr = [torch.zeros_like(tr) for _ in range(world)]
        dist.all_gather(allr, tr)
        tr_all = torch.stack(allr).cpu().numpy()                     # (rank, round)
        durations = np.diff(np.concatenate([np.zeros((world, 1), np.float32), tr_all], 1), axis=1)
        trace = {"rounds": int(tr_all.shape[1]), "last_round_done_ms": [float(x) for x in tr_all[:, -1]],
                 "skew_ms_mean": float((tr_all.max(0) - tr_all.min(0)).mean()), "skew_ms_max": float((tr_all.max(0) - tr_all.min(0)).max()),
                 "note": "ms after the run's start at which a rank had finished round k (compute stream); skew = max - min over ranks per round"}
        if tr_all.shape[1] <= 64:
            trace["round_done_ms_rank0"] = [round(float(x), 2) for x in tr_all[0]]
            trace["round_ms_max_over_ranks"] = [round(float(x), 2) for x in durations.max(0)]
            trace["round_ms_min_over_ranks"] = [round(float(x), 2) for x in durations.min(0)]
    launches = scan.timing()["launches"] - l0
    hs = scan.host_stats()              # host side of the planar prior in the timed region: triangulation time, and how long the stream waited for it
    level_tm_timed = {lv: dict(t) for lv, t in level_tm.items()}
    tm = {k: sum(t[k] for t in level_tm_timed.values()) for k in ("init_ms", "sweeps_ms", "tail_ms", "total_ms")}
    step(True)                          # e2e warm-up (pinned staging paths)
    ms_e2e = timed(args.steps, True)
    sizes = []
    scan.reset()
    for lvl in range(scan.num_levels()):
        scan.begin_level(lvl)
        sizes.append(scan.view_size(0))
    scan.sync()
    # executed work: one extra, untimed step with the device-side NCC counter switched on
    scan.count_ncc(True)
    scan.ncc_count()
    ex0 = n_exchanges[0]
    ncc_level = {}

    def count_level(level):
        ncc_level[level] = scan.ncc_count()

    scan.reset()
    for v in range(n):
        scan.set_image(v, imgs_dev[v].data_ptr(), on_device=True)
    if dataflow:
        scan.run_dataflow(1234)
        scan.sync()
        for lv in range(n_levels):
            t = scan.level_stats(lv)
            ncc_level[lv] = {"init": t["init"], "sweeps": t["sweeps"]}
        exchanges_per_step = scan.dataflow_stats()["rounds"]
    else:
        scan.run_all(seed=1234, exchange=exchange, level_done=count_level, stages=stages)
        scan.sync()
        exchanges_per_step = n_exchanges[0] - ex0
    ncc = {k: sum(c[k] for c in ncc_level.values()) for k in ("init", "sweeps")}
    scan.count_ncc(False)
    fp32_tflops, mufu_gops = host.measure_fp32_peak(local)
    tex_coherent, tex_grouped = host.measure_tex_peak(local)
    tex8_coherent, tex8_grouped = host.measure_tex_peak(local, u8=True)
    # dominant kernel = the sweep kernel of the finest level (R8 texels there when the images allow it)
    top = len(sizes) - 1
    top_u8 = bool(level_u8.get(top, False))
    top_sweep_s = level_tm_timed[top]["sweeps_ms"] * 1e-3
    top_fetches = ncc_level[top]["sweeps"] * 36 * args.steps
    top_launches = runs_rank[top] * 4 * args.steps   # 4 half-sweeps per RunPatchMatch
    top_peak, top_grouped = (tex8_coherent, tex8_grouped) if top_u8 else (tex_coherent, tex_grouped)
    top_achieved = top_fetches / top_sweep_s / 1e9
    per_level = []
    for lv in range(len(sizes)):
        s_ms = level_tm_timed[lv]["sweeps_ms"]
        per_level.append({"level": lv, "size": list(sizes[lv]), "texels": "R8 unorm" if level_u8.get(lv) else "R32F",
                          "sweeps_ms_per_step": s_ms / args.steps, "init_ms_per_step": level_tm_timed[lv]["init_ms"] / args.steps,
                          "gfetch_per_s": ncc_level[lv]["sweeps"] * 36 * args.steps / (s_ms * 1e-3) / 1e9})
    hpv = hyp_evals_per_view(sizes, runs_per_level)
    views_per_s = n * args.steps / (ms_dev * 1e-3)
    e2e_views_per_s = n * args.steps / (ms_e2e * 1e-3)
    # dominant kernel: sweep_kernel (rank 0's launches in the timed region).
    #  reference formulation: per run P * 28 hypotheses * nsrc views, 1538 flop each
    #  executed: counted NCC evaluations x 36 bilinear fetches (the kernel scores the current plane and
    #  the 5 refinement candidates against the sampled views only)
    ncc_sweeps = sum(runs_rank[lv] * w * h * 28 * args.nsrc for lv, (w, h) in enumerate(sizes)) * args.steps
    sweep_s = tm["sweeps_ms"] * 1e-3
    fp32_achieved = ncc_sweeps * FLOP_PER_NCC / sweep_s / 1e12
    fetches = ncc["sweeps"] * 36 * args.steps
    tex_achieved = fetches / sweep_s / 1e9
    n_sweep_launches = 4 * sum(runs_rank) * args.steps   # runs x 4 half-sweeps
    views_rank = sum(runs_rank) / float(runs_per_level * len(sizes))   # this rank's RunPatchMatch calls, in whole-view equivalents
    h2d = sum(int(np.prod(im.shape)) for im in imgs_np)   # uint8
    d2h = len(owned) * H * W * 5 * 4
    line = {
        "metric": CONFIGS[args.config]["metric"],
        "value": views_per_s, "unit": "views/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_dev / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": CONFIGS[args.config]["name"].format(W=args.width, H=args.height, n=n, ns=args.nsrc, nl=len(sizes)),
                   "name": args.config,
                   "geom1_order": "in place (the reference's Gauss-Seidel order)" if world == 1 else "Jacobi (sharded scan)",
                   "levels": [list(s) for s in sizes],
                   "schedule": ("dataflow: %d (stage, view) tasks in %d rounds of %d, lag %d (acmmp_scan_run_dataflow)" % (len(task_rank), n_rounds, world, df_lag)) if dataflow
                               else ("stage by stage, view %% %d" % world),
                   "sharding": "tasks dealt to ranks round by round" if dataflow else f"view % {world}",
                   "exchange": ("grouped ncclBroadcast of the records [depth|normal|cost] a round wrote, issued by the library on its second stream, overlapped with the next round" if dataflow
                                else "none (no stage reads another view's maps)" if photo_only
                                else "ncclAllGather of the depth maps per stage, issued by the library on the scan's stream"),
                   "nccl_version": host.nccl_version(),
                   "texels": [p["texels"] for p in per_level],
                   "l2": "inputs (%d images of %.1f MB + per-view state) exceed the 126 MB L2; no explicit flush" % (n, args.width * args.height * 4 / 1e6),
                   "hyp_evals_per_view": hpv},
        "mpix_hyp_per_s": views_per_s * hpv / 1e6,
        "e2e": {"value": e2e_views_per_s, "unit": "views/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": ms_e2e / args.steps, "mpix_hyp_per_s": e2e_views_per_s * hpv / 1e6},
        "gpu_launches": int(launches),
        "kernel_ms_per_step": {k: tm[k] / args.steps for k in ("init_ms", "sweeps_ms", "tail_ms", "total_ms")},
        # the RunPatchMatch kernels alone (CUDA events around init / sweeps / tail of every run, this rank): what the
        # reference arm's `kernel_only` is compared with - the rest of `value` is schedule (prior construction, JBU, copies)
        "kernel_only": {"value": views_rank * args.steps / (tm["total_ms"] * 1e-3), "unit": "views/s (this rank's share of the scan's RunPatchMatch calls, in views / its kernel time)",
                        "ms_per_view": tm["total_ms"] / (views_rank * args.steps)},
        "roofline": {"bound": "tex", "kernel": "sweep_kernel<%s> at %dx%d" % ("R8" if top_u8 else "R32F", sizes[top][0], sizes[top][1]),
                     "achieved": top_achieved, "peak": top_peak, "unit": "Gfetch/s",
                     "frac": top_achieved / top_peak,
                     "traffic": (SWEEP_DRAM_BYTES_PER_LAUNCH_U8 if top_u8 else SWEEP_DRAM_BYTES_PER_LAUNCH) if (tuple(sizes[top]) == (1600, 1200) and args.nsrc == 10) else None,
                     "launches": top_launches, "avg_launch_ms": level_tm_timed[top]["sweeps_ms"] / top_launches,
                     "fetches_per_launch": top_fetches / top_launches, "peak_grouped_access": top_grouped,
                     "share_of_kernel_time": level_tm_timed[top]["sweeps_ms"] / tm["total_ms"],
                     "all_levels": {"achieved": tex_achieved, "launches": n_sweep_launches, "avg_launch_ms": tm["sweeps_ms"] / n_sweep_launches,
                                    "per_level": per_level,
                                    "peaks": {"R32F": {"coherent": tex_coherent, "grouped": tex_grouped},
                                              "R8": {"coherent": tex8_coherent, "grouped": tex8_grouped}}},
                     "fp32": {"achieved": fp32_achieved, "peak": fp32_tflops, "unit": "TFLOP/s", "frac": fp32_achieved / fp32_tflops,
                              "mufu_gops_measured": mufu_gops},
                     "ncc_evals_per_step": {"init": ncc["init"], "sweeps": ncc["sweeps"], "reference_formulation": ncc_sweeps // args.steps},
                     # not measured by this run: the binding units' utilisation of the same kernel under ncu --set full
                     "ncu": {"source": "profiles/r2ae_sweep_full_summary.txt (one 1600x1200 R8 launch on converged state)",
                             "l1tex_data_stage_busy": 0.856, "issue_slots_busy": 0.758, "t_stage_wavefronts_per_request": 1.22,
                             "tex_l1_hit_rate": 0.96, "warps_active_of_peak": 0.468},
                     "note": "texture-pipe bound (ncu: l1tex data-stage is the top unit; HBM traffic is ~0.7 % of peak, no tensor-shaped work): "
                             "achieved = bilinear fetches issued by sweep_kernel (device-counted NCC evaluations x 36) / its time from CUDA events "
                             "on the launching stream, for the launches of the finest level (the dominant kernel; all_levels has the rest); "
                             "peak = fetch rate of tex_peak_kernel measured live on this GPU with warp-coherent "
                             "coordinates on the same texel format (peak_grouped_access: same with the kernel's 8-hypotheses-per-pixel shape). fp32: SURVEY.md 8(d)'s "
                             "reference-formulation flop (NCC x 1538) against the live FP32 FMA peak. traffic = dram bytes per launch from the "
                             "ncu --set full capture in profiles/ (1600x1200 level)"},
        "clocks": clocks.summary(),
        "dataflow_trace": trace,
        "exchanges_per_step": exchanges_per_step,
        "planar_prior_host": {"delaunay_ms_per_step": hs["prior_host_ms"] / args.steps, "stream_wait_ms_per_step": hs["wait_ms"] / args.steps,
                              "note": "host triangulation of this rank's views (worker threads, overlapped with device work of the next views); "
                                      "stream_wait = time the issuing thread was blocked on a triangulation"},
    }
    # bit-level checksum of the final maps of every view (depth from this rank's exchange buffer, i.e.
    # through the NCCL gathers; normals and costs of the owned views summed over ranks): equal numbers
    # at different --gpus mean the sharded run reproduces the single-GPU maps bit for bit
    # (geom-1 in Jacobi order at every N for this pass: the unsharded default is the reference's in-place order,
    # which a sharded scan cannot follow)
    scan.set_sequential_geom(False)
    scan.reset()
    if dataflow:
        scan.run_dataflow(1234)
    else:
        scan.run_all(seed=1234, exchange=exchange, stages=stages)
        if photo_only:
            gather(*scan.exchange_info())
    scan.sync()
    csum = torch.zeros(1, dtype=torch.int64, device=dev)
    for v in range(n):
        d_, _, _ = scan.view_device(v) if (v in owned or dataflow) else (scan.exchange_view_ptr(v), None, None)
        w_, h_ = scan.view_size(v)
        csum += torch.as_tensor(DevPtr(d_, w_ * h_), device=dev).view(torch.int32).to(torch.int64).sum() * (v + 1)
    own = torch.zeros(1, dtype=torch.int64, device=dev)
    for v in owned:
        _, n_, c_ = scan.view_device(v)
        w_, h_ = scan.view_size(v)
        own += torch.as_tensor(DevPtr(n_, 3 * w_ * h_), device=dev).view(torch.int32).to(torch.int64).sum() * (v + 1)
        own += torch.as_tensor(DevPtr(c_, w_ * h_), device=dev).view(torch.int32).to(torch.int64).sum() * (v + 1)
    if dist is not None:
        dist.all_reduce(own, op=dist.ReduceOp.SUM)
    line["result_checksum"] = {"depth_all_views": int(csum.item()), "normals_costs": int(own.item())}
    # next row of the scope table: depth-map fusion of this scan's final maps on the
    # device, beside the CPU restatement of the reference's RunFusion on the same maps (outside the
    # timed region; not part of `value`).  Sharded runs: fusion is sequential over views, so the normals
    # are sent to rank 0 (depths are there already, in the exchange buffer) and rank 0 fuses
    # ("replicas only" for this step).
    gathered = {}
    if world > 1 and not dataflow:                  # owner -> rank 0 through the library (ncclSend / ncclRecv on the scan's stream)
        for v in range(n):
            w_, h_ = scan.view_size(v)
            if rank == 0 and v % world != 0:
                gathered[v] = (torch.empty(3 * w_ * h_, dtype=torch.float32, device=dev), torch.empty(w_ * h_, dtype=torch.float32, device=dev))
                scan.collect_view(v, 0, gathered[v][0].data_ptr(), gathered[v][1].data_ptr())
            elif v % world == rank and rank != 0:
                scan.collect_view(v, 0)
        scan.sync()
    if rank == 0:
        fcams = []
        for v in range(n):                      # RescaleImageAndCamera (ACMMP.cpp:179-203): cameras at the size of the maps
            c = sc["cams"][v].copy()
            w_, h_ = scan.view_size(v)
            if (w_, h_) != (c.width, c.height):
                sx, sy = np.float32(w_) / np.float32(c.width), np.float32(h_) / np.float32(c.height)
                c.K[0] = float(np.float32(c.K[0]) * sx); c.K[2] = float(np.float32(c.K[2]) * sx)
                c.K[4] = float(np.float32(c.K[4]) * sy); c.K[5] = float(np.float32(c.K[5]) * sy)
                c.width, c.height = w_, h_
            fcams.append(c)
        fus = host.Fusion(fcams, sc["pairs"], device=local)
        for v in range(n):
            if v in gathered:
                fus.set_view(v, scan.exchange_view_ptr(v), gathered[v][0].data_ptr(), on_device=True)
            else:
                d_, n_, _ = scan.view_device(v)
                fus.set_view(v, d_, n_, on_device=True)
        fus.run(0.3, 1)
        t0 = time.time()
        npts = fus.run(0.3, 1)
        fwall = time.time() - t0
        fst = fus.stats()
        line["fusion"] = {"points": int(npts), "ms": fwall * 1e3, "device_ms": fst["score_ms"] + fst["greedy_ms"] + fst["emit_ms"],
                          "fixed_point_rounds_max": max(fst["rounds"]), "consistency_scalar": 0.3, "num_consistent_thresh": 1,
                          "where": "rank 0" + ((" (holds every view's record after the dataflow run)" if dataflow else " after gathering the normals over NCCL") if world > 1 else "")}
        if world == 1 and not args.no_cpu_baseline and args.config == "full":
            import oraclebind as ob
            maps = [scan.get_view(v, want_cost=False) for v in range(n)]
            t0 = time.time()
            cnt = ob.fusion(sc["cams"], [m[0] for m in maps], [m[1] for m in maps], sc["pairs"], 0.3, 1)
            line["fusion"].update({"cpu_port_s": time.time() - t0, "cpu_port_points": int(cnt), "cpu_cores": 1})
        fus.close()
    if rank == 0 and not args.no_cpu_baseline:
        if args.config in ("full", "photometric"):
            line["planar_prior"] = planar_prior_baseline(args, sc)
        line["cpu_baseline"] = cpu_baseline(args, sc, hpv)
    if rank == 0:
        print(json.dumps(line), flush=True)
    scan.close()
    if dist is not None:
        dist.destroy_process_group()


def planar_prior_baseline(args, sc):
    """The second baseline north_star names: the reference's CPU planar-prior code (GetSupportPoints ->
    Delaunay -> triangle raster + plane fit + range check, ACMMP.cpp:868-958, acmmp_definitions.cpp:306-378),
    restated in oracle/prior.c, on one host core for ONE view, next to the product's construction of the
    same prior (support cells, raster, plane fit and range check as kernels; Delaunay on the host).
    Outside the timed region; the scan's 98 prior constructions per step are inside `value`."""
    import oraclebind as ob
    from acmmp_b200 import host, lib as L
    v = args.views // 2
    ids = [v] + sc["pairs"][v]
    to_np = lambda t: t.detach().float().cpu().numpy() if hasattr(t, "detach") else np.asarray(t, np.float32)
    images = [to_np(sc["images"][i]) for i in ids]
    cams = [sc["cams"][i] for i in ids]
    m = host.ACMMP(0)
    m.SetTextureFormat(L.TEX_AUTO)
    m.InputInitialization(images, cams)
    m.RunPatchMatch(1234)
    depth, _, cost = m.GetMaps()
    pr = m.params()
    dev_ms = []
    for _ in range(3):
        t0 = time.time()
        ns, nt = m.BuildPlanarPrior()
        m.sync()
        dev_ms.append((time.time() - t0) * 1e3)
    _, mask = m.GetPlanarPrior()
    lib = L.load()
    lib.acmmp_set_delaunay_threads(1)   # the reference's cv::Subdiv2D is single-threaded
    try:
        t0 = time.time()
        pts = ob.support_points(cost)
        t1 = time.time()
        tris = host.DelaunayTriangulation(pts)
        t2 = time.time()
        _, mask_o, _ = ob.prior_from_triangles(cams[0], depth, tris.reshape(-1, 6), pr.depth_min, pr.depth_max)
        t3 = time.time()
    finally:
        lib.acmmp_set_delaunay_threads(0)
    m.close()
    return {"what": "planar-prior construction of one %dx%d view" % (args.width, args.height), "support_points": int(ns), "triangles": int(nt),
            "ours_ms": min(dev_ms), "cpu_port_ms": (t3 - t0) * 1e3, "cpu_cores": 1,
            "cpu_port_breakdown_ms": {"support_points": (t1 - t0) * 1e3,
                                      "delaunay (the product's host code on 1 thread; the reference calls cv::Subdiv2D)": (t2 - t1) * 1e3,
                                      "raster + plane fit + range check": (t3 - t2) * 1e3},
            "labels_equal_frac": float((mask == mask_o).mean())}


def cpu_baseline(args, sc, hyp_per_view):
    """The CPU port of the path (oracle/*.c, single thread) on a bounded sample of the
    same workload: one reference view of the scan with its sources at full
    resolution, random init on a (inner + 24 px margin) window, then the four
    half-sweeps on the inner window."""
    import oraclebind as ob
    ref = args.views // 2
    ids = [ref] + sc["pairs"][ref]
    images = [sc["images"][i].cpu().numpy() for i in ids]
    cams = [sc["cams"][i] for i in ids]
    H, W = images[0].shape
    iw, ih, m = 96, 64, 24
    x0, y0 = W // 2 - iw // 2, H // 2 - ih // 2
    pr = ob.Problem(images, cams)
    pr.rng_init(1234)
    t0 = time.time()
    pr.init(ob.INIT_RANDOM, (x0 - m, y0 - m, x0 + iw + m, y0 + ih + m))
    for it in range(2):
        for color in range(2):
            pr.sweep(it, color, (x0, y0, x0 + iw, y0 + ih))
    dt = time.time() - t0
    hyp = (iw + 2 * m) * (ih + 2 * m) * 1 + iw * ih * 28
    return {"value": hyp / dt / hyp_per_view, "unit": "views/s", "cores": 1, "kind": "port",
            "mpix_hyp_per_s": hyp / dt / 1e6,
            "sample": f"view {ref} + {len(ids) - 1} sources at {W}x{H}: random init on a {iw + 2 * m}x{ih + 2 * m} window + 2x(black,red) sweeps "
                      f"on the inner {iw}x{ih} ({hyp} hypothesis evaluations, {dt:.1f} s), scaled by {hyp_per_view} evaluations per view",
            "note": "the reference has no CPU PatchMatch; this is the oracle restatement (exact libm, no SIMD)"}


def level_sizes(W, H, single_scale=False, cap=3200, bound=1000):
    """(w, h) of every pyramid level of a W x H view, coarse to fine (ComputeMultiScaleSettings + InputInitialization)"""
    m = min(max(W, H), cap)
    k = 0
    while m > bound and not single_scale:
        m //= 2
        k += 1
    out = []
    for lvl in range(k, -1, -1):
        size = int(min(max(W, H), cap) / 2 ** lvl)
        if W <= size and H <= size:
            out.append((W, H))
        else:
            f = min(np.float32(size) / np.float32(W), np.float32(size) / np.float32(H))
            out.append((int(round(float(np.float32(W) * f))), int(round(float(np.float32(H) * f)))))
    return out


def run_reference(args):
    """Baseline arm, 1 GPU (the reference has no multi-GPU mode and no CPU PatchMatch): the reference's own code on
    this box, compiled from where it lies by `make ref`.
    --config photometric: its kernels in the order of ACMMP::RunPatchMatch, per view its own problem set-up
    (textures, buffers) and result copy (oracle/_ref/libacmmp_ref.so), every view of the scan per step.
    other configs: the reference PROGRAM (oracle/_ref/libacmmp_ref_full.so: kernels, host class, ProcessProblem,
    JointBilateralUpsampling, RunFusion, run in the order of its main() on its own on-disk layout, .dmb round trips
    included).  Each step = the whole schedule over a bounded sub-scan: the centre view of the rig and its sources
    (nsrc + 1 views, every view paired with all the others) = the per-view work of the full scan; --ref-full-scan
    runs all views.  Beside `value`: `p2` = one more step of the same program built without `#define DEBUG`
    (BASELINE.md patch P2), `kernel_only` = its RunPatchMatch kernels alone (CUDA events; the kernels harness in the
    reference's per-view flow on --ref-views views), `delaunay_standin_share`."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import refbind
    if args.ref_kernel_only_child:
        import torch
        torch.cuda.set_device(0)
        print(json.dumps(reference_kernel_only(args, make_scan_inputs(args, torch.device("cuda", 0)))), flush=True)
        return
    if args.config == "photometric":
        return run_reference_photometric(args)
    if not refbind.full_available():
        return run_reference_kernels(args)
    import shutil
    import tempfile
    import torch
    torch.cuda.set_device(0)
    sc = make_scan_inputs(args, torch.device("cuda", 0))
    n = args.views
    centre = n // 2
    sample = list(range(n)) if args.ref_full_scan else sorted(set([centre] + sc["pairs"][centre][:args.nsrc]))
    images = [sc["images"][i].cpu().numpy() for i in sample]
    cams = [sc["cams"][i] for i in sample]
    if args.ref_full_scan:
        pairs = sc["pairs"]
    else:
        pairs = [[j for j in range(len(sample)) if j != i] for i in range(len(sample))]
    H, W = args.height, args.width
    sizes = level_sizes(W, H)
    hpv = hyp_evals_per_view(sizes)
    # same number of warm-up passes as the product's arm for the config the driver runs; the long configs keep one
    warm = args.warmup if (args.config == "full" and not args.ref_full_scan) else min(args.warmup, 1)
    dense = tempfile.mkdtemp(prefix="acmmp_ref_")
    p2 = None
    t_arm = time.time()
    try:
        refbind.write_dense_folder(dense, cams, images, pairs)
        for _ in range(warm):
            refbind.full_run(dense, seed=1234, fuse=False)
        t_pm = 0.0
        t_fu = 0.0
        refbind.full_delaunay_seconds()
        with ClockSampler(0) as clocks:
            for k in range(args.steps):
                a, b = refbind.full_run(dense, seed=1234 + k, fuse=(k == args.steps - 1))
                t_pm += a
                t_fu = b
        t_del = refbind.full_delaunay_seconds()
        npts = len(refbind.read_ply(os.path.join(dense, "ACMMP", "ACMMP_model.ply"))[0])
        # (the P2 variant costs two more steps: skipped when the arm has already run for ten minutes, so that a
        # --steps 20 --warmup 5 invocation on a slow day still ends inside the driver's limit)
        if refbind.full_available("p2") and not args.no_ref_p2 and time.time() - t_arm > 600.0:
            p2 = {"skipped": "time budget: the arm had run for %.0f s" % (time.time() - t_arm)}
        elif refbind.full_available("p2") and not args.no_ref_p2:
            refbind.full_run(dense, seed=1234, fuse=False, variant="p2")
            a, _ = refbind.full_run(dense, seed=1234, fuse=False, variant="p2")
            p2 = {"value": len(sample) / a, "unit": "views/s", "ms_per_step": a * 1e3, "steps": 1, "warmup": 1,
                  "what": "the same program built without `#define DEBUG` (ACMMP.cu:3; BASELINE.md patch P2): no per-run debug copies of the plane map"}
    finally:
        shutil.rmtree(dense, ignore_errors=True)
    v = len(sample) * args.steps / t_pm
    kernel_only = None
    if args.config == "full" and refbind.available():
        # in a process of its own: libacmmp_ref.so and libacmmp_ref_full.so both define the reference's kernels (same
        # symbols), and the dynamic linker would bind the second library's launches to the first one's stubs
        cmd = [sys.executable, os.path.abspath(__file__), "--impl", "reference", "--ref-kernel-only-child", "--config", args.config,
               "--views", str(args.views), "--width", str(args.width), "--height", str(args.height), "--nsrc", str(args.nsrc),
               "--seed", str(args.seed), "--ref-views", str(args.ref_views)]
        out = subprocess.run(cmd, capture_output=True, text=True, timeout=1200)
        lines = [l for l in out.stdout.splitlines() if l.startswith("{")]
        kernel_only = json.loads(lines[-1]) if (out.returncode == 0 and lines) else {"unavailable": (out.stderr or "no output")[-300:]}
    line = {
        "impl": "reference",
        "metric": CONFIGS[args.config]["metric"],
        "value": v, "unit": "views/s", "n_gpus": 1, "steps": args.steps, "warmup": warm,
        "ms_per_step": t_pm / args.steps * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": CONFIGS[args.config]["name"].format(W=W, H=H, n=n, ns=args.nsrc, nl=len(sizes)), "name": args.config,
                   "levels": [list(x) for x in sizes], "sample_views": sample, "hyp_evals_per_view": hpv,
                   "what": "the reference program (ACMMP.cu / ACMMP.cpp / acmmp_definitions.cpp compiled from where they lie, sm_100, its own flags, "
                           "`#define DEBUG` as shipped) in the order of its main(): per view and stage its own image read + resize, texture set-up, "
                           "RunPatchMatch, host planar prior and .dmb round trips; stand-ins: raw images instead of JPEG decode, the product's "
                           "Delaunay for cv::Subdiv2D (share of the step: delaunay_standin_share), inert imshow / imwrite, fixed seed"},
        "mpix_hyp_per_s": v * hpv / 1e6,
        "delaunay_standin_share": t_del / t_pm,
        "fusion": {"points": npts, "ms": t_fu * 1e3, "what": "its own RunFusion (CPU, 1 thread) on the sample"},
        "cpu_baseline": {"value": v, "unit": "views/s", "cores": 1, "kind": "reference",
                         "sample": (f"all {n} views per step" if args.ref_full_scan else
                                    f"{len(sample)}-view sub-scan per step (view {centre} and its {args.nsrc} sources, every view paired with all the others)")
                                   + ", all scales and stages; 1 B200 for the reference's CUDA kernels, 1 host thread for its CPU steps"},
        "e2e": {"value": v, "unit": "views/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "clocks": clocks.summary(),
    }
    if p2 is not None:
        line["p2"] = p2
    if kernel_only is not None:
        line["kernel_only"] = kernel_only
    print(json.dumps(line), flush=True)


def reference_kernel_only(args, sc):
    """RunPatchMatch kernels of the reference alone: oracle/_ref/libacmmp_ref.so (its __global__ functions launched in the
    order of ACMMP::RunPatchMatch, CUDA events around each run) in the reference's per-view / per-stage flow
    (oracle/ref_pipeline.py) on --ref-views views; views outside the sample contribute ground-truth depth maps to the
    geometric-consistency stages, so a view's kernel work is that of the full scan."""
    import cv2
    import ref_pipeline
    import torch
    images = [im.cpu().numpy() for im in sc["images"]]
    gts = [d.cpu().numpy() for d in sc["depths"]]
    n = args.views
    centre = n // 2
    sample = sorted(set([centre] + sc["pairs"][centre][:max(0, args.ref_views - 1)]))
    rs = ref_pipeline.RefScan(images, sc["cams"], sc["pairs"], 0)

    def external(i, w, h):
        g = gts[i]
        return g if g.shape == (h, w) else cv2.resize(g, (w, h), interpolation=cv2.INTER_NEAREST)

    rs.run_all(1234, sample[:1], external)       # warm-up
    torch.cuda.synchronize()
    rs.kernel_ms, rs.prior_s, rs.runs = 0.0, 0.0, 0
    rs.run_all(1234, sample, external)
    torch.cuda.synchronize()
    return {"value": len(sample) / (rs.kernel_ms * 1e-3), "unit": "views/s (views / RunPatchMatch kernel time)",
            "ms_per_view": rs.kernel_ms / len(sample), "runs": rs.runs, "views": sample,
            "how": "reference kernels (libacmmp_ref.so) in the reference's per-view flow, CUDA events around each RunPatchMatch"}


def run_reference_photometric(args):
    """configs[1]: one RunPatchMatch per view, single scale, no prior, no geometric consistency - the reference's kernels in the
    order of ACMMP::RunPatchMatch, with its per-view problem set-up (one texture per image, buffers: ACMMP.cpp:638-708) and the
    result copy (ACMMP.cu:1453-1455).  Every view of the scan per step (same config as the product's arm)."""
    import refbind
    import torch
    torch.cuda.set_device(0)
    sc = make_scan_inputs(args, torch.device("cuda", 0))
    n = args.views
    images = [im.cpu().numpy() for im in sc["images"]]
    ref = refbind.RefACMMP(0)
    W, H = args.width, args.height
    hpv = hyp_evals_per_view([(W, H)], 1)

    def step(seed):
        kms = 0.0
        for v in range(n):
            ids = [v] + sc["pairs"][v]
            ref.set_problem([images[i] for i in ids], [sc["cams"][i] for i in ids])
            kms += ref.run(seed + v)
            ref.get_result()
        return kms

    for _ in range(args.warmup):
        step(1234)
    torch.cuda.synchronize()
    kernel_ms = 0.0
    with ClockSampler(0) as clocks:
        t0 = time.time()
        for k in range(args.steps):
            kernel_ms += step(1234 + k)
        torch.cuda.synchronize()
        dt = time.time() - t0
    ref.close()
    v = n * args.steps / dt
    line = {
        "impl": "reference", "metric": CONFIGS[args.config]["metric"],
        "value": v, "unit": "views/s", "n_gpus": 1, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": CONFIGS[args.config]["name"].format(W=W, H=H, n=n, ns=args.nsrc, nl=1), "name": args.config, "levels": [[W, H]],
                   "hyp_evals_per_view": hpv,
                   "what": "reference ACMMP.cu kernels (sm_100, -O3 --use_fast_math --maxrregcount=128) launched in the order of ACMMP::RunPatchMatch "
                           "with a device-wide synchronisation after each (ACMMP.cu:1415-1448), per view its texture / buffer set-up from host "
                           "images and the D2H of planes + costs; wall clock"},
        "mpix_hyp_per_s": v * hpv / 1e6,
        "kernel_only": {"value": n * args.steps / (kernel_ms * 1e-3), "unit": "views/s (views / RunPatchMatch kernel time)",
                        "ms_per_view": kernel_ms / (n * args.steps), "how": "CUDA events around the kernel sequence of each run"},
        "cpu_baseline": {"value": v, "unit": "views/s", "cores": 1, "kind": "reference",
                         "sample": f"all {n} views per step; 1 B200 for the reference's CUDA kernels, 1 host thread"},
        "e2e": {"value": v, "unit": "views/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "clocks": clocks.summary(),
    }
    print(json.dumps(line), flush=True)


def run_reference_kernels(args):
    """Fallback baseline arm (libacmmp_ref_full.so absent): the reference's own CUDA kernels (oracle/_ref, compiled from
    /root/reference/src/ACMMP.cu for sm_100 with the reference's flags) in the
    reference's own per-view / per-stage flow (oracle/ref_pipeline.py), 1 GPU - the
    reference has no multi-GPU mode and no CPU PatchMatch.  Each step = a bounded
    sample of the scan: --ref-views reference views go through the whole schedule
    with all their sources; views outside the sample contribute ground-truth depth
    maps to the geometric-consistency stages, so the work per view is that of the
    full scan."""
    import torch
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import refbind
    if not refbind.available():
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/libacmmp_ref.so was not built (needs /root/reference at build time)"}))
        return
    import cv2
    import ref_pipeline
    torch.cuda.set_device(0)
    sc = make_scan_inputs(args, torch.device("cuda", 0))
    images = [im.cpu().numpy() for im in sc["images"]]
    gts = [d.cpu().numpy() for d in sc["depths"]]
    n = args.views
    centre = n // 2
    sample = sorted(set([centre] + sc["pairs"][centre][:max(0, args.ref_views - 1)]))
    rs = ref_pipeline.RefScan(images, sc["cams"], sc["pairs"], 0)

    def external(i, w, h):
        g = gts[i]
        return g if g.shape == (h, w) else cv2.resize(g, (w, h), interpolation=cv2.INTER_NEAREST)

    sizes = []
    mis, nds, mx = ref_pipeline.multi_scale_settings(sc["cams"])
    for lvl in range(mx + 1):
        w, h, _ = ref_pipeline.level_geometry(sc["cams"][0], int(mis[0] / 2 ** (nds[0] - lvl)))
        sizes.append((w, h))
    hpv = hyp_evals_per_view(sizes)
    for _ in range(min(args.warmup, 1)):       # one warm-up pass is enough for a synchronous flow
        rs.run_all(1234, sample, external)
    torch.cuda.synchronize()
    rs.kernel_ms, rs.prior_s, rs.runs = 0.0, 0.0, 0
    with ClockSampler(0) as clocks:
        t0 = time.time()
        for _ in range(args.steps):
            rs.run_all(1234, sample, external)
        torch.cuda.synchronize()
        dt = time.time() - t0
    v = len(sample) * args.steps / dt
    H, W = args.height, args.width
    line = {
        "impl": "reference",
        "metric": "views/s, DTU-shape 1600x1200 49-view full ACMMP (multi-scale + planar prior + geometric consistency)",
        "value": v, "unit": "views/s", "n_gpus": 1, "steps": args.steps, "warmup": min(args.warmup, 1),
        "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"configs[2]: DTU-shape {W}x{H} x {n} views, {args.nsrc} sources, full ACMMP", "levels": [list(s) for s in sizes],
                   "sample_views": sample, "hyp_evals_per_view": hpv,
                   "what": "reference ACMMP.cu kernels (sm_100, -O3 --use_fast_math --maxrregcount=128) in the reference's per-view/"
                           "per-stage flow; JPEG decode, .dmb disk I/O and imshow left out (in its favour); wall clock incl. its "
                           "uploads, syncs, D2H and CPU planar prior"},
        "mpix_hyp_per_s": v * hpv / 1e6,
        "kernel_only": {"views_per_s": len(sample) * args.steps / (rs.kernel_ms * 1e-3), "kernel_ms": rs.kernel_ms, "runs": rs.runs},
        "cpu_prior_s_per_view": rs.prior_s / (len(sample) * args.steps),
        "cpu_baseline": {"value": v, "unit": "views/s", "cores": 1, "kind": "reference",
                         "sample": f"{len(sample)} of {n} views per step (view {centre} and its nearest sources), all stages; "
                                   "device = 1 B200 for the reference's CUDA kernels, 1 host thread for its CPU steps"},
        "e2e": {"value": v, "unit": "views/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "clocks": clocks.summary(),
    }
    print(json.dumps(line), flush=True)


def main():
    # stdout carries the JSON line and nothing else: native libraries write to file descriptor 1 behind Python's back
    # (NCCL prints "NCCL version ..." there when the box sets NCCL_DEBUG), so descriptor 1 is pointed at stderr and
    # Python's own stdout keeps the original descriptor
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(real_stdout, "w")
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
