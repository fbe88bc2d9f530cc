"""Python host-side mirror of the reference's operator interface for the hot path:
`class ACMMP` (ACMMP.h:58-90) and `RunJBU` (ACMMP.h:18-20), calling the CUDA
library through the C ABI.  Method names, call order and argument meaning follow
the reference; cv::Mat becomes numpy, file names become in-memory arrays (the
file-based driver lives in pipeline.py).  Host logic only - no arithmetic of the
path is done in Python."""
import ctypes as C

import numpy as np

from . import lib as L


def _check(ctx, rc, what):
    if rc != 0:
        msg = L.load().acmmp_last_error(ctx).decode() if ctx else L.load().acmmp_last_error(None).decode()
        raise L.AcmmpError(f"{what} failed ({rc}): {msg}")


class ACMMP:
    """One per (device); reusable across problems (the reference constructs one
    per view per stage, acmmp_definitions.cpp:260)."""

    def __init__(self, device=0):
        self._lib = L.load()
        self._ctx = C.c_void_p()
        rc = self._lib.acmmp_create(C.byref(self._ctx), device)
        if rc != 0:
            raise L.AcmmpError(f"acmmp_create failed ({rc}): {self._lib.acmmp_last_error(None).decode()}")
        self.device = device
        self._flags = []

    def close(self):
        if self._ctx:
            self._lib.acmmp_destroy(self._ctx)
            self._ctx = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- mode setters (ACMMP.cpp:447-464); applied at InputInitialization -------
    def SetGeomConsistencyParams(self, multi_geometry=False):
        _check(self._ctx, self._lib.acmmp_set_geom_consistency(self._ctx, int(multi_geometry)), "SetGeomConsistencyParams")

    def SetHierarchyParams(self):
        _check(self._ctx, self._lib.acmmp_set_hierarchy(self._ctx), "SetHierarchyParams")

    def SetPlanarPriorParams(self):
        _check(self._ctx, self._lib.acmmp_set_planar_prior_params(self._ctx), "SetPlanarPriorParams")

    def SetCompatMode(self, mode):
        _check(self._ctx, self._lib.acmmp_set_compat_mode(self._ctx, int(mode)), "SetCompatMode")

    def SetWaveEmulation(self, resident_blocks):
        """parity mode: run each half-sweep as the reference launch's block waves (148 resident blocks on a B200;
        0 = off, the deterministic default).  include/acmmp_b200.h: acmmp_set_wave_emulation."""
        _check(self._ctx, self._lib.acmmp_set_wave_emulation(self._ctx, int(resident_blocks)), "SetWaveEmulation")

    def SetBoxStaging(self, on):
        """TMA box staging of the sweep kernel on R8 problems: True / False, None = the process default
        (include/acmmp_b200.h: acmmp_set_box_staging); results do not depend on it"""
        _check(self._ctx, self._lib.acmmp_set_box_staging(self._ctx, -1 if on is None else int(bool(on))), "SetBoxStaging")

    def SetTextureFormat(self, fmt):
        """L.TEX_R32F (default) or L.TEX_AUTO: R8 texels for the next problems whose images all hold
        integer grey levels (bit-identical samples, a quarter of the texture footprint)."""
        _check(self._ctx, self._lib.acmmp_set_texture_format(self._ctx, int(fmt)), "SetTextureFormat")

    def TextureIsU8(self):
        return bool(self._lib.acmmp_texture_is_u8(self._ctx))

    # ---- inputs ---------------------------------------------------------------------
    def InputInitialization(self, images, cams, depths=None):
        """images[0] / cams[0]: reference view (float32 HxW at this scale); the
        rest are the sources.  `depths`: per-view depth maps for geometric
        consistency (ACMMP.cpp:608-635)."""
        n = len(images)
        imgs = [L.f32(im) for im in images]
        for im, cam in zip(imgs, cams):
            if im.shape != (cam.height, cam.width):
                raise L.AcmmpError(f"image {im.shape} does not match camera {cam.height}x{cam.width}")
        cam_arr = (L.Camera * n)(*[c.copy() for c in cams])
        ptrs = (C.c_void_p * n)(*[im.ctypes.data for im in imgs])
        _check(self._ctx, self._lib.acmmp_set_problem(self._ctx, n, cam_arr, ptrs, 0), "InputInitialization")
        if depths is not None:
            for i, d in enumerate(depths):
                if d is None:
                    continue
                d = L.f32(d)
                _check(self._ctx, self._lib.acmmp_set_depth_map(self._ctx, i, d.ctypes.data, d.shape[1], d.shape[0], 0),
                       "set_depth_map")

    def CudaSpaceInitialization(self, warm_start=None, hierarchy=None):
        """warm_start = (depth, normal[H,W,3], cost) of this view's previous stage
        (ACMMP.cpp:710-742); hierarchy = (coarse_normal, coarse_cost, fine_depth)
        (ACMMP.cpp:745-808)."""
        if warm_start is not None:
            d, n, c = (L.f32(a) for a in warm_start)
            _check(self._ctx, self._lib.acmmp_set_warm_start(self._ctx, L.fptr(d), L.fptr(n), L.fptr(c)), "set_warm_start")
        if hierarchy is not None:
            cn, cc, fd = (L.f32(a) for a in hierarchy)
            _check(self._ctx, self._lib.acmmp_set_hierarchy_inputs(self._ctx, L.fptr(cn), L.fptr(cc), cc.shape[1], cc.shape[0], L.fptr(fd)),
                   "set_hierarchy_inputs")

    def SetPlanarPrior(self, planes):
        """fixed per-pixel plane hypotheses = the reference's `seeded` mode."""
        p = L.f32(planes)
        _check(self._ctx, self._lib.acmmp_set_seeded_planes(self._ctx, L.fptr(p)), "SetPlanarPrior")

    def CudaPlanarPriorInitialization(self, planes, masks):
        p = L.f32(planes)
        m = np.ascontiguousarray(masks, dtype=np.uint32)
        _check(self._ctx, self._lib.acmmp_set_planar_prior(self._ctx, L.fptr(p), L.uptr(m)), "CudaPlanarPriorInitialization")

    def BuildPlanarPrior(self):
        """GetSupportPoints -> DelaunayTriangulation -> raster -> plane fit -> upload."""
        ns, nt = C.c_int(), C.c_int()
        _check(self._ctx, self._lib.acmmp_build_planar_prior(self._ctx, C.byref(ns), C.byref(nt)), "BuildPlanarPrior")
        return ns.value, nt.value

    def GetPlanarPrior(self):
        """prior planes (H,W,4) and triangle labels (H,W) as uploaded."""
        h, w = self._shape()
        pl = np.empty((h, w, 4), np.float32)
        m = np.empty((h, w), np.uint32)
        _check(self._ctx, self._lib.acmmp_get_planar_prior(self._ctx, L.fptr(pl), L.uptr(m)), "GetPlanarPrior")
        return pl, m

    # ---- run ----------------------------------------------------------------------------
    def RunPatchMatch(self, seed=0):
        _check(self._ctx, self._lib.acmmp_run(self._ctx, int(seed)), "RunPatchMatch")

    def run_init(self, seed=0):
        _check(self._ctx, self._lib.acmmp_run_init(self._ctx, int(seed)), "run_init")

    def run_sweep(self, it, color):
        _check(self._ctx, self._lib.acmmp_run_sweep(self._ctx, int(it), int(color)), "run_sweep")

    def run_sweep_snapshot(self, it, color, snap_planes, snap_costs):
        """diagnostics: neighbour reads from the given snapshot, own pixel from the live state (in place)."""
        sp, sc = L.f32(snap_planes), L.f32(snap_costs)
        _check(self._ctx, self._lib.acmmp_run_sweep_snapshot(self._ctx, int(it), int(color), L.fptr(sp), L.fptr(sc)), "run_sweep_snapshot")

    def run_tail(self):
        _check(self._ctx, self._lib.acmmp_run_tail(self._ctx), "run_tail")

    def sync(self):
        _check(self._ctx, self._lib.acmmp_sync(self._ctx), "sync")

    # ---- outputs ------------------------------------------------------------------------
    def GetReferenceImageWidth(self):
        return self._lib.acmmp_width(self._ctx)

    def GetReferenceImageHeight(self):
        return self._lib.acmmp_height(self._ctx)

    def params(self):
        p = L.Params()
        _check(self._ctx, self._lib.acmmp_get_params(self._ctx, C.byref(p)), "get_params")
        return p

    def GetMinDepth(self):
        return self.params().depth_min

    def GetMaxDepth(self):
        return self.params().depth_max

    def _shape(self):
        return self.GetReferenceImageHeight(), self.GetReferenceImageWidth()

    def GetPlaneHypotheses(self):
        """(H,W,4): world normal + depth after RunPatchMatch (ACMMP.cpp:848-851)."""
        h, w = self._shape()
        out = np.empty((h, w, 4), np.float32)
        _check(self._ctx, self._lib.acmmp_get_result(self._ctx, L.fptr(out), None, None), "GetPlaneHypothesis")
        return out

    def GetCosts(self):
        h, w = self._shape()
        out = np.empty((h, w), np.float32)
        _check(self._ctx, self._lib.acmmp_get_result(self._ctx, None, L.fptr(out), None), "GetCost")
        return out

    def GetMaps(self):
        """depth (H,W), normal (H,W,3), cost (H,W): the three .dmb payloads."""
        h, w = self._shape()
        d = np.empty((h, w), np.float32)
        n = np.empty((h, w, 3), np.float32)
        c = np.empty((h, w), np.float32)
        _check(self._ctx, self._lib.acmmp_get_maps(self._ctx, L.fptr(d), L.fptr(n), L.fptr(c)), "get_maps")
        return d, n, c

    def get_state(self, with_rng=True):
        h, w = self._shape()
        planes = np.empty((h, w, 4), np.float32)
        costs = np.empty((h, w), np.float32)
        sel = np.empty((h, w), np.uint32)
        rng = np.empty((h, w, 6), np.uint32) if with_rng else None
        _check(self._ctx, self._lib.acmmp_get_state(self._ctx, L.fptr(planes), L.fptr(costs), L.uptr(sel), L.uptr(rng)), "get_state")
        return planes, costs, sel, rng

    def set_state(self, planes=None, costs=None, sel=None, rng=None):
        planes = L.f32(planes) if planes is not None else None
        costs = L.f32(costs) if costs is not None else None
        sel = np.ascontiguousarray(sel, np.uint32) if sel is not None else None
        rng = np.ascontiguousarray(rng, np.uint32) if rng is not None else None
        _check(self._ctx, self._lib.acmmp_set_state(self._ctx, L.fptr(planes), L.fptr(costs), L.uptr(sel), L.uptr(rng)), "set_state")

    def launch_count(self):
        return int(self._lib.acmmp_launch_count(self._ctx))

    def last_timing(self):
        out = (C.c_float * 4)()
        _check(self._ctx, self._lib.acmmp_last_timing(self._ctx, C.byref(out)), "last_timing")
        return {"init_ms": out[0], "sweeps_ms": out[1], "tail_ms": out[2], "total_ms": out[3]}


def RunJBU(image, src_depth, device=0):
    """Joint bilateral upsample of `src_depth` to the size of `image`
    (RunJBU, ACMMP.cpp:1022-1087).  Returns None when the integer scale is 1."""
    lib = L.load()
    img = L.f32(image)
    dep = L.f32(src_depth)
    out = np.empty_like(img)
    rc = lib.acmmp_jbu_depth(device, L.fptr(img), img.shape[1], img.shape[0], L.fptr(dep), dep.shape[1], dep.shape[0], L.fptr(out))
    if rc == 1:
        return None
    _check(None, rc, "RunJBU")
    return out


def resize_linear(image, new_w, new_h, device=0):
    """cv::resize(INTER_LINEAR) of a float image on the device (ACMMP.cpp:589)."""
    lib = L.load()
    img = L.f32(image)
    out = np.empty((new_h, new_w), np.float32)
    _check(None, lib.acmmp_resize_linear(device, L.fptr(img), img.shape[1], img.shape[0], L.fptr(out), new_w, new_h), "resize_linear")
    return out


def prior_planes_from_maps(cam, depth16, normal16, rows, cols, device=0):
    """pSampler::GetPriorPlaneEstimate (acmmp_definitions.cpp:98-177) on the device -> (rows, cols, 4) planes"""
    lib = L.load()
    d = np.ascontiguousarray(depth16, np.uint16)
    n = np.ascontiguousarray(normal16, np.uint16)
    out = np.empty((rows, cols, 4), np.float32)
    u16p = C.POINTER(C.c_uint16)
    c = cam.copy()
    rc = lib.acmmp_prior_planes_from_maps(device, C.byref(c), d.ctypes.data_as(u16p), n.ctypes.data_as(u16p), d.shape[0], d.shape[1],
                                          rows, cols, L.fptr(out))
    if rc != 0:
        raise L.AcmmpError(f"acmmp_prior_planes_from_maps failed ({rc}): {lib.acmmp_seeded_last_error().decode()}")
    return out


def tex_sample(image, xs, ys, device=0):
    """Bilinear fetches of `image` at (xs, ys) through the views' texture set-up (ACMMP.cpp:646-663)."""
    lib = L.load()
    img, x, y = L.f32(image), L.f32(xs).ravel(), L.f32(ys).ravel()
    out = np.empty(len(x), np.float32)
    _check(None, lib.acmmp_tex_sample(device, L.fptr(img), img.shape[1], img.shape[0], L.fptr(x), L.fptr(y), len(x), L.fptr(out)),
           "tex_sample")
    return out


def tex_sample_u8(image, xs, ys, device=0):
    """The same through an R8 unorm texture and the rounding of the kU8 kernels (ncc.cuh); `image` must hold
    integer grey levels 0..255."""
    lib = L.load()
    img, x, y = L.f32(image), L.f32(xs).ravel(), L.f32(ys).ravel()
    out = np.empty(len(x), np.float32)
    _check(None, lib.acmmp_tex_sample_u8(device, L.fptr(img), img.shape[1], img.shape[0], L.fptr(x), L.fptr(y), len(x), L.fptr(out)),
           "tex_sample_u8")
    return out


def DelaunayTriangulation(points):
    """points (n,2) int -> triangles (m,3,2) int (ACMMP.cpp:896-918); host only."""
    lib = L.load()
    pts = np.ascontiguousarray(points, np.int32).reshape(-1, 2)
    cap = max(2 * len(pts), 1)
    out = np.zeros((cap, 6), np.int32)
    i32 = C.POINTER(C.c_int32)
    nt = lib.acmmp_delaunay(pts.ctypes.data_as(i32), len(pts), out.ctypes.data_as(i32), cap)
    if nt < 0:
        raise L.AcmmpError(f"acmmp_delaunay failed ({nt})")
    return out[:nt].reshape(-1, 3, 2)


def rng_init_states(seed, w, h, device=0):
    lib = L.load()
    out = np.empty((h, w, 6), np.uint32)
    _check(None, lib.acmmp_rng_init_states(device, int(seed), w, h, L.uptr(out)), "rng_init_states")
    return out


def measure_fp32_peak(device=0):
    lib = L.load()
    a, b = C.c_double(), C.c_double()
    _check(None, lib.acmmp_measure_fp32_peak(device, C.byref(a), C.byref(b)), "measure_fp32_peak")
    return a.value, b.value


def measure_tex_peak(device=0, u8=False):
    """(coherent, grouped) bilinear fetch rate in G fetches/s: the propagation kernel's roofline
    (u8: on R8 unorm texels incl. the rounding of the kU8 kernels)."""
    lib = L.load()
    a, b = C.c_double(), C.c_double()
    fn = lib.acmmp_measure_tex_peak_u8 if u8 else lib.acmmp_measure_tex_peak
    _check(None, fn(device, C.byref(a), C.byref(b)), "measure_tex_peak")
    return a.value, b.value


STAGE_PHOTO, STAGE_GEOM0, STAGE_GEOM1 = 0, 1, 2


def plan_dataflow(pairs, n_levels, world, photometric_only=False, level_cost=None):
    """host only: (task_round, task_rank, n_rounds, lag) of the dataflow schedule; task = stage * n_views + view"""
    lib = L.load()
    n = len(pairs)
    off = np.zeros(n + 1, np.int32)
    off[1:] = np.cumsum([len(p) for p in pairs])
    ids = np.asarray([i for p in pairs for i in p], np.int32)
    T = n_levels * (1 if photometric_only else 3) * n
    rounds, ranks = np.zeros(T, np.int32), np.zeros(T, np.int32)
    nr, lag = C.c_int32(0), C.c_int32(0)
    I32 = C.POINTER(C.c_int32)
    cost = None if level_cost is None else (C.c_float * n_levels)(*[float(c) for c in level_cost])
    rc = lib.acmmp_plan_dataflow(n, off.ctypes.data_as(I32), ids.ctypes.data_as(I32), n_levels, int(photometric_only), world, cost,
                                 rounds.ctypes.data_as(I32), ranks.ctypes.data_as(I32), C.byref(nr), C.byref(lag))
    if rc:
        raise RuntimeError("acmmp_plan_dataflow failed (%d)" % rc)
    return rounds, ranks, nr.value, lag.value


def nccl_version():
    return int(L.load().acmmp_nccl_version())


def nccl_unique_id():
    """128 bytes to hand to every rank's Scan.comm_init (rank 0 calls this)"""
    buf = (C.c_uint8 * 128)()
    rc = L.load().acmmp_nccl_get_unique_id(buf)
    if rc != 0:
        raise L.AcmmpError(f"acmmp_nccl_get_unique_id failed ({rc}): {L.load().acmmp_last_error(None).decode()}")
    return bytes(buf)


class Scan:
    """The stage graph of the reference's main() (main_ACMMP.cpp:89-176) over a set of
    views with all data resident on the device; `exchange` is called after every
    stage with (device pointer, floats per rank segment) so that a multi-GPU caller
    can all-gather the depth maps (no-op on one GPU)."""

    def __init__(self, cams, pairs, device=0, rank=0, world=1, sequential_geom=None, tex_format=None):
        self._lib = L.load()
        self._s = C.c_void_p()
        n = len(cams)
        cam_arr = (L.Camera * n)(*[c.copy() for c in cams])
        offs = np.zeros(n + 1, np.int32)
        for i, p in enumerate(pairs):
            offs[i + 1] = offs[i] + len(p)
        ids = np.ascontiguousarray([j for p in pairs for j in p], np.int32)
        i32 = C.POINTER(C.c_int32)
        rc = self._lib.acmmp_scan_create(C.byref(self._s), device, n, cam_arr, offs.ctypes.data_as(i32), ids.ctypes.data_as(i32))
        if rc != 0:
            raise L.AcmmpError(f"acmmp_scan_create failed ({rc}): {self._lib.acmmp_last_error(None).decode()}")
        self.n, self.rank, self.world = n, rank, world
        self._ck(self._lib.acmmp_scan_set_shard(self._s, rank, world), "set_shard")
        # None: the library's default = the reference's in-place (Gauss-Seidel) geom-1 order on an unsharded scan,
        # Jacobi on a sharded one; True / False force either
        if sequential_geom is not None:
            self._ck(self._lib.acmmp_scan_set_sequential_geom(self._s, int(bool(sequential_geom))), "set_sequential_geom")
        if tex_format is not None:   # default of the library: L.TEX_AUTO
            self._ck(self._lib.acmmp_scan_set_texture_format(self._s, int(tex_format)), "set_texture_format")

    def _ck(self, rc, what):
        if rc != 0:
            raise L.AcmmpError(f"{what} failed ({rc}): {self._lib.acmmp_scan_last_error(self._s).decode()}")

    def close(self):
        if self._s:
            self._lib.acmmp_scan_destroy(self._s)
            self._s = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def owned(self):
        return [v for v in range(self.n) if v % self.world == self.rank]

    def set_image(self, view, gray, on_device=False, sync=True):
        """gray: float32 HxW numpy array (host, may be pinned) or an int device pointer.
        sync=False: the caller keeps the host array alive until the next sync()."""
        if on_device:
            self._ck(self._lib.acmmp_scan_set_image(self._s, view, C.c_void_p(int(gray)), 1), "set_image")
        else:
            g = L.f32(gray)
            self._ck(self._lib.acmmp_scan_set_image(self._s, view, C.c_void_p(g.ctypes.data), 0), "set_image")
            if sync:
                self.sync()   # the host array may be released after this returns

    def set_image_u8(self, view, gray, on_device=False, sync=True):
        """gray: uint8 HxW numpy array (what cv::imread(..., IMREAD_GRAYSCALE) returns) or an int device
        pointer; converted to float on the device.  sync=False: keep the host array alive until sync()."""
        if on_device:
            self._ck(self._lib.acmmp_scan_set_image_u8(self._s, view, C.c_void_p(int(gray)), 1), "set_image_u8")
        else:
            g = np.ascontiguousarray(gray, np.uint8)
            self._ck(self._lib.acmmp_scan_set_image_u8(self._s, view, C.c_void_p(g.ctypes.data), 0), "set_image_u8")
            if sync:
                self.sync()

    def num_levels(self):
        return self._lib.acmmp_scan_num_levels(self._s)

    # ---- multi-GPU: NCCL inside the library (include/acmmp_b200.h) ---------------------------------------------
    def comm_init(self, unique_id):
        """collective over the world given at construction; unique_id: the 128 bytes of nccl_unique_id() of rank 0"""
        buf = (C.c_uint8 * 128).from_buffer_copy(bytes(unique_id))
        self._ck(self._lib.acmmp_scan_comm_init(self._s, buf), "comm_init")

    @staticmethod
    def comm_init_all(scans):
        """one process, one scan (and one thread) per GPU: scans[r] = rank r of len(scans) on distinct devices (ncclCommInitAll)"""
        arr = (C.c_void_p * len(scans))(*[s._s for s in scans])
        rc = scans[0]._lib.acmmp_scan_comm_init_all(arr, len(scans))
        if rc != 0:
            raise L.AcmmpError(f"comm_init_all failed ({rc}): {scans[0]._lib.acmmp_scan_last_error(scans[0]._s).decode()}")

    def all_gather(self):
        """depth maps of the owned views -> every rank, on the scan's stream (no host sync); no-op when unsharded"""
        self._ck(self._lib.acmmp_scan_all_gather(self._s), "all_gather")

    def run_dataflow(self, seed):
        """the whole scan (all levels, all stages) as a dataflow of (stage, view) tasks; collective on a sharded scan.
        Afterwards every rank holds every view's final maps (get_view / view_device)."""
        self._ck(self._lib.acmmp_scan_run_dataflow(self._s, C.c_uint64(seed)), "run_dataflow")

    def set_box_staging(self, on):
        """TMA box staging of the sweep kernel on R8 levels: True / False, None = the process default"""
        self._ck(self._lib.acmmp_scan_set_box_staging(self._s, -1 if on is None else int(bool(on))), "set_box_staging")

    def keep_photometric(self, on=True):
        self._ck(self._lib.acmmp_scan_keep_photometric(self._s, int(bool(on))), "keep_photometric")

    def photometric_depth(self, view):
        """depths.dmb of the reference: the finest level's photometric (+ prior) depth, kept by keep_photometric()"""
        w, h = self.view_size(view)
        out = np.empty((h, w), np.float32)
        self._ck(self._lib.acmmp_scan_get_photometric_depth(self._s, view, out.ctypes.data_as(C.c_void_p)), "photometric_depth")
        return out

    def level_stats(self, level):
        """kernel time by phase (+ NCC counts of a counted dataflow run, texel format) of one pyramid level since the last call"""
        ms, ncc, u8 = (C.c_double * 4)(), (C.c_uint64 * 2)(), C.c_int(0)
        self._ck(self._lib.acmmp_scan_level_stats(self._s, level, ms, ncc, C.byref(u8)), "level_stats")
        return {"init_ms": ms[0], "sweeps_ms": ms[1], "tail_ms": ms[2], "total_ms": ms[3], "init": int(ncc[0]), "sweeps": int(ncc[1]), "u8": bool(u8.value)}

    def dataflow_trace(self):
        """ms after the run's start at which this rank finished each round (sharded scans; empty otherwise)"""
        cap = max(1, self.dataflow_stats()["rounds"])
        buf = (C.c_float * cap)()
        k = self._lib.acmmp_scan_dataflow_trace(self._s, buf, cap)
        if k < 0:
            self._ck(k, "dataflow_trace")
        return [float(buf[i]) for i in range(k)]

    def dataflow_stats(self):
        r, lag = C.c_int(0), C.c_int(0)
        self._ck(self._lib.acmmp_scan_dataflow_stats(self._s, C.byref(r), C.byref(lag)), "dataflow_stats")
        return {"rounds": r.value, "lag": lag.value}

    def collect_view(self, view, root, d_normal3=None, d_cost=None):
        """owner -> root: normals + costs of `view` into the root's device buffers (int pointers)"""
        self._ck(self._lib.acmmp_scan_collect_view(self._s, view, root, C.c_void_p(d_normal3), C.c_void_p(d_cost)), "collect_view")

    def set_sequential_geom(self, on):
        """geom-1 order: True = the reference's in-place order, False = Jacobi, None = the default (in place when unsharded)"""
        self._ck(self._lib.acmmp_scan_set_sequential_geom(self._s, -1 if on is None else int(bool(on))), "set_sequential_geom")

    def set_schedule(self, single_scale=False, photometric_only=False):
        """reduced schedules (before the first level): no pyramid / photometric stage without the planar prior"""
        self._ck(self._lib.acmmp_scan_set_schedule(self._s, int(bool(single_scale)), int(bool(photometric_only))), "set_schedule")

    def begin_level(self, level):
        self._ck(self._lib.acmmp_scan_begin_level(self._s, level), "begin_level")

    def level_is_u8(self):
        """True when the current level's views sit in R8 texels (no resize, integer grey levels)."""
        return bool(self._lib.acmmp_scan_level_is_u8(self._s))

    def run_stage(self, stage, seed=0):
        self._ck(self._lib.acmmp_scan_run_stage(self._s, stage, int(seed)), "run_stage")

    def exchange_info(self):
        ptr, stride, slots = C.c_void_p(), C.c_int64(), C.c_int()
        self._ck(self._lib.acmmp_scan_exchange(self._s, C.byref(ptr), C.byref(stride), C.byref(slots)), "exchange")
        return ptr.value, stride.value, slots.value

    def view_size(self, view):
        w, h = C.c_int(), C.c_int()
        self._ck(self._lib.acmmp_scan_view_size(self._s, view, C.byref(w), C.byref(h)), "view_size")
        return w.value, h.value

    def get_view(self, view, want_normal=True, want_cost=True, out=None):
        w, h = self.view_size(view)
        if out is not None:
            d, n, c = out
        else:
            d = np.empty((h, w), np.float32)
            n = np.empty((h, w, 3), np.float32) if want_normal else None
            c = np.empty((h, w), np.float32) if want_cost else None
        self._ck(self._lib.acmmp_scan_get_view(self._s, view, L.fptr(d), L.fptr(n), L.fptr(c)), "get_view")
        return d, n, c

    def view_device(self, view):
        """(depth, normal3, cost) device pointers of an owned view (ints)"""
        d, n, c = C.c_void_p(), C.c_void_p(), C.c_void_p()
        self._ck(self._lib.acmmp_scan_view_device(self._s, view, C.byref(d), C.byref(n), C.byref(c)), "view_device")
        return d.value, n.value, c.value

    def exchange_view_ptr(self, view):
        """device pointer of any view's depth map inside this rank's exchange buffer (after the gather)"""
        d = C.c_void_p()
        self._ck(self._lib.acmmp_scan_view_device(self._s, view, C.byref(d), None, None), "view_device")
        return d.value

    def timing(self):
        out, runs, launches = (C.c_double * 4)(), C.c_int(), C.c_uint64()
        self._ck(self._lib.acmmp_scan_timing(self._s, C.byref(out), C.byref(runs), C.byref(launches)), "timing")
        return {"init_ms": out[0], "sweeps_ms": out[1], "tail_ms": out[2], "total_ms": out[3], "runs": runs.value,
                "launches": int(launches.value)}

    def sync(self):
        self._ck(self._lib.acmmp_scan_sync(self._s), "sync")

    def host_stats(self):
        a, b = C.c_double(), C.c_double()
        self._ck(self._lib.acmmp_scan_host_stats(self._s, C.byref(a), C.byref(b)), "host_stats")
        return {"prior_host_ms": a.value, "wait_ms": b.value}

    def set_prior_maps(self, view, depth16=None, normal16=None):
        """seeded mode (-p): the 16-bit prior images of a view as cv::imread(path, -1) returns them
        (pSampler, acmmp_definitions.cpp:98-131); None, None removes them"""
        u16p = C.POINTER(C.c_uint16)
        if depth16 is None and normal16 is None:
            self._ck(self._lib.acmmp_scan_set_prior_maps(self._s, view, None, None, 0, 0), "set_prior_maps")
            return
        d = np.ascontiguousarray(depth16, np.uint16)
        n = np.ascontiguousarray(normal16, np.uint16)
        if n.shape != d.shape + (3,):
            raise ValueError("normal16 must be (rows, cols, 3) for depth16 (rows, cols)")
        self._ck(self._lib.acmmp_scan_set_prior_maps(self._s, view, d.ctypes.data_as(u16p), n.ctypes.data_as(u16p), d.shape[0], d.shape[1]),
                 "set_prior_maps")

    def count_ncc(self, enable=True):
        """switch the device-side count of NCC evaluations on / off (instrumentation)"""
        self._ck(self._lib.acmmp_scan_count_ncc(self._s, int(bool(enable))), "count_ncc")

    def ncc_count(self):
        a, b = C.c_uint64(), C.c_uint64()
        self._ck(self._lib.acmmp_scan_ncc_count(self._s, C.byref(a), C.byref(b)), "ncc_count")
        return {"init": a.value, "sweeps": b.value}

    def reset(self):
        self._ck(self._lib.acmmp_scan_reset(self._s), "reset")

    def mark(self, i):
        self._ck(self._lib.acmmp_scan_mark(self._s, i), "mark")

    def elapsed_ms(self, i, j):
        ms = C.c_float()
        self._ck(self._lib.acmmp_scan_elapsed(self._s, i, j, C.byref(ms)), "elapsed")
        return float(ms.value)

    def run_all(self, seed=0, exchange=None, stages=(STAGE_PHOTO, STAGE_GEOM0, STAGE_GEOM1), levels=None, level_done=None):
        """the full-ACMMP schedule: for every level, coarse to fine: photometric +
        planar prior, geom-0, geom-1.  level_done(level): called after the last stage of a level."""
        nl = self.num_levels() if levels is None else levels
        for level in range(nl):
            self.begin_level(level)
            for st in stages:
                self.run_stage(st, seed)
                if exchange is not None:
                    exchange(*self.exchange_info())
            if level_done is not None:
                level_done(level)


class Fusion:
    """RunFusion (acmmp_definitions.cpp:828-1043) on the device, without the file I/O:
    `cams` are the cameras rescaled to the depth maps, `pairs[v]` the source-view indices."""

    def __init__(self, cams, pairs, device=0):
        self._lib = L.load()
        self._f = C.c_void_p()
        n = len(cams)
        cam_arr = (L.Camera * n)(*[c.copy() for c in cams])
        offs = np.zeros(n + 1, np.int32)
        for i, p in enumerate(pairs):
            offs[i + 1] = offs[i] + len(p)
        ids = np.ascontiguousarray([j for p in pairs for j in p], np.int32)
        i32 = C.POINTER(C.c_int32)
        rc = self._lib.acmmp_fusion_create(C.byref(self._f), device, n, cam_arr, offs.ctypes.data_as(i32), ids.ctypes.data_as(i32))
        if rc != 0:
            raise L.AcmmpError(f"acmmp_fusion_create failed ({rc}): {self._lib.acmmp_fusion_last_error(None).decode()}")
        self.n = n
        self.sizes = [(c.height, c.width) for c in cams]

    def _ck(self, rc, what):
        if rc != 0:
            raise L.AcmmpError(f"{what} failed ({rc}): {self._lib.acmmp_fusion_last_error(self._f).decode()}")

    def close(self):
        if self._f:
            self._lib.acmmp_fusion_destroy(self._f)
            self._f = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_view(self, view, depth, normal, color=None, mask=None, on_device=False):
        """numpy arrays, or raw device pointers (ints) with on_device=True"""
        if on_device:
            ptrs = [C.c_void_p(int(p)) if p else None for p in (depth, normal, color, mask)]
        else:
            keep = [L.f32(depth), L.f32(normal),
                    None if color is None else np.ascontiguousarray(color, np.uint8),
                    None if mask is None else np.ascontiguousarray(mask, np.uint8)]
            ptrs = [None if a is None else C.c_void_p(a.ctypes.data) for a in keep]
        self._ck(self._lib.acmmp_fusion_set_view(self._f, view, ptrs[0], ptrs[1], ptrs[2], ptrs[3], int(on_device)), "fusion_set_view")

    def run(self, consistency_scalar=0.3, con_num_thresh=1):
        n = C.c_int64()
        self._ck(self._lib.acmmp_fusion_run(self._f, float(consistency_scalar), int(con_num_thresh), C.byref(n)), "fusion_run")
        self.num_points = n.value
        return n.value

    def set_view_second(self, view, depth2, normal2, on_device=False):
        """the second (depth, normal) set of a view, for run_prior_aware"""
        if on_device:
            ptrs = [C.c_void_p(int(depth2)), C.c_void_p(int(normal2))]
        else:
            keep = [L.f32(depth2), L.f32(normal2)]
            ptrs = [C.c_void_p(a.ctypes.data) for a in keep]
        self._ck(self._lib.acmmp_fusion_set_view_second(self._f, view, ptrs[0], ptrs[1], int(on_device)), "fusion_set_view_second")

    def run_prior_aware(self, consistency_scalar=0.3, num_consistent_thresh=1, single_match_penalty=1):
        """RunPriorAwareFusion (acmmp_definitions.cpp:575-826) on the device"""
        n = C.c_int64()
        self._ck(self._lib.acmmp_fusion_run_prior_aware(self._f, float(consistency_scalar), int(num_consistent_thresh),
                                                        int(single_match_penalty), C.byref(n)), "fusion_run_prior_aware")
        self.num_points = n.value
        return n.value

    def points(self):
        pts = np.empty((self.num_points, 9), np.float32)
        per = (C.c_int64 * self.n)()
        self._ck(self._lib.acmmp_fusion_get_points(self._f, C.c_void_p(pts.ctypes.data), self.num_points, per), "fusion_get_points")
        return pts, list(per)

    def mask(self, view):
        h, w = self.sizes[view]
        out = np.empty((h, w), np.uint8)
        self._ck(self._lib.acmmp_fusion_get_mask(self._f, view, C.c_void_p(out.ctypes.data)), "fusion_get_mask")
        return out

    def stats(self):
        ms = (C.c_double * 3)()
        rounds = (C.c_int32 * self.n)()
        self._ck(self._lib.acmmp_fusion_stats(self._f, ms, rounds), "fusion_stats")
        return {"score_ms": ms[0], "greedy_ms": ms[1], "emit_ms": ms[2], "rounds": list(rounds)}


def RunFusion(cams, pairs, depths, normals, colors=None, masks=None, consistency_scalar=0.3, con_num_thresh=1, device=0):
    """RunFusion with in-memory maps: returns (points[N, 9], per-view accepted counts)."""
    f = Fusion(cams, pairs, device)
    try:
        for v in range(len(cams)):
            f.set_view(v, depths[v], normals[v], None if colors is None else colors[v], None if masks is None else masks[v])
        f.run(consistency_scalar, con_num_thresh)
        return f.points()
    finally:
        f.close()


def RunPriorAwareFusion(cams, pairs, depths, normals, depths2, normals2, colors=None, masks=None, consistency_scalar=0.3,
                        num_consistent_thresh=1, single_match_penalty=1, device=0):
    """RunPriorAwareFusion with in-memory maps (set 0 = the reference's fusion-folder maps, set 1 = the run's
    own): returns (points[N, 9], per-view accepted counts)."""
    f = Fusion(cams, pairs, device)
    try:
        for v in range(len(cams)):
            f.set_view(v, depths[v], normals[v], None if colors is None else colors[v], None if masks is None else masks[v])
            f.set_view_second(v, depths2[v], normals2[v])
        f.run_prior_aware(consistency_scalar, num_consistent_thresh, single_match_penalty)
        return f.points()
    finally:
        f.close()


def decode_jpeg(data, color=False, device=0):
    """JPEG bytes -> (H, W) uint8 luma, or (H, W, 3) uint8 BGR with color=True, decoded on the
    device by nvJPEG: what cv::imread(IMREAD_GRAYSCALE / IMREAD_COLOR) hands the reference."""
    lib = L.load()
    buf = np.frombuffer(bytes(data), np.uint8)
    w, h = C.c_int(), C.c_int()
    ptr = C.c_void_p(buf.ctypes.data)
    if lib.acmmp_decode_jpeg(device, ptr, buf.size, int(color), None, 0, C.byref(w), C.byref(h)) != 0:
        raise L.AcmmpError(lib.acmmp_jpeg_last_error().decode())
    out = np.empty((h.value, w.value, 3) if color else (h.value, w.value), np.uint8)
    if lib.acmmp_decode_jpeg(device, ptr, buf.size, int(color), C.c_void_p(out.ctypes.data), out.nbytes, C.byref(w), C.byref(h)) != 0:
        raise L.AcmmpError(lib.acmmp_jpeg_last_error().decode())
    return out
