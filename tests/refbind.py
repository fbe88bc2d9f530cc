"""ctypes binding of oracle/_ref/libacmmp_ref.so - the REFERENCE's own CUDA kernels
behind a C ABI (oracle/ref_harness.cu).  Test infrastructure: used by tests/,
tools/gpu_parity.py and bench.py --impl reference only."""
import ctypes as C
import os

import numpy as np

from acmmp_b200 import lib as L

REF_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "oracle", "_ref", "libacmmp_ref.so")


def available():
    return os.path.exists(REF_PATH)


_ref = None


def load():
    global _ref
    if _ref is None:
        lib = C.CDLL(REF_PATH)
        vp, fp, up = C.c_void_p, C.POINTER(C.c_float), C.POINTER(C.c_uint32)
        lib.ref_create.argtypes = [C.POINTER(vp), C.c_int]
        lib.ref_destroy.argtypes = [vp]
        lib.ref_destroy.restype = None
        lib.ref_last_error.argtypes = [vp]
        lib.ref_last_error.restype = C.c_char_p
        lib.ref_set_problem.argtypes = [vp, C.c_int, C.POINTER(L.Camera), C.POINTER(vp)]
        lib.ref_set_flags.argtypes = [vp] + [C.c_int] * 6 + [C.c_float, C.c_float]
        lib.ref_set_depth_map.argtypes = [vp, C.c_int, fp, C.c_int, C.c_int]
        lib.ref_set_planes.argtypes = [vp, C.c_int, fp, C.c_int]
        lib.ref_set_scalars.argtypes = [vp, C.c_int, vp]
        lib.ref_set_rng.argtypes = [vp, up]
        lib.ref_get_state.argtypes = [vp, fp, fp, up, up, fp]
        lib.ref_run_init.argtypes = [vp, C.c_uint64]
        lib.ref_run_sweep.argtypes = [vp, C.c_int, C.c_int]
        lib.ref_run_tail.argtypes = [vp]
        lib.ref_run.argtypes = [vp, C.c_uint64]
        lib.ref_last_ms.argtypes = [vp]
        lib.ref_last_ms.restype = C.c_float
        lib.ref_ncc.argtypes = [vp, fp, C.c_int, fp]
        lib.ref_homography.argtypes = [vp, fp, C.c_int, C.c_int, fp]
        lib.ref_geom_cost.argtypes = [vp, fp, C.c_int, fp]
        lib.ref_jbu.argtypes = [fp, C.c_int, C.c_int, fp, C.c_int, C.c_int, fp]
        _ref = lib
    return _ref


class RefACMMP:
    """The reference kernels driven like the reference's class ACMMP."""

    def __init__(self, device=0):
        self.lib = load()
        self.ctx = C.c_void_p()
        if self.lib.ref_create(C.byref(self.ctx), device) != 0:
            raise RuntimeError("ref_create failed")
        self.flags = dict(geom=0, planar_prior=0, multi_geometry=0, hierarchy=0, upsample=0, seeded=0,
                          scaled_cols=0.0, scaled_rows=0.0)

    def close(self):
        if self.ctx:
            self.lib.ref_destroy(self.ctx)
            self.ctx = C.c_void_p()

    def _ck(self, rc, what):
        if rc != 0:
            raise RuntimeError(f"{what}: {self.lib.ref_last_error(self.ctx).decode()}")

    def set_problem(self, images, cams):
        n = len(images)
        self.imgs = [L.f32(i) for i in images]
        self.h, self.w = self.imgs[0].shape
        cam_arr = (L.Camera * n)(*[c.copy() for c in cams])
        ptrs = (C.c_void_p * n)(*[im.ctypes.data for im in self.imgs])
        self._ck(self.lib.ref_set_problem(self.ctx, n, cam_arr, ptrs), "ref_set_problem")
        self._push_flags()

    def _push_flags(self):
        f = self.flags
        self.lib.ref_set_flags(self.ctx, f["geom"], f["planar_prior"], f["multi_geometry"], f["hierarchy"], f["upsample"],
                               f["seeded"], f["scaled_cols"], f["scaled_rows"])

    def set_flags(self, **kw):
        self.flags.update(kw)
        self._push_flags()

    def set_depth_map(self, slot, depth):
        d = L.f32(depth)
        self._ck(self.lib.ref_set_depth_map(self.ctx, slot, L.fptr(d), d.shape[1], d.shape[0]), "ref_set_depth_map")

    def set_planes(self, which, planes):
        p = L.f32(planes)
        self._ck(self.lib.ref_set_planes(self.ctx, which, L.fptr(p), p.size // 4), "ref_set_planes")

    def set_scalars(self, which, arr):
        a = np.ascontiguousarray(arr)
        assert a.itemsize == 4
        self._ck(self.lib.ref_set_scalars(self.ctx, which, a.ctypes.data), "ref_set_scalars")

    def set_rng(self, rng6):
        r = np.ascontiguousarray(rng6, np.uint32)
        self._ck(self.lib.ref_set_rng(self.ctx, L.uptr(r)), "ref_set_rng")

    def get_state(self):
        planes = np.empty((self.h, self.w, 4), np.float32)
        costs = np.empty((self.h, self.w), np.float32)
        sel = np.empty((self.h, self.w), np.uint32)
        rng = np.empty((self.h, self.w, 6), np.uint32)
        pre = np.empty((self.h, self.w), np.float32)
        self._ck(self.lib.ref_get_state(self.ctx, L.fptr(planes), L.fptr(costs), L.uptr(sel), L.uptr(rng), L.fptr(pre)), "ref_get_state")
        return planes, costs, sel, rng, pre

    def get_result(self):
        """planes + costs only = the D2H of ACMMP::RunPatchMatch (ACMMP.cu:1453-1455)."""
        planes = np.empty((self.h, self.w, 4), np.float32)
        costs = np.empty((self.h, self.w), np.float32)
        self._ck(self.lib.ref_get_state(self.ctx, L.fptr(planes), L.fptr(costs), None, None, None), "ref_get_state")
        return planes, costs

    def run_init(self, seed):
        self._ck(self.lib.ref_run_init(self.ctx, seed), "ref_run_init")

    def run_sweep(self, it, color):
        self._ck(self.lib.ref_run_sweep(self.ctx, it, color), "ref_run_sweep")

    def run_tail(self):
        self._ck(self.lib.ref_run_tail(self.ctx), "ref_run_tail")

    def run(self, seed):
        self._ck(self.lib.ref_run(self.ctx, seed), "ref_run")
        return float(self.lib.ref_last_ms(self.ctx))

    def ncc(self, planes, view):
        p = L.f32(planes)
        out = np.empty((self.h, self.w), np.float32)
        self._ck(self.lib.ref_ncc(self.ctx, L.fptr(p), view, L.fptr(out)), "ref_ncc")
        return out

    def homography(self, planes, view):
        p = L.f32(planes).reshape(-1, 4)
        out = np.empty((p.shape[0], 9), np.float32)
        self._ck(self.lib.ref_homography(self.ctx, L.fptr(p), p.shape[0], view, L.fptr(out)), "ref_homography")
        return out

    def geom_cost(self, planes, view):
        p = L.f32(planes)
        out = np.empty((self.h, self.w), np.float32)
        self._ck(self.lib.ref_geom_cost(self.ctx, L.fptr(p), view, L.fptr(out)), "ref_geom_cost")
        return out


def ref_jbu(image, depth):
    lib = load()
    img, dep = L.f32(image), L.f32(depth)
    out = np.empty_like(img)
    rc = lib.ref_jbu(L.fptr(img), img.shape[1], img.shape[0], L.fptr(dep), dep.shape[1], dep.shape[0], L.fptr(out))
    if rc == 1:
        return None
    if rc != 0:
        raise RuntimeError("ref_jbu failed")
    return out


# ---- the reference's HOST code (RunFusion) ---------------------------------------------
_REF_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "oracle", "_ref")
REF_HOST_PATHS = {"reference_flags": os.path.join(_REF_DIR, "libacmmp_ref_host.so"),     # -O3 -ffast-math
                  "ieee": os.path.join(_REF_DIR, "libacmmp_ref_host_ieee.so")}           # -O2
_ref_host = {}


def host_available():
    return all(os.path.exists(p) for p in REF_HOST_PATHS.values())


def _write_dmb(path, arr):
    a = np.ascontiguousarray(arr, np.float32)
    h, w = a.shape[:2]
    nb = 1 if a.ndim == 2 else a.shape[2]
    with open(path, "wb") as f:
        np.array([1, h, w, nb], np.int32).tofile(f)     # writeDepthDmb / writeNormalDmb, ACMMP.cpp:294-380
        a.tofile(f)


def _write_cam(path, cam):
    """cams/%08d_cam.txt as ReadCamera parses it (ACMMP.cpp:149-177)"""
    R, t, K = cam.R[:], cam.t[:], cam.K[:]
    with open(path, "w") as f:
        f.write("extrinsic\n")
        for i in range(3):
            f.write("%.9g %.9g %.9g %.9g\n" % (R[3 * i], R[3 * i + 1], R[3 * i + 2], t[i]))
        f.write("0 0 0 1\n\nintrinsic\n")
        for i in range(3):
            f.write("%.9g %.9g %.9g\n" % (K[3 * i], K[3 * i + 1], K[3 * i + 2]))
        f.write("\n%.9g 1 192 %.9g\n" % (cam.depth_min, cam.depth_max))


def host_fusion(workdir, cams, depths, normals, colors, pairs, consistency_scalar=0.3, con_num_thresh=1, geom=True,
                build="reference_flags", timing=None):
    """The reference's own RunFusion on its own on-disk layout under `workdir`; returns the
    points of ACMMP_model.ply as (N, 9) floats: x y z nx ny nz + the three colour bytes in the
    order they sit in the file."""
    if build not in _ref_host:
        lib = C.CDLL(REF_HOST_PATHS[build])
        lib.ref_host_run_fusion.argtypes = [C.c_char_p, C.c_char_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int),
                                            C.POINTER(C.c_int), C.c_int, C.c_float, C.c_int]
        _ref_host[build] = lib
    n = len(cams)
    dense = os.path.join(workdir, "dense")
    out = os.path.join(workdir, "dense", "ACMMP")
    os.makedirs(os.path.join(dense, "images"), exist_ok=True)
    os.makedirs(os.path.join(dense, "cams"), exist_ok=True)
    for i in range(n):
        h, w = depths[i].shape
        col = np.zeros((h, w, 3), np.uint8) if colors is None else np.ascontiguousarray(colors[i], np.uint8)
        with open(os.path.join(dense, "images", "%08d.jpg" % i), "wb") as f:      # raw container, see oracle/ref_stub
            f.write(b"ACMRAW1 %d %d 3\n" % (h, w))
            f.write(col.tobytes())
        _write_cam(os.path.join(dense, "cams", "%08d_cam.txt" % i), cams[i])
        vd = os.path.join(out, "2333_%08d" % i)
        os.makedirs(vd, exist_ok=True)
        _write_dmb(os.path.join(vd, "depths_geom.dmb" if geom else "depths.dmb"), depths[i])
        _write_dmb(os.path.join(vd, "normals.dmb"), normals[i])
    ids = (C.c_int * n)(*range(n))
    offs = [0]
    flat = []
    for p in pairs:
        flat += list(p)
        offs.append(len(flat))
    import time
    t0 = time.time()
    rc = _ref_host[build].ref_host_run_fusion(dense.encode(), out.encode(), n, ids, (C.c_int * (n + 1))(*offs),
                                              (C.c_int * max(1, len(flat)))(*flat), int(geom), float(consistency_scalar), int(con_num_thresh))
    if timing is not None:
        timing["run_fusion_s"] = time.time() - t0     # includes its file reads and the PLY write, as in the reference
    assert rc == 0
    with open(os.path.join(out, "ACMMP_model.ply"), "rb") as f:
        blob = f.read()
    end = blob.index(b"end_header\n") + len(b"end_header\n")
    header = blob[:end].decode()
    count = int([l for l in header.splitlines() if l.startswith("element vertex")][0].split()[-1])
    rec = np.frombuffer(blob[end:end + 27 * count], dtype=np.dtype([("f", "<f4", 6), ("c", "u1", 3)]))
    pts = np.zeros((count, 9), np.float32)
    pts[:, :6] = rec["f"]
    pts[:, 6:] = rec["c"]
    return pts


def _read_ply_points(path):
    with open(path, "rb") as f:
        blob = f.read()
    end = blob.index(b"end_header\n") + len(b"end_header\n")
    count = int([l for l in blob[:end].decode().splitlines() if l.startswith("element vertex")][0].split()[-1])
    rec = np.frombuffer(blob[end:end + 27 * count], dtype=np.dtype([("f", "<f4", 6), ("c", "u1", 3)]))
    pts = np.zeros((count, 9), np.float32)
    pts[:, :6] = rec["f"]
    pts[:, 6:] = rec["c"]
    return pts


def host_prior_fusion(workdir, cams, depths, normals, depths2, normals2, colors, pairs, consistency_scalar=0.3,
                      num_consistent_thresh=1, single_match_penalty=1, geom=True, build="ieee"):
    """The reference's own RunPriorAwareFusion (acmmp_definitions.cpp:575-826) on its own on-disk layout:
    map set 0 under <dense>/FUSION (its `fusion_folder`), set 1 under <dense>/ACMMP (its `outfolder`);
    returns the points of ACMMP_prior_model.ply as (N, 9) floats."""
    key = "prior_" + build
    if key not in _ref_host:
        lib = C.CDLL(REF_HOST_PATHS[build])
        lib.ref_host_run_prior_fusion.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int),
                                                  C.POINTER(C.c_int), C.c_int, C.c_float, C.c_int, C.c_int]
        _ref_host[key] = lib
    n = len(cams)
    dense = os.path.join(workdir, "dense")
    out = os.path.join(dense, "ACMMP")
    fus = os.path.join(dense, "FUSION")
    os.makedirs(os.path.join(dense, "images"), exist_ok=True)
    os.makedirs(os.path.join(dense, "cams"), exist_ok=True)
    for i in range(n):
        h, w = depths[i].shape
        col = np.zeros((h, w, 3), np.uint8) if colors is None else np.ascontiguousarray(colors[i], np.uint8)
        with open(os.path.join(dense, "images", "%08d.jpg" % i), "wb") as f:      # raw container, see oracle/ref_stub
            f.write(b"ACMRAW1 %d %d 3\n" % (h, w))
            f.write(col.tobytes())
        _write_cam(os.path.join(dense, "cams", "%08d_cam.txt" % i), cams[i])
        for folder, d, nm in ((fus, depths[i], normals[i]), (out, depths2[i], normals2[i])):
            vd = os.path.join(folder, "2333_%08d" % i)
            os.makedirs(vd, exist_ok=True)
            _write_dmb(os.path.join(vd, "depths_geom.dmb" if geom else "depths.dmb"), d)
            _write_dmb(os.path.join(vd, "normals.dmb"), nm)
    ids = (C.c_int * n)(*range(n))
    offs, flat = [0], []
    for p in pairs:
        flat += list(p)
        offs.append(len(flat))
    rc = _ref_host[key].ref_host_run_prior_fusion(dense.encode(), out.encode(), fus.encode(), n, ids, (C.c_int * (n + 1))(*offs),
                                                  (C.c_int * max(1, len(flat)))(*flat), int(geom), float(consistency_scalar),
                                                  int(num_consistent_thresh), int(single_match_penalty))
    assert rc == 0
    return _read_ply_points(os.path.join(out, "ACMMP_prior_model.ply"))


def write_prior_images(dense, index, depth16, normal16):
    """<dense>/priors/{depths,normals}/%08d.png as the 16-bit raw container of oracle/ref_stub"""
    for sub, arr in (("depths", depth16), ("normals", normal16)):
        folder = os.path.join(dense, "priors", sub)
        os.makedirs(folder, exist_ok=True)
        a = np.ascontiguousarray(arr, np.uint16)
        cn = 1 if a.ndim == 2 else a.shape[2]
        with open(os.path.join(folder, "%08d.png" % index), "wb") as f:
            f.write(b"ACMRAW2 %d %d %d\n" % (a.shape[0], a.shape[1], cn) + a.astype("<u2").tobytes())


def host_prior_planes(dense, ncams, index, cam, rows, cols, build="ieee"):
    """The reference's own pSampler::GetPriorPlaneEstimate on <dense>/priors (written by write_prior_images)."""
    key = "planes_" + build
    if key not in _ref_host:
        lib = C.CDLL(REF_HOST_PATHS[build])
        lib.ref_host_prior_planes.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_float)]
        _ref_host[key] = lib
    out = np.zeros((rows, cols, 4), np.float32)
    c = cam.copy() if hasattr(cam, "copy") else cam
    rc = _ref_host[key].ref_host_prior_planes((dense.rstrip("/") + "/").encode(), ncams, index, C.byref(c), rows, cols,
                                              out.ctypes.data_as(C.POINTER(C.c_float)))
    assert rc == 0, rc
    return out


# ---- the reference PROGRAM (kernels + host class + ProcessProblem + JBU + RunFusion) -----------
REF_FULL_PATH = os.path.join(_REF_DIR, "libacmmp_ref_full.so")
_ref_full = None


def full_available(variant=""):
    return os.path.exists(REF_FULL_PATH if not variant else REF_FULL_PATH.replace(".so", f"_{variant}.so"))


def write_dense_folder(dense, cams, images, pairs):
    """the reference's input layout: images/%08d.jpg (raw container, see oracle/ref_stub), cams/, pair.txt"""
    os.makedirs(os.path.join(dense, "images"), exist_ok=True)
    os.makedirs(os.path.join(dense, "cams"), exist_ok=True)
    n = len(cams)
    with open(os.path.join(dense, "pair.txt"), "w") as f:
        f.write("%d\n" % n)
        for i in range(n):
            f.write("%d\n%d " % (i, len(pairs[i])) + " ".join("%d 1.0" % j for j in pairs[i]) + "\n")
    for i in range(n):
        img = np.clip(np.rint(np.asarray(images[i])), 0, 255).astype(np.uint8)
        with open(os.path.join(dense, "images", "%08d.jpg" % i), "wb") as f:
            f.write(b"ACMRAW1 %d %d 1\n" % img.shape + img.tobytes())
        _write_cam(os.path.join(dense, "cams", "%08d_cam.txt" % i), cams[i])


def read_dmb(path):
    hdr = np.fromfile(path, np.int32, 4)
    a = np.fromfile(path, np.float32, offset=16)
    return a.reshape(hdr[1], hdr[2]) if hdr[3] == 1 else a.reshape(hdr[1], hdr[2], hdr[3])


def read_ply(path):
    with open(path, "rb") as f:
        blob = f.read()
    end = blob.index(b"end_header\n") + len(b"end_header\n")
    count = int([l for l in blob[:end].decode().splitlines() if l.startswith("element vertex")][0].split()[-1])
    rec = np.frombuffer(blob[end:end + 27 * count], dtype=np.dtype([("f", "<f4", 6), ("c", "u1", 3)]))
    return rec["f"].copy(), rec["c"].copy()


_ref_full_variants = {}


def _full_lib(variant=""):
    """variant "": the program as shipped (#define DEBUG on); "p2": built without that line (BASELINE.md patch P2)"""
    if variant not in _ref_full_variants:
        lib = C.CDLL(REF_FULL_PATH if not variant else REF_FULL_PATH.replace(".so", f"_{variant}.so"))
        lib.ref_full_run.argtypes = [C.c_char_p, C.c_char_p, C.c_uint64, C.c_int, C.c_float, C.c_int,
                                     C.POINTER(C.c_double), C.POINTER(C.c_double)]
        lib.ref_full_delaunay_seconds.argtypes = [C.c_int]
        lib.ref_full_delaunay_seconds.restype = C.c_double
        _ref_full_variants[variant] = lib
    return _ref_full_variants[variant]


def full_delaunay_seconds(variant="", reset=True):
    """seconds the reference program spent in the stand-in for cv::Subdiv2D (the product's triangulation behind a hook)"""
    return float(_full_lib(variant).ref_full_delaunay_seconds(int(reset)))


def full_run(dense, seed=1234, fuse=True, consistency_scalar=0.3, num_consistent_thresh=1, output_dir="/ACMMP", quiet=True, variant=""):
    """The reference program on `dense` (written by write_dense_folder).  Returns seconds spent in
    the PatchMatch schedule (all ProcessProblem + JBU calls, file round trips included) and in RunFusion."""
    _ref_full = _full_lib(variant)
    a, b = C.c_double(), C.c_double()
    saved = None
    if quiet:   # the reference prints a few lines per view
        import sys
        sys.stdout.flush()
        saved = os.dup(1)
        devnull = os.open(os.devnull, os.O_WRONLY)
        os.dup2(devnull, 1)
        os.close(devnull)
    try:
        rc = _ref_full.ref_full_run(dense.encode(), output_dir.encode(), seed, int(fuse), consistency_scalar, num_consistent_thresh,
                                    C.byref(a), C.byref(b))
    finally:
        if saved is not None:
            os.dup2(saved, 1)
            os.close(saved)
    assert rc == 0
    return a.value, b.value


# ---- the drop-in proof: the reference's unchanged ACMMP.h + acmmp_definitions.cpp over libacmmp_b200.so --------------
DROPIN_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "oracle", "_ref", "libacmmp_dropin.so")
_dropin = None


def dropin_available():
    return os.path.exists(DROPIN_PATH)


def dropin_run(dense, seed=1234, fuse=True, consistency_scalar=0.3, num_consistent_thresh=1, output_dir="/ACMMP", quiet=True):
    """main() of the reference on top of oracle/dropin/ACMMP_dropin.cpp (class ACMMP over the C ABI)"""
    global _dropin
    if _dropin is None:
        _dropin = C.CDLL(DROPIN_PATH)
        _dropin.dropin_run.argtypes = [C.c_char_p, C.c_char_p, C.c_uint64, C.c_int, C.c_float, C.c_int]
    saved = None
    if quiet:
        import sys
        sys.stdout.flush()
        saved = os.dup(1)
        devnull = os.open(os.devnull, os.O_WRONLY)
        os.dup2(devnull, 1)
        os.close(devnull)
    try:
        rc = _dropin.dropin_run(dense.encode(), output_dir.encode(), seed, int(fuse), consistency_scalar, num_consistent_thresh)
    finally:
        if saved is not None:
            os.dup2(saved, 1)
            os.close(saved)
    assert rc == 0


# ---- the reference's own planar-prior functions (ACMMP.cpp:868-958) and what its ProcessProblem builds -------------
def full_support_points(cam, costs):
    """ACMMP::GetSupportPoints of the reference on a cost map (host code; no GPU needed)"""
    lib = _full_lib()
    c = np.ascontiguousarray(costs, np.float32)
    h, w = c.shape
    cc = cam.copy()
    cc.width, cc.height = w, h
    cap = ((w + 4) // 5) * ((h + 4) // 5)
    xy = np.zeros((cap, 2), np.int32)
    lib.ref_full_support_points.argtypes = [C.POINTER(L.Camera), C.POINTER(C.c_float), C.POINTER(C.c_int), C.c_int]
    n = lib.ref_full_support_points(C.byref(cc), L.fptr(c), xy.ctypes.data_as(C.POINTER(C.c_int)), cap)
    return xy[:n]


def full_prior_planes(cam, depth, tris):
    """ACMMP::GetPriorPlaneParams of the reference for every triangle (cv::SVD::solveZ: the stand-in's cross product)"""
    lib = _full_lib()
    d = np.ascontiguousarray(depth, np.float32)
    cc = cam.copy()
    cc.height, cc.width = d.shape
    t = np.ascontiguousarray(tris, np.int32).reshape(-1, 6)
    out = np.zeros((len(t), 4), np.float32)
    lib.ref_full_prior_planes.argtypes = [C.POINTER(L.Camera), C.POINTER(C.c_float), C.POINTER(C.c_int), C.c_int, C.POINTER(C.c_float)]
    lib.ref_full_prior_planes(C.byref(cc), L.fptr(d), t.ctypes.data_as(C.POINTER(C.c_int)), len(t), L.fptr(out))
    return out


def full_depths_from_planes(cam, planes, xy):
    """ACMMP::GetDepthFromPlaneParam of the reference at pixels xy (n, 2) for planes (n, 4)"""
    lib = _full_lib()
    p = np.ascontiguousarray(planes, np.float32).reshape(-1, 4)
    q = np.ascontiguousarray(xy, np.int32).reshape(-1, 2)
    out = np.zeros(len(p), np.float32)
    cc = cam.copy()
    lib.ref_full_depths_from_planes.argtypes = [C.POINTER(L.Camera), C.POINTER(C.c_float), C.POINTER(C.c_int), C.c_int, C.POINTER(C.c_float)]
    lib.ref_full_depths_from_planes(C.byref(cc), L.fptr(p), q.ctypes.data_as(C.POINTER(C.c_int)), len(p), L.fptr(out))
    return out


def full_capture_prior(dense, idx=0, seed=1234, output_dir="/ACMMP_CAPTURE"):
    """One ProcessProblem (photometric + planar prior) of the reference program on view `idx`; returns what its driver built:
    dict(depth (h, w) the prior was built from, points (n, 2), tris (m, 6), planes (k, 4) = planeParams_tri,
    mask (h, w) float = mask_tri after the range check).  Needs a GPU (RunPatchMatch)."""
    lib = _full_lib()
    ip = C.POINTER(C.c_int)
    lib.ref_full_capture_prior.argtypes = [C.c_char_p, C.c_char_p, C.c_int, C.c_uint64, ip, ip, ip, ip, ip]
    lib.ref_full_captured.argtypes = [C.POINTER(C.c_float), ip, ip, C.POINTER(C.c_float), C.POINTER(C.c_float)]
    w, h, ns, nt, npl = C.c_int(), C.c_int(), C.c_int(), C.c_int(), C.c_int()
    import sys
    sys.stdout.flush()
    saved = os.dup(1)
    devnull = os.open(os.devnull, os.O_WRONLY)
    os.dup2(devnull, 1)
    os.close(devnull)
    try:
        rc = lib.ref_full_capture_prior(dense.encode(), output_dir.encode(), idx, seed, C.byref(w), C.byref(h), C.byref(ns), C.byref(nt), C.byref(npl))
    finally:
        os.dup2(saved, 1)
        os.close(saved)
    assert rc == 0
    depth = np.zeros((h.value, w.value), np.float32)
    mask = np.zeros((h.value, w.value), np.float32)
    pts = np.zeros((ns.value, 2), np.int32)
    tris = np.zeros((nt.value, 6), np.int32)
    planes = np.zeros((npl.value, 4), np.float32)
    assert lib.ref_full_captured(L.fptr(depth), pts.ctypes.data_as(ip), tris.ctypes.data_as(ip), L.fptr(planes), L.fptr(mask)) == 0
    return {"depth": depth, "points": pts, "tris": tris, "planes": planes, "mask": mask}
