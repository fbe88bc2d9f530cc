"""ctypes binding of oracle/liboracle.so - the CPU restatement of the reference's
PatchMatch path (oracle/*.c).  TEST INFRASTRUCTURE: imported only by tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs."""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_PATH = os.path.join(ROOT, "oracle", "liboracle.so")


class OrcCamera(C.Structure):
    _fields_ = [("K", C.c_float * 9), ("R", C.c_float * 9), ("t", C.c_float * 3),
                ("height", C.c_int32), ("width", C.c_int32),
                ("depth_min", C.c_float), ("depth_max", C.c_float)]


fp = C.POINTER(C.c_float)
up = C.POINTER(C.c_uint32)


class OrcProblem(C.Structure):
    _fields_ = [("n", C.c_int32), ("cams", C.POINTER(OrcCamera)),
                ("images", C.POINTER(fp)), ("depths", C.POINTER(fp)),
                ("depth_min", C.c_float), ("depth_max", C.c_float),
                ("sigma_spatial", C.c_float), ("sigma_color", C.c_float),
                ("top_k", C.c_int32), ("geom", C.c_int32), ("planar_prior", C.c_int32), ("hierarchy", C.c_int32),
                ("compat", C.c_int32), ("scaled_cols", C.c_float), ("scaled_rows", C.c_float),
                ("planes", fp), ("costs", fp), ("pre_costs", fp), ("sel", up), ("rng6", up),
                ("prior_planes", fp), ("masks", up), ("scaled_planes", fp), ("seeded", fp)]


_lib = None


def available():
    return os.path.exists(ORACLE_PATH)


def load():
    global _lib
    if _lib is not None:
        return _lib
    lib = C.CDLL(ORACLE_PATH)
    u6 = C.c_uint32 * 6
    lib.orc_rng_seed.argtypes = [C.c_uint64, C.c_uint64, C.c_uint64, C.POINTER(u6)]
    lib.orc_rng_seed.restype = None
    lib.orc_rng_next.argtypes = [C.POINTER(u6)]
    lib.orc_rng_next.restype = C.c_uint32
    lib.orc_rng_uniform.argtypes = [C.POINTER(u6)]
    lib.orc_rng_uniform.restype = C.c_float
    lib.orc_rng_init_image.argtypes = [C.c_uint64, C.c_int, C.c_int, up]
    lib.orc_rng_init_image.restype = None
    lib.orc_tex_bilinear.argtypes = [fp, C.c_int, C.c_int, C.c_float, C.c_float]
    lib.orc_tex_bilinear.restype = C.c_float
    lib.orc_homography.argtypes = [C.POINTER(OrcCamera), C.POINTER(OrcCamera), fp, fp]
    lib.orc_homography.restype = None
    pp = C.POINTER(OrcProblem)
    lib.orc_ncc.argtypes = [pp, C.c_int, C.c_int, fp, C.c_int]
    lib.orc_ncc.restype = C.c_float
    lib.orc_geom_cost.argtypes = [pp, C.c_int, C.c_int, fp, C.c_int]
    lib.orc_geom_cost.restype = C.c_float
    lib.orc_plane_depth.argtypes = [C.POINTER(OrcCamera), fp, C.c_int, C.c_int]
    lib.orc_plane_depth.restype = C.c_float
    lib.orc_plane_offset.argtypes = [C.POINTER(OrcCamera), C.c_int, C.c_int, C.c_float, fp]
    lib.orc_plane_offset.restype = C.c_float
    lib.orc_initial_cost.argtypes = [pp, C.c_int, C.c_int, fp, up, fp]
    lib.orc_initial_cost.restype = C.c_float
    lib.orc_init.argtypes = [pp, C.c_int] + [C.c_int] * 4
    lib.orc_init.restype = None
    lib.orc_sweep.argtypes = [pp, C.c_int, C.c_int] + [C.c_int] * 4
    lib.orc_sweep.restype = None
    lib.orc_depth_normal.argtypes = [pp]
    lib.orc_depth_normal.restype = None
    lib.orc_median_filter.argtypes = [pp, C.c_int]
    lib.orc_median_filter.restype = None
    lib.orc_run.argtypes = [pp, C.c_int, C.c_int]
    lib.orc_run.restype = None
    lib.orc_jbu_depth.argtypes = [fp, C.c_int, C.c_int, fp, C.c_int, C.c_int, fp]
    lib.orc_resize_linear.argtypes = [fp, C.c_int, C.c_int, fp, C.c_int, C.c_int]
    lib.orc_resize_linear.restype = None
    lib.orc_fusion.argtypes = [C.c_int, C.POINTER(OrcCamera), C.POINTER(fp), C.POINTER(fp),
                               C.POINTER(C.POINTER(C.c_uint8)), C.POINTER(C.c_int), C.POINTER(C.c_int),
                               C.c_float, C.c_int, fp, C.c_long]
    lib.orc_fusion.restype = C.c_long
    u8pp = C.POINTER(C.POINTER(C.c_uint8))
    lib.orc_prior_fusion.argtypes = [C.c_int, C.POINTER(OrcCamera), C.POINTER(fp), C.POINTER(fp), C.POINTER(fp), C.POINTER(fp),
                                     u8pp, u8pp, C.POINTER(C.c_int), C.POINTER(C.c_int), C.c_float, C.c_int, C.c_int, fp, C.c_long]
    lib.orc_prior_fusion.restype = C.c_long
    u16p = C.POINTER(C.c_uint16)
    lib.orc_prior_planes_from_maps.argtypes = [C.POINTER(OrcCamera), u16p, u16p, C.c_int, C.c_int, C.c_int, C.c_int, fp]
    lib.orc_prior_planes_from_maps.restype = None
    i32p = C.POINTER(C.c_int32)
    lib.orc_support_points.argtypes = [fp, C.c_int, C.c_int, i32p]
    lib.orc_prior_from_triangles.argtypes = [C.POINTER(OrcCamera), fp, C.c_int, C.c_int, i32p, C.c_int, C.c_float, C.c_float, fp, fp, up]
    lib.orc_prior_from_triangles.restype = None
    _lib = lib
    return lib


def _f32(a):
    return np.ascontiguousarray(a, np.float32)


def _fp(a):
    return a.ctypes.data_as(fp) if a is not None else None


def _up(a):
    return a.ctypes.data_as(up) if a is not None else None


def to_orc_camera(cam):
    """accepts acmmp_b200.lib.Camera (same 100-byte layout)."""
    o = OrcCamera()
    C.memmove(C.byref(o), C.byref(cam), C.sizeof(OrcCamera))
    return o


INIT_RANDOM, INIT_SEEDED, INIT_PRIOR, INIT_UPSAMPLE, INIT_WARM = range(5)


class Problem:
    """One (reference view, sources) problem with its per-pixel state held in numpy."""

    def __init__(self, images, cams, depths=None, geom=0, planar_prior=0, hierarchy=0, compat=0):
        self.lib = load()
        self.n = len(images)
        self.images = [_f32(i) for i in images]
        self.cams = (OrcCamera * self.n)(*[to_orc_camera(c) for c in cams])
        self.h, self.w = self.images[0].shape
        P = self.h * self.w
        self.depths = [None if d is None else _f32(d) for d in depths] if depths is not None else None
        self.planes = np.zeros((self.h, self.w, 4), np.float32)
        self.costs = np.zeros((self.h, self.w), np.float32)
        self.pre_costs = np.zeros((self.h, self.w), np.float32)
        self.sel = np.zeros((self.h, self.w), np.uint32)
        self.rng = np.zeros((self.h, self.w, 6), np.uint32)
        self.prior_planes = np.zeros((self.h, self.w, 4), np.float32)
        self.masks = np.zeros((self.h, self.w), np.uint32)
        self.scaled_planes = None
        self.seeded = None
        p = OrcProblem()
        p.n = self.n
        p.cams = self.cams
        self._img_ptrs = (fp * self.n)(*[_fp(i) for i in self.images])
        p.images = self._img_ptrs
        if self.depths is not None:
            self._dep_ptrs = (fp * self.n)(*[_fp(d) if d is not None else None for d in self.depths])
            p.depths = self._dep_ptrs
        p.depth_min = np.float32(cams[0].depth_min) * np.float32(0.6)   # ACMMP.cpp:600-601
        p.depth_max = np.float32(cams[0].depth_max) * np.float32(1.2)
        p.sigma_spatial, p.sigma_color, p.top_k = 5.0, 3.0, 4
        p.geom, p.planar_prior, p.hierarchy, p.compat = geom, planar_prior, hierarchy, compat
        self.p = p
        self._sync()

    def _sync(self):
        p = self.p
        p.planes, p.costs, p.pre_costs = _fp(self.planes), _fp(self.costs), _fp(self.pre_costs)
        p.sel, p.rng6 = _up(self.sel), _up(self.rng)
        p.prior_planes, p.masks = _fp(self.prior_planes), _up(self.masks)
        p.scaled_planes = _fp(self.scaled_planes)
        p.seeded = _fp(self.seeded)

    def set_seeded(self, planes):
        self.seeded = _f32(planes)
        self._sync()

    def set_prior(self, planes, masks):
        self.prior_planes = _f32(planes)
        self.masks = np.ascontiguousarray(masks, np.uint32)
        self._sync()

    def set_scaled(self, planes4, cols, rows):
        self.scaled_planes = _f32(planes4)
        self.p.scaled_cols, self.p.scaled_rows = float(cols), float(rows)
        self._sync()

    def set_state(self, planes=None, costs=None, sel=None, rng=None):
        if planes is not None:
            self.planes = _f32(planes).copy()
        if costs is not None:
            self.costs = _f32(costs).copy()
        if sel is not None:
            self.sel = np.ascontiguousarray(sel, np.uint32).copy()
        if rng is not None:
            self.rng = np.ascontiguousarray(rng, np.uint32).copy()
        self._sync()

    def rng_init(self, seed):
        self.lib.orc_rng_init_image(int(seed), self.w, self.h, _up(self.rng))

    def region(self, region):
        return region if region is not None else (0, 0, self.w, self.h)

    def init(self, mode, region=None):
        self.lib.orc_init(C.byref(self.p), mode, *self.region(region))

    def sweep(self, it, color, region=None):
        self.lib.orc_sweep(C.byref(self.p), it, color, *self.region(region))

    def tail(self):
        self.lib.orc_depth_normal(C.byref(self.p))
        self.lib.orc_median_filter(C.byref(self.p), 0)
        self.lib.orc_median_filter(C.byref(self.p), 1)

    def run(self, mode, seed, iterations=2):
        self.rng_init(seed)
        self.lib.orc_run(C.byref(self.p), mode, iterations)

    def ncc(self, px, py, plane, view):
        pl = _f32(plane)
        return float(self.lib.orc_ncc(C.byref(self.p), px, py, _fp(pl), view))

    def ncc_map(self, planes, view, region=None):
        x0, y0, x1, y1 = self.region(region)
        pl = _f32(planes)
        out = np.zeros((self.h, self.w), np.float32)
        for y in range(y0, y1):
            for x in range(x0, x1):
                out[y, x] = self.lib.orc_ncc(C.byref(self.p), x, y, _fp(pl[y, x]), view)
        return out

    def geom_cost(self, px, py, plane, view):
        pl = _f32(plane)
        return float(self.lib.orc_geom_cost(C.byref(self.p), px, py, _fp(pl), view))


def rng_state(seed, subsequence, offset):
    st = (C.c_uint32 * 6)()
    load().orc_rng_seed(int(seed), int(subsequence), int(offset), C.byref(st))
    return st


def rng_init_image(seed, w, h):
    out = np.zeros((h, w, 6), np.uint32)
    load().orc_rng_init_image(int(seed), w, h, _up(out))
    return out


def homography(cam_r, cam_s, plane):
    H = np.zeros(9, np.float32)
    pl = _f32(plane)
    load().orc_homography(C.byref(to_orc_camera(cam_r)), C.byref(to_orc_camera(cam_s)), _fp(pl), _fp(H))
    return H


def tex_bilinear(img, x, y):
    im = _f32(img)
    return float(load().orc_tex_bilinear(_fp(im), im.shape[1], im.shape[0], x, y))


def jbu_depth(image, depth):
    img, dep = _f32(image), _f32(depth)
    out = np.empty_like(img)
    rc = load().orc_jbu_depth(_fp(img), img.shape[1], img.shape[0], _fp(dep), dep.shape[1], dep.shape[0], _fp(out))
    return None if rc == 1 else out


def resize_linear(image, new_w, new_h):
    img = _f32(image)
    out = np.empty((new_h, new_w), np.float32)
    load().orc_resize_linear(_fp(img), img.shape[1], img.shape[0], _fp(out), new_w, new_h)
    return out


def fusion(cams, depths, normals, pairs, consistency_scalar=0.3, con_num_thresh=1, colors=None, want_points=False):
    """RunFusion restatement; pairs[i] = list of source view indices of view i."""
    n = len(cams)
    cam_arr = (OrcCamera * n)(*[to_orc_camera(c) for c in cams])
    d = [_f32(x) for x in depths]
    nm = [_f32(x) for x in normals]
    dp = (fp * n)(*[_fp(x) for x in d])
    npz = (fp * n)(*[_fp(x) for x in nm])
    offs = np.zeros(n + 1, np.int32)
    for i, p in enumerate(pairs):
        offs[i + 1] = offs[i] + len(p)
    ids = np.ascontiguousarray([j for p in pairs for j in p], np.int32)
    cp = None
    if colors is not None:
        cols = [np.ascontiguousarray(c, np.uint8) for c in colors]
        cp = (C.POINTER(C.c_uint8) * n)(*[c.ctypes.data_as(C.POINTER(C.c_uint8)) for c in cols])
    pts, cap = None, 0
    if want_points:
        cap = int(sum(x.size for x in d))
        pts = np.zeros((cap, 9), np.float32)
    cnt = load().orc_fusion(n, cam_arr, dp, npz, cp, offs.ctypes.data_as(C.POINTER(C.c_int)),
                            ids.ctypes.data_as(C.POINTER(C.c_int)), consistency_scalar, con_num_thresh, _fp(pts), cap)
    return (int(cnt), pts[:cnt]) if want_points else int(cnt)


def prior_fusion(cams, depths, normals, depths2, normals2, pairs, consistency_scalar=0.3, num_consistent_thresh=1,
                 single_match_penalty=1, colors=None, masks=None):
    """RunPriorAwareFusion restatement -> (count, points (count, 9))."""
    n = len(cams)
    cam_arr = (OrcCamera * n)(*[to_orc_camera(c) for c in cams])
    keep = [[_f32(x) for x in arrs] for arrs in (depths, normals, depths2, normals2)]
    ptrs = [(fp * n)(*[_fp(x) for x in arrs]) for arrs in keep]
    offs = np.zeros(n + 1, np.int32)
    for i, p in enumerate(pairs):
        offs[i + 1] = offs[i] + len(p)
    ids = np.ascontiguousarray([j for p in pairs for j in p], np.int32)
    u8p = C.POINTER(C.c_uint8)

    def bytes_of(arrs):
        if arrs is None:
            return None, None
        k = [np.ascontiguousarray(a, np.uint8) for a in arrs]
        return k, (u8p * n)(*[a.ctypes.data_as(u8p) for a in k])
    kc, cp = bytes_of(colors)
    km, mp = bytes_of(masks)
    cap = int(sum(x.size for x in keep[0]))
    pts = np.zeros((cap, 9), np.float32)
    cnt = load().orc_prior_fusion(n, cam_arr, ptrs[0], ptrs[1], ptrs[2], ptrs[3], cp, mp, offs.ctypes.data_as(C.POINTER(C.c_int)),
                                  ids.ctypes.data_as(C.POINTER(C.c_int)), consistency_scalar, num_consistent_thresh,
                                  single_match_penalty, _fp(pts), cap)
    return int(cnt), pts[:cnt]


def prior_planes_from_maps(cam, depth16, normal16, rows, cols):
    """pSampler::GetPriorPlaneEstimate restatement -> (rows, cols, 4) planes (camera-frame normal, distance)."""
    d = np.ascontiguousarray(depth16, np.uint16)
    n = np.ascontiguousarray(normal16, np.uint16)
    assert n.shape == d.shape + (3,)
    out = np.zeros((rows, cols, 4), np.float32)
    oc = to_orc_camera(cam)
    u16p = C.POINTER(C.c_uint16)
    load().orc_prior_planes_from_maps(C.byref(oc), d.ctypes.data_as(u16p), n.ctypes.data_as(u16p), d.shape[0], d.shape[1], rows, cols,
                                      _fp(out))
    return out


def support_points(costs):
    c = _f32(costs)
    h, w = c.shape
    xy = np.zeros((((w + 4) // 5) * ((h + 4) // 5), 2), np.int32)
    n = load().orc_support_points(_fp(c), w, h, xy.ctypes.data_as(C.POINTER(C.c_int32)))
    return xy[:n]


def prior_from_triangles(cam, depth, tris, depth_min, depth_max):
    d = _f32(depth)
    h, w = d.shape
    t = np.ascontiguousarray(tris, np.int32).reshape(-1, 6)
    tri_planes = np.zeros((max(len(t), 1), 4), np.float32)
    prior = np.zeros((h, w, 4), np.float32)
    mask = np.zeros((h, w), np.uint32)
    oc = to_orc_camera(cam)
    load().orc_prior_from_triangles(C.byref(oc), _fp(d), w, h, t.ctypes.data_as(C.POINTER(C.c_int32)), len(t),
                                    float(depth_min), float(depth_max), _fp(tri_planes), _fp(prior), _up(mask))
    return prior, mask, tri_planes[:len(t)]
