"""The reference's own flow for the path, driven stage by stage like its main():
ProcessProblem / JointBilateralUpsampling (acmmp_definitions.cpp:245-440) with the
reference's UNMODIFIED CUDA kernels (oracle/_ref/libacmmp_ref.so, built from
/root/reference/src/ACMMP.cu for sm_100 with the reference's flags) and the
reference's host steps restated (oracle/prior.c) or delegated to the reference's
real dependency (cv2.resize, cv2.Subdiv2D).

TEST INFRASTRUCTURE: used by bench.py --impl reference (the baseline arm) and by
tests/ (full-pipeline parity).  Nothing under acmmp_b200/ imports it.

What it keeps from the reference's design, because it is the design being
compared: a fresh problem set-up per view per stage (cudaMallocArray + H2D of
the reference image and every source image, texture objects, per-pixel buffers:
ACMMP.cpp:638-809), the per-view resize of every image (ACMMP.cpp:566-598), a
device-wide synchronisation after every kernel (ACMMP.cu:1421-1448), D2H of the
plane map and costs after every run (ACMMP.cu:1453-1455), the CPU planar-prior
construction, and stage-to-stage hand-over through host memory.  What it leaves
out, in the reference's favour: JPEG decoding (ACMMP.cpp:539-557, per view per
stage), the .dmb disk round trips (acmmp_definitions.cpp:392-401) and the
cv::imshow debug windows (#define DEBUG, ACMMP.cu:3).
"""
import math
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

import oraclebind as ob  # noqa: E402
import refbind  # noqa: E402


def multi_scale_settings(cams, max_image_size=3200, size_bound=1000):
    """ComputeMultiScaleSettings, acmmp_definitions.cpp:207-243."""
    mis, nds = [], []
    for c in cams:
        m = min(max(c.height, c.width), max_image_size)
        mis.append(m)
        k = 0
        while m > size_bound:
            m //= 2
            k += 1
        nds.append(k)
    return mis, nds, max(nds)


def level_geometry(cam, max_size):
    """InputInitialization's rescale rule, ACMMP.cpp:566-598 -> (w, h, scaled camera)."""
    c = cam.copy()
    if cam.width <= max_size and cam.height <= max_size:
        return cam.width, cam.height, c
    f = min(np.float32(max_size) / np.float32(cam.width), np.float32(max_size) / np.float32(cam.height))
    w = int(round(float(np.float32(cam.width) * f)))
    h = int(round(float(np.float32(cam.height) * f)))
    sx, sy = np.float32(w) / np.float32(cam.width), np.float32(h) / np.float32(cam.height)
    c.K[0] = float(np.float32(c.K[0]) * sx)
    c.K[2] = float(np.float32(c.K[2]) * sx)
    c.K[4] = float(np.float32(c.K[4]) * sy)
    c.K[5] = float(np.float32(c.K[5]) * sy)
    c.width, c.height = w, h
    return w, h, c


class RefScan:
    def __init__(self, images, cams, pairs, device=0):
        import cv2
        self.cv2 = cv2
        self.images = [np.ascontiguousarray(i, np.float32) for i in images]
        self.cams, self.pairs, self.n = cams, pairs, len(cams)
        self.max_image_size, self.num_downscale, self.max_num_downscale = multi_scale_settings(cams)
        self.cur_size = list(self.max_image_size)
        self.ref = refbind.RefACMMP(device)
        self.depths = [None] * self.n        # depths.dmb
        self.depths_geom = [None] * self.n   # depths_geom.dmb
        self.normals = [None] * self.n       # normals.dmb
        self.costs = [None] * self.n         # costs.dmb
        self.kernel_ms = 0.0
        self.prior_s = 0.0
        self.runs = 0
        self.stats = {}

    def num_levels(self):
        return self.max_num_downscale + 1

    def begin_level(self, first, views=None):
        """main_ACMMP.cpp:99-105 (+ :142-145 JBU on finer scales)."""
        for i in range(self.n):
            if self.num_downscale[i] >= 0:
                self.cur_size[i] = int(self.max_image_size[i] / math.pow(2, self.num_downscale[i]))
                self.num_downscale[i] -= 1
        if not first:
            for i in (range(self.n) if views is None else views):
                self.joint_bilateral_upsampling(i)

    def _scaled(self, v):
        w, h, c = level_geometry(self.cams[v], self.cur_size[v])
        img = self.images[v]
        if (w, h) != (img.shape[1], img.shape[0]):
            img = self.cv2.resize(img, (w, h), interpolation=self.cv2.INTER_LINEAR)
        return img, c

    def joint_bilateral_upsampling(self, v):
        """acmmp_definitions.cpp:405-440 + RunJBU ACMMP.cpp:1022-1087."""
        img, _ = self._scaled(v)
        out = refbind.ref_jbu(img, self.depths_geom[v])
        if out is not None:
            self.depths[v] = out

    def process_problem(self, v, geom, planar_prior, hierarchy, multi_geometry, seed):
        """ProcessProblem, acmmp_definitions.cpp:245-403."""
        ref = self.ref
        ids = [v] + list(self.pairs[v])
        scaled = [self._scaled(i) for i in ids]
        images = [s[0] for s in scaled]
        cams = [s[1] for s in scaled]
        ref.set_problem(images, cams)
        h, w = images[0].shape
        ref.set_flags(geom=int(geom), planar_prior=0, multi_geometry=int(multi_geometry), hierarchy=int(hierarchy), upsample=0,
                      seeded=0, scaled_cols=0.0, scaled_rows=0.0)
        if geom:
            src = self.depths_geom if multi_geometry else self.depths
            for k, i in enumerate(ids):
                ref.set_depth_map(k, src[i])
            planes = np.concatenate([self.normals[v], src[v][..., None]], -1)
            ref.set_planes(0, planes)
            ref.set_scalars(0, self.costs[v])
        if hierarchy:
            cn, cc = self.normals[v], self.costs[v]
            ch, cw = cc.shape
            ref.set_flags(upsample=1, scaled_cols=float(cw), scaled_rows=float(ch))   # ACMMP.cpp:766 (always true for w != h)
            ref.set_planes(3, np.concatenate([cn, cc[..., None]], -1))
            fp = np.zeros((h, w, 4), np.float32)
            fp[..., 3] = self.depths[v]
            ref.set_planes(0, fp)
        self.kernel_ms += ref.run(seed)
        self.runs += 1
        planes, costs = ref.get_result()
        if planar_prior:
            t0 = time.time()
            dmin = np.float32(cams[0].depth_min) * np.float32(0.6)
            dmax = np.float32(cams[0].depth_max) * np.float32(1.2)
            pts = ob.support_points(costs)
            tris = np.zeros((0, 6), np.int32)
            if len(pts):
                sub = self.cv2.Subdiv2D((0, 0, w, h))
                sub.insert(pts.astype(np.float32))
                tl = sub.getTriangleList().astype(np.int32)   # (int) truncation, ACMMP.cpp:912-914
                x, y = tl[:, 0::2], tl[:, 1::2]
                inside = ((x >= 0) & (x < w) & (y >= 0) & (y < h)).all(1)
                tris = np.ascontiguousarray(tl[inside])
            prior, mask, _ = ob.prior_from_triangles(cams[0], planes[..., 3], tris, dmin, dmax)
            self.prior_s += time.time() - t0
            self.stats.setdefault("support", []).append(len(pts))
            self.stats.setdefault("triangles", []).append(len(tris))
            ref.set_flags(planar_prior=1)
            ref.set_planes(1, prior)
            ref.set_scalars(3, mask)
            self.kernel_ms += ref.run(seed + 1)
            self.runs += 1
            planes, costs = ref.get_result()
        depth = np.ascontiguousarray(planes[..., 3])
        if geom:
            self.depths_geom[v] = depth
        else:
            self.depths[v] = depth
        self.normals[v] = np.ascontiguousarray(planes[..., :3])
        self.costs[v] = costs

    def run_level(self, level, views, seed, external=None):
        """one iteration of the scale loop, main_ACMMP.cpp:96-175, over `views`.
        external(i, w, h): depth map standing in for a view outside `views` (bounded
        samples of a scan: the work per processed view stays that of the full scan)."""
        self.begin_level(level == 0, views)
        if external is not None:
            for i in range(self.n):
                if i not in views:
                    w, h, _ = level_geometry(self.cams[i], self.cur_size[i])
                    self.depths[i] = self.depths_geom[i] = np.ascontiguousarray(external(i, w, h), np.float32)
        for v in views:
            self.process_problem(v, False, True, level > 0, False, self._seed(seed, v, level, 0))
        for it in range(2):
            for v in views:
                self.process_problem(v, True, False, False, it > 0, self._seed(seed, v, level, 1 + it))

    @staticmethod
    def _seed(seed, v, level, stage):
        return (seed * 1000003 + v * 8191 + level * 131 + stage * 17) & 0xFFFFFFFFFFFFFFFF

    def run_all(self, seed=1234, views=None, external=None):
        views = list(range(self.n)) if views is None else views
        self.max_image_size, self.num_downscale, self.max_num_downscale = multi_scale_settings(self.cams)
        for level in range(self.num_levels()):
            self.run_level(level, views, seed, external)
