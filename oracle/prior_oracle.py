"""CPU restatement of the reference's planar-prior construction, for tests only.

TEST INFRASTRUCTURE (see oracle/oracle.h): imported by tests/ only.

Follows ACMMP::GetSupportPoints (ACMMP.cpp:868-894), DelaunayTriangulation
(ACMMP.cpp:896-918), the raster + range check of ProcessProblem
(acmmp_definitions.cpp:332-374), GetPriorPlaneParams (ACMMP.cpp:920-953) and
GetDepthFromPlaneParam (ACMMP.cpp:955-958).  The triangulation itself is the
reference's real third-party dependency, cv::Subdiv2D, reached through the
Python binding (cv2 4.13 here; the reference's Docker image has OpenCV 3.2 -
Subdiv2D's algorithm has not changed); cv::SVD::solveZ is replaced by
numpy.linalg.svd in float32.
"""
import numpy as np


def support_points(costs):
    """best pixel of each 5x5 cell (column-major cell order) if its cost < 0.1."""
    h, w = costs.shape
    pts = []
    for col in range(0, w, 5):
        for row in range(0, h, 5):
            blk = costs[row:min(h, row + 5), col:min(w, col + 5)]
            best, bc, br = np.float32(2.0), -1, -1
            for c in range(blk.shape[1]):          # column outer, row inner; strict `>`
                for r in range(blk.shape[0]):
                    v = blk[r, c]
                    if v < 2.0 and best > v:
                        best, bc, br = v, col + c, row + r
            if best < 0.1:
                pts.append((bc, br))
    return np.array(pts, np.int32).reshape(-1, 2)


def triangulate_cv2(points, w, h):
    """cv::Subdiv2D insert + getTriangleList, vertices truncated to int, triangles
    with a vertex outside the image dropped (acmmp_definitions.cpp:335)."""
    import cv2
    sub = cv2.Subdiv2D((0, 0, w, h))
    for x, y in points:
        sub.insert((float(x), float(y)))
    tl = sub.getTriangleList().reshape(-1, 3, 2).astype(np.int64)   # (int) truncation
    inside = ((tl[..., 0] >= 0) & (tl[..., 0] < w) & (tl[..., 1] >= 0) & (tl[..., 1] < h)).all(1)
    return tl[inside].astype(np.int32)


def plane_of_triangle(tri, depth, K):
    X = []
    for x, y in tri:
        d = np.float32(depth[y, x])
        X.append([d * (np.float32(x) - K[2]) / K[0], d * (np.float32(y) - K[5]) / K[4], d, np.float32(1.0)])
    A = np.array(X, np.float32)
    _, _, vt = np.linalg.svd(A.astype(np.float64))
    n4 = vt[-1]
    nrm = np.sqrt(n4[0] ** 2 + n4[1] ** 2 + n4[2] ** 2)
    if n4[3] < 0:
        nrm = -nrm
    return (n4 / nrm).astype(np.float32)


def raster_triangle(mask, tri, label):
    """barycentric stepping with float loop counters (acmmp_definitions.cpp:336-353)."""
    (x1, y1), (x2, y2), (x3, y3) = [(int(a), int(b)) for a, b in tri]
    L01 = np.float32(np.sqrt(float((x1 - x2) ** 2 + (y1 - y2) ** 2)))
    L02 = np.float32(np.sqrt(float((x1 - x3) ** 2 + (y1 - y3) ** 2)))
    L12 = np.float32(np.sqrt(float((x2 - x3) ** 2 + (y2 - y3) ** 2)))
    step = np.float32(1.0 / float(max(L01, L02, L12)))
    p = np.float32(0.0)
    f = np.float32
    while float(p) < 1.0:
        q = np.float32(0.0)
        while float(q) < 1.0 - float(p):
            r = 1.0 - float(p) - float(q)
            x = int(float(f(f(p * f(x1)) + f(q * f(x2)))) + r * x3)
            y = int(float(f(f(p * f(y1)) + f(q * f(y2)))) + r * y3)
            mask[y, x] = label
            q = f(q + step)
        p = f(p + step)


def build_prior(costs, depth, K, depth_min, depth_max, triangles=None):
    """returns (prior_planes (H,W,4), labels (H,W) uint32, support points, triangles)."""
    h, w = costs.shape
    K = np.asarray(K, np.float32)
    pts = support_points(costs)
    if triangles is None:
        triangles = triangulate_cv2(pts, w, h) if len(pts) else np.zeros((0, 3, 2), np.int32)
    mask = np.zeros((h, w), np.uint32)
    planes = []
    for i, tri in enumerate(triangles):
        raster_triangle(mask, tri, i + 1)
        planes.append(plane_of_triangle(tri, depth, K))
    planes = np.array(planes, np.float32).reshape(-1, 4)
    prior = np.zeros((h, w, 4), np.float32)
    ys, xs = np.nonzero(mask)
    for y, x in zip(ys, xs):
        pl = planes[mask[y, x] - 1]
        with np.errstate(all="ignore"):
            d = -pl[3] * K[0] / ((np.float32(x) - K[2]) * pl[0] + (K[0] / K[4]) * (np.float32(y) - K[5]) * pl[1] + K[0] * pl[2])
        if d <= depth_max and d >= depth_min:
            prior[y, x] = pl
        else:
            mask[y, x] = 0
    return prior, mask, pts, triangles
