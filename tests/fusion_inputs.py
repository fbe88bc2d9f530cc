"""Inputs shared by the fusion tests (CPU pinning and GPU parity)."""
import numpy as np


def noisy_maps(sc, seed, noise=0.002, holes=0.02, outliers=0.03):
    """depth maps like PatchMatch leaves them: small noise, some holes (0) and gross errors;
    world-frame normals around one direction with a few degrees of scatter"""
    rng = np.random.default_rng(seed)
    depths, normals, colors = [], [], []
    for d in sc["depths"]:
        h, w = d.shape
        x = d * (1.0 + noise * rng.standard_normal((h, w)))
        bad = rng.random((h, w)) < outliers
        x[bad] *= rng.uniform(0.8, 1.2, size=int(bad.sum()))
        x[rng.random((h, w)) < holes] = 0.0
        depths.append(np.ascontiguousarray(x, np.float32))
        n = rng.normal(size=(h, w, 3)) * 0.06 + np.array([0.0, 0.0, 1.0])
        n /= np.linalg.norm(n, axis=-1, keepdims=True)
        normals.append(np.ascontiguousarray(n, np.float32))
        colors.append(rng.integers(0, 256, size=(h, w, 3), dtype=np.uint8))
    return depths, normals, colors


def _world_point(cam, x, y, d):
    K, R, t = np.float32(cam.K[:]), np.float32(cam.R[:]).reshape(3, 3), np.float32(cam.t[:])
    X = np.array([d * (x - K[2]) / K[0], d * (y - K[5]) / K[4], d], np.float32)
    return R.T @ X - R.T @ t


def _project(cam, W):
    K, R, t = np.float32(cam.K[:]).reshape(3, 3), np.float32(cam.R[:]).reshape(3, 3), np.float32(cam.t[:])
    p = K @ (R @ W + t)
    return p[0] / p[2], p[1] / p[2], p[2]


def fusion_no_stale(cams, depths, normals, pairs, scalar, thresh):
    """RunFusion with used_list reset for every pixel (what upstream ACMMP does); plain python,
    float32 math close to the reference - only used to show that the quirk changes the count."""
    n = len(cams)
    masks = [np.zeros(d.shape, bool) for d in depths]
    count = 0
    for i in range(n):
        h, w = depths[i].shape
        for r in range(h):
            for c in range(w):
                if masks[i][r, c]:
                    continue
                d = depths[i][r, c]
                if d <= 0 or d >= cams[i].depth_max:
                    continue
                X = _world_point(cams[i], c, r, d)
                used, num, dyn = [], 0, 0.0
                for s in pairs[i]:
                    qx, qy, _ = _project(cams[s], X)
                    sr, scol = int(qy + np.float32(0.5)), int(qx + np.float32(0.5))
                    if not (0 <= scol < depths[s].shape[1] and 0 <= sr < depths[s].shape[0]) or masks[s][sr, scol]:
                        continue
                    sd = depths[s][sr, scol]
                    if sd <= 0:
                        continue
                    T = _world_point(cams[s], scol, sr, sd)
                    tx, ty, pd = _project(cams[i], T)
                    reproj = np.sqrt((c - tx) ** 2 + (r - ty) ** 2)
                    rel = abs(pd - d) / d
                    ang = np.arccos(np.clip(np.dot(normals[i][r, c], normals[s][sr, scol]), -1, 1))
                    if reproj < 2.0 and rel < 0.01 and ang < 0.174533:
                        used.append((s, sr, scol))
                        dyn += np.exp(-(reproj + 200 * rel + 10 * ang))
                        num += 1
                if num >= thresh and dyn > scalar * num:
                    count += 1
                    for s, sr, scol in used:
                        masks[s][sr, scol] = True
    return count
