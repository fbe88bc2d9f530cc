"""16-bit prior images for the seeded mode (-p): what an external prior generator would hand to the
reference under <dense>/priors/{depths,normals}/%08d.png."""
import numpy as np


def encode_priors(cam, depth, normal_cam, noise=0.0, seed=0):
    """depth (H, W) and camera-frame normals (H, W, 3) -> (depth16, normal16) in the reference's coding:
    depth = v * (dmax - dmin) / 65535 + dmin, normal = v * 2 / 65536 - 1 (acmmp_definitions.cpp:118-131)"""
    rng = np.random.default_rng(seed)
    d = np.asarray(depth, np.float64) * (1.0 + noise * rng.standard_normal(depth.shape))
    v = np.rint((d - cam.depth_min) / (cam.depth_max - cam.depth_min) * 65535.0)
    depth16 = np.clip(v, 0, 65535).astype(np.uint16)
    n = np.asarray(normal_cam, np.float64) + noise * 5 * rng.standard_normal(normal_cam.shape)
    n /= np.linalg.norm(n, axis=-1, keepdims=True)
    normal16 = np.clip(np.rint((n + 1.0) * 32768.0), 0, 65535).astype(np.uint16)
    return depth16, normal16
