"""Synthetic MVS scenes with DTU / ETH3D / Tanks&Temples *shape* (SURVEY 8d): a
procedurally textured height field seen by pinhole cameras on a spherical cap,
rendered by ray/height-field intersection.  Nothing is downloaded; the same seed
gives the same scene.  torch is used only as an array library so the renderer can
run on the GPU box's device for the 49 x 1600 x 1200 bench scan and on the CPU for
the small test scenes."""
import math

import numpy as np
import torch

from .lib import Camera


def _look_at(center):
    """world->camera rotation with the optical axis through the origin."""
    c = np.asarray(center, np.float64)
    f = -c / np.linalg.norm(c)
    up = np.array([0.0, 1.0, 0.0])
    r = np.cross(up, f)
    r /= np.linalg.norm(r)
    d = np.cross(f, r)
    R = np.stack([r, d, f])
    t = -R @ c
    return R, t


def make_rig(n_views, width, height, focal=None, distance=600.0, span_deg=(24.0, 18.0), layout="grid"):
    """Cameras on a cap around the +z axis looking at the origin.
    layout 'grid': ceil(sqrt(n)) x ... grid (49 -> 7 x 7, DTU-like);
    layout 'ring': a ring / helix (Tanks&Temples-like)."""
    if focal is None:
        focal = 2892.0 * width / 1600.0   # DTU focal length scaled to the width
    cams = []
    if layout == "grid":
        nx = int(math.ceil(math.sqrt(n_views)))
        ny = int(math.ceil(n_views / nx))
        k = 0
        for j in range(ny):
            for i in range(nx):
                if k >= n_views:
                    break
                a = math.radians(span_deg[0] * (2.0 * i / max(nx - 1, 1) - 1.0)) if nx > 1 else 0.0
                b = math.radians(span_deg[1] * (2.0 * j / max(ny - 1, 1) - 1.0)) if ny > 1 else 0.0
                c = distance * np.array([math.sin(a) * math.cos(b), math.sin(b), math.cos(a) * math.cos(b)])
                cams.append(c)
                k += 1
    else:
        for k in range(n_views):
            a = 2.0 * math.pi * k / n_views
            tilt = math.radians(span_deg[1]) * (0.6 + 0.4 * math.sin(3 * a))
            c = distance * np.array([math.sin(tilt) * math.cos(a), math.sin(tilt) * math.sin(a), math.cos(tilt)])
            cams.append(c)
    out = []
    K = np.array([[focal, 0, width / 2.0], [0, focal, height / 2.0], [0, 0, 1.0]])
    for c in cams:
        R, t = _look_at(c)
        out.append((K.copy(), R, t))
    return out


class HeightField:
    """z = h(x, y): smooth bumps + a raised slab (depth discontinuity), textured by
    multi-octave value noise quantised to 8 bits like a decoded JPEG."""

    def __init__(self, seed, device="cpu", amplitude=25.0, slab=True, texel_mm=0.2):
        g = torch.Generator().manual_seed(int(seed))
        self.device = torch.device(device)
        self.amp = amplitude
        self.slab = slab
        self.phase = (torch.rand(6, generator=g) * 6.2831853).tolist()
        self.grids = [torch.rand(256, 256, generator=g).to(self.device) for _ in range(4)]
        # noise cell sizes in mm, tied to the pixel footprint so that every
        # resolution sees the same texture frequencies in pixels (4.5 .. 200 px)
        self.cells = [4.5 * texel_mm, 13.5 * texel_mm, 45.0 * texel_mm, 200.0 * texel_mm]
        self.gain = [0.45, 0.30, 0.17, 0.08]

    def height(self, x, y):
        p = self.phase
        h = self.amp * (torch.sin(x / 70.0 + p[0]) * torch.cos(y / 90.0 + p[1]) + 0.5 * torch.sin(x / 31.0 + y / 47.0 + p[2]))
        if self.slab:
            # smooth-edged slab, 30 mm high
            sx = torch.sigmoid((x + 40.0) * 2.0) * torch.sigmoid((110.0 - x) * 2.0)
            sy = torch.sigmoid((y + 60.0) * 2.0) * torch.sigmoid((50.0 - y) * 2.0)
            h = h + 30.0 * sx * sy
        return h

    def _noise(self, grid, u, v):
        n = grid.shape[0]
        u0 = torch.floor(u)
        v0 = torch.floor(v)
        fu = u - u0
        fv = v - v0
        i0 = torch.remainder(u0.long(), n)
        j0 = torch.remainder(v0.long(), n)
        i1 = torch.remainder(i0 + 1, n)
        j1 = torch.remainder(j0 + 1, n)
        fu = fu * fu * (3 - 2 * fu)
        fv = fv * fv * (3 - 2 * fv)
        a = grid[j0, i0] * (1 - fu) + grid[j0, i1] * fu
        b = grid[j1, i0] * (1 - fu) + grid[j1, i1] * fu
        return a * (1 - fv) + b * fv

    def texture(self, x, y):
        t = torch.zeros_like(x)
        for grid, cell, gain in zip(self.grids, self.cells, self.gain):
            t = t + gain * self._noise(grid, x / cell + 1000.0, y / cell + 1000.0)
        t = (t - 0.5) * 1.9 + 0.5
        return torch.clamp(torch.round(t * 255.0), 0.0, 255.0)


def render(field, K, R, t, width, height, iters=24):
    """image (float32, integer valued 0..255) and ground-truth depth of one view."""
    dev = field.device
    Kt = torch.tensor(K, dtype=torch.float64, device=dev)
    Rt = torch.tensor(R, dtype=torch.float64, device=dev)
    tt = torch.tensor(t, dtype=torch.float64, device=dev)
    C = -(Rt.T @ tt)
    ys, xs = torch.meshgrid(torch.arange(height, device=dev, dtype=torch.float64),
                            torch.arange(width, device=dev, dtype=torch.float64), indexing="ij")
    dc = torch.stack([(xs - Kt[0, 2]) / Kt[0, 0], (ys - Kt[1, 2]) / Kt[1, 1], torch.ones_like(xs)], -1)
    dw = dc @ Rt          # R^T applied to every ray (row-vector form)
    s = torch.full_like(xs, float(torch.linalg.norm(C)))
    for _ in range(iters):
        px = C[0] + s * dw[..., 0]
        py = C[1] + s * dw[..., 1]
        s = (field.height(px, py) - C[2]) / dw[..., 2]
    px = C[0] + s * dw[..., 0]
    py = C[1] + s * dw[..., 1]
    img = field.texture(px, py)
    return img.to(torch.float32), s.to(torch.float32)


def select_pairs(rig, n_src):
    """pair.txt-style neighbour lists: n_src nearest views by optical-axis angle
    (the rule of python_scripts/select_dtu_cams.py:38-56 with score 1)."""
    axes = np.stack([R[2] for (_, R, _) in rig])
    ang = np.arccos(np.clip(axes @ axes.T, -1, 1))
    pairs = []
    for i in range(len(rig)):
        order = [int(j) for j in np.argsort(ang[i], kind="stable") if j != i]
        pairs.append(order[:n_src])
    return pairs


def make_scene(n_views=5, width=640, height=480, n_src=4, seed=0, device="cpu", layout="grid",
               span_deg=None, depth_range=(450.0, 800.0), slab=True, as_numpy=True):
    """Returns dict(images, depths, cams (lib.Camera), pairs, rig)."""
    if span_deg is None:
        span_deg = (24.0, 18.0) if n_views > 9 else (8.0, 6.0)
    rig = make_rig(n_views, width, height, span_deg=span_deg, layout=layout)
    field = HeightField(seed, device=device, slab=slab, texel_mm=600.0 / rig[0][0][0, 0])
    images, depths, cams = [], [], []
    for (K, R, t) in rig:
        img, dep = render(field, K, R, t, width, height)
        if as_numpy:
            img = img.cpu().numpy()
            dep = dep.cpu().numpy()
        images.append(img)
        depths.append(dep)
        cams.append(Camera.make(K, R, t, width, height, depth_range[0], depth_range[1]))
    return {"images": images, "depths": depths, "cams": cams, "pairs": select_pairs(rig, n_src), "rig": rig,
            "width": width, "height": height}
