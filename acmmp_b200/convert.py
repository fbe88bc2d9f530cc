"""Input converter: COLMAP sparse model -> `cams/%08d_cam.txt`, `images/%08d.jpg`, `pair.txt`.

Host-side mirror of the reference's converter
(/root/reference/python_scripts/colmap2mvsnet_acm.py): same command line,
same output files.  The reference's slow step - scoring all V(V-1)/2 image pairs with a Python
`in` over numpy arrays, O(n_i n_j) per pair and run twice (:408-409) - is one call into the CUDA
library here (`acmmp_pair_scores`, csrc/pairs.cu).  There is no CPU fallback: without the library
and a device `pair_scores` raises.

    python -m acmmp_b200.convert --dense_folder <colmap dense dir> --save_folder <out>
"""
import argparse
import ctypes as C
import os
import shutil
import struct
from collections import namedtuple

import numpy as np

from . import lib as L

# name, number of parameters, index of fx/fy/cx/cy in the parameter list   [reference :32-44, :322-334]
CAMERA_MODELS = {
    0: ("SIMPLE_PINHOLE", 3, (0, 0, 1, 2)),
    1: ("PINHOLE", 4, (0, 1, 2, 3)),
    2: ("SIMPLE_RADIAL", 4, (0, 0, 1, 2)),
    3: ("RADIAL", 5, (0, 0, 1, 2)),
    4: ("OPENCV", 8, (0, 1, 2, 3)),
    5: ("OPENCV_FISHEYE", 8, (0, 1, 2, 3)),
    6: ("FULL_OPENCV", 12, (0, 1, 2, 3)),
    7: ("FOV", 5, (0, 1, 2, 3)),
    8: ("SIMPLE_RADIAL_FISHEYE", 4, (0, 0, 1, 2)),
    9: ("RADIAL_FISHEYE", 5, (0, 0, 1, 2)),
    10: ("THIN_PRISM_FISHEYE", 12, (0, 1, 2, 3)),
}
_MODEL_BY_NAME = {v[0]: v for v in CAMERA_MODELS.values()}

CameraRec = namedtuple("CameraRec", ["id", "model", "width", "height", "params"])


class SparseModel:
    """A COLMAP sparse model in flat arrays.  Images are ordered by image id - the reference
    renumbers them the same way (:343-346) - and keypoints keep their order."""

    def __init__(self, cameras, image_ids, qvecs, tvecs, camera_ids, names, obs_offsets, obs_point_ids, point_ids, xyz):
        order = np.argsort(np.asarray(image_ids, np.int64), kind="stable")
        self.cameras = cameras
        self.image_ids = np.asarray(image_ids, np.int64)[order]
        self.qvecs = np.asarray(qvecs, np.float64).reshape(-1, 4)[order]
        self.tvecs = np.asarray(tvecs, np.float64).reshape(-1, 3)[order]
        self.camera_ids = np.asarray(camera_ids, np.int64)[order]
        self.names = [names[k] for k in order]
        off = np.asarray(obs_offsets, np.int64)
        ids = np.asarray(obs_point_ids, np.int64)
        chunks = [ids[off[k]:off[k + 1]] for k in order]
        self.obs_offsets = np.concatenate([[0], np.cumsum([len(c) for c in chunks])]).astype(np.int64)
        self.obs_point_ids = np.concatenate(chunks) if chunks else np.zeros(0, np.int64)
        porder = np.argsort(np.asarray(point_ids, np.int64), kind="stable")
        self.point_ids = np.asarray(point_ids, np.int64)[porder]
        self.xyz = np.asarray(xyz, np.float64).reshape(-1, 3)[porder]

    @property
    def num_images(self):
        return len(self.image_ids)

    def obs_point_index(self):
        """Keypoint -> row of `xyz`, -1 where the keypoint has no 3-D point.  A keypoint that names a
        point missing from points3D is an error (the reference dies with a KeyError, :293/:366)."""
        ids = self.obs_point_ids
        idx = np.full(len(ids), -1, np.int32)
        has = ids != -1
        if has.any():
            pos = np.searchsorted(self.point_ids, ids[has])
            pos = np.minimum(pos, max(len(self.point_ids) - 1, 0))
            if len(self.point_ids) == 0 or (self.point_ids[pos] != ids[has]).any():
                raise KeyError("images refer to 3-D points that points3D does not hold")
            idx[has] = pos
        return idx


# ---- readers: COLMAP's published text and binary formats --------------------------------------

def _data_lines(path):
    with open(path, "r") as f:
        for line in f:
            line = line.strip()
            if line and not line.startswith("#"):
                yield line


def read_cameras_text(path):
    cams = {}
    for line in _data_lines(path):
        e = line.split()
        cams[int(e[0])] = CameraRec(int(e[0]), e[1], int(e[2]), int(e[3]), np.array([float(x) for x in e[4:]]))
    return cams


def read_cameras_binary(path):
    cams = {}
    with open(path, "rb") as f:
        (n,) = struct.unpack("<Q", f.read(8))
        for _ in range(n):
            cid, model_id, w, h = struct.unpack("<iiQQ", f.read(24))
            name, npar, _ = CAMERA_MODELS[model_id]
            cams[cid] = CameraRec(cid, name, w, h, np.array(struct.unpack("<%dd" % npar, f.read(8 * npar))))
    return cams


def _read_images_text(path):
    ids, q, t, cam, names, off, obs = [], [], [], [], [], [0], []
    with open(path, "r") as f:
        while True:
            line = f.readline()
            if not line:
                break
            line = line.strip()
            if not line or line.startswith("#"):
                continue
            e = line.split()
            ids.append(int(e[0]))
            q.append([float(x) for x in e[1:5]])
            t.append([float(x) for x in e[5:8]])
            cam.append(int(e[8]))
            names.append(e[9])
            pts = f.readline().split()          # the keypoint line follows, possibly empty
            p3 = np.array(pts[2::3], dtype=np.int64) if pts else np.zeros(0, np.int64)
            obs.append(p3)
            off.append(off[-1] + len(p3))
    return ids, q, t, cam, names, off, (np.concatenate(obs) if obs else np.zeros(0, np.int64))


def _read_images_binary(path):
    ids, q, t, cam, names, off, obs = [], [], [], [], [], [0], []
    with open(path, "rb") as f:
        buf = f.read()
    (n,) = struct.unpack_from("<Q", buf, 0)
    pos = 8
    rec = np.dtype([("x", "<f8"), ("y", "<f8"), ("id", "<i8")])
    for _ in range(n):
        v = struct.unpack_from("<idddddddi", buf, pos)
        pos += 64
        end = buf.index(b"\x00", pos)
        names.append(buf[pos:end].decode("utf-8"))
        pos = end + 1
        (m,) = struct.unpack_from("<Q", buf, pos)
        pos += 8
        p = np.frombuffer(buf, rec, m, pos)
        pos += 24 * m
        ids.append(v[0])
        q.append(v[1:5])
        t.append(v[5:8])
        cam.append(v[8])
        obs.append(p["id"].astype(np.int64))
        off.append(off[-1] + m)
    return ids, q, t, cam, names, off, (np.concatenate(obs) if obs else np.zeros(0, np.int64))


def _read_points_text(path):
    ids, xyz = [], []
    for line in _data_lines(path):
        e = line.split(None, 4)
        ids.append(int(e[0]))
        xyz.append((float(e[1]), float(e[2]), float(e[3])))
    return ids, xyz


def _read_points_binary(path):
    ids, xyz = [], []
    with open(path, "rb") as f:
        buf = f.read()
    (n,) = struct.unpack_from("<Q", buf, 0)
    pos = 8
    for _ in range(n):
        pid, x, y, z = struct.unpack_from("<Qddd", buf, pos)
        (track,) = struct.unpack_from("<Q", buf, pos + 43)
        pos += 51 + 8 * track
        ids.append(pid)
        xyz.append((x, y, z))
    return ids, xyz


def read_model(path, ext=".txt"):
    """cameras / images / points3D of `path` with extension `ext` (".txt" or ".bin")  [reference :240-250]"""
    if ext == ".txt":
        cams = read_cameras_text(os.path.join(path, "cameras.txt"))
        img = _read_images_text(os.path.join(path, "images.txt"))
        pts = _read_points_text(os.path.join(path, "points3D.txt"))
    elif ext == ".bin":
        cams = read_cameras_binary(os.path.join(path, "cameras.bin"))
        img = _read_images_binary(os.path.join(path, "images.bin"))
        pts = _read_points_binary(os.path.join(path, "points3D.bin"))
    else:
        raise ValueError("model_ext must be .txt or .bin")
    return SparseModel(cams, *img, *pts)


# ---- cameras ------------------------------------------------------------------------------------

def intrinsic_matrix(cam):
    """3x3 K; distortion parameters are dropped as in the reference (:336-341)."""
    ifx, ify, icx, icy = _MODEL_BY_NAME[cam.model][2]
    p = cam.params
    return np.array([[p[ifx], 0, p[icx]], [0, p[ify], p[icy]], [0, 0, 1]], dtype=np.float64)


def rotation_from_quaternion(q):
    """COLMAP (w, x, y, z) -> R; same roundings as the reference's qvec2rotmat (:252-262) - scaling by
    two is exact, so only the products and the left-to-right sums round."""
    w, x, y, z = (np.float64(v) for v in q)
    xx, yy, zz = x * x, y * y, z * z
    xy, zx, yz = x * y, z * x, y * z
    wx, wy, wz = w * x, w * y, w * z
    return np.array([[1 - 2 * yy - 2 * zz, 2 * xy - 2 * wz, 2 * zx + 2 * wy],
                     [2 * xy + 2 * wz, 1 - 2 * xx - 2 * zz, 2 * yz - 2 * wx],
                     [2 * zx - 2 * wy, 2 * yz + 2 * wx, 1 - 2 * xx - 2 * yy]], dtype=np.float64)


def extrinsic_matrices(model):
    E = np.zeros((model.num_images, 4, 4))
    for v in range(model.num_images):
        E[v, :3, :3] = rotation_from_quaternion(model.qvecs[v])
        E[v, :3, 3] = model.tvecs[v]
        E[v, 3, 3] = 1
    return E


def camera_centers(E):
    """-R^T t per image (calc_score :285-286)."""
    return np.stack([-np.matmul(e[:3, :3].transpose(), e[:3, 3:4])[:, 0] for e in E]) if len(E) else np.zeros((0, 3))


def depth_ranges(model, E, K_of_image, max_d=192, interval_scale=1.0):
    """(depth_min, depth_interval, depth_num, depth_max) per image               [reference :358-389]
    from the 1 % / 99 % ranks of the camera-frame depths of the image's 3-D points (each keypoint
    counts), relaxed by 0.75 / 1.25; max_d == 0 selects the inverse-depth sampling count."""
    idx = model.obs_point_index()
    out = []
    for v in range(model.num_images):
        sl = idx[model.obs_offsets[v]:model.obs_offsets[v + 1]]
        sl = sl[sl >= 0]
        if len(sl) == 0:
            raise IndexError("image %d observes no 3-D point" % v)          # the reference fails at zs_sorted[0]
        P = np.concatenate([model.xyz[sl], np.ones((len(sl), 1))], axis=1)
        zs = np.sort(np.matmul(P, E[v][2]))
        dmin = zs[int(len(zs) * .01)] * 0.75
        dmax = zs[int(len(zs) * .99)] * 1.25
        if max_d == 0:
            K, R, t = K_of_image(v), E[v][:3, :3], E[v][:3, 3]
            Kinv, Rinv = np.linalg.inv(K), np.linalg.inv(R)
            a = np.matmul(Rinv, np.matmul(Kinv, [K[0, 2], K[1, 2], 1]) * dmin - t)
            b = np.matmul(Rinv, np.matmul(Kinv, [K[0, 2] + 1, K[1, 2], 1]) * dmin - t)
            num = (1 / dmin - 1 / dmax) / (1 / dmin - 1 / (dmin + np.linalg.norm(b - a)))
        else:
            num = max_d
        out.append((dmin, (dmax - dmin) / (num - 1) / interval_scale, num, dmax))
    return out


# ---- view selection -----------------------------------------------------------------------------

def pair_scores(model, E, device=0, min_angle_deg=1.0, quantile=0.75, return_ms=False):
    """V x V score matrix on the device (csrc/pairs.cu) - calc_score over all pairs (:280-302, :399-413)."""
    lib = L.load()
    V = model.num_images
    idx = np.ascontiguousarray(model.obs_point_index(), np.int32)
    off = np.ascontiguousarray(model.obs_offsets, np.int64)
    xyz = np.ascontiguousarray(model.xyz, np.float64)
    cen = np.ascontiguousarray(camera_centers(E), np.float64)
    score = np.zeros((V, V), np.float64)
    ms = C.c_double(0.0)
    dp = C.POINTER(C.c_double)
    rc = lib.acmmp_pair_scores(device, V, off.ctypes.data_as(C.POINTER(C.c_int64)), idx.ctypes.data_as(C.POINTER(C.c_int32)),
                               len(xyz), xyz.ctypes.data_as(dp), cen.ctypes.data_as(dp), min_angle_deg, quantile,
                               score.ctypes.data_as(dp), C.byref(ms))
    if rc != 0:
        raise RuntimeError("acmmp_pair_scores failed (%d): %s" % (rc, lib.acmmp_pairs_last_error().decode()))
    return (score, ms.value) if return_ms else score


def select_views(score, max_views=20):
    """Per image the `max_views` best partners, best first - numpy's own argsort, reversed, as the
    reference does (:417-421), so ties fall the same way."""
    n = min(max_views, len(score) - 1)
    sel = []
    for i in range(len(score)):
        order = np.argsort(score[i])[::-1]
        sel.append([(int(k), score[i, k]) for k in order[:n]])
    return sel


# ---- writers ------------------------------------------------------------------------------------

def write_cam_file(path, E, K, rng):
    """cam text in the reference's formatting (:430-443): str() of every float64, %f depth line."""
    with open(path, "w") as f:
        f.write("extrinsic\n")
        for j in range(4):
            f.write("".join(str(E[j, k]) + " " for k in range(4)) + "\n")
        f.write("\nintrinsic\n")
        for j in range(3):
            f.write("".join(str(K[j, k]) + " " for k in range(3)) + "\n")
        f.write("\n%f %f %f %f\n" % (rng[0], rng[1], rng[2], rng[3]))


def write_pair_file(path, view_sel):
    """pair.txt (:444-450): count, then per image its index and `n id score id score ...`."""
    with open(path, "w") as f:
        f.write("%d\n" % len(view_sel))
        for i, row in enumerate(view_sel):
            f.write("%d\n%d " % (i, len(row)))
            for k, s in row:
                f.write("%d %d " % (k, s))
            f.write("\n")


def convert_images(model, image_dir, out_dir):
    """images/%08d.jpg: JPEGs are copied, anything else is re-encoded (:452-459)."""
    for i, name in enumerate(model.names):
        src = os.path.join(image_dir, name)
        dst = os.path.join(out_dir, "%08d.jpg" % i)
        if src.endswith(".jpg"):
            shutil.copyfile(src, dst)
        else:
            import cv2  # only needed for non-JPEG inputs
            cv2.imwrite(dst, cv2.imread(src))


def processing_single_scene(args, device=0, copy_images=True):
    """The converter's whole job, same files as the reference's function of this name (:304-459)."""
    image_dir = os.path.join(args.dense_folder, "images")
    model_dir = os.path.join(args.dense_folder, "sparse")
    cam_dir = os.path.join(args.save_folder, "cams")
    out_images = os.path.join(args.save_folder, "images")
    for d in (out_images, cam_dir):
        if os.path.exists(d):
            shutil.rmtree(d)
    os.makedirs(out_images)
    os.makedirs(cam_dir)

    model = read_model(model_dir, args.model_ext)
    E = extrinsic_matrices(model)
    Ks = {cid: intrinsic_matrix(cam) for cid, cam in model.cameras.items()}
    ranges = depth_ranges(model, E, lambda v: Ks[int(model.camera_ids[v])], args.max_d, args.interval_scale)
    score = pair_scores(model, E, device=device)
    view_sel = select_views(score)
    for v in range(model.num_images):
        write_cam_file(os.path.join(cam_dir, "%08d_cam.txt" % v), E[v], Ks[int(model.camera_ids[v])], ranges[v])
    write_pair_file(os.path.join(args.save_folder, "pair.txt"), view_sel)
    if copy_images:
        convert_images(model, image_dir, out_images)
    return model, score, view_sel


def build_parser():
    p = argparse.ArgumentParser(description="Convert colmap camera")
    p.add_argument("--dense_folder", required=True, type=str, help="dense_folder.")
    p.add_argument("--save_folder", required=True, type=str, help="save_folder.")
    p.add_argument("--max_d", type=int, default=192)
    p.add_argument("--interval_scale", type=float, default=1)
    p.add_argument("--theta0", type=float, default=5)        # accepted and unused, as in the reference
    p.add_argument("--sigma1", type=float, default=1)
    p.add_argument("--sigma2", type=float, default=10)
    p.add_argument("--model_ext", type=str, default=".txt", choices=[".txt", ".bin"], help="sparse model ext")
    p.add_argument("--device", type=int, default=0)
    return p


def main(argv=None):
    args = build_parser().parse_args(argv)
    os.makedirs(args.save_folder, exist_ok=True)
    processing_single_scene(args, device=args.device)


if __name__ == "__main__":
    main()
