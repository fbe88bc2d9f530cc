"""Synthetic COLMAP sparse models for the converter tests (text and binary writers of COLMAP's
published formats).  Test infrastructure."""
import os
import struct

import numpy as np

MODEL_IDS = {"SIMPLE_PINHOLE": 0, "PINHOLE": 1, "SIMPLE_RADIAL": 2, "RADIAL": 3, "OPENCV": 4}


def _quat_from_R(R):
    # Shepperd, scalar first
    t = np.trace(R)
    if t > 0:
        s = np.sqrt(t + 1.0) * 2
        q = [0.25 * s, (R[2, 1] - R[1, 2]) / s, (R[0, 2] - R[2, 0]) / s, (R[1, 0] - R[0, 1]) / s]
    else:
        i = int(np.argmax(np.diag(R)))
        j, k = (i + 1) % 3, (i + 2) % 3
        s = np.sqrt(1.0 + R[i, i] - R[j, j] - R[k, k]) * 2
        q = [0.0] * 4
        q[0] = (R[k, j] - R[j, k]) / s
        q[1 + i] = 0.25 * s
        q[1 + j] = (R[j, i] + R[i, j]) / s
        q[1 + k] = (R[k, i] + R[i, k]) / s
    q = np.array(q)
    return q / np.linalg.norm(q)


def make_model(seed=0, n_images=12, n_points=600, visibility=0.5, keypoints_without_point=40, duplicates=6,
               twin_images=True, scattered_ids=True):
    """Cameras on an arc looking at a point cloud around the origin.  Image ids are scattered and
    listed out of order, a few keypoints observe the same point twice, some have no point, and (with
    twin_images) the last image sits a hair beside the first so that their triangulation angles
    fall below one degree."""
    rng = np.random.default_rng(seed)
    pts = rng.normal(0.0, 1.0, (n_points, 3)) * np.array([2.0, 1.5, 1.0])
    point_ids = np.sort(rng.choice(np.arange(1, 5 * n_points), n_points, replace=False)) if scattered_ids else np.arange(1, n_points + 1)
    image_ids = rng.permutation(np.arange(1, 3 * n_images))[:n_images] if scattered_ids else np.arange(1, n_images + 1)
    cameras = {1: ("PINHOLE", 640, 480, [520.0, 515.5, 319.5, 239.5]),
               2: ("SIMPLE_RADIAL", 800, 600, [701.25, 400.0, 300.0, -0.0123])}
    images = []
    for v in range(n_images):
        ang = -0.6 + 1.2 * v / max(n_images - 1, 1)
        if twin_images and v == n_images - 1:
            ang = -0.6 + 2e-4
        C = np.array([9.0 * np.sin(ang), 0.4 * np.cos(3 * ang), -9.0 * np.cos(ang)])
        z = -C / np.linalg.norm(C)
        x = np.cross([0.0, 1.0, 0.0], z)
        x /= np.linalg.norm(x)
        y = np.cross(z, x)
        R = np.stack([x, y, z])
        t = -R @ C
        cam_id = 1 + (v % 2)
        K = cameras[cam_id][3]
        seen = np.flatnonzero(rng.random(n_points) < visibility)
        if v < duplicates:
            seen = np.concatenate([seen, seen[:3]])          # the same point under two keypoints
        Xc = (R @ pts[seen].T).T + t
        f = K[0]
        cx, cy = (K[2], K[3]) if cam_id == 1 else (K[1], K[2])
        xy = np.stack([f * Xc[:, 0] / Xc[:, 2] + cx, f * Xc[:, 1] / Xc[:, 2] + cy], 1)
        ids = point_ids[seen].astype(np.int64)
        extra = rng.uniform(0, 600, (keypoints_without_point, 2))
        xy = np.concatenate([xy, extra])
        ids = np.concatenate([ids, -np.ones(keypoints_without_point, np.int64)])
        perm = rng.permutation(len(ids))
        ext = ".jpg" if v % 3 else ".png"
        images.append(dict(id=int(image_ids[v]), qvec=_quat_from_R(R), tvec=t, camera_id=cam_id, name="view_%02d%s" % (v, ext),
                           xys=xy[perm], point3D_ids=ids[perm]))
    return dict(cameras=cameras, images=images, point_ids=point_ids.astype(np.int64), xyz=pts,
                rgb=rng.integers(0, 256, (n_points, 3)), error=rng.random(n_points))


def _tracks(model):
    tr = {int(p): [] for p in model["point_ids"]}
    for im in model["images"]:
        for k, p in enumerate(im["point3D_ids"]):
            if p != -1 and int(p) in tr:
                tr[int(p)].append((im["id"], k))
    return tr


def write_text(model, folder):
    os.makedirs(folder, exist_ok=True)
    with open(os.path.join(folder, "cameras.txt"), "w") as f:
        f.write("# Camera list with one line of data per camera:\n#   CAMERA_ID, MODEL, WIDTH, HEIGHT, PARAMS[]\n")
        for cid, (m, w, h, p) in model["cameras"].items():
            f.write("%d %s %d %d %s\n" % (cid, m, w, h, " ".join(repr(float(x)) for x in p)))
    with open(os.path.join(folder, "images.txt"), "w") as f:
        f.write("# Image list with two lines of data per image:\n")
        for im in model["images"]:
            f.write("%d %s %s %d %s\n" % (im["id"], " ".join(repr(float(x)) for x in im["qvec"]),
                                          " ".join(repr(float(x)) for x in im["tvec"]), im["camera_id"], im["name"]))
            f.write(" ".join("%r %r %d" % (float(x), float(y), p) for (x, y), p in zip(im["xys"], im["point3D_ids"])) + "\n")
    tr = _tracks(model)
    with open(os.path.join(folder, "points3D.txt"), "w") as f:
        f.write("# 3D point list with one line of data per point:\n")
        for pid, X, c, e in zip(model["point_ids"], model["xyz"], model["rgb"], model["error"]):
            f.write("%d %s %d %d %d %r %s\n" % (pid, " ".join(repr(float(x)) for x in X), c[0], c[1], c[2], float(e),
                                                " ".join("%d %d" % t for t in tr[int(pid)])))


def write_binary(model, folder):
    os.makedirs(folder, exist_ok=True)
    with open(os.path.join(folder, "cameras.bin"), "wb") as f:
        f.write(struct.pack("<Q", len(model["cameras"])))
        for cid, (m, w, h, p) in model["cameras"].items():
            f.write(struct.pack("<iiQQ", cid, MODEL_IDS[m], w, h))
            f.write(struct.pack("<%dd" % len(p), *p))
    with open(os.path.join(folder, "images.bin"), "wb") as f:
        f.write(struct.pack("<Q", len(model["images"])))
        for im in model["images"]:
            f.write(struct.pack("<idddddddi", im["id"], *[float(x) for x in im["qvec"]], *[float(x) for x in im["tvec"]], im["camera_id"]))
            f.write(im["name"].encode() + b"\x00")
            f.write(struct.pack("<Q", len(im["point3D_ids"])))
            for (x, y), p in zip(im["xys"], im["point3D_ids"]):
                f.write(struct.pack("<ddq", float(x), float(y), int(p)))
    tr = _tracks(model)
    with open(os.path.join(folder, "points3D.bin"), "wb") as f:
        f.write(struct.pack("<Q", len(model["point_ids"])))
        for pid, X, c, e in zip(model["point_ids"], model["xyz"], model["rgb"], model["error"]):
            f.write(struct.pack("<QdddBBBd", int(pid), float(X[0]), float(X[1]), float(X[2]), int(c[0]), int(c[1]), int(c[2]), float(e)))
            t = tr[int(pid)]
            f.write(struct.pack("<Q", len(t)))
            for a, b in t:
                f.write(struct.pack("<ii", a, b))


def write_images(model, folder, seed=0):
    """tiny image files under the names the model lists (needs cv2)."""
    import cv2
    os.makedirs(folder, exist_ok=True)
    rng = np.random.default_rng(seed)
    for im in model["images"]:
        cv2.imwrite(os.path.join(folder, im["name"]), rng.integers(0, 256, (12, 16, 3)).astype(np.uint8))
