"""CPU restatement of the converter's view-pair score.  TEST INFRASTRUCTURE: only tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this.

Follows calc_score, /root/reference/python_scripts/colmap2mvsnet_acm.py:280-302, per pair and in
the reference's arithmetic (numpy float64, the same dot / norm / norm / arccos expression), with
the membership test `it in id_j` done by np.isin instead of a linear scan.  Pinned against the
reference's own function and output files: tests/golden/convert/ (made by
tests/golden/make_convert_golden.py, which imports the reference module in the build container).
"""
import numpy as np


def calc_score(ids_i, ids_j, center_i, center_j, xyz_of):
    """ids_*: point3D id per keypoint (-1 = none); xyz_of: id -> (3,) float64.   [:280-302]"""
    shared = ids_i[np.isin(ids_i, ids_j)]           # keeps keypoint order and repeats of image i  (:284)
    score = 0
    angles = []
    for pid in shared:
        if pid == -1:                               # (:290-291) - also met when both images hold a -1
            continue
        p = xyz_of(int(pid))
        a, b = center_i - p, center_j - p
        theta = (180 / np.pi) * np.arccos(np.dot(a, b) / np.linalg.norm(a) / np.linalg.norm(b))   # (:293)
        angles.append(theta)
        score += 1
    if len(angles) > 0:                             # (:297-301)
        if sorted(angles)[int(len(angles) * 0.75)] < 1:
            score = 0.0
    return score


def score_matrix(obs_offsets, obs_point_ids, centers, point_ids, xyz, pairs=None):
    """Symmetric V x V matrix of calc_score over i < j (:399-413); `pairs` restricts the work to a
    sample of (i, j) tuples (the rest stays 0)."""
    V = len(obs_offsets) - 1
    lookup = {int(p): k for k, p in enumerate(point_ids)}
    xyz = np.asarray(xyz, np.float64)
    xyz_of = lambda pid: xyz[lookup[pid]]           # noqa: E731
    ids = [np.asarray(obs_point_ids[obs_offsets[v]:obs_offsets[v + 1]], np.int64) for v in range(V)]
    score = np.zeros((V, V))
    if pairs is None:
        pairs = [(i, j) for i in range(V) for j in range(i + 1, V)]
    for i, j in pairs:
        s = calc_score(ids[i], ids[j], centers[i], centers[j], xyz_of)
        score[i, j] = s
        score[j, i] = s
    return score
