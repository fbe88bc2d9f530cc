"""Shared helpers of the converter tests."""
import argparse
import filecmp
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden", "convert")
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import pair_scores as oracle_pairs  # noqa: E402

CASES = {"a": 192, "b": 0}     # tag -> --max_d used when the golden files were made


def args_for(tag, save_folder, dense=None, ext=".txt"):
    return argparse.Namespace(dense_folder=dense or os.path.join(GOLDEN, tag, "dense"), save_folder=save_folder,
                              max_d=CASES[tag], interval_scale=1, theta0=5, sigma1=1, sigma2=10, model_ext=ext)


def oracle_scores(model, E, pairs=None):
    from acmmp_b200 import convert
    return oracle_pairs.score_matrix(model.obs_offsets, model.obs_point_ids, convert.camera_centers(E), model.point_ids,
                                     model.xyz, pairs=pairs)


def assert_same_outputs(tag, save_folder):
    exp = os.path.join(GOLDEN, tag, "expected")
    assert filecmp.cmp(os.path.join(exp, "pair.txt"), os.path.join(save_folder, "pair.txt"), shallow=False)
    names = sorted(os.listdir(os.path.join(exp, "cams")))
    assert names == sorted(os.listdir(os.path.join(save_folder, "cams")))
    for n in names:
        a = open(os.path.join(exp, "cams", n)).read()
        b = open(os.path.join(save_folder, "cams", n)).read()
        assert a == b, (n, a, b)


def expected_image_sizes(tag):
    out = {}
    for line in open(os.path.join(GOLDEN, tag, "expected_image_sizes.txt")):
        n, s = line.split()
        out[n] = int(s)
    return out


def reference_scores(tag):
    return np.load(os.path.join(GOLDEN, tag, "reference_scores.npy"))
