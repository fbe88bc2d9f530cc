"""Makes tests/golden/convert/: a small synthetic COLMAP model (text), the files the REFERENCE's
converter writes for it, and the reference's own pair scores.  Runs only where /root/reference
exists (the build container); the committed outputs are what the tests read.

    python tests/golden/make_convert_golden.py
"""
import argparse
import importlib.util
import os
import shutil
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import colmap_model as cm  # noqa: E402

REF = "/root/reference/python_scripts/colmap2mvsnet_acm.py"


def reference_module():
    spec = importlib.util.spec_from_file_location("ref_colmap2mvsnet_acm", REF)
    mod = importlib.util.module_from_spec(spec)
    sys.modules[spec.name] = mod          # its multiprocessing pool pickles calc_score by module name
    spec.loader.exec_module(mod)
    return mod


def main():
    out = os.path.join(HERE, "convert")
    shutil.rmtree(out, ignore_errors=True)
    ref = reference_module()
    for tag, kw, max_d in (("a", dict(seed=3, n_images=12, n_points=500), 192),
                           ("b", dict(seed=8, n_images=7, n_points=160, visibility=0.7, scattered_ids=False), 0)):
        dense = os.path.join(out, tag, "dense")
        save = os.path.join(out, tag, "expected")
        model = cm.make_model(**kw)
        cm.write_text(model, os.path.join(dense, "sparse"))
        cm.write_images(model, os.path.join(dense, "images"), seed=kw["seed"])
        os.makedirs(save)
        args = argparse.Namespace(dense_folder=dense, save_folder=save, max_d=max_d, interval_scale=1, theta0=5, sigma1=1,
                                  sigma2=10, model_ext=".txt")
        ref.processing_single_scene(args)
        # the reference's own score per pair, straight from calc_score
        cameras, images, points3d = ref.read_model(os.path.join(dense, "sparse"), ".txt")
        images = {i + 1: images[k] for i, k in enumerate(sorted(images.keys()))}
        extrinsic = {}
        for k, im in images.items():
            e = np.zeros((4, 4))
            e[:3, :3] = ref.qvec2rotmat(im.qvec)
            e[:3, 3] = im.tvec
            e[3, 3] = 1
            extrinsic[k] = e
        V = len(images)
        score = np.zeros((V, V))
        for i in range(V):
            for j in range(i + 1, V):
                _, _, s = ref.calc_score((i, j), images, points3d, extrinsic, args)
                score[i, j] = score[j, i] = s
        np.save(os.path.join(out, tag, "reference_scores.npy"), score)
        # converted images are byte copies / re-encodes: keep their sizes only
        sizes = {n: os.path.getsize(os.path.join(save, "images", n)) for n in sorted(os.listdir(os.path.join(save, "images")))}
        with open(os.path.join(out, tag, "expected_image_sizes.txt"), "w") as f:
            for n, s in sizes.items():
                f.write("%s %d\n" % (n, s))
        shutil.rmtree(os.path.join(save, "images"))


if __name__ == "__main__":
    main()
