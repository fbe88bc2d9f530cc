"""Input converter on the device (acmmp_pair_scores, csrc/pairs.cu) against the REFERENCE converter's
own scores and output files (tests/golden/convert) and, at sizes the reference would need hours
for, against the oracle restatement on a sample of pairs.  pytest -m gpu."""
import os
import subprocess
import sys

import numpy as np
import pytest

import colmap_model as cm
import convert_util as cu

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
if not torch.cuda.is_available():
    pytest.skip("no CUDA device", allow_module_level=True)

from acmmp_b200 import convert  # noqa: E402


@pytest.mark.parametrize("tag", sorted(cu.CASES))
def test_device_scores_equal_reference(tag):
    model = convert.read_model(os.path.join(cu.GOLDEN, tag, "dense", "sparse"), ".txt")
    E = convert.extrinsic_matrices(model)
    got = convert.pair_scores(model, E)
    assert (got == cu.reference_scores(tag)).all()


@pytest.mark.parametrize("tag,ext", [("a", ".txt"), ("b", ".txt"), ("a", ".bin")])
def test_command_line_writes_the_reference_files(tag, ext, tmp_path):
    dense = os.path.join(cu.GOLDEN, tag, "dense")
    if ext == ".bin":                    # same model, binary container
        import shutil
        dense = str(tmp_path / "dense")
        shutil.copytree(os.path.join(cu.GOLDEN, tag, "dense", "images"), os.path.join(dense, "images"))
        seeds = {"a": dict(seed=3, n_images=12, n_points=500)}
        cm.write_binary(cm.make_model(**seeds[tag]), os.path.join(dense, "sparse"))
    save = str(tmp_path / "out")
    r = subprocess.run([sys.executable, "-m", "acmmp_b200.convert", "--dense_folder", dense, "--save_folder", save,
                        "--max_d", str(cu.CASES[tag]), "--model_ext", ext], cwd=cu.ROOT, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    cu.assert_same_outputs(tag, save)
    sizes = cu.expected_image_sizes(tag)
    for n, s in sizes.items():
        assert os.path.getsize(os.path.join(save, "images", n)) == s


def test_large_model_against_oracle_sample(tmp_path):
    """60 images x ~3 k keypoints: every pair's count checked in closed form, 40 pairs by the oracle."""
    m = cm.make_model(seed=21, n_images=60, n_points=6000, visibility=0.45, keypoints_without_point=300, duplicates=20)
    cm.write_binary(m, str(tmp_path / "m"))
    model = convert.read_model(str(tmp_path / "m"), ".bin")
    E = convert.extrinsic_matrices(model)
    got, ms = convert.pair_scores(model, E, return_ms=True)
    assert ms > 0 and (got == got.T).all() and (np.diag(got) == 0).all()
    rng = np.random.default_rng(0)
    V = model.num_images
    pairs = [(0, 1), (0, V - 1), (V - 2, V - 1)]
    # include the twin pair (scattered ids move it): the only zero among otherwise populated pairs
    zi, zj = np.argwhere(np.triu(got == 0, 1))[0]
    pairs.append((int(zi), int(zj)))
    while len(pairs) < 40:
        i, j = sorted(rng.choice(V, 2, replace=False))
        pairs.append((int(i), int(j)))
    ref = cu.oracle_scores(model, E, pairs=pairs)
    for i, j in pairs:
        assert got[i, j] == ref[i, j], (i, j, got[i, j], ref[i, j])
    # counts of all pairs: keypoints of image i whose point image j also sees
    ids = [model.obs_point_ids[model.obs_offsets[v]:model.obs_offsets[v + 1]] for v in range(V)]
    ids = [a[a != -1] for a in ids]
    for i in range(0, V, 7):
        for j in range(i + 1, V):
            n = int(np.isin(ids[i], ids[j]).sum())
            assert got[i, j] in (0.0, float(n))


def test_edge_cases():
    one = convert.SparseModel({}, [7], [[1, 0, 0, 0]], [[0, 0, 0]], [1], ["a.jpg"], [0, 3], [-1, -1, -1], [], np.zeros((0, 3)))
    assert (convert.pair_scores(one, convert.extrinsic_matrices(one)) == 0).all()
    # three images, the middle one sees nothing; images 0 and 2 share point 5 (twice in image 0) and point 9
    m = convert.SparseModel({}, [1, 2, 3], [[1, 0, 0, 0]] * 3, [[0, 0, 0], [-1, 0, 0], [-2, 0, 0]], [1, 1, 1], ["a", "b", "c"],
                            [0, 4, 4, 7], [5, 9, 5, -1, 9, -1, 5], [5, 9], [[0.5, 0.2, 4.0], [1.0, -0.3, 5.0]])
    s = convert.pair_scores(m, convert.extrinsic_matrices(m))
    assert s[0, 2] == 3 and s[2, 0] == 3 and (s[1] == 0).all() and (s[:, 1] == 0).all()
    assert (s == cu.oracle_scores(m, convert.extrinsic_matrices(m))).all()
    # a higher angle threshold zeroes the pair: all three angles are well below 45 degrees
    assert convert.pair_scores(m, convert.extrinsic_matrices(m), min_angle_deg=45.0)[0, 2] == 0
