"""Input converter, CPU side: the oracle restatement of calc_score against
the REFERENCE's own scores (tests/golden/convert, made by importing the reference converter), and
the host logic (model readers, cameras, depth ranges, view selection, writers) against the files
the reference wrote - with the pair scores, which the product computes on the device only, taken
from the oracle here."""
import os

import numpy as np
import pytest

import colmap_model as cm
import convert_util as cu
from acmmp_b200 import convert


@pytest.mark.parametrize("tag", sorted(cu.CASES))
def test_oracle_scores_equal_reference_calc_score(tag):
    model = convert.read_model(os.path.join(cu.GOLDEN, tag, "dense", "sparse"), ".txt")
    E = convert.extrinsic_matrices(model)
    mine = cu.oracle_scores(model, E)
    ref = cu.reference_scores(tag)
    assert (mine == ref).all()
    assert (ref == ref.T).all() and (np.diag(ref) == 0).all()
    # exactly one pair - the twin images - is zeroed by the one-degree rule
    off = ref[np.triu_indices(len(ref), 1)]
    assert (off == 0).sum() == 1 and (off > 0).sum() == len(off) - 1


@pytest.mark.parametrize("tag", sorted(cu.CASES))
def test_host_logic_reproduces_reference_files(tag, tmp_path, monkeypatch):
    monkeypatch.setattr(convert, "pair_scores", lambda model, E, device=0: cu.oracle_scores(model, E))
    save = str(tmp_path / "out")
    os.makedirs(save)
    model, score, sel = convert.processing_single_scene(cu.args_for(tag, save))
    cu.assert_same_outputs(tag, save)
    sizes = cu.expected_image_sizes(tag)
    assert sorted(os.listdir(os.path.join(save, "images"))) == sorted(sizes)
    for v, name in enumerate(model.names):
        out = os.path.join(save, "images", "%08d.jpg" % v)
        if name.endswith(".jpg"):       # JPEGs are byte copies
            assert open(out, "rb").read() == open(os.path.join(cu.GOLDEN, tag, "dense", "images", name), "rb").read()
        assert os.path.getsize(out) == sizes["%08d.jpg" % v]


def test_text_and_binary_models_agree(tmp_path):
    m = cm.make_model(seed=11, n_images=9, n_points=200)
    cm.write_text(m, str(tmp_path / "t"))
    cm.write_binary(m, str(tmp_path / "b"))
    a = convert.read_model(str(tmp_path / "t"), ".txt")
    b = convert.read_model(str(tmp_path / "b"), ".bin")
    assert a.names == b.names and (a.image_ids == b.image_ids).all() and (np.diff(a.image_ids) > 0).all()
    for f in ("qvecs", "tvecs", "camera_ids", "obs_offsets", "obs_point_ids", "point_ids", "xyz"):
        assert (getattr(a, f) == getattr(b, f)).all(), f
    for cid in a.cameras:
        assert a.cameras[cid].model == b.cameras[cid].model and (a.cameras[cid].params == b.cameras[cid].params).all()
    assert (convert.intrinsic_matrix(a.cameras[2]) == np.array([[701.25, 0, 400.0], [0, 701.25, 300.0], [0, 0, 1]])).all()
    idx = a.obs_point_index()
    has = a.obs_point_ids != -1
    assert (idx[~has] == -1).all() and (a.point_ids[idx[has]] == a.obs_point_ids[has]).all()


def test_rotation_and_centres():
    m = cm.make_model(seed=2, n_images=5, n_points=50)
    for im in m["images"]:
        R = convert.rotation_from_quaternion(im["qvec"])
        assert np.abs(R @ R.T - np.eye(3)).max() < 1e-12 and abs(np.linalg.det(R) - 1) < 1e-12
    E = np.zeros((1, 4, 4))
    E[0, :3, :3] = convert.rotation_from_quaternion(m["images"][0]["qvec"])
    E[0, :3, 3] = m["images"][0]["tvec"]
    c = convert.camera_centers(E)[0]
    assert np.abs(E[0, :3, :3] @ c + E[0, :3, 3]).max() < 1e-12


def test_view_selection_order_and_missing_points(tmp_path):
    score = np.array([[0, 5, 9, 5], [5, 0, 1, 0], [9, 1, 0, 2], [5, 0, 2, 0]], float)
    sel = convert.select_views(score, max_views=2)
    assert [k for k, _ in sel[0]] == [int(k) for k in np.argsort(score[0])[::-1][:2]] and sel[0][0] == (2, 9.0)
    convert.write_pair_file(str(tmp_path / "pair.txt"), sel)
    txt = open(str(tmp_path / "pair.txt")).read().split("\n")
    assert txt[0] == "4" and txt[1] == "0" and txt[2].startswith("2 2 9 ")
    # a keypoint naming a point that points3D lacks is an error, as in the reference (KeyError)
    m = cm.make_model(seed=5, n_images=3, n_points=30)
    m["images"][0]["point3D_ids"][0] = 10 ** 6
    cm.write_text(m, str(tmp_path / "bad"))
    with pytest.raises(KeyError):
        convert.read_model(str(tmp_path / "bad"), ".txt").obs_point_index()


def test_pair_scores_needs_the_device_library(monkeypatch):
    """no CPU fallback: without a CUDA device the call fails loudly."""
    torch = pytest.importorskip("torch")
    if torch.cuda.is_available():
        pytest.skip("a device is present")
    model = convert.read_model(os.path.join(cu.GOLDEN, "b", "dense", "sparse"), ".txt")
    with pytest.raises(RuntimeError):
        convert.pair_scores(model, convert.extrinsic_matrices(model))
