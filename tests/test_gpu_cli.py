"""The C++ command line (cli/acmmp_main.cpp, build/acmmp_main) on the reference's on-disk layout:
pair.txt, cams/*_cam.txt, images/*  ->  2333_*/{depths,depths_geom,normals,costs}.dmb + ACMMP_model.ply;
the files it writes are then fed to the reference's OWN RunFusion.  pytest -m gpu."""
import os
import shutil
import subprocess

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
if not torch.cuda.is_available():
    pytest.skip("no CUDA device", allow_module_level=True)

import refbind  # noqa: E402
from acmmp_b200.scene import make_scene  # noqa: E402

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "build", "acmmp_main")


def read_dmb(path):
    hdr = np.fromfile(path, np.int32, 4)
    assert hdr[0] == 1
    a = np.fromfile(path, np.float32, offset=16)
    return a.reshape(hdr[1], hdr[2]) if hdr[3] == 1 else a.reshape(hdr[1], hdr[2], hdr[3])


def read_ply(path):
    blob = open(path, "rb").read()
    end = blob.index(b"end_header\n") + 11
    n = int([l for l in blob[:end].decode().splitlines() if l.startswith("element vertex")][0].split()[-1])
    rec = np.frombuffer(blob[end:end + 27 * n], dtype=np.dtype([("f", "<f4", 6), ("c", "u1", 3)]))
    return rec["f"].copy(), rec["c"].copy()


@pytest.fixture(scope="module")
def cli_run(tmp_path_factory):
    if not os.path.exists(EXE):
        pytest.skip("build/acmmp_main not built (make cli)")
    sc = make_scene(5, 320, 240, 4, seed=6)
    dense = str(tmp_path_factory.mktemp("cli") / "dense")
    os.makedirs(os.path.join(dense, "images"))
    os.makedirs(os.path.join(dense, "cams"))
    n = 5
    ids = [10 * i + 3 for i in range(n)]                      # image ids are not indices
    with open(os.path.join(dense, "pair.txt"), "w") as f:
        f.write("%d\n" % n)
        for i in range(n):
            f.write("%d\n%d " % (ids[i], len(sc["pairs"][i])) + " ".join("%d 1.0" % ids[j] for j in sc["pairs"][i]) + "\n")
    for i in range(n):
        img = np.clip(np.rint(sc["images"][i]), 0, 255).astype(np.uint8)
        with open(os.path.join(dense, "images", "%08d.jpg" % ids[i]), "wb") as f:   # raw container under the reference's name
            f.write(b"ACMRAW1 %d %d 1\n" % img.shape + img.tobytes())
        refbind._write_cam(os.path.join(dense, "cams", "%08d_cam.txt" % ids[i]), sc["cams"][i])
    out = subprocess.run([EXE, dense, "--seed", "5"], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr + out.stdout
    return sc, dense, ids, out.stdout


def test_cli_writes_the_reference_layout(cli_run):
    sc, dense, ids, stdout = cli_run
    n = len(ids)
    outdir = os.path.join(dense, "ACMMP")
    for i in range(n):
        folder = os.path.join(outdir, "2333_%08d" % ids[i])
        depth = read_dmb(os.path.join(folder, "depths_geom.dmb"))
        normal = read_dmb(os.path.join(folder, "normals.dmb"))
        cost = read_dmb(os.path.join(folder, "costs.dmb"))
        photo = read_dmb(os.path.join(folder, "depths.dmb"))
        assert depth.shape == (240, 320) and normal.shape == (240, 320, 3) and cost.shape == (240, 320) and photo.shape == (240, 320)
        gt = sc["depths"][i]
        assert (np.abs(depth - gt)[8:-8, 8:-8] / gt[8:-8, 8:-8] < 0.01).mean() >= 0.9
        assert np.abs(np.linalg.norm(normal, axis=-1) - 1).max() < 1e-3
    xyz_n, col = read_ply(os.path.join(outdir, "ACMMP_model.ply"))
    assert len(xyz_n) > 50000 and np.isfinite(xyz_n).all()
    assert "approved" in stdout


@pytest.mark.skipif(not refbind.host_available(), reason="oracle/_ref/libacmmp_ref_host*.so not built")
def test_reference_runfusion_reads_the_cli_output(cli_run):
    """the reference's own RunFusion on the files the CLI wrote (same ids, same layout)"""
    sc, dense, ids, _ = cli_run
    n = len(ids)
    outdir = os.path.join(dense, "ACMMP")
    xyz_n, col = read_ply(os.path.join(outdir, "ACMMP_model.ply"))
    shutil.copy(os.path.join(outdir, "ACMMP_model.ply"), os.path.join(outdir, "ours.ply"))
    import ctypes as C
    lib = C.CDLL(refbind.REF_HOST_PATHS["ieee"])
    lib.ref_host_run_fusion.argtypes = [C.c_char_p, C.c_char_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int),
                                        C.c_int, C.c_float, C.c_int]
    offs, flat = [0], []
    for p in sc["pairs"]:
        flat += [ids[j] for j in p]
        offs.append(len(flat))
    rc = lib.ref_host_run_fusion(dense.encode(), outdir.encode(), n, (C.c_int * n)(*ids), (C.c_int * (n + 1))(*offs),
                                 (C.c_int * len(flat))(*flat), 1, 0.3, 1)
    assert rc == 0
    ref_xyz_n, ref_col = read_ply(os.path.join(outdir, "ACMMP_model.ply"))
    assert abs(len(ref_xyz_n) - len(xyz_n)) <= max(1, int(1e-3 * len(ref_xyz_n))), (len(ref_xyz_n), len(xyz_n))
    if len(ref_xyz_n) == len(xyz_n):
        same = (ref_xyz_n.view(np.uint32) == xyz_n.view(np.uint32)).all(1).mean()
        assert same >= 0.999, same
        assert (ref_col == col).all()


def test_cli_reads_real_jpegs(tmp_path):
    """images/%08d.jpg as real JPEG files (decoded by nvJPEG inside the library)"""
    cv2 = pytest.importorskip("cv2")
    if not os.path.exists(EXE):
        pytest.skip("build/acmmp_main not built (make cli)")
    sc = make_scene(5, 320, 240, 4, seed=6)
    dense = str(tmp_path / "dense")
    os.makedirs(os.path.join(dense, "images"))
    os.makedirs(os.path.join(dense, "cams"))
    n = 5
    with open(os.path.join(dense, "pair.txt"), "w") as f:
        f.write("%d\n" % n)
        for i in range(n):
            f.write("%d\n%d " % (i, len(sc["pairs"][i])) + " ".join("%d 1.0" % j for j in sc["pairs"][i]) + "\n")
    for i in range(n):
        img = np.clip(np.rint(sc["images"][i]), 0, 255).astype(np.uint8)
        assert cv2.imwrite(os.path.join(dense, "images", "%08d.jpg" % i), cv2.cvtColor(img, cv2.COLOR_GRAY2BGR), [cv2.IMWRITE_JPEG_QUALITY, 98])
        refbind._write_cam(os.path.join(dense, "cams", "%08d_cam.txt" % i), sc["cams"][i])
    out = subprocess.run([EXE, dense, "--seed", "5", "--output_dir", "/OUT"], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr + out.stdout
    depth = read_dmb(os.path.join(dense, "OUT", "2333_%08d" % 2, "depths_geom.dmb"))
    gt = sc["depths"][2]
    assert (np.abs(depth - gt)[8:-8, 8:-8] / gt[8:-8, 8:-8] < 0.01).mean() >= 0.9
    xyz_n, col = read_ply(os.path.join(dense, "OUT", "ACMMP_model.ply"))
    assert len(xyz_n) > 50000 and col.max() > 0


@pytest.mark.skipif(not refbind.host_available(), reason="oracle/_ref/libacmmp_ref_host*.so not built")
def test_cli_prior_aware_fusion_against_reference(cli_run):
    """--force_fusion --multi_fusion /ACMMP: a second run (another seed, output /ACMMP_B) fused with the
    first run's maps; the reference's own RunPriorAwareFusion then reads the same two folders."""
    sc, dense, ids, _ = cli_run
    n = len(ids)
    out = subprocess.run([EXE, dense, "--seed", "9", "--output_dir", "/ACMMP_B", "--force_fusion", "--multi_fusion", "/ACMMP",
                          "--single_match_penalty", "1", "--num_consistent_thresh", "2"], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr + out.stdout
    outdir = os.path.join(dense, "ACMMP_B")
    ply = os.path.join(outdir, "ACMMP_prior_model.ply")
    xyz_n, col = read_ply(ply)
    assert len(xyz_n) > 50000 and not os.path.exists(os.path.join(outdir, "ACMMP_model.ply"))
    shutil.copy(ply, os.path.join(outdir, "ours_prior.ply"))
    import ctypes as C
    lib = C.CDLL(refbind.REF_HOST_PATHS["ieee"])
    lib.ref_host_run_prior_fusion.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int),
                                              C.POINTER(C.c_int), C.c_int, C.c_float, C.c_int, C.c_int]
    offs, flat = [0], []
    for p in sc["pairs"]:
        flat += [ids[j] for j in p]
        offs.append(len(flat))
    rc = lib.ref_host_run_prior_fusion(dense.encode(), outdir.encode(), os.path.join(dense, "ACMMP").encode(), n, (C.c_int * n)(*ids),
                                       (C.c_int * (n + 1))(*offs), (C.c_int * len(flat))(*flat), 1, 0.3, 2, 1)
    assert rc == 0
    ref_xyz_n, ref_col = read_ply(ply)
    assert abs(len(ref_xyz_n) - len(xyz_n)) <= max(1, int(1e-3 * len(ref_xyz_n))), (len(ref_xyz_n), len(xyz_n))
    if len(ref_xyz_n) == len(xyz_n):
        same = (ref_xyz_n.view(np.uint32) == xyz_n.view(np.uint32)).all(1).mean()
        assert same >= 0.999, same
        assert (ref_col == col).all()


def test_cli_seeded_mode_reads_prior_pngs(cli_run):
    """-p: <dense>/priors/{depths,normals}/%08d.png (16-bit PNG, numbered by problem index) seed the first
    scale; the maps written under /ACMMP_PRIOR equal a scan seeded through the Python host with the arrays
    cv2.imread(path, -1) returns."""
    cv2 = pytest.importorskip("cv2")
    from acmmp_b200 import host
    from seeded_inputs import encode_priors
    sc, dense, ids, _ = cli_run
    n = len(ids)
    priors = []
    for v in range(n):
        nrm = np.zeros(sc["depths"][v].shape + (3,))
        nrm[..., 2] = -1.0
        nrm[..., 0] = 0.1
        d16, n16 = encode_priors(sc["cams"][v], sc["depths"][v], nrm, noise=0.01, seed=v)
        os.makedirs(os.path.join(dense, "priors", "depths"), exist_ok=True)
        os.makedirs(os.path.join(dense, "priors", "normals"), exist_ok=True)
        assert cv2.imwrite(os.path.join(dense, "priors", "depths", "%08d.png" % v), d16)
        assert cv2.imwrite(os.path.join(dense, "priors", "normals", "%08d.png" % v), n16)   # cv2 stores channel 0 as blue
        back = cv2.imread(os.path.join(dense, "priors", "normals", "%08d.png" % v), -1)
        assert (back == n16).all()
        priors.append((d16, n16))
    out = subprocess.run([EXE, dense, "-p", "--seed", "5", "--no_fusion"], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr + out.stdout
    s = host.Scan(sc["cams"], sc["pairs"])
    for v in range(n):
        s.set_image(v, np.clip(np.rint(sc["images"][v]), 0, 255).astype(np.float32))
        s.set_prior_maps(v, *priors[v])
    s.run_all(seed=5)
    s.sync()
    for v in range(n):
        depth, normal, cost = s.get_view(v)
        folder = os.path.join(dense, "ACMMP_PRIOR", "2333_%08d" % ids[v])
        assert (read_dmb(os.path.join(folder, "depths_geom.dmb")).view(np.uint32) == depth.view(np.uint32)).all()
        assert (read_dmb(os.path.join(folder, "normals.dmb")).view(np.uint32) == normal.view(np.uint32)).all()
    s.close()
    # without the prior images the mode refuses to start, as the reference does (main_ACMMP.cpp:73-77)
    import shutil as sh
    sh.rmtree(os.path.join(dense, "priors"))
    out = subprocess.run([EXE, dense, "-p", "--no_fusion"], capture_output=True, text=True)
    assert out.returncode != 0 and "no suitable priors" in out.stderr


def test_cli_reads_png_masks_in_the_reference_format(cli_run):
    """--mask_dir: <dense>/<mask_dir>/%08d.png read like cv::imread(-1), resized to the depth map with
    cv::resize(INTER_LINEAR) and thresholded `< 128` (acmmp_definitions.cpp:884-899).  The masks are written at
    twice the depth-map size, so the CLI's own 8-bit resize (io.hpp ResizeLinear8) is on the path; the expected
    cloud comes from the device fusion fed with masks that cv2 made from the same files."""
    cv2 = pytest.importorskip("cv2")
    from acmmp_b200 import host
    sc, dense, ids, _ = cli_run
    n = len(ids)
    rng = np.random.default_rng(3)
    os.makedirs(os.path.join(dense, "masks"), exist_ok=True)
    masks = []
    for i in range(n):
        big = cv2.GaussianBlur((rng.random((480, 640)) * 255).astype(np.uint8), (0, 0), 9)
        big = cv2.normalize(big, None, 0, 255, cv2.NORM_MINMAX)
        assert cv2.imwrite(os.path.join(dense, "masks", "%08d.png" % ids[i]), big)
        back = cv2.imread(os.path.join(dense, "masks", "%08d.png" % ids[i]), -1)
        masks.append((cv2.resize(back, (320, 240)) < 128).astype(np.uint8))
        assert 0.2 < masks[-1].mean() < 0.8
    out = subprocess.run([EXE, dense, "--seed", "5", "--output_dir", "/MASKED", "--mask_dir", "masks"], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr + out.stdout
    xyz_n, col = read_ply(os.path.join(dense, "MASKED", "ACMMP_model.ply"))
    depths, normals = [], []
    for i in range(n):
        folder = os.path.join(dense, "MASKED", "2333_%08d" % ids[i])
        depths.append(read_dmb(os.path.join(folder, "depths_geom.dmb")))
        normals.append(read_dmb(os.path.join(folder, "normals.dmb")))
    pts, per = host.RunFusion(sc["cams"], sc["pairs"], depths, normals, masks=masks)
    assert len(pts) == len(xyz_n), (len(pts), len(xyz_n))
    assert (pts[:, :6].astype(np.float32).view(np.uint32) == xyz_n.view(np.uint32)).all()
    unmasked, _ = read_ply(os.path.join(dense, "ACMMP", "ours.ply")) if os.path.exists(os.path.join(dense, "ACMMP", "ours.ply")) else (None, None)
    if unmasked is not None:
        assert len(xyz_n) < len(unmasked)
    # a missing mask is an error, not a silent no-mask run
    os.remove(os.path.join(dense, "masks", "%08d.png" % ids[0]))
    out = subprocess.run([EXE, dense, "--seed", "5", "--output_dir", "/MASKED", "--mask_dir", "masks"], capture_output=True, text=True)
    assert out.returncode != 0 and "mask of image" in out.stderr


@pytest.mark.skipif(not refbind.host_available(), reason="oracle/_ref/libacmmp_ref_host*.so not built")
def test_fuse_data_refuses_the_output_folder_like_the_reference(cli_run):
    """build/fuse_data = main_fusion.cpp: re-fuses <dense>/ACMMP from depths.dmb (geom_consistency = false there,
    main_fusion.cpp:75) + normals.dmb; the reference's own RunFusion with the same flag reads the same folder."""
    exe = os.path.join(ROOT, "build", "fuse_data")
    if not os.path.exists(exe):
        pytest.skip("build/fuse_data not built (make cli)")
    sc, dense, ids, _ = cli_run
    n = len(ids)
    outdir = os.path.join(dense, "ACMMP")
    out = subprocess.run([exe, dense, "-f", "0.3", "--num_consistent_thresh", "2"], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr + out.stdout
    assert "approved" in out.stdout
    xyz_n, col = read_ply(os.path.join(outdir, "ACMMP_model.ply"))
    shutil.copy(os.path.join(outdir, "ACMMP_model.ply"), os.path.join(outdir, "fuse_data.ply"))
    import ctypes as C
    lib = C.CDLL(refbind.REF_HOST_PATHS["ieee"])
    lib.ref_host_run_fusion.argtypes = [C.c_char_p, C.c_char_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int),
                                        C.c_int, C.c_float, C.c_int]
    offs, flat = [0], []
    for p in sc["pairs"]:
        flat += [ids[j] for j in p]
        offs.append(len(flat))
    rc = lib.ref_host_run_fusion(dense.encode(), outdir.encode(), n, (C.c_int * n)(*ids), (C.c_int * (n + 1))(*offs),
                                 (C.c_int * len(flat))(*flat), 0, 0.3, 2)
    assert rc == 0
    ref_xyz_n, ref_col = read_ply(os.path.join(outdir, "ACMMP_model.ply"))
    assert len(ref_xyz_n) > 20000
    assert abs(len(ref_xyz_n) - len(xyz_n)) <= max(1, int(1e-3 * len(ref_xyz_n))), (len(ref_xyz_n), len(xyz_n))
    if len(ref_xyz_n) == len(xyz_n):
        assert (ref_xyz_n.view(np.uint32) == xyz_n.view(np.uint32)).all(1).mean() >= 0.999
        assert (ref_col == col).all()
    # a missing map is an error
    os.rename(os.path.join(outdir, "2333_%08d" % ids[1], "normals.dmb"), os.path.join(outdir, "2333_%08d" % ids[1], "normals.bak"))
    out = subprocess.run([exe, dense], capture_output=True, text=True)
    os.rename(os.path.join(outdir, "2333_%08d" % ids[1], "normals.bak"), os.path.join(outdir, "2333_%08d" % ids[1], "normals.dmb"))
    assert out.returncode != 0 and "cannot read the maps" in out.stderr


def test_cli_reference_flow_through_the_cpp_class(cli_run):
    """--reference_flow: the reference's own flow - ProcessProblem per view and stage with .dmb hand-over, host planar
    prior, JBU between scales - written against `class ACMMP` with the reference's method signatures
    (acmmp_b200/host/acmmp.hpp, driver.hpp).  Same scene as the resident run: the two flows must agree like two seeds."""
    sc, dense, ids, _ = cli_run
    n = len(ids)
    out = subprocess.run([EXE, dense, "--seed", "5", "--output_dir", "/FLOW", "--reference_flow"], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr + out.stdout
    assert out.stdout.count("Run Planar Prior Assisted PatchMatch MVS") == n and out.stdout.count("Processing image") == 2 * 3 * n
    inner = (slice(8, -8), slice(8, -8))
    for i in range(n):
        flow = read_dmb(os.path.join(dense, "FLOW", "2333_%08d" % ids[i], "depths_geom.dmb"))
        resident = read_dmb(os.path.join(dense, "ACMMP", "2333_%08d" % ids[i], "depths_geom.dmb"))
        gt = sc["depths"][i]
        acc_flow = (np.abs(flow - gt)[inner] / gt[inner] < 0.01).mean()
        acc_res = (np.abs(resident - gt)[inner] / gt[inner] < 0.01).mean()
        assert acc_flow >= 0.9 and abs(acc_flow - acc_res) <= 0.01, (acc_flow, acc_res)
        assert (np.abs(flow - resident)[inner] / gt[inner] < 0.001).mean() >= 0.93
        for name in ("depths.dmb", "normals.dmb", "costs.dmb"):
            assert os.path.exists(os.path.join(dense, "FLOW", "2333_%08d" % ids[i], name))
    xyz_flow, _ = read_ply(os.path.join(dense, "FLOW", "ACMMP_model.ply"))
    xyz_res, _ = read_ply(os.path.join(dense, "ACMMP", "ours.ply")) if os.path.exists(os.path.join(dense, "ACMMP", "ours.ply")) else (None, None)
    assert len(xyz_flow) > 50000
    if xyz_res is not None:
        assert abs(len(xyz_flow) - len(xyz_res)) <= 0.01 * len(xyz_res), (len(xyz_flow), len(xyz_res))
