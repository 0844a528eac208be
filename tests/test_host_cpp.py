"""C++17 host layer (visualslam_b200/host): the reference's own names and executables above the C ABI.

CPU part: the layer builds with g++ against include/vslam.h, the executables exist, print their usage, and FAIL LOUDLY
without a CUDA device (no CPU fallback).  GPU part: ./DoG, ./Harris and tests/Pyramid_Test run on the bundled images
(pre-decoded to PGM from tests/golden/images.npz) and print / write exactly the golden keypoints and corner counts.
"""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST = os.path.join(ROOT, "visualslam_b200", "host")
BIN = os.path.join(HOST, "build")


def _build():
    from visualslam_b200 import build as b
    b.build()
    subprocess.check_call(["make", "-C", HOST, "-s"])


def _pgm(path, img):
    with open(path, "wb") as f:
        f.write(b"P5\n%d %d\n255\n" % (img.shape[1], img.shape[0]))
        f.write(np.ascontiguousarray(img, np.uint8).tobytes())


def _run(exe, *args):
    return subprocess.run([os.path.join(BIN, exe), *args], capture_output=True, text=True, timeout=300)


def test_host_layer_builds_and_prints_usage():
    _build()
    for exe in ("DoG", "Harris", os.path.join("tests", "Pyramid_Test"), os.path.join("tests", "RotateImgTest"), "libvslam_host.a"):
        assert os.path.exists(os.path.join(BIN, exe)), exe
    for exe in ("DoG", "Harris"):
        r = _run(exe, "--help")
        assert r.returncode == 0 and "usage" in r.stdout


def test_host_layer_has_no_cpu_fallback(tmp_path, images):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    _build()
    p = str(tmp_path / "blox.pgm")
    _pgm(p, images["blox"])
    for exe in ("DoG", "Harris", os.path.join("tests", "Pyramid_Test")):
        r = _run(exe, p)
        assert r.returncode != 0
        assert "CUDA" in r.stderr or "cuda" in r.stderr


def _last_int(stdout, prefix):
    for line in stdout.splitlines():
        if line.startswith(prefix):
            return int(line[len(prefix):].strip())
    raise AssertionError(f"{prefix!r} not printed:\n{stdout}")


@pytest.mark.gpu
@pytest.mark.parametrize("stem", ["home", "building", "blox"])
def test_dog_executable_matches_golden(tmp_path, images, golden_lists, stem):
    from conftest import image_for
    _build()
    p = str(tmp_path / f"{stem}.pgm")
    _pgm(p, image_for(images, stem, "cvt"))
    want = golden_lists[f"dog_{stem}"]
    for flags in ([], ["--fused"]):
        out = str(tmp_path / f"kp_{stem}_{len(flags)}.txt")
        r = _run("DoG", p, "--out", out, *flags)
        assert r.returncode == 0, r.stderr
        assert _last_int(r.stdout, "Number of keypoints after edge response:") == len(want)
        got = np.loadtxt(out, dtype=np.int64).reshape(-1, 6)
        assert np.array_equal(got, want)
        if not flags:       # the reference's own stdout lines (Diff_of_Gauss.cpp:784, 786)
            assert "Pyramid octave: 0" in r.stdout and "Pyramid octave: 3" in r.stdout
            h, w = image_for(images, stem, "cvt").shape
            assert f"image size: width.height ({2 * w}, {2 * h})" in r.stdout


@pytest.mark.gpu
@pytest.mark.parametrize("flags", [[], ["--fused"]])
def test_dog_executable_writes_the_descriptor_file(tmp_path, images, flags):
    """./DoG after the hot path: oriented keypoints and featureDescriptors.dat (Diff_of_Gauss.cpp:787-791, 815, 834, 837-863)
    against the cv2 fixtures of tools/make_golden_descriptors.py."""
    from conftest import image_for
    _build()
    with np.load(os.path.join(ROOT, "tests", "golden", "descriptors.npz")) as z:
        gd = {k: z[k] for k in z.files}
    p = str(tmp_path / "home.pgm")
    _pgm(p, image_for(images, "home", "cvt"))
    ori, dat = str(tmp_path / "oriented.txt"), str(tmp_path / "featureDescriptors.dat")
    r = _run("DoG", p, "--oriented", ori, "--desc", dat, *flags)
    assert r.returncode == 0, r.stderr
    want = gd["home_oriented"]
    assert _last_int(r.stdout, "Number of features found in image:") == len(want)
    assert _last_int(r.stdout, "Number of keypoints after edge orientation:") == len(want)
    assert np.array_equal(np.loadtxt(ori, dtype=np.int64).reshape(-1, 6), want)
    raw = open(dat, "rb").read()
    assert np.frombuffer(raw[:12], "<i4").tolist() == [len(want), 128, 24]
    desc = np.frombuffer(raw[12:], "<f4").reshape(-1, 128)
    ok = gd["home_valid"].astype(bool)
    assert np.nanmax(np.abs(desc[ok] - gd["home_desc"][ok])) <= 1e-4


@pytest.mark.gpu
def test_dog_executable_on_colour_input(tmp_path, golden_lists):
    """./DoG on a colour PPM: imread(IMREAD_COLOR) + cvtColor on the GPU (Diff_of_Gauss.cpp:730-739), then the golden keypoints."""
    _build()
    with np.load(os.path.join(ROOT, "tests", "golden", "home_bgr.npz")) as z:
        bgr = z["home_bgr"]
    p = str(tmp_path / "home.ppm")
    with open(p, "wb") as f:
        f.write(b"P6\n%d %d\n255\n" % (bgr.shape[1], bgr.shape[0]))
        f.write(np.ascontiguousarray(bgr[..., ::-1]).tobytes())          # PPM is RGB
    out = str(tmp_path / "kp.txt")
    r = _run("DoG", p, "--out", out, "--fused")
    assert r.returncode == 0, r.stderr
    assert np.array_equal(np.loadtxt(out, dtype=np.int64).reshape(-1, 6), golden_lists["dog_home"])


@pytest.mark.gpu
@pytest.mark.parametrize("stem", ["blox", "chessboard", "building"])
def test_harris_executable_matches_golden(tmp_path, images, golden_lists, stem):
    from conftest import image_for
    _build()
    p = str(tmp_path / f"{stem}.pgm")
    img = image_for(images, stem, "gray")
    _pgm(p, img)
    want = golden_lists[f"harris_{stem}"]
    vals = golden_lists[f"harris_val_{stem}"]
    outs = []
    for flags in ([], ["--fused"]):
        r = _run("Harris", p, "--dump", str(tmp_path / f"{stem}{len(flags)}"), *flags)
        assert r.returncode == 0, r.stderr
        assert _last_int(r.stdout, "Number of corners (NMS2):") == len(want)
        assert f"Response rows: {img.shape[0]}" in r.stdout and f"Response cols: {img.shape[1]}" in r.stdout
        outs.append(r.stdout)
        resp = np.fromfile(str(tmp_path / f"{stem}{len(flags)}_response.pfm"), dtype=np.float32, offset=len(f"Pf\n{img.shape[1]} {img.shape[0]}\n-1.0\n"))
        resp = resp.reshape(img.shape)
        assert np.array_equal(resp[want[:, 0], want[:, 1]], vals)
    assert outs[0] == outs[1]       # function-by-function and fused pipelines print the same numbers


@pytest.mark.gpu
def test_rotate_img_test_executable(tmp_path, images):
    """tests/RotateImgTest (rotate_image_test.cpp:55-84): the timing line, and the dumped gathers equal the oracle's."""
    from oracle import c_oracle as O
    _build()
    img = images["blox"]
    p = str(tmp_path / "blox.pgm")
    _pgm(p, img)
    r = _run(os.path.join("tests", "RotateImgTest"), p, "--dump", str(tmp_path / "r"))
    assert r.returncode == 0, r.stderr
    assert "Time rotating 16x16 ROI (averaged for 100 runs):" in r.stdout
    cs = O.cos_sin_of_angle(45.0)
    center = (img.shape[1] // 2, img.shape[0] // 2)

    def pgm(path):
        raw = open(path, "rb").read()
        parts = raw.split(b"\n", 3)
        w, h = map(int, parts[1].split())
        return np.frombuffer(parts[3], np.uint8).reshape(h, w)
    assert np.array_equal(pgm(str(tmp_path / "r_rotated.pgm")), O.rotate_mat_ccw(img, center, cs))
    assert np.array_equal(pgm(str(tmp_path / "r_roi.pgm")), O.rotated_window(img, center, 100, cs))
    # the double-crop comparison (rotate_image_test.cpp:85-94): its timing line, and the crop equals the oracle's (= cv2's, = the
    # reference's own compiled doubleCrop: tests/test_oracle_ref.py)
    assert "Time rotating 16x16 ROI with the double crop method (averaged for 100 runs):" in r.stdout
    assert np.array_equal(pgm(str(tmp_path / "r_roi2.pgm")), O.double_crop(img, center, 100, 45.0)[0])


@pytest.mark.gpu
def test_pyramid_test_executable(tmp_path, images):
    from conftest import image_for
    _build()
    p = str(tmp_path / "building.pgm")
    _pgm(p, image_for(images, "building", "cvt"))
    r = _run(os.path.join("tests", "Pyramid_Test"), p)
    assert r.returncode == 0, r.stdout + r.stderr
    # the values the reference's own test prints (tests/GaussPyramid_Test.cpp:88-110)
    assert "Number of expected octaves: 4" in r.stdout
    assert "Number of expected levels per octave: 6" in r.stdout
    assert "image size: width.height (1736, 1200)" in r.stdout and "image size: width.height (217, 150)" in r.stdout
    assert "First sigma value of pyramid : 1.6" in r.stdout
    assert "Last sigma value of 1st octave : 5.07968" in r.stdout
    assert "double sigma value : 3.2" in r.stdout
    assert "Number of sigma values per octave: 6" in r.stdout
