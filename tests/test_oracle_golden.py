"""CPU tests: the plain-C oracle (oracle/vslam_oracle.c) against the fixtures generated from the reference's own
OpenCV calls (tools/make_golden.py -> tests/golden/), and against cv2 live where it is importable."""
import hashlib
import os

import numpy as np
import pytest

from conftest import GOLDEN, image_for
from oracle import c_oracle as O


def md5(a):
    return hashlib.md5(np.ascontiguousarray(a).tobytes()).hexdigest()


def test_reference_pins_pyramid_test(golden):
    """The only values the reference's own tests print (tests/GaussPyramid_Test.cpp:88-110) on building.jpg."""
    L = O.lib()
    assert O.pyramid_sizes(600, 868, 4) == [(1200, 1736), (600, 868), (300, 434), (150, 217)]
    assert L.vo_calculate_sigma(0, 0, 1.6, 3) == 1.6
    assert abs(L.vo_calculate_sigma(0, 5, 1.6, 3) - 1.6 * 2 ** (5 / 3)) < 1e-12
    assert abs(L.vo_calculate_sigma(0, 3, 1.6, 3) - 3.2) < 1e-12
    assert L.vo_calculate_num_octaves(600, 868) == 5      # floor(log2(600)) - 4, GaussPyramid.cpp:151
    assert golden["pyramid"]["building"]["sizes"] == [[1200, 1736], [600, 868], [300, 434], [150, 217]]


def test_bgr_to_gray_matches_golden(images):
    """Diff_of_Gauss.cpp:730,739: the colour decode of home.jpg (fixture) through the oracle's cvtColor is the golden grey image."""
    import os
    with np.load(os.path.join(os.path.dirname(__file__), "golden", "home_bgr.npz")) as z:
        bgr = z["home_bgr"]
    assert np.array_equal(O.bgr_to_gray(bgr), images["home_cvt"])
    v = np.arange(256, dtype=np.uint8)
    assert np.array_equal(O.bgr_to_gray(np.stack([v, v, v], -1)[None])[0], v)       # grey in, grey out


def test_rotation_known_answers():
    """SLAM::Rotation (rotation.cpp:9-28, 66-108), parity unpinned (no reference vectors): known answers of the literal
    float arithmetic.  0 degrees is the identity except row 0 / column 0 (the strict '> 0' tests of rotate_mat_CCW)."""
    c0 = O.cos_sin_of_angle(0.0)
    assert c0 == (1.0, 0.0)
    c, s = O.cos_sin_of_angle(90.0)
    assert s == 1.0 and c == np.float32(np.cos(np.float64(np.float32(90.0 * (np.pi / np.float64(np.float32(180.0)))))))
    img = np.arange(1, 8 * 9 + 1, dtype=np.uint8).reshape(8, 9)
    want = img.copy(); want[0, :] = 0; want[:, 0] = 0
    assert np.array_equal(O.rotate_mat_ccw(img, (4, 4), c0), want)
    assert np.array_equal(O.rotated_window(img, (4, 4), 4, c0), img[2:6, 2:6])
    # literal numpy-float32 transcription of rotate_pt_CW for an arbitrary angle
    cs = O.cos_sin_of_angle(33.0)
    got = O.rotate_mat_ccw(img, (3, 5), cs)
    cf, sf = np.float32(cs[0]), np.float32(cs[1])
    for y in range(8):
        for x in range(9):
            rx, ry = np.float32(x - 3), np.float32(y - 5)
            xr = int(np.float32(rx * cf) - np.float32(ry * sf)) + 3
            yr = int(np.float32(rx * sf) + np.float32(ry * cf)) + 5
            assert got[y, x] == (img[yr, xr] if 0 < xr < 9 and 0 < yr < 8 else 0)


def test_taps_match_cv2_impulse_tables(golden):
    for key, e in golden["taps"].items():
        t = O.gauss_taps(e["sigma"])
        assert len(t) == e["n"], key
        assert t.tolist() == e["taps"], key
        assert t.sum() == 256
    for e in golden["taps_sweep"]:
        assert O.gauss_taps(e["sigma"]).tolist() == e["taps"], e["sigma"]
    assert O.gauss_taps(0.0, 3).tolist() == [64, 128, 64]


@pytest.mark.parametrize("stem", ["home", "building", "blox"])
def test_pyramid_and_keypoints_match_golden(stem, golden, images, golden_lists):
    g = golden["pyramid"][stem]
    img = image_for(images, stem, "cvt")
    kp, counts, pyr = O.dog_keypoints(img, 4, 1.6, 3)
    assert [list(b.shape) for b in pyr["base"]] == g["sizes"]
    assert [md5(b) for b in pyr["base"]] == g["base"]
    assert [[md5(m) for m in o] for o in pyr["gauss"]] == g["gauss"]
    assert [[md5(m) for m in o] for o in pyr["dog"]] == g["dog"]
    assert counts == g["counts"]
    assert len(kp) == g["n_keypoints"]
    assert np.array_equal(kp, golden_lists["dog_" + stem])


def test_known_counts(golden):
    """Totals the survey recorded from the cv2 run of the reference's calls."""
    assert golden["pyramid"]["home"]["n_keypoints"] == 91
    assert golden["pyramid"]["building"]["n_keypoints"] == 735
    assert [c["edge"] for c in golden["pyramid"]["home"]["counts"]] == [58, 32, 1, 0]
    assert golden["harris"]["blox"]["n_corners"] == 1328
    assert golden["harris"]["chessboard"]["n_corners"] == 283


def test_gradients_match_golden(golden, images, golden_lists):
    img = image_for(images, "home", "cvt")
    pyr = O.build_pyramid(img, 4, 1.6, gradients=True, mag_orient=True)
    g = golden["pyramid"]["home"]
    assert [[md5(m) for m in o] for o in pyr["gx"]] == g["gx"]
    assert [[md5(m) for m in o] for o in pyr["gy"]] == g["gy"]
    # magnitude: <= 1 ulp of sqrtf; phase: OpenCV fastAtan polynomial, SIMD rounding differs by < 1e-3 deg
    assert np.allclose(pyr["mag"][3][1], golden_lists["mag_o3l1_home"], rtol=2e-7, atol=0)
    assert np.abs(pyr["orient"][3][1] - golden_lists["orient_o3l1_home"]).max() < 1e-3


def test_localize_known_answers(golden):
    L = O.lib()
    for dx, dy, ds, val, want in golden["localize"]:
        got = L.vo_localize(dx, dy, ds, val)
        assert np.float32(got) == np.float32(want) or (np.isnan(got) and np.isnan(want)), (dx, dy, ds, val)


@pytest.mark.parametrize("stem", ["blox", "chessboard", "building", "home"])
def test_harris_matches_golden(stem, golden, images, golden_lists):
    g = golden["harris"][stem]
    h = O.harris(image_for(images, stem, "gray"))
    assert md5(h["blurred"]) == g["blurred"]
    assert md5(h["response"]) == g["response"]
    assert md5(h["r8"]) == g["r8"]
    assert md5(h["mask3"]) == g["mask3"]
    assert int((h["mask3"] > 0).sum()) == g["mask3_count"]
    assert len(h["corners"]) == g["n_corners"]
    assert np.array_equal(h["corners"], golden_lists["harris_" + stem])
    assert np.array_equal(h["corner_values"], golden_lists["harris_val_" + stem])
    assert h["true_max"] == g["true_max"]


def test_oracle_vs_cv2_live_on_awkward_sizes():
    """The C restatement against cv2 itself on seeded images: odd sizes, kernel wider than the image, 1-pixel rows."""
    cv2 = pytest.importorskip("cv2")
    from oracle import cv2_reference as R
    rng = np.random.default_rng(5)
    for (h, w) in [(1, 37), (37, 1), (2, 2), (5, 7), (33, 64), (48, 31), (97, 131)]:
        img = rng.integers(0, 256, (h, w), dtype=np.uint8)
        assert np.array_equal(O.upsample2x(img), cv2.resize(img, None, fx=2, fy=2, interpolation=cv2.INTER_LINEAR)), (h, w)
        for s in (1.6, 3.2, 12.8, 40.63):
            assert np.array_equal(O.gaussian_blur(img, s), cv2.GaussianBlur(img, (0, 0), s, sigmaY=0)), (h, w, s)
        hc, hr = O.harris(img), R.harris(img)
        for k in ("blurred", "response", "r8", "mask3", "corners", "corner_values"):
            assert np.array_equal(hc[k], hr[k]), (h, w, k)
    for (h, w) in [(24, 40), (33, 47), (64, 64)]:
        img = (rng.integers(0, 256, (h, w)) // 3 + np.linspace(0, 120, w)[None, :]).astype(np.uint8)
        kc, cc, _ = O.dog_keypoints(img, 3, 1.6, 3)
        kr, cr, _ = R.dog_keypoints(img, 3, 1.6, 3)
        assert cc == cr and np.array_equal(kc, kr), (h, w)


def test_localize_numpy_vs_literal_cv2():
    pytest.importorskip("cv2")
    from oracle import cv2_reference as R
    rng = np.random.default_rng(11)
    L = O.lib()
    for _ in range(300):
        dx, dy, ds = (int(v) for v in rng.integers(-60, 61, 3))
        val = int(rng.integers(0, 80))
        lit = np.float32(R.localize_literal_cv2(dx, dy, ds, val))
        vec = np.float32(R.localize_numpy(dx, dy, ds, val))
        c = np.float32(L.vo_localize(dx, dy, ds, val))
        assert (lit == vec == c) or (np.isnan(lit) and np.isnan(vec) and np.isnan(c)), (dx, dy, ds, val)


# ---- the stage after the hot path: orientation assignment and SIFT descriptors (tests/golden/descriptors.npz, generated
# ---- by tools/make_golden_descriptors.py through cv2) -------------------------------------------------------------------
FP32_TOL = 1e-4          # BASELINE.json north_star: max abs err <= 1e-4 on [0, 1]-normalised values


@pytest.fixture(scope="module")
def golden_desc():
    with np.load(os.path.join(os.path.dirname(__file__), "golden", "descriptors.npz")) as z:
        return {k: z[k] for k in z.files}


def test_fp32_gauss_taps_match_cv2():
    cv2 = pytest.importorskip("cv2")
    for o in range(4):
        for l in range(6):
            s = 1.5 * O.lib().vo_calculate_sigma(o, l, 1.6, 3)
            n = O.lib().vo_gauss_ksize_32f(s)
            assert np.array_equal(O.gauss_taps_32f(s), cv2.getGaussianKernel(n, s, cv2.CV_32F).ravel()), (o, l)
    assert O.lib().vo_gauss_ksize_32f(1.5 * 1.6 * 8 * 2 ** (5 / 3)) == 489


@pytest.mark.parametrize("stem", ["home", "building"])
def test_oracle_orientations_and_descriptors_match_cv2_golden(stem, images, golden_lists, golden_desc):
    img = image_for(images, stem, "cvt")
    kp = golden_lists[f"dog_{stem}"]
    pyr = O.build_pyramid(img)
    hist = np.stack([O.orientation_histogram(pyr["gauss"][k[4]][k[5]], int(k[0]), int(k[1]), 1.5 * pyr["sigma"][k[4]][k[5]])
                     for k in kp], 0)
    want = golden_desc[f"{stem}_hist"]
    peak = want.max(1, keepdims=True)
    assert np.abs(hist - want).max() <= FP32_TOL * max(1.0, float(peak.max()))
    assert (np.abs(hist - want) / peak).max() <= FP32_TOL          # peak-normalised, like the 0.8 * peak decision
    ori = np.concatenate([O.assign_orientations(pyr["gauss"][o], pyr["sigma"][o], kp[kp[:, 4] == o]) for o in range(4)], 0)
    assert np.array_equal(ori, golden_desc[f"{stem}_oriented"])    # no bin of these images sits within tolerance of 0.8 * peak
    desc, valid = [], []
    for o in range(4):
        d, v = O.sift(pyr["gauss"][o], pyr["sigma"][o], ori[ori[:, 4] == o])
        desc.append(d); valid.append(v)
    desc, valid = np.concatenate(desc, 0), np.concatenate(valid, 0)
    assert np.array_equal(valid, golden_desc[f"{stem}_valid"])
    ok = valid.astype(bool)
    assert ok.sum() > 50
    a, b = desc[ok], golden_desc[f"{stem}_desc"][ok]
    assert np.array_equal(np.isnan(a), np.isnan(b))
    assert np.nanmax(np.abs(a - b)) <= FP32_TOL
    assert np.nanmax(a) == 1.0 and np.nanmin(a) >= 0.0


def test_oracle_descriptor_known_answers():
    """A flat image has zero gradients: every histogram is empty, no orientation is assigned, and SIFT divides 0 by 0."""
    g = [np.full((64, 80), 77, np.uint8)] * 6
    sig = [1.6 * 2 ** (l / 3) for l in range(6)]
    kp = np.array([[30, 40, 9, 1, 0, 2]], np.int32)
    assert np.array_equal(O.orientation_histogram(g[2], 30, 40, 1.5 * sig[2]), np.zeros(36, np.float32))
    assert len(O.assign_orientations(g, sig, kp)) == 0
    d, v = O.sift(g, sig, np.array([[30, 40, 90, 0, 0, 2]], np.int32))
    assert np.isnan(d).all() and v[0] == 1
    # a horizontal ramp has gx = 2, gy = 0 away from the border: orientation 0 degrees, everything lands in bin 0
    ramp = np.tile((np.arange(80) * 2).astype(np.uint8), (64, 1))
    h = O.orientation_histogram(ramp, 30, 40, 1.5 * sig[1])
    assert h[0] > 0 and not h[1:].any()
    ori = O.assign_orientations([ramp] * 6, sig, np.array([[30, 40, 9, 1, 0, 1]], np.int32))
    assert ori.tolist() == [[30, 40, 0, 0, 0, 1]]
    d, v = O.sift([ramp] * 6, sig, ori)
    assert v[0] == 1 and set(np.nonzero(d[0])[0]) == set(range(0, 128, 8)) and d[0].max() == 1.0


def test_oracle_descriptor_stage_vs_cv2_live_on_a_tiny_pyramid():
    """Kernels several times wider than the map (octave 2 of a 48 x 64 frame is 24 x 32; up to 155 taps): OpenCV's repeated
    reflect-101 bouncing, live through cv2 (oracle/cv2_reference.py), against the C restatement."""
    pytest.importorskip("cv2")
    from oracle import cv2_reference as R
    from visualslam_b200 import synth
    img = synth.frame(5, 48, 64)
    pyr = O.build_pyramid(img, 3, 1.6)
    ref = R.build_pyramid(img, 3, 1.6)
    for o in range(3):
        rows, cols = pyr["gauss"][o][0].shape
        pts = [(0, 0), (rows - 1, cols - 1), (rows // 2, cols // 3), (1, cols - 2), (rows - 2, 3)]
        for l in (1, 2, 3):
            sig = 1.5 * pyr["sigma"][o][l]
            kp = np.array([[y, x, 0, 1, o, l] for (y, x) in pts], np.int32)
            want = R.orientation_histograms(ref["mag"][o][l], ref["orient"][o][l], kp, sig)
            got = np.stack([O.orientation_histogram(pyr["gauss"][o][l], y, x, sig) for (y, x) in pts], 0)
            assert np.abs(got - want).max() <= FP32_TOL * max(1.0, float(want.max())), (o, l)
            ori = O.assign_orientations(pyr["gauss"][o], pyr["sigma"][o], kp)
            desc, valid = O.sift(pyr["gauss"][o], pyr["sigma"][o], ori)
            for n, k in enumerate(ori):
                d, ok = R.sift_descriptor(ref["mag"][o][l], ref["orient"][o][l], int(k[0]), int(k[1]), int(k[2]), sig)
                assert ok == bool(valid[n]), (o, l, n)
                if ok:
                    assert np.array_equal(np.isnan(d), np.isnan(desc[n]))
                    if not np.isnan(d).any():
                        assert np.abs(d - desc[n]).max() <= FP32_TOL, (o, l, n)


def test_rotation_matrix_warp_affine_and_double_crop_match_cv2():
    """Rotation::doubleCrop's OpenCV calls (rotation.cpp:152,185,198) and the whole function (:147-205) against cv2 4.13 vectors
    (tools/make_golden_rotation.py): matrices bit for bit, warped images bit for bit, crops and ROIs identical."""
    with np.load(os.path.join(GOLDEN, "warp_affine.npz")) as z:
        g = {k: z[k] for k in z.files}
    for i, (h, w, ang, cx, cy, dh, dw, scale, seed) in enumerate(g["warp_cases"]):
        img = np.random.default_rng(int(seed)).integers(0, 256, (int(h), int(w)), dtype=np.uint8)
        m = O.get_rotation_matrix_2d((cx, cy), ang, scale)
        assert m.tobytes() == g[f"warp_m_{i}"].tobytes(), (i, m, g[f"warp_m_{i}"])
        assert np.array_equal(O.warp_affine(img, m, (int(dw), int(dh))), g[f"warp_out_{i}"]), i
    with np.load(os.path.join(GOLDEN, "images.npz")) as z:
        blox = z["blox"]
    for k, (cx, cy, window, ang, x0, y0, rw, rh) in enumerate(g["dcrop_cases"]):
        crop, roi = O.double_crop(blox, (int(cx), int(cy)), int(window), float(ang))
        assert roi == (int(x0), int(y0), int(rw), int(rh)), (k, roi)
        assert crop.shape == g[f"dcrop_{k}"].shape and np.array_equal(crop, g[f"dcrop_{k}"]), (k, crop.shape)
    with pytest.raises(ValueError):
        O.double_crop(blox, (10, 10), 100, 45.0)          # the first crop leaves the image: cv::Mat(src, Rect) throws
