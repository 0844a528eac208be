"""CPU tests: the plain-C oracle (oracle/vslam_oracle.c) against outputs of the REFERENCE'S OWN code.

tests/golden/ref_golden.npz was written by tools/make_ref_golden.py from oracle/_ref -- Harris_corners.cpp,
Diff_of_Gauss.cpp, GaussPyramid.cpp and rotation.cpp of /root/reference compiled unmodified against the OpenCV stand-in of
oracle/ref_shim (whose numerics are the cv2-pinned primitives).  These vectors pin the reference's loops (StructureMatrix,
HarrisCorner, NonMaximumSuppression, NMS2, initialKeypointDetection, FeaturePointLocalization, computeEdgeResponse,
filterKeypoints, SIFT, Rotation gathers) to the oracle the GPU path is compared with.  Where oracle/_ref itself is present
(this container; it also travels to the GPU box as a built file) a few small cases are compared live as well.
"""
import hashlib
import os

import numpy as np
import pytest

from conftest import GOLDEN
from oracle import c_oracle as O
from oracle import ref as R


def md5(a):
    return hashlib.md5(np.ascontiguousarray(a).tobytes()).hexdigest()


@pytest.fixture(scope="module")
def refg():
    with np.load(os.path.join(GOLDEN, "ref_golden.npz")) as z:
        return {k: z[k] for k in z.files}


def _squares(images):
    return {"blox": images["blox"], "chessboard1240": images["chessboard"][:, :1240], "home384": images["home_gray"][:, :384],
            "building600": images["building_gray"][:, 100:700]}


@pytest.mark.parametrize("name", ["blox", "chessboard1240", "home384", "building600"])
def test_harris_matches_reference_code(images, refg, name):
    """Harris_corners.cpp:10-129,158-179 run as written (square images: the only inputs it is defined on)."""
    img = np.ascontiguousarray(_squares(images)[name])
    o = O.harris(img)
    assert [md5(o["response"]), md5(o["r8"]), md5(o["mask3"])] == list(refg[f"harris_{name}_md5"])
    assert np.array_equal(o["corners"], refg[f"harris_{name}_corners"])            # NMS2 marks, row-major
    assert np.array_equal(o["corner_values"], refg[f"harris_{name}_values"])        # the value NMS2 writes (:118)


@pytest.mark.parametrize("name", ["home", "building", "home_odd"])
def test_dog_keypoints_match_reference_code(images, refg, name):
    """GaussPyramid{img, 4, 1.6} + initialKeypointDetection + computeEdgeResponse < 12.1 (Diff_of_Gauss.cpp:254-297, 223-251,
    79-109, 326-336), lists in the reference's emission order."""
    img = {"home": images["home_cvt"], "building": images["building_cvt"], "home_odd": images["home_cvt"][3:140, 5:206]}[name]
    k, cs, _ = O.dog_keypoints(np.ascontiguousarray(img))
    assert np.array_equal(k, refg[f"dog_{name}_edge"])
    assert [[c["contrast"], c["edge"]] for c in cs] == refg[f"dog_{name}_counts"].tolist()
    if name == "home":
        assert refg["dog_home_counts"].tolist() == [[368, 58], [253, 32], [11, 1], [0, 0]]   # SURVEY.md 8(a) golden counts


def test_pyramid_maps_match_reference_constructor(images, refg):
    """createPyramid / GaussVector / Diff_of_Gauss / processGradients (GaussPyramid.cpp:65-200) as orchestrated by the
    reference: which level is blurred from what, with which sigma, which level seeds the next octave."""
    pyr = O.build_pyramid(np.ascontiguousarray(images["home_cvt"]), 4, 1.6, gradients=True)
    g = [md5(pyr["gauss"][o][l]) for o in range(4) for l in range(6)]
    d = [md5(pyr["dog"][o][l]) for o in range(4) for l in range(5)]
    assert g == list(refg["pyr_home_gauss_md5"])
    assert d == list(refg["pyr_home_dog_md5"])
    assert np.array_equal(np.array([pyr["sigma"][o][l] for o in range(4) for l in range(6)]), refg["pyr_home_sigma"])
    assert [md5(pyr["gx"][1][2]), md5(pyr["gy"][1][2])] == list(refg["pyr_home_gx_o1l2_md5"])


def test_localization_matches_reference_code(refg):
    """FeaturePointLocalization (:223-251) on 6000 seeded derivative triples: decision and the new value (:245-246)."""
    L = O.lib()
    for (dx, dy, ds, v), (kept, nv) in zip(refg["loc_in"], refg["loc_out"]):
        z = np.float32(L.vo_localize(int(dx), int(dy), int(ds), int(v)))
        ok = bool(z > np.float32(0.03))
        assert ok == bool(kept), (dx, dy, ds, v)
        if ok:
            assert int(np.float32(z) * np.float32(255.0)) == nv, (dx, dy, ds, v)


def test_descriptors_match_reference_code(images, refg):
    """filterKeypoints' orientation assignment and SIFT (:340-369, 561-693) on home.jpg: oriented keypoints identical;
    descriptors bit-identical wherever the reference's swapped indexing (:549-554) stays inside its matrices."""
    ori, desc, valid, _ = O.dog_features(np.ascontiguousarray(images["home_cvt"]))
    assert np.array_equal(ori, refg["desc_home_oriented"])
    ok = refg["desc_home_oob"] == 0
    assert np.array_equal(valid.astype(bool), ok)
    assert np.array_equal(desc[ok], refg["desc_home_desc"][ok], equal_nan=True)


def test_rotation_matches_reference_code(images, refg):
    """SLAM::Rotation::rotate_mat_CCW / getRotatedWindow (rotation.cpp:66-108)."""
    for th in (0, 30, 45, 90, 200, 333):
        cs = O.cos_sin_of_angle(float(th))
        assert md5(O.rotate_mat_ccw(images["blox"], (128, 128), cs)) == refg[f"rot_ccw_{th}_md5"][0]
        assert np.array_equal(O.rotated_window(images["blox"], (100, 120), 16, cs), refg[f"rot_win_{th}"])


def test_double_crop_matches_reference_code(images, refg):
    """SLAM::Rotation::doubleCrop (rotation.cpp:147-205) compiled from the reference: crops identical, and it throws where the
    first crop leaves the image."""
    assert bool(refg["dcrop_throws"][0])
    for k, (cx, cy, win, ang) in enumerate(refg["dcrop_cases"]):
        crop, _ = O.double_crop(images["blox"], (int(cx), int(cy)), int(win), float(ang))
        assert crop.shape == refg[f"dcrop_{k}"].shape and np.array_equal(crop, refg[f"dcrop_{k}"]), k


@pytest.mark.skipif(not R.available(), reason="oracle/_ref is built only where /root/reference exists")
def test_live_reference_code_on_seeded_inputs():
    """oracle/_ref executed here: seeded square noise + structure (not a bundled image), Harris and DoG."""
    rng = np.random.default_rng(11)
    img = (rng.integers(0, 256, (96, 96)) // 64 * 64).astype(np.uint8)
    img[20:60, 30:70] = 200
    r, o = R.harris(img), O.harris(img)
    assert np.array_equal(r["response"], o["response"]) and np.array_equal(r["mask3"], o["mask3"])
    assert np.array_equal(r["corners"], o["corners"]) and np.array_equal(r["corner_values"], o["corner_values"])
    # StructureMatrix at a border position of the padded maps (:10-29)
    resp_sm = R.structure_matrix(img, 1, 1)
    assert resp_sm[1] == resp_sm[2]
    smooth = np.ascontiguousarray(np.clip(np.add.outer(np.arange(70), np.arange(90)) + rng.integers(0, 40, (70, 90)), 0, 255).astype(np.uint8))
    a, b, c = R.dog_keypoints(smooth, num_octaves=3)
    k, cs, _ = O.dog_keypoints(smooth, num_octaves=3)
    assert np.array_equal(b, k) and [[x["contrast"], x["edge"]] for x in cs] == c.tolist()
