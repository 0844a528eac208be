"""Writes tests/golden/ref_golden.npz: outputs of the REFERENCE'S OWN code (oracle/_ref: Harris_corners.cpp,
Diff_of_Gauss.cpp, GaussPyramid.cpp, rotation.cpp compiled unmodified against oracle/ref_shim) on the bundled images.
Run in the container that has /root/reference:   make -C oracle ref && python tools/make_ref_golden.py
tests/test_oracle_ref.py checks the plain-C oracle against these vectors (and live against oracle/_ref where it exists).
"""
import hashlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref as R  # noqa: E402


def md5(a):
    return hashlib.md5(np.ascontiguousarray(a).tobytes()).hexdigest()


def main():
    assert R.available(), "build oracle/_ref first (make -C oracle ref)"
    z = np.load(os.path.join(ROOT, "tests", "golden", "images.npz"))
    out = {}
    # Harris: square inputs only (HarrisCorner swaps rows / columns, Harris_corners.cpp:40,47-48)
    squares = {"blox": z["blox"], "chessboard1240": z["chessboard"][:, :1240], "home384": z["home_gray"][:, :384],
               "building600": z["building_gray"][:, 100:700]}
    for name, img in squares.items():
        h = R.harris(np.ascontiguousarray(img))
        out[f"harris_{name}_md5"] = np.array([md5(h["response"]), md5(h["r8"]), md5(h["mask3"])])
        out[f"harris_{name}_corners"] = h["corners"]
        out[f"harris_{name}_values"] = h["corner_values"]
    # DoG: the two images the reference's executables read, plus a small odd-sized crop (kernel wider than octave 3)
    dogs = {"home": z["home_cvt"], "building": z["building_cvt"], "home_odd": z["home_cvt"][3:140, 5:206]}
    for name, img in dogs.items():
        a, b, c = R.dog_keypoints(np.ascontiguousarray(img))
        out[f"dog_{name}_contrast"] = a
        out[f"dog_{name}_edge"] = b
        out[f"dog_{name}_counts"] = c
    # maps of the reference's GaussPyramid constructor on home.jpg: every gauss / dog level as md5, sizes, sigmas
    img = np.ascontiguousarray(z["home_cvt"])
    g, d, sg = [], [], []
    for o in range(4):
        for l in range(6):
            g.append(md5(R.pyramid_map(img, "gauss", o, l)))
            sg.append(R.sigma_at(img, o, l))
        for l in range(5):
            d.append(md5(R.pyramid_map(img, "dog", o, l)))
    out["pyr_home_gauss_md5"] = np.array(g); out["pyr_home_dog_md5"] = np.array(d); out["pyr_home_sigma"] = np.array(sg)
    out["pyr_home_gx_o1l2_md5"] = np.array([md5(R.pyramid_map(img, "gx", 1, 2)), md5(R.pyramid_map(img, "gy", 1, 2))])
    # FeaturePointLocalization as written, on seeded derivative triples (a quarter with a zero derivative)
    rng = np.random.default_rng(7)
    q = rng.integers(-255, 256, (6000, 3)).astype(np.int32)
    q[::4, rng.integers(0, 3)] = 0
    v = rng.integers(0, 256, 6000).astype(np.int32)
    res = np.array([R.localize(int(a0), int(a1), int(a2), int(vv)) for (a0, a1, a2), vv in zip(q, v)], np.int32)
    out["loc_in"] = np.concatenate([q, v[:, None]], 1); out["loc_out"] = res
    # the stage after the hot path on home.jpg
    ori, desc, oob = R.describe(img)
    out["desc_home_oriented"] = ori; out["desc_home_oob"] = oob
    out["desc_home_desc"] = np.where((oob == 0)[:, None], desc, 0).astype(np.float32)
    # SLAM::Rotation gathers on blox.jpg
    for th in (0, 30, 45, 90, 200, 333):
        out[f"rot_ccw_{th}_md5"] = np.array([md5(R.rotate_mat_ccw(z["blox"], (128, 128), float(th)))])
        out[f"rot_win_{th}"] = R.rotated_window(z["blox"], (100, 120), 16, float(th))[0]
    # SLAM::Rotation::doubleCrop (rotation.cpp:147-205; warpAffine / getRotationMatrix2D through the stand-in)
    DC = ((128, 128, 100, 45.0), (100, 120, 64, 30.0), (128, 128, 100, 0.0), (90, 150, 41, -17.5), (128, 128, 50, 90.0), (140, 100, 80, 200.0),
          (128, 128, 16, 333.0), (60, 200, 33, 12.25))
    out["dcrop_cases"] = np.array(DC, np.float64)
    for k, (cx, cy, win, ang) in enumerate(DC):
        out[f"dcrop_{k}"] = R.double_crop(z["blox"], (cx, cy), win, ang)
    out["dcrop_throws"] = np.array([R.double_crop(z["blox"], (10, 10), 100, 45.0) is None])
    path = os.path.join(ROOT, "tests", "golden", "ref_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes;", {k: (v.shape if hasattr(v, "shape") else v) for k, v in list(out.items())[:6]})


if __name__ == "__main__":
    main()
