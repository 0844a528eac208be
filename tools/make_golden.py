#!/usr/bin/env python
"""Generate tests/golden/* from the reference's own inputs, run through the reference's OpenCV calls (cv2).

Run in the build container only (needs /root/reference and cv2):
    python tools/make_golden.py
Writes
    tests/golden/images.npz      the four bundled images decoded exactly as each reference entry point decodes
                                 them (Diff_of_Gauss.cpp:730,739 / GaussPyramid_Test.cpp:78-81: COLOR+cvtColor;
                                 Harris_corners.cpp:149 / rotate_image_test.cpp:25: IMREAD_GRAYSCALE)
    tests/golden/golden.json     md5 of every pyramid map, per-octave detection counts, Harris md5s / counts,
                                 the 8.8 fixed-point tap tables of the 24 default sigmas + a sweep
    tests/golden/keypoints.npz   hot-path keypoint lists (row,col,value,padding,octave,level) and Harris corners
The GPU box has no /root/reference: tests read only these files.
"""
import hashlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import cv2_reference as R  # noqa: E402

IMG_DIR = "/root/reference/KeyPointDetection/images/"
OUT = os.path.join(ROOT, "tests", "golden")


def md5(a):
    return hashlib.md5(np.ascontiguousarray(a).tobytes()).hexdigest()


def main():
    import cv2
    os.makedirs(OUT, exist_ok=True)
    names = ["blox.jpg", "home.jpg", "building.jpg", "chessboard.png"]
    images = {}
    for n in names:
        stem = n.split(".")[0]
        images[stem + "_cvt"] = R.decode_color_then_gray(IMG_DIR + n)
        images[stem + "_gray"] = R.decode_grayscale(IMG_DIR + n)
    # drop duplicates (PNG and the greyscale JPEG decode identically both ways)
    for stem in [n.split(".")[0] for n in names]:
        if np.array_equal(images[stem + "_cvt"], images[stem + "_gray"]):
            images[stem] = images.pop(stem + "_cvt")
            images.pop(stem + "_gray")
    np.savez_compressed(os.path.join(OUT, "images.npz"), **images)

    def get(stem, mode):
        return images[stem] if stem in images else images[stem + "_" + mode]

    gold = {"cv2_version": cv2.__version__, "pyramid": {}, "harris": {}, "taps": {}, "image_md5": {}}
    for k, v in images.items():
        gold["image_md5"][k] = {"shape": list(v.shape), "md5": md5(v)}
    kp_store = {}
    for stem in ["home", "building", "blox"]:
        img = get(stem, "cvt")
        kp, counts, pyr = R.dog_keypoints(img)
        full = R.build_pyramid(img, 4, 1.6, gradients=True, mag_orient=True)
        gold["pyramid"][stem] = {
            "num_octaves": 4, "sigma0": 1.6,
            "sizes": [list(b.shape) for b in pyr["base"]],
            "sigma": pyr["sigma"],
            "base": [md5(b) for b in pyr["base"]],
            "gauss": [[md5(g) for g in o] for o in pyr["gauss"]],
            "dog": [[md5(g) for g in o] for o in pyr["dog"]],
            "gx": [[md5(g) for g in o] for o in full["gx"]],
            "gy": [[md5(g) for g in o] for o in full["gy"]],
            "counts": counts, "n_keypoints": int(len(kp)), "keypoints_md5": md5(kp),
        }
        kp_store["dog_" + stem] = kp
        # magnitude / phase are tolerance-compared: keep octave 3 (small) as arrays
        kp_store["mag_o3l1_" + stem] = full["mag"][3][1]
        kp_store["orient_o3l1_" + stem] = full["orient"][3][1]
    for stem in ["blox", "chessboard", "building", "home"]:
        img = get(stem, "gray")
        h = R.harris(img)
        gold["harris"][stem] = {
            "shape": list(img.shape), "blurred": md5(h["blurred"]), "ix": md5(h["ix"]), "iy": md5(h["iy"]),
            "response": md5(h["response"]), "r8": md5(h["r8"]), "mask3": md5(h["mask3"]),
            "mask3_count": int((h["mask3"] > 0).sum()), "n_corners": int(len(h["corners"])),
            "corners_md5": md5(h["corners"]), "true_max": h["true_max"],
            "positive_frac": float((h["response"] > 0).mean()),
        }
        kp_store["harris_" + stem] = h["corners"]
        kp_store["harris_val_" + stem] = h["corner_values"]
    # fixed-point tap tables read from cv2 by impulse response
    for o in range(4):
        for l in range(6):
            s = R.calculate_sigma(o, l, 1.6)
            gold["taps"][f"o{o}l{l}"] = {"sigma": s, "n": R.gaussian_ksize_8u(s),
                                         "taps": R.gaussian_taps_by_impulse(s).tolist()}
    rng = np.random.default_rng(7)
    sweep = []
    for s in np.concatenate([rng.uniform(0.35, 12.0, 40), rng.uniform(12.0, 45.0, 10)]):
        t = R.gaussian_taps_by_impulse(float(s))
        if (t >= 0).all():
            sweep.append({"sigma": float(s), "n": int(len(t)), "taps": t.tolist()})
    gold["taps_sweep"] = sweep
    # localisation known answers straight from cv2.gemm / cv2.invert (literal MatExpr lowering)
    loc = []
    for _ in range(400):
        dx, dy, ds = (int(v) for v in rng.integers(-40, 41, 3))
        val = int(rng.integers(0, 60))
        loc.append([dx, dy, ds, val, float(np.float32(R.localize_literal_cv2(dx, dy, ds, val)))])
    for dx, dy, ds, val in [(0, 0, 0, 7), (0, 0, 0, 8), (0, 3, 4, 8), (255, -255, 255, 255), (1, 1, 1, 1), (-1, 2, -3, 20)]:
        loc.append([dx, dy, ds, val, float(np.float32(R.localize_literal_cv2(dx, dy, ds, val)))])
    gold["localize"] = loc
    with open(os.path.join(OUT, "golden.json"), "w") as f:
        json.dump(gold, f, indent=1)
    np.savez_compressed(os.path.join(OUT, "keypoints.npz"), **kp_store)
    for fn in os.listdir(OUT):
        print(fn, os.path.getsize(os.path.join(OUT, fn)))


if __name__ == "__main__":
    main()
