#!/usr/bin/env python
"""Generate tests/golden/descriptors.npz: the stage after the hot path in ./DoG (orientation assignment,
Diff_of_Gauss.cpp:340-369, and SIFT descriptors, :561-693) run through the reference's OpenCV calls (cv2) on the
bundled images (tests/golden/images.npz, decoded by tools/make_golden.py).

Run in the build container (needs cv2):   python tools/make_golden_descriptors.py
Per image stem (home, building):
    <stem>_hist      n x 36 float32   orientation histograms of the hot-path keypoints (tests/golden/keypoints.npz order)
    <stem>_oriented  m x 6 int32      {row, col, angle, 0, octave, level} per bin above 0.8 of the peak
    <stem>_desc      m x 128 float32  SIFT descriptors
    <stem>_valid     m uint8          0 where rotateImageSection's swapped indexing leaves the padded map (undefined
                                      in the reference; the stored descriptor is then not compared)
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import cv2_reference as R  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")


def main():
    with np.load(os.path.join(OUT, "images.npz")) as z:
        images = {k: z[k] for k in z.files}
    store = {}
    for stem in ("home", "building"):
        img = images[stem] if stem in images else images[stem + "_cvt"]
        kp, _, _ = R.dog_keypoints(img)
        kp = np.asarray(kp, np.int32).reshape(-1, 6)
        pyr = R.build_pyramid(img, 4, 1.6, gradients=True, mag_orient=True)
        hist = np.zeros((len(kp), 36), np.float32)
        for o in range(4):
            for l in range(6):
                sel = np.nonzero((kp[:, 4] == o) & (kp[:, 5] == l))[0]
                if len(sel):
                    hist[sel] = R.orientation_histograms(pyr["mag"][o][l], pyr["orient"][o][l], kp[sel],
                                                         1.5 * R.calculate_sigma(o, l, 1.6))
        oriented = []
        for k, h in zip(kp, hist):
            thr = h.max() * np.float32(0.8)
            for b in np.nonzero(h > thr)[0]:
                oriented.append([k[0], k[1], int(b) * 10, 0, k[4], k[5]])
        oriented = np.asarray(oriented, np.int32).reshape(-1, 6)
        desc = np.zeros((len(oriented), 128), np.float32)
        valid = np.zeros(len(oriented), np.uint8)
        for n, (y, x, a, _, o, l) in enumerate(oriented.tolist()):
            desc[n], ok = R.sift_descriptor(pyr["mag"][o][l], pyr["orient"][o][l], y, x, a, 1.5 * R.calculate_sigma(o, l, 1.6))
            valid[n] = ok
        store.update({f"{stem}_hist": hist, f"{stem}_oriented": oriented, f"{stem}_desc": desc, f"{stem}_valid": valid})
        print(stem, img.shape, "keypoints", len(kp), "oriented", len(oriented), "valid", int(valid.sum()))
    np.savez_compressed(os.path.join(OUT, "descriptors.npz"), **store)


if __name__ == "__main__":
    main()
