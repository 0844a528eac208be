#!/usr/bin/env python
"""Fixture for the BGR -> grey ingest step (Diff_of_Gauss.cpp:730,739): tests/golden/home_bgr.npz holds home.jpg decoded
with cv2.imread(IMREAD_COLOR) (the decoder is not on the GPU path; its output is the fixture), so that
cvtColor(home_bgr) == images.npz['home_cvt'] pins the oracle's and the kernel's arithmetic.  Also checks the 15-bit
formula against cv2 over all 2^24 (B, G, R) triples.  Run in the build container (needs cv2 and /root/reference)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import cv2  # noqa: E402

IMG = "/root/reference/KeyPointDetection/images/home.jpg"
bgr = cv2.imread(IMG, cv2.IMREAD_COLOR)
with np.load(os.path.join(ROOT, "tests", "golden", "images.npz")) as z:
    assert np.array_equal(cv2.cvtColor(bgr, cv2.COLOR_BGR2GRAY), z["home_cvt"])
B, G, R = np.meshgrid(np.arange(256), np.arange(256), np.arange(256), indexing="ij")
full = np.stack([B, G, R], -1).astype(np.uint8).reshape(256, -1, 3)
want = cv2.cvtColor(full, cv2.COLOR_BGR2GRAY)
got = ((full[..., 0].astype(np.int64) * 3735 + full[..., 1].astype(np.int64) * 19235 + full[..., 2].astype(np.int64) * 9798 + 16384) >> 15)
assert np.array_equal(want, got.astype(np.uint8)), "15-bit BGR2GRAY formula differs from cv2"
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "home_bgr.npz"), home_bgr=bgr)
print("home_bgr", bgr.shape, "all 2^24 triples agree with cv2", cv2.__version__)
