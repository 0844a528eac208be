#!/usr/bin/env python
"""Golden vectors for Rotation::doubleCrop's OpenCV calls (rotation.cpp:152,185,198), generated with cv2 in this container:
cv2.getRotationMatrix2D matrices (bit patterns), cv2.warpAffine outputs (INTER_LINEAR, BORDER_CONSTANT 0) on the bundled blox
image and seeded noise, and the whole doubleCrop sequence re-enacted with cv2 calls (gemm for drawRotated's product).
Writes tests/golden/warp_affine.npz.  The C oracle (vo_get_rotation_matrix_2d, vo_warp_affine_u8, vo_affine_point,
vo_double_crop) is held to these by tests/test_oracle_golden.py."""
import hashlib, os
import cv2
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def draw_rotated(pt, m):                      # rotation.cpp:133-145 with cv2's own gemm
    pos = np.array([[pt[0]], [pt[1]], [1.0]], np.float64)
    r = cv2.gemm(m, pos, 1.0, None, 0.0)
    return int(r[0, 0]), int(r[1, 0])         # (int) truncates toward zero, like Python's int()


def double_crop_cv2(src, center, window, angle):
    pad = window // 2
    m = cv2.getRotationMatrix2D((float(center[0]), float(center[1])), angle, 1.0)
    pts = [draw_rotated(p, m) for p in ((center[0] - pad, center[1] - pad), (center[0] + pad, center[1] - pad),
                                        (center[0] - pad, center[1] + pad), (center[0] + pad, center[1] + pad))]
    xs, ys = [p[0] for p in pts], [p[1] for p in pts]
    x0, y0, w, h = min(xs), min(ys), max(xs) - min(xs), max(ys) - min(ys)
    roi = src[y0:y0 + h, x0:x0 + w]
    assert roi.shape == (h, w) and x0 >= 0 and y0 >= 0
    m2 = cv2.getRotationMatrix2D((float(w // 2), float(h // 2)), angle, 1.0)
    p1, p4 = draw_rotated((w // 2, 0), m2), draw_rotated((w // 2, h), m2)
    rot = cv2.warpAffine(np.ascontiguousarray(roi), m2, (w, h))
    cx0, cy0 = min(p1[0], p4[0]), min(p1[1], p4[1])
    cw, ch = max(p1[0], p4[0]) - cx0, max(p1[1], p4[1]) - cy0
    assert cx0 >= 0 and cy0 >= 0 and cx0 + cw <= w and cy0 + ch <= h
    return rot[cy0:cy0 + ch, cx0:cx0 + cw].copy(), (x0, y0, w, h)


def main():
    z = np.load(os.path.join(ROOT, "tests", "golden", "images.npz"))
    blox = z["blox"]
    out = {}
    rng = np.random.default_rng(77)
    cases = []
    for i in range(24):
        h, w = int(rng.integers(3, 160)), int(rng.integers(3, 160))
        ang = float(rng.uniform(-360, 360)) if i % 3 else float(rng.choice([0.0, 45.0, 90.0, -90.0, 180.0, 30.0]))
        c = (float(rng.integers(0, w)), float(rng.integers(0, h))) if i % 2 else (float(w // 2), float(h // 2))
        dh, dw = (h, w) if i % 4 else (int(rng.integers(1, 200)), int(rng.integers(1, 200)))
        scale = 1.0 if i % 5 else float(rng.uniform(0.5, 2.0))
        cases.append((h, w, ang, c[0], c[1], dh, dw, scale, 1000 + i))
    out["warp_cases"] = np.array(cases, np.float64)
    for i, (h, w, ang, cx, cy, dh, dw, scale, seed) in enumerate(cases):
        img = np.random.default_rng(int(seed)).integers(0, 256, (int(h), int(w)), dtype=np.uint8)
        m = cv2.getRotationMatrix2D((cx, cy), ang, scale)
        out[f"warp_m_{i}"] = m
        out[f"warp_out_{i}"] = cv2.warpAffine(img, m, (int(dw), int(dh)))
    dc = []
    for (cx, cy, window, ang) in ((128, 128, 100, 45.0), (100, 120, 64, 30.0), (128, 128, 100, 0.0), (90, 150, 41, -17.5), (128, 128, 50, 90.0),
                                  (140, 100, 80, 200.0), (128, 128, 16, 333.0)):
        crop, roi = double_crop_cv2(blox, (cx, cy), window, ang)
        k = len(dc)
        dc.append((cx, cy, window, ang) + roi)
        out[f"dcrop_{k}"] = crop
    out["dcrop_cases"] = np.array(dc, np.float64)
    path = os.path.join(ROOT, "tests", "golden", "warp_affine.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes;", len(cases), "warps,", len(dc), "double crops", [out[f"dcrop_{k}"].shape for k in range(len(dc))])


if __name__ == "__main__":
    main()
