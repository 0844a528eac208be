"""Synthetic frame batches for the bench and the large-size parity tests.

8 unique frames per resolution, cycled to fill a batch, seeded with default_rng(1234 + frame_id):
  ids 0-3  natural-tile: a 600 x 868 grey photograph (visualslam_b200/data/synth_tile.npz; the package's own copy, so that the
           bench does not reach into tests/) mirror-tiled 2x2, cropped,
           circularly shifted by a seeded offset (realistic DoG sparsity / keypoint density)
  ids 4-6  procedural: random filled rectangles and discs over a linear ramp + N(0,2) noise
  id  7    uniform noise U{0..255}
"""
from __future__ import annotations

import os

import numpy as np

_TILE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data", "synth_tile.npz")
_tile_cache = {}


def _natural_tile():
    if "t" not in _tile_cache:
        with np.load(_TILE) as z:
            b = z["building_cvt"]
        top = np.concatenate([b, b[:, ::-1]], 1)
        _tile_cache["t"] = np.concatenate([top, top[::-1]], 0)
    return _tile_cache["t"]


def frame(frame_id: int, h: int, w: int) -> np.ndarray:
    rng = np.random.default_rng(1234 + frame_id)
    kind = frame_id % 8
    if kind < 4:
        t = _natural_tile()
        th, tw = t.shape
        oy, ox = int(rng.integers(0, th)), int(rng.integers(0, tw))
        ys = (np.arange(h) + oy) % th
        xs = (np.arange(w) + ox) % tw
        return np.ascontiguousarray(t[np.ix_(ys, xs)])
    if kind < 7:
        img = np.linspace(30, 200, w, dtype=np.float32)[None, :].repeat(h, 0)
        n = max(4, int(200 * h * w / 1e6))
        yy, xx = None, None
        for _ in range(n):
            g = float(rng.integers(0, 256))
            cy, cx = int(rng.integers(0, h)), int(rng.integers(0, w))
            s = int(rng.integers(4, 60))
            y0, y1, x0, x1 = max(cy - s, 0), min(cy + s, h), max(cx - s, 0), min(cx + s, w)
            if rng.random() < 0.5:
                img[y0:y1, x0:x1] = g
            else:
                yy, xx = np.ogrid[y0:y1, x0:x1]
                m = (yy - cy) ** 2 + (xx - cx) ** 2 <= s * s
                img[y0:y1, x0:x1][m] = g
        img += rng.normal(0, 2, img.shape).astype(np.float32)
        return np.clip(np.rint(img), 0, 255).astype(np.uint8)
    return rng.integers(0, 256, (h, w), dtype=np.uint8)


def batch(n: int, h: int, w: int, unique: int = 8) -> np.ndarray:
    u = [frame(i, h, w) for i in range(min(unique, n))]
    return np.stack([u[i % len(u)] for i in range(n)], 0)
