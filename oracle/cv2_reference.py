"""TEST INFRASTRUCTURE ONLY -- the reference's hot path issued through cv2.

This file is part of the *oracle*: it is imported by tests/, tools/make_golden.py,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs, and by
nothing on the product path (visualslam_b200/ never imports oracle/).

The reference (JacobYoung115/VisualSLAM, KeyPointDetection/) delegates all arithmetic to
OpenCV, a third-party dependency that is not vendored under /root/reference (README.md:31
names "4.7.0-dev, 3.5+ should work"; there is no lock file).  OpenCV C++ headers do not
exist in this image, so the reference cannot be compiled; its Python binding cv2 4.13.0
can be imported and runs the very same imgproc code.  This module therefore

  * makes the reference's OpenCV calls verbatim (same function, same arguments), citing
    the call site, and
  * restates the reference's own C++ loops (which are not OpenCV) in vectorised numpy with
    the exact IEEE rounding sequence of SURVEY.md section 8(a).

Parity status: the reference ships no golden vectors for this path; maps
and keypoints are pinned by executing cv2 here, and the C restatement oracle/vslam_oracle.c
is pinned against this module (tests/test_oracle_vs_cv2.py) and against fixtures generated
by tools/make_golden.py.
"""
from __future__ import annotations

import numpy as np

try:  # cv2 is present in this image and on the GPU box; gate so the import never hides an error
    import cv2
except Exception as exc:  # pragma: no cover
    cv2 = None
    _cv2_error = exc

NUM_LEVELS = 6          # GaussPyramid.hpp:65-66  scaleSamples_ + 3
SCALE_SAMPLES = 3
K = float(np.power(2.0, 1.0 / 3.0))   # GaussPyramid.hpp:69  pow(2.0f, 1.0f/double(3)) in double


def require_cv2():
    if cv2 is None:  # pragma: no cover
        raise RuntimeError(f"cv2 is required for the cv2 reference path: {_cv2_error!r}")
    return cv2


# ----------------------------------------------------------------------------------------------
# decode -- the two ways the reference entry points read an image
# ----------------------------------------------------------------------------------------------
def decode_color_then_gray(path: str) -> np.ndarray:
    """Diff_of_Gauss.cpp:730,739 and tests/GaussPyramid_Test.cpp:78-81: imread COLOR + cvtColor."""
    cv = require_cv2()
    col = cv.imread(path, cv.IMREAD_COLOR)
    return cv.cvtColor(col, cv.COLOR_BGR2GRAY)


def decode_grayscale(path: str) -> np.ndarray:
    """Harris_corners.cpp:149 and tests/rotate_image_test.cpp:25: imread GRAYSCALE."""
    cv = require_cv2()
    return cv.imread(path, cv.IMREAD_GRAYSCALE)


# ----------------------------------------------------------------------------------------------
# GaussPyramid (include/src/GaussPyramid/GaussPyramid.cpp)
# ----------------------------------------------------------------------------------------------
def calculate_sigma(octave: int, level: int, sigma0: float) -> float:
    """GaussPyramid.cpp:156-162: pow(2,octave)*sigma_*pow(k_,level), all double."""
    return float(np.power(2.0, octave) * sigma0 * np.power(K, level))


def calculate_num_octaves(rows: int, cols: int) -> int:
    """GaussPyramid.cpp:150-152."""
    return int(np.floor(np.log2(min(rows, cols)))) - 4


def build_pyramid(img: np.ndarray, num_octaves: int = 4, sigma0: float = 1.6,
                  gradients: bool = True, mag_orient: bool = True) -> dict:
    """GaussPyramid::createPyramid (GaussPyramid.cpp:106-131) through cv2, call for call."""
    cv = require_cv2()
    assert img.dtype == np.uint8 and img.ndim == 2
    out = {"base": [], "sigma": [], "gauss": [], "dog": [], "gx": [], "gy": [], "mag": [], "orient": []}
    base = cv.resize(img, None, fx=2, fy=2, interpolation=cv.INTER_LINEAR)            # :110
    for o in range(num_octaves):                                                        # :112
        sig = [calculate_sigma(o, l, sigma0) for l in range(NUM_LEVELS)]                # :173
        g = [cv.GaussianBlur(base, (0, 0), s, sigmaY=0, borderType=cv.BORDER_DEFAULT) for s in sig]  # :177
        d = [cv.subtract(g[l], g[l - 1]) for l in range(1, NUM_LEVELS)]                # :197 (saturating 8U)
        out["base"].append(base); out["sigma"].append(sig); out["gauss"].append(g); out["dog"].append(d)
        if gradients:                                                                    # :86-98
            gx = [cv.Sobel(m, cv.CV_32F, 1, 0, ksize=1, scale=1, delta=0, borderType=cv.BORDER_DEFAULT) for m in g]
            gy = [cv.Sobel(m, cv.CV_32F, 0, 1, ksize=1, scale=1, delta=0, borderType=cv.BORDER_DEFAULT) for m in g]
            out["gx"].append(gx); out["gy"].append(gy)
            if mag_orient:
                out["mag"].append([cv.magnitude(a, b) for a, b in zip(gx, gy)])
                out["orient"].append([cv.phase(a, b, angleInDegrees=True) for a, b in zip(gx, gy)])
        base = cv.resize(g[SCALE_SAMPLES], None, fx=0.5, fy=0.5, interpolation=cv.INTER_NEAREST)  # :123-126
    return out


def pad_octave(padding: int, images):
    """GaussPyramid::padOctave (GaussPyramid.cpp:133-141)."""
    cv = require_cv2()
    return [cv.copyMakeBorder(m, padding, padding, padding, padding, cv.BORDER_REPLICATE).copy() for m in images]


# ----------------------------------------------------------------------------------------------
# FeaturePointLocalization (Diff_of_Gauss.cpp:223-251) -- two forms
# ----------------------------------------------------------------------------------------------
def localize_literal_cv2(dx: int, dy: int, ds: int, value: int) -> float:
    """One candidate through the very cv2 calls cv::MatExpr lowers to (gemm / invert).  Slow; used by
    tests to pin localize_numpy on every distinct (dx,dy,ds,value) an image produces."""
    cv = require_cv2()
    f = np.float32
    A = np.array([[f(dx) / f(255.0)], [f(dy) / f(255.0)], [f(ds) / f(255.0)]], dtype=np.float32)     # :233
    A_T = cv.transpose(A)                                                                            # :236
    B = cv.gemm(A, A_T, 1.0, None, 0.0)                                                              # :238
    ok, Binv = cv.invert(B)                                                                          # :239 (DECOMP_LU; zeros if singular)
    if Binv is None:
        Binv = np.zeros((3, 3), np.float32)
    Bneg = -Binv
    z = cv.gemm(Bneg, A, 1.0, None, 0.0)                                                             # :240
    q = cv.gemm(A_T, z, 0.5, None, 0.0)                                                              # :241
    return float(f(f(value) / f(255.0)) + q[0, 0])


def localize_numpy(dx, dy, ds, value):
    """Vectorised exact restatement of FeaturePointLocalization's arithmetic.
    Every operation below is a single IEEE-754 operation in the stated precision (numpy never fuses)."""
    f32, f64 = np.float32, np.float64
    with np.errstate(all="ignore"):
        a = [np.asarray(d, dtype=f32) / f32(255.0) for d in (dx, dy, ds)]
        A64 = [x.astype(f64) for x in a]
        B = [[(A64[i] * A64[j]).astype(f32) for j in range(3)] for i in range(3)]
        b = [[B[i][j].astype(f64) for j in range(3)] for i in range(3)]
        # det3 (OpenCV matrix_operations / lapack.cpp), double, left to right
        d = (b[0][0] * (b[1][1] * b[2][2] - b[1][2] * b[2][1])
             - b[0][1] * (b[1][0] * b[2][2] - b[1][2] * b[2][0])
             + b[0][2] * (b[1][0] * b[2][1] - b[1][1] * b[2][0]))
        nz = d != 0.0
        inv = np.where(nz, 1.0 / np.where(nz, d, 1.0), 0.0)
        t = [
            (b[1][1] * b[2][2] - b[1][2] * b[2][1]) * inv,
            (b[0][2] * b[2][1] - b[0][1] * b[2][2]) * inv,
            (b[0][1] * b[1][2] - b[0][2] * b[1][1]) * inv,
            (b[1][2] * b[2][0] - b[1][0] * b[2][2]) * inv,
            (b[0][0] * b[2][2] - b[0][2] * b[2][0]) * inv,
            (b[0][2] * b[1][0] - b[0][0] * b[1][2]) * inv,
            (b[1][0] * b[2][1] - b[1][1] * b[2][0]) * inv,
            (b[0][1] * b[2][0] - b[0][0] * b[2][1]) * inv,
            (b[0][0] * b[1][1] - b[0][1] * b[1][0]) * inv,
        ]
        M = [-(np.where(nz, x, 0.0).astype(f32)) for x in t]      # M = -(B.inv())
        z = []
        for i in range(3):  # small-matrix gemm path: float accumulate
            p0 = (M[3 * i + 0] * a[0]).astype(f32)
            p1 = (M[3 * i + 1] * a[1]).astype(f32)
            p2 = (M[3 * i + 2] * a[2]).astype(f32)
            z.append(((p0 + p1).astype(f32) + p2).astype(f32))
        acc = (A64[0] * z[0].astype(f64) + A64[1] * z[1].astype(f64)) + A64[2] * z[2].astype(f64)
        q = (0.5 * acc).astype(f32)
        v = np.asarray(value, dtype=f32) / f32(255.0)
        return (v + q).astype(f32)


# ----------------------------------------------------------------------------------------------
# initialKeypointDetection + filterKeypoints' edge test (Diff_of_Gauss.cpp:254-297, 79-109, 326-338)
# ----------------------------------------------------------------------------------------------
def edge_response_numpy(gx: np.ndarray, gy: np.ndarray, rows: np.ndarray, cols: np.ndarray, padding: int = 1):
    """computeEdgeResponse (Diff_of_Gauss.cpp:79-109) on unpadded gradients with padded coordinates."""
    f32, f64 = np.float32, np.float64
    ix2 = np.zeros(rows.shape, f32); iy2 = np.zeros(rows.shape, f32); ixy = np.zeros(rows.shape, f32)
    for du in range(-padding, padding):
        for dv in range(-padding, padding):
            # padding 1 stays inside the maps; wider windows leave them in the reference (undefined): repeat the border sample
            uu = np.clip(rows + du, 0, gx.shape[0] - 1); vv = np.clip(cols + dv, 0, gx.shape[1] - 1)
            a = gx[uu, vv]; b = gy[uu, vv]
            ix2 = (ix2 + (a * a).astype(f32)).astype(f32)
            iy2 = (iy2 + (b * b).astype(f32)).astype(f32)
            ixy = (ixy + (a * b).astype(f32)).astype(f32)
    with np.errstate(all="ignore"):
        det = (ix2.astype(f64) * iy2.astype(f64) - ixy.astype(f64) * ixy.astype(f64)).astype(f32)
        tr = (ix2 + iy2).astype(f32)
        return ((tr * tr).astype(f32) / det).astype(f32)


def detect_octave(dogs, gx, gy, octave: int, window: int = 3, contrast_thr: float = 0.03, edge_r: float = 10.0):
    """Hot-path keypoints of one octave, in the reference's emission order.

    Returns (keypoints int32[n,6] = row,col,value,padding,octave,level ; counts dict).
    row/col are PADDED coordinates exactly as SLAM::point carries them (Diff_of_Gauss.cpp:289).
    """
    f32 = np.float32
    pad = (window - 1) // 2
    P = [np.pad(d, pad, mode="edge").astype(np.int32) for d in dogs]       # padOctave(padding, dogs) :260
    R, C = dogs[0].shape
    ii = np.arange(pad, R, window); jj = np.arange(pad, C, window)          # :267-268
    I, J = np.meshgrid(ii, jj, indexing="ij")
    kps = []
    counts = {"tested": 0, "candidates": 0, "contrast": 0, "edge": 0}
    thr_edge = f32((f32(edge_r) + f32(1.0)) * (f32(edge_r) + f32(1.0))) / f32(edge_r)    # :332-333 (float)
    for level in range(1, len(dogs) - 1):                                   # :264
        c = P[level][I, J]
        mn = c.copy(); mx = c.copy()
        for du in range(-pad, pad):                                         # :273 half-open
            for dv in range(-pad, pad):                                     # :274
                for l in (level - 1, level, level + 1):
                    v = P[l][I + du, J + dv]
                    mn = np.minimum(mn, v); mx = np.maximum(mx, v)
        if pad == 0:   # degenerate window: the neighbour list is empty in the reference (UB on min_element); no candidates
            cand = np.zeros_like(c, dtype=bool)
        else:
            cand = (c == mn) | (c == mx)                                    # :287
        counts["tested"] += c.size; counts["candidates"] += int(cand.sum())
        ci = I[cand]; cj = J[cand]; cv_ = c[cand]
        dx = P[level][ci, cj - 1] - P[level][ci, cj + 1]                    # :226
        dy = P[level][ci - 1, cj] - P[level][ci + 1, cj]                    # :227
        ds = P[level - 1][ci, cj] - P[level + 1][ci, cj]                    # :228
        zhat = localize_numpy(dx, dy, ds, cv_)
        keep = zhat > f32(contrast_thr)                                     # :245
        counts["contrast"] += int(keep.sum())
        ci = ci[keep]; cj = cj[keep]
        with np.errstate(all="ignore"):
            newv = (zhat[keep] * f32(255.0)).astype(f32)
        newv = np.trunc(newv).astype(np.int64).astype(np.int32)            # (int)(dog_zhat*255.0f) :246
        resp = edge_response_numpy(gx[level], gy[level], ci, cj, pad)       # :329-331 (unpadded maps, padded coords)
        ek = resp < thr_edge                                                # :336
        counts["edge"] += int(ek.sum())
        n = int(ek.sum())
        if n:
            k = np.empty((n, 6), np.int32)
            k[:, 0] = ci[ek]; k[:, 1] = cj[ek]; k[:, 2] = newv[ek]; k[:, 3] = pad; k[:, 4] = octave; k[:, 5] = level
            kps.append(k)
    kp = np.concatenate(kps, 0) if kps else np.zeros((0, 6), np.int32)
    return kp, counts


def dog_keypoints(img: np.ndarray, num_octaves: int = 4, sigma0: float = 1.6, window: int = 3):
    """./DoG up to the edge test (Diff_of_Gauss.cpp:746-787 / :336): all octaves, emission order."""
    pyr = build_pyramid(img, num_octaves, sigma0, gradients=True, mag_orient=False)
    all_k, all_c = [], []
    for o in range(num_octaves):
        k, c = detect_octave(pyr["dog"][o], pyr["gx"][o], pyr["gy"][o], o, window)
        all_k.append(k); all_c.append(c)
    return np.concatenate(all_k, 0), all_c, pyr


# ----------------------------------------------------------------------------------------------
# Harris (Harris_corners.cpp)
# ----------------------------------------------------------------------------------------------
def harris_opencv_part(img: np.ndarray):
    """Harris_corners.cpp:158-164: 3x3 pre-blur (sigma 0) and the two ksize=1 Sobels."""
    cv = require_cv2()
    blurred = cv.GaussianBlur(img, (3, 3), 0, sigmaY=0, borderType=cv.BORDER_DEFAULT)
    ix = cv.Sobel(blurred, cv.CV_32F, 1, 0, ksize=1, scale=1, delta=0, borderType=cv.BORDER_DEFAULT)
    iy = cv.Sobel(blurred, cv.CV_32F, 0, 1, ksize=1, scale=1, delta=0, borderType=cv.BORDER_DEFAULT)
    return blurred, ix, iy


def harris_response(ix: np.ndarray, iy: np.ndarray, window: int = 3, k: float = 0.04) -> np.ndarray:
    """HarrisCorner + StructureMatrix (Harris_corners.cpp:10-68), intended (row,col) semantics for every
    pixel; identical to the literal code on square images."""
    cv = require_cv2()
    f32, f64 = np.float32, np.float64
    p = (window - 1) // 2
    pix = cv.copyMakeBorder(ix, p, p, p, p, cv.BORDER_REPLICATE)            # :42
    piy = cv.copyMakeBorder(iy, p, p, p, p, cv.BORDER_REPLICATE)            # :43
    H, W = ix.shape
    ix2 = np.zeros((H, W), f32); iy2 = np.zeros((H, W), f32); ixy = np.zeros((H, W), f32)
    for u in range(0, 2 * p + 1):                                           # :16-17 closed window
        for v in range(0, 2 * p + 1):
            a = pix[u:u + H, v:v + W]; b = piy[u:u + H, v:v + W]
            ix2 = (ix2 + (a * a).astype(f32)).astype(f32)
            iy2 = (iy2 + (b * b).astype(f32)).astype(f32)
            ixy = (ixy + (a * b).astype(f32)).astype(f32)
    det = (ix2.astype(f64) * iy2.astype(f64) - ixy.astype(f64) * ixy.astype(f64)).astype(f32)   # :54 det2 in double
    tr = (ix2 + iy2).astype(f32)                                                                  # :55
    resp = (det - (f32(k) * (tr * tr).astype(f32)).astype(f32)).astype(f32)                      # :57
    return np.where(resp > 0, resp, f32(0)).astype(f32)                                           # :60-62


def harris_nms3(resp: np.ndarray, window: int = 3):
    """convertScaleAbs + NonMaximumSuppression (Harris_corners.cpp:176,70-81) verbatim through cv2."""
    cv = require_cv2()
    r8 = cv.convertScaleAbs(resp)
    c = (window - 1) // 2
    ker = np.ones((window, window), np.uint8); ker[c, c] = 0
    dil = cv.dilate(r8, ker)
    return r8, ((r8 > dil).astype(np.uint8) * 255)


def harris_nms2(resp: np.ndarray, window: int = 5):
    """NMS2 (Harris_corners.cpp:83-129): corner set, row-major.
    Returns (corners int32[n,2] (row, col), values float32[n], true_max)."""
    p = (window - 1) // 2
    H, W = resp.shape
    if H <= 2 * p or W <= 2 * p:
        return np.zeros((0, 2), np.int32), np.zeros((0,), np.float32), 0.0
    m = np.zeros((H - 2 * p, W - 2 * p), np.float32)
    for du in range(-p, p):
        for dv in range(-p, p):
            m = np.maximum(m, resp[p + du:H - p + du, p + dv:W - p + dv])
    centre = resp[p:H - p, p:W - p]
    keep = (centre >= m) & (centre > 0)
    rr, cc = np.nonzero(keep)
    corners = np.stack([rr + p, cc + p], 1).astype(np.int32)
    return corners, centre[keep].astype(np.float32), float(m.max()) if m.size else 0.0


def harris(img: np.ndarray):
    """./Harris up to :181 -- returns a dict of every intermediate the tests compare."""
    blurred, ix, iy = harris_opencv_part(img)
    resp = harris_response(ix, iy)
    r8, mask3 = harris_nms3(resp, 3)
    corners, values, true_max = harris_nms2(resp, 5)
    return {"blurred": blurred, "ix": ix, "iy": iy, "response": resp, "r8": r8, "mask3": mask3,
            "corners": corners, "corner_values": values, "true_max": true_max}


# ----------------------------------------------------------------------------------------------
# fixed-point Gaussian taps by impulse response (known-answer vectors for the tap generator)
# ----------------------------------------------------------------------------------------------
def gaussian_ksize_8u(sigma: float) -> int:
    """OpenCV createGaussianKernels for CV_8U: cvRound(sigma*3*2+1)|1 (round-half-even)."""
    return int(np.rint(sigma * 3 * 2 + 1)) | 1


def gaussian_taps_by_impulse(sigma: float, n: int | None = None) -> np.ndarray:
    """Read cv2's 8.8 fixed-point taps as the response of a (n x 1) GaussianBlur to a single 255 pixel
    (valid while every tap < 128)."""
    cv = require_cv2()
    if n is None:
        n = gaussian_ksize_8u(sigma)
    w = 2 * n + 1
    row = np.zeros((1, w), np.uint8); row[0, n] = 255
    # out = (255*tap*256 + 32768)>>16 with a vertical 1-tap kernel of 256  ->  recover tap exactly
    resp = cv.GaussianBlur(row, (n, 1), sigma, sigmaY=sigma, borderType=cv.BORDER_CONSTANT).astype(np.int64)[0]
    taps = resp[n - n // 2:n + n // 2 + 1]
    # invert y = (255*256*t + 32768) >> 16  for small t: y = floor((65280 t + 32768)/65536)
    cand = np.arange(0, 257)
    lut = (65280 * cand + 32768) >> 16
    out = np.empty(n, np.int32)
    for i, y in enumerate(taps):
        m = np.nonzero(lut == y)[0]
        out[i] = m[0] if len(m) == 1 else -1
    return out


# ----------------------------------------------------------------------------------------------
# "next" rows: orientation assignment (Diff_of_Gauss.cpp:340-369) and SIFT (:561-693) through cv2
# ----------------------------------------------------------------------------------------------
def orientation_histograms(mag: np.ndarray, orient: np.ndarray, kps, sigma: float) -> np.ndarray:
    """36-bin histograms of the keypoints (rows of (row, col, ...)) of ONE level.  A cv::Mat ROI keeps its parent, so
    GaussianBlur(windowMag, ...) (:348) filters with the padded map's pixels around the window; cv2 cannot express a
    parented ROI, but filtering the whole padded map and cropping is the same arithmetic per output pixel."""
    cv = require_cv2()
    pm = pad_octave(8, [mag])[0]
    po = pad_octave(8, [orient])[0]
    blurred = cv.GaussianBlur(pm, (0, 0), sigma, sigmaY=0, borderType=cv.BORDER_DEFAULT)
    coeff = np.float32(36) / np.float32(360.0)
    out = np.zeros((len(kps), 36), np.float32)
    for n, k in enumerate(kps):
        y, x = int(k[0]), int(k[1])
        w = blurred[y:y + 16, x:x + 16]
        idx = (po[y:y + 16, x:x + 16] * coeff).astype(np.int32)
        for i in range(16):
            for j in range(16):
                out[n, idx[i, j]] += w[i, j]
    return out


def sift_descriptor(mag: np.ndarray, orient: np.ndarray, y: int, x: int, angle: int, sigma: float):
    """SIFT() for one oriented keypoint of a level; returns (descriptor[128], in_bounds)."""
    cv = require_cv2()
    f = np.float32
    pm = pad_octave(20, [mag])[0]
    po = pad_octave(20, [orient])[0]
    cx, cy = x + 20, y + 20
    ang = f(np.float64(f(angle)) * (np.pi / np.float64(f(180.0))))
    c, s = f(np.cos(np.float64(ang))), f(np.sin(np.float64(ang)))
    pts = []
    for i in range(cy - 8, cy + 9):                     # rotation.cpp:122-127 ('<=': 17 x 17 points)
        for j in range(cx - 8, cx + 9):
            rx, ry = f(j - cx), f(i - cy)
            pts.append((int(f(rx * c) - f(ry * s)) + cx, int(f(rx * s) + f(ry * c)) + cy))
    m = np.zeros((16, 16), np.float32)
    o = np.zeros((16, 16), np.float32)
    ok = True
    fm, fo = pm.reshape(-1), po.reshape(-1)
    for i in range(16):
        for j in range(16):
            px, py = pts[i * 16 + j]                    # :545
            lin = px * pm.shape[1] + py                 # :549-554 at<>(x, y)
            if not (0 <= lin < fm.size):
                ok = False
                lin = min(max(lin, 0), fm.size - 1)
            m[i, j] = fm[lin]; o[i, j] = fo[lin]
    w = cv.GaussianBlur(m, (0, 0), sigma, sigmaY=0, borderType=cv.BORDER_DEFAULT)    # :611
    coeff = np.float32(8) / np.float32(360.0)
    d = np.zeros(128, np.float32)
    idx = (o * coeff).astype(np.int32)
    for cell in range(16):
        r0, c0 = 4 * (cell // 4), 4 * (cell % 4)
        for i in range(r0, r0 + 4):
            for j in range(c0, c0 + 4):
                d[8 * cell + idx[i, j]] += w[i, j]
    with np.errstate(all="ignore"):
        d = d / d.max()
        d = np.where(f(0.2) < d, f(0.2), d)
        d = d / d.max()
    return d.astype(np.float32), ok
