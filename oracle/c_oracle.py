"""TEST INFRASTRUCTURE ONLY -- ctypes loader for oracle/vslam_oracle.c (the plain-C restatement).

Imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs only.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libvslam_oracle.so")
_lib = None

u8p = C.POINTER(C.c_uint8)
f32p = C.POINTER(C.c_float)
i32p = C.POINTER(C.c_int32)


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "vslam_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _SO


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_SO)
        L.vo_calculate_sigma.restype = C.c_double
        L.vo_calculate_sigma.argtypes = [C.c_int, C.c_int, C.c_double, C.c_int]
        L.vo_calculate_num_octaves.argtypes = [C.c_int, C.c_int]
        L.vo_gauss_ksize_8u.argtypes = [C.c_double]
        L.vo_gauss_taps_8u.argtypes = [C.c_double, C.c_int, C.POINTER(C.c_int)]
        L.vo_upsample2x.argtypes = [u8p, C.c_int, C.c_int, u8p]
        L.vo_upsample2x.restype = None
        L.vo_blur_8u.argtypes = [u8p, C.c_int, C.c_int, C.POINTER(C.c_int), C.c_int, u8p]
        L.vo_gaussian_blur_8u.argtypes = [u8p, C.c_int, C.c_int, C.c_double, u8p]
        L.vo_bgr_to_gray.argtypes = [u8p, C.c_int, C.c_int, u8p]
        L.vo_cos_sin_of_angle.argtypes = [C.c_float, C.POINTER(C.c_float), C.POINTER(C.c_float)]; L.vo_cos_sin_of_angle.restype = None
        L.vo_rotate_mat_ccw.argtypes = [u8p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float, u8p]; L.vo_rotate_mat_ccw.restype = None
        L.vo_rotated_window.argtypes = [u8p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float, u8p]; L.vo_rotated_window.restype = None
        f64p = C.POINTER(C.c_double); i32p_ = C.POINTER(C.c_int)
        L.vo_get_rotation_matrix_2d.argtypes = [C.c_float, C.c_float, C.c_double, C.c_double, f64p]; L.vo_get_rotation_matrix_2d.restype = None
        L.vo_warp_affine_u8.argtypes = [u8p, C.c_int, C.c_int, C.c_size_t, f64p, u8p, C.c_int, C.c_int]
        L.vo_affine_point.argtypes = [f64p, C.c_int, C.c_int, i32p_, i32p_]; L.vo_affine_point.restype = None
        L.vo_double_crop.argtypes = [u8p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_double, u8p, C.c_int, i32p_, i32p_, i32p_]
        L.vo_dog.argtypes = [u8p, u8p, C.c_size_t, u8p]
        L.vo_dog.restype = None
        L.vo_half_size.argtypes = [C.c_int, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.vo_half_size.restype = None
        L.vo_downsample_nearest.argtypes = [u8p, C.c_int, C.c_int, u8p]
        L.vo_downsample_nearest.restype = None
        L.vo_sobel1.argtypes = [u8p, C.c_int, C.c_int, f32p, f32p]
        L.vo_sobel1.restype = None
        L.vo_mag_orient.argtypes = [f32p, f32p, C.c_size_t, f32p, f32p]
        L.vo_mag_orient.restype = None
        L.vo_localize.argtypes = [C.c_int] * 4
        L.vo_localize.restype = C.c_float
        L.vo_edge_response.argtypes = [u8p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]
        L.vo_edge_response.restype = C.c_float
        L.vo_detect_octave.argtypes = [C.POINTER(u8p), C.POINTER(u8p), C.c_int, C.c_int, C.c_int, C.c_int,
                                       C.c_float, C.c_float, i32p, C.c_long, C.POINTER(C.c_long)]
        L.vo_detect_octave.restype = C.c_long
        L.vo_pyramid_sizes.argtypes = [C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int)]
        L.vo_pyramid_sizes.restype = None
        L.vo_build_pyramid.argtypes = [u8p, C.c_int, C.c_int, C.c_int, C.c_double, C.POINTER(u8p),
                                       C.POINTER(u8p), C.POINTER(u8p)]
        L.vo_harris_preblur.argtypes = [u8p, C.c_int, C.c_int, u8p]
        L.vo_harris_preblur.restype = None
        L.vo_harris_response.argtypes = [u8p, C.c_int, C.c_int, C.c_float, f32p]
        L.vo_harris_response.restype = None
        L.vo_harris_nms3.argtypes = [f32p, C.c_int, C.c_int, u8p, u8p]
        L.vo_harris_nms3.restype = None
        L.vo_harris_nms2.argtypes = [f32p, C.c_int, C.c_int, C.c_int, i32p, f32p, C.c_long, f32p]
        L.vo_harris_nms2.restype = C.c_long
        L.vo_gauss_ksize_32f.argtypes = [C.c_double]
        L.vo_gauss_taps_32f.argtypes = [C.c_double, C.c_int, f32p]
        L.vo_orientation_histogram.argtypes = [u8p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_double, f32p]
        L.vo_assign_orientations.argtypes = [C.POINTER(u8p), C.c_int, C.c_int, C.POINTER(C.c_double), i32p, C.c_long, i32p, C.c_long]
        L.vo_assign_orientations.restype = C.c_long
        L.vo_sift.argtypes = [C.POINTER(u8p), C.c_int, C.c_int, C.POINTER(C.c_double), i32p, C.c_long, f32p, u8p]
        _lib = L
    return _lib


def _u8(a):
    return a.ctypes.data_as(u8p)


def _f32(a):
    return a.ctypes.data_as(f32p)


def _ptr_array(arrs):
    return (u8p * len(arrs))(*[_u8(a) for a in arrs])


def gauss_taps(sigma: float, n: int | None = None) -> np.ndarray:
    L = lib()
    if n is None:
        n = L.vo_gauss_ksize_8u(sigma)
    t = (C.c_int * n)()
    rc = L.vo_gauss_taps_8u(sigma, n, t)
    if rc:
        raise ValueError(f"vo_gauss_taps_8u rc={rc}")
    return np.array(list(t), np.int32)


def upsample2x(img: np.ndarray) -> np.ndarray:
    img = np.ascontiguousarray(img, np.uint8)
    h, w = img.shape
    out = np.empty((2 * h, 2 * w), np.uint8)
    lib().vo_upsample2x(_u8(img), h, w, _u8(out))
    return out


def cos_sin_of_angle(theta_deg: float):
    c, s = C.c_float(), C.c_float()
    lib().vo_cos_sin_of_angle(theta_deg, C.byref(c), C.byref(s))
    return c.value, s.value


def rotate_mat_ccw(img: np.ndarray, center, cos_sin) -> np.ndarray:
    img = np.ascontiguousarray(img, np.uint8)
    out = np.empty_like(img)
    lib().vo_rotate_mat_ccw(_u8(img), img.shape[0], img.shape[1], int(center[0]), int(center[1]), cos_sin[0], cos_sin[1], _u8(out))
    return out


def rotated_window(img: np.ndarray, center, window: int, cos_sin) -> np.ndarray:
    img = np.ascontiguousarray(img, np.uint8)
    out = np.empty((window, window), np.uint8)
    lib().vo_rotated_window(_u8(img), img.shape[0], img.shape[1], int(center[0]), int(center[1]), window, cos_sin[0], cos_sin[1], _u8(out))
    return out


def get_rotation_matrix_2d(center, angle_deg: float, scale: float = 1.0) -> np.ndarray:
    """cv::getRotationMatrix2D(Point2f(center), angle, scale) -> 2 x 3 float64"""
    m = np.zeros(6, np.float64)
    lib().vo_get_rotation_matrix_2d(float(center[0]), float(center[1]), float(angle_deg), float(scale), m.ctypes.data_as(C.POINTER(C.c_double)))
    return m.reshape(2, 3)


def warp_affine(img: np.ndarray, m: np.ndarray, dsize) -> np.ndarray:
    """cv::warpAffine(img, M, dsize = (width, height)): INTER_LINEAR, BORDER_CONSTANT 0, 8UC1"""
    img = np.ascontiguousarray(img, np.uint8)
    m = np.ascontiguousarray(m, np.float64).reshape(6)
    out = np.empty((int(dsize[1]), int(dsize[0])), np.uint8)
    rc = lib().vo_warp_affine_u8(_u8(img), img.shape[0], img.shape[1], img.shape[1], m.ctypes.data_as(C.POINTER(C.c_double)), _u8(out), out.shape[0], out.shape[1])
    if rc:
        raise ValueError(f"vo_warp_affine_u8 -> {rc}")
    return out


def affine_point(m: np.ndarray, x: int, y: int):
    """Rotation::drawRotated's arithmetic: (int) of M * [x y 1]^T"""
    m = np.ascontiguousarray(m, np.float64).reshape(6)
    xr, yr = C.c_int(), C.c_int()
    lib().vo_affine_point(m.ctypes.data_as(C.POINTER(C.c_double)), int(x), int(y), C.byref(xr), C.byref(yr))
    return xr.value, yr.value


def double_crop(img: np.ndarray, center, window: int, angle: float):
    """Rotation::doubleCrop (rotation.cpp:147-205) -> (cropped image, first ROI (x, y, w, h)); raises where OpenCV would"""
    img = np.ascontiguousarray(img, np.uint8)
    cap = 16 * (window + 4) * (window + 4)
    out = np.empty(cap, np.uint8)
    oh, ow = C.c_int(), C.c_int()
    roi = (C.c_int * 4)()
    rc = lib().vo_double_crop(_u8(img), img.shape[0], img.shape[1], int(center[0]), int(center[1]), int(window), float(angle), _u8(out), cap,
                              C.byref(oh), C.byref(ow), roi)
    if rc:
        raise ValueError(f"vo_double_crop -> {rc} (roi {list(roi)})")
    return out[:oh.value * ow.value].reshape(oh.value, ow.value).copy(), tuple(roi)


def bgr_to_gray(bgr: np.ndarray) -> np.ndarray:
    bgr = np.ascontiguousarray(bgr, np.uint8)
    h, w, _ = bgr.shape
    out = np.empty((h, w), np.uint8)
    assert lib().vo_bgr_to_gray(_u8(bgr), h, w, _u8(out)) == 0
    return out


def gaussian_blur(img: np.ndarray, sigma: float) -> np.ndarray:
    img = np.ascontiguousarray(img, np.uint8)
    h, w = img.shape
    out = np.empty_like(img)
    rc = lib().vo_gaussian_blur_8u(_u8(img), h, w, sigma, _u8(out))
    assert rc == 0
    return out


def pyramid_sizes(h: int, w: int, num_octaves: int):
    s = (C.c_int * (2 * num_octaves))()
    lib().vo_pyramid_sizes(h, w, num_octaves, s)
    return [(s[2 * o], s[2 * o + 1]) for o in range(num_octaves)]


def build_pyramid(img: np.ndarray, num_octaves: int = 4, sigma0: float = 1.6, gradients: bool = False,
                  mag_orient: bool = False) -> dict:
    img = np.ascontiguousarray(img, np.uint8)
    h, w = img.shape
    sizes = pyramid_sizes(h, w, num_octaves)
    base = [np.empty(s, np.uint8) for s in sizes]
    gauss = [[np.empty(s, np.uint8) for _ in range(6)] for s in sizes]
    dog = [[np.empty(s, np.uint8) for _ in range(5)] for s in sizes]
    rc = lib().vo_build_pyramid(_u8(img), h, w, num_octaves, sigma0, _ptr_array(base),
                                _ptr_array([g for o in gauss for g in o]), _ptr_array([d for o in dog for d in o]))
    assert rc == 0
    out = {"base": base, "gauss": gauss, "dog": dog,
           "sigma": [[lib().vo_calculate_sigma(o, l, sigma0, 3) for l in range(6)] for o in range(num_octaves)]}
    if gradients:
        out["gx"], out["gy"], out["mag"], out["orient"] = [], [], [], []
        for o in range(num_octaves):
            gxs, gys, ms, os_ = [], [], [], []
            for g in gauss[o]:
                gx = np.empty(g.shape, np.float32); gy = np.empty(g.shape, np.float32)
                lib().vo_sobel1(_u8(g), g.shape[0], g.shape[1], _f32(gx), _f32(gy))
                gxs.append(gx); gys.append(gy)
                if mag_orient:
                    m = np.empty(g.shape, np.float32); a = np.empty(g.shape, np.float32)
                    lib().vo_mag_orient(_f32(gx), _f32(gy), g.size, _f32(m), _f32(a))
                    ms.append(m); os_.append(a)
            out["gx"].append(gxs); out["gy"].append(gys); out["mag"].append(ms); out["orient"].append(os_)
    return out


def detect_octave(dogs, gauss, octave: int, window: int = 3, contrast_thr: float = 0.03, edge_r: float = 10.0,
                  capacity: int | None = None):
    rows, cols = dogs[0].shape
    if capacity is None:
        capacity = 3 * ((rows + window - 1) // window) * ((cols + window - 1) // window) + 16
    out = np.zeros((capacity, 6), np.int32)
    counts = (C.c_long * 4)()
    n = lib().vo_detect_octave(_ptr_array(dogs), _ptr_array(gauss), rows, cols, octave, window,
                               contrast_thr, edge_r, out.ctypes.data_as(i32p), capacity, counts)
    c = {"tested": counts[0], "candidates": counts[1], "contrast": counts[2], "edge": counts[3]}
    return out[:min(n, capacity)].copy(), c


def dog_keypoints(img: np.ndarray, num_octaves: int = 4, sigma0: float = 1.6, window: int = 3):
    pyr = build_pyramid(img, num_octaves, sigma0)
    ks, cs = [], []
    for o in range(num_octaves):
        k, c = detect_octave(pyr["dog"][o], pyr["gauss"][o], o, window)
        ks.append(k); cs.append(c)
    return np.concatenate(ks, 0), cs, pyr


def harris(img: np.ndarray, k: float = 0.04, nms_window: int = 5) -> dict:
    img = np.ascontiguousarray(img, np.uint8)
    h, w = img.shape
    L = lib()
    blurred = np.empty_like(img)
    L.vo_harris_preblur(_u8(img), h, w, _u8(blurred))
    resp = np.empty((h, w), np.float32)
    L.vo_harris_response(_u8(blurred), h, w, k, _f32(resp))
    r8 = np.empty((h, w), np.uint8); mask = np.empty((h, w), np.uint8)
    L.vo_harris_nms3(_f32(resp), h, w, _u8(r8), _u8(mask))
    cap = h * w
    corners = np.zeros((cap, 2), np.int32); vals = np.zeros(cap, np.float32)
    tmax = C.c_float(0)
    n = L.vo_harris_nms2(_f32(resp), h, w, nms_window, corners.ctypes.data_as(i32p), _f32(vals), cap, C.byref(tmax))
    return {"blurred": blurred, "response": resp, "r8": r8, "mask3": mask, "corners": corners[:n].copy(),
            "corner_values": vals[:n].copy(), "true_max": float(tmax.value)}


# ---- "next" rows: orientation assignment and SIFT descriptors (Diff_of_Gauss.cpp:340-369, 561-693) ----------------
def gauss_taps_32f(sigma: float, n: int | None = None) -> np.ndarray:
    n = lib().vo_gauss_ksize_32f(sigma) if n is None else n
    t = np.empty(n, np.float32)
    assert lib().vo_gauss_taps_32f(sigma, n, _f32(t)) == 0
    return t


def orientation_histogram(g: np.ndarray, y: int, x: int, sigma: float) -> np.ndarray:
    h = np.empty(36, np.float32)
    assert lib().vo_orientation_histogram(_u8(g), g.shape[0], g.shape[1], y, x, sigma, _f32(h)) == 0
    return h


def assign_orientations(gauss, sigmas, kps: np.ndarray) -> np.ndarray:
    """filterKeypoints' second half over one octave's keypoints (n x 6 int32) -> oriented keypoints (m x 6)."""
    kps = np.ascontiguousarray(kps, np.int32).reshape(-1, 6)
    rows, cols = gauss[0].shape
    out = np.zeros((36 * max(len(kps), 1), 6), np.int32)
    sg = (C.c_double * 6)(*[float(v) for v in sigmas])
    m = lib().vo_assign_orientations(_ptr_array(gauss), rows, cols, sg, kps.ctypes.data_as(i32p), len(kps),
                                     out.ctypes.data_as(i32p), len(out))
    assert m >= 0
    return out[:m].copy()


def sift(gauss, sigmas, oriented: np.ndarray):
    """SIFT() over one octave's oriented keypoints -> (m x 128 float32 descriptors, m uint8 valid flags)."""
    oriented = np.ascontiguousarray(oriented, np.int32).reshape(-1, 6)
    rows, cols = gauss[0].shape
    desc = np.zeros((len(oriented), 128), np.float32)
    valid = np.zeros(len(oriented), np.uint8)
    sg = (C.c_double * 6)(*[float(v) for v in sigmas])
    rc = lib().vo_sift(_ptr_array(gauss), rows, cols, sg, oriented.ctypes.data_as(i32p), len(oriented), _f32(desc), _u8(valid))
    assert rc == 0, rc
    return desc, valid


def dog_features(img: np.ndarray, num_octaves: int = 4, sigma0: float = 1.6, window: int = 3):
    """./DoG's per-octave loop (Diff_of_Gauss.cpp:781-790): keypoints -> oriented keypoints -> descriptors."""
    pyr = build_pyramid(img, num_octaves, sigma0)
    ori, des, val = [], [], []
    for o in range(num_octaves):
        k, _ = detect_octave(pyr["dog"][o], pyr["gauss"][o], o, window)
        r = assign_orientations(pyr["gauss"][o], pyr["sigma"][o], k)
        d, v = sift(pyr["gauss"][o], pyr["sigma"][o], r)
        ori.append(r); des.append(d); val.append(v)
    return np.concatenate(ori, 0), np.concatenate(des, 0), np.concatenate(val, 0), pyr
