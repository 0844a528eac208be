"""TEST INFRASTRUCTURE ONLY -- ctypes loader for oracle/_ref: the reference's own translation units compiled unmodified
against the OpenCV stand-in (oracle/Makefile target `ref`, oracle/ref_driver.cpp).  The libraries are built in the
container that has /root/reference; they travel to the GPU box as built files.  available() says whether they exist.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_H = os.path.join(_HERE, "_ref", "libvslam_ref_harris.so")
_D = os.path.join(_HERE, "_ref", "libvslam_ref_dog.so")
_libs: dict = {}

u8p = C.POINTER(C.c_uint8)
f32p = C.POINTER(C.c_float)
i32p = C.POINTER(C.c_int32)


def available() -> bool:
    return os.path.exists(_H) and os.path.exists(_D)


def _lib(which: str):
    if which not in _libs:
        L = C.CDLL(_H if which == "harris" else _D)
        if which == "harris":
            L.ref_harris.argtypes = [u8p, C.c_int, f32p, u8p, u8p, f32p]
            L.ref_harris.restype = C.c_long
            L.ref_structure_matrix.argtypes = [u8p, C.c_int, C.c_int, C.c_int, f32p]
            L.ref_structure_matrix.restype = None
        else:
            L.ref_dog_keypoints.argtypes = [u8p, C.c_int, C.c_int, C.c_int, C.c_double, C.c_int, i32p, C.c_long, i32p, C.c_long, C.POINTER(C.c_long)]
            L.ref_pyramid_map.argtypes = [u8p, C.c_int, C.c_int, C.c_int, C.c_double, C.c_int, C.c_int, C.c_int, C.c_void_p, C.POINTER(C.c_int)]
            L.ref_sigma_at.argtypes = [u8p, C.c_int, C.c_int, C.c_int, C.c_double, C.c_int, C.c_int]
            L.ref_sigma_at.restype = C.c_double
            L.ref_localize.argtypes = [C.c_int] * 4 + [C.POINTER(C.c_int)]
            L.ref_describe.argtypes = [u8p, C.c_int, C.c_int, C.c_int, C.c_double, i32p, f32p, i32p, C.c_long]
            L.ref_describe.restype = C.c_long
            L.ref_rotate_mat_ccw.argtypes = [u8p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, u8p]
            L.ref_rotate_mat_ccw.restype = None
            L.ref_rotated_window.argtypes = [u8p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, u8p]
            L.ref_rotated_window.restype = C.c_long
            L.ref_double_crop.argtypes = [u8p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_double, u8p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]
            L.ref_rotate_pt.argtypes = [C.c_int] * 4 + [C.c_float, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]
            L.ref_rotate_pt.restype = None
        _libs[which] = L
    return _libs[which]


def _u8(a):
    return a.ctypes.data_as(u8p)


def harris(img: np.ndarray) -> dict:
    """Harris_corners.cpp main() (:158-179) on a SQUARE image: response, r8, mask3 and the NMS2 plane."""
    img = np.ascontiguousarray(img, np.uint8)
    n = img.shape[0]
    assert img.shape == (n, n), "the reference's HarrisCorner is defined on square images only"
    resp = np.empty((n, n), np.float32); nms2 = np.empty((n, n), np.float32)
    r8 = np.empty((n, n), np.uint8); mask = np.empty((n, n), np.uint8)
    oob = _lib("harris").ref_harris(_u8(img), n, resp.ctypes.data_as(f32p), _u8(r8), _u8(mask), nms2.ctypes.data_as(f32p))
    assert oob == 0, f"{oob} out-of-buffer accesses"
    ij = np.argwhere(nms2 > 0).astype(np.int32)          # zero marks are no-ops (responses are >= 0)
    return {"response": resp, "r8": r8, "mask3": mask, "nms2": nms2, "corners": ij, "corner_values": nms2[nms2 > 0]}


def structure_matrix(img: np.ndarray, i: int, j: int) -> np.ndarray:
    img = np.ascontiguousarray(img, np.uint8)
    m = np.empty(4, np.float32)
    _lib("harris").ref_structure_matrix(_u8(img), img.shape[0], i, j, m.ctypes.data_as(f32p))
    return m


def dog_keypoints(img: np.ndarray, num_octaves: int = 4, sigma0: float = 1.6, window: int = 3):
    """GaussPyramid + initialKeypointDetection + the edge test, per octave -> (after contrast, after edge, counts)."""
    img = np.ascontiguousarray(img, np.uint8)
    h, w = img.shape
    cap = 3 * (2 * h // window + 2) * (2 * w // window + 2) * 2
    a = np.zeros((cap, 6), np.int32); b = np.zeros((cap, 6), np.int32)
    counts = (C.c_long * (2 * num_octaves))()
    rc = _lib("dog").ref_dog_keypoints(_u8(img), h, w, num_octaves, sigma0, window, a.ctypes.data_as(i32p), cap, b.ctypes.data_as(i32p), cap, counts)
    assert rc == 0
    c = np.array(list(counts), np.int64).reshape(num_octaves, 2)
    return a[:c[:, 0].sum()].copy(), b[:c[:, 1].sum()].copy(), c


_KINDS = {"image": 0, "gauss": 1, "dog": 2, "gx": 3, "gy": 4, "mag": 5, "orient": 6}


def pyramid_map(img: np.ndarray, kind: str, octave: int, level: int = 0, num_octaves: int = 4, sigma0: float = 1.6) -> np.ndarray:
    img = np.ascontiguousarray(img, np.uint8)
    h, w = img.shape
    dims = (C.c_int * 2)()
    L = _lib("dog")
    assert L.ref_pyramid_map(_u8(img), h, w, num_octaves, sigma0, _KINDS[kind], octave, level, None, dims) == 0
    out = np.empty((dims[0], dims[1]), np.uint8 if _KINDS[kind] < 3 else np.float32)
    assert L.ref_pyramid_map(_u8(img), h, w, num_octaves, sigma0, _KINDS[kind], octave, level, out.ctypes.data_as(C.c_void_p), dims) == 0
    return out


def sigma_at(img: np.ndarray, octave: int, level: int, num_octaves: int = 4, sigma0: float = 1.6) -> float:
    img = np.ascontiguousarray(img, np.uint8)
    return _lib("dog").ref_sigma_at(_u8(img), img.shape[0], img.shape[1], num_octaves, sigma0, octave, level)


def localize(dx: int, dy: int, ds: int, value: int):
    nv = C.c_int(0)
    kept = _lib("dog").ref_localize(dx, dy, ds, value, C.byref(nv))
    return bool(kept), nv.value


def describe(img: np.ndarray, num_octaves: int = 4, sigma0: float = 1.6, cap: int = 200000):
    """filterKeypoints + SIFT as main() runs them -> (oriented keypoints, descriptors, out-of-buffer reads per keypoint)."""
    img = np.ascontiguousarray(img, np.uint8)
    h, w = img.shape
    ori = np.zeros((cap, 6), np.int32); desc = np.zeros((cap, 128), np.float32); oob = np.zeros(cap, np.int32)
    n = _lib("dog").ref_describe(_u8(img), h, w, num_octaves, sigma0, ori.ctypes.data_as(i32p), desc.ctypes.data_as(f32p), oob.ctypes.data_as(i32p), cap)
    assert 0 <= n <= cap, n
    return ori[:n].copy(), desc[:n].copy(), oob[:n].copy()


def rotate_mat_ccw(img: np.ndarray, center, theta_deg: float) -> np.ndarray:
    img = np.ascontiguousarray(img, np.uint8)
    out = np.empty_like(img)
    _lib("dog").ref_rotate_mat_ccw(_u8(img), img.shape[0], img.shape[1], int(center[0]), int(center[1]), theta_deg, _u8(out))
    return out


def double_crop(img: np.ndarray, center, window: int, angle: float):
    """SLAM::Rotation::doubleCrop (rotation.cpp:147-205) as compiled from the reference; None where it throws"""
    img = np.ascontiguousarray(img, np.uint8)
    cap = 16 * (window + 4) * (window + 4)
    out = np.empty(cap, np.uint8)
    oh, ow = C.c_int(), C.c_int()
    rc = _lib("dog").ref_double_crop(_u8(img), img.shape[0], img.shape[1], int(center[0]), int(center[1]), int(window), float(angle), _u8(out), cap,
                                     C.byref(oh), C.byref(ow))
    if rc:
        return None
    return out[:oh.value * ow.value].reshape(oh.value, ow.value).copy()


def rotated_window(img: np.ndarray, center, window: int, theta_deg: float):
    img = np.ascontiguousarray(img, np.uint8)
    out = np.empty((window, window), np.uint8)
    oob = _lib("dog").ref_rotated_window(_u8(img), img.shape[0], img.shape[1], int(center[0]), int(center[1]), window, theta_deg, _u8(out))
    return out, int(oob)


def rotate_pt(pt, center, theta_deg: float, ccw: bool):
    x = C.c_int(0); y = C.c_int(0)
    _lib("dog").ref_rotate_pt(int(pt[0]), int(pt[1]), int(center[0]), int(center[1]), theta_deg, int(ccw), C.byref(x), C.byref(y))
    return x.value, y.value
