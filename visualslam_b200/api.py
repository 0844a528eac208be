"""ctypes binding of libvslam_b200.so (include/vslam.h) -- the same calls a C++ / cgo / JNI host would make.

Host-side mirror of the reference's surface for Python callers (tests, bench):
    Context.build_pyramid(img)   <-> GaussPyramid(img, numOctaves, sigma)            GaussPyramid.hpp:17
    Pyramid.octave_blur(o) ...   <-> GaussPyramid::octaveBlur / octaveDiff / ...     GaussPyramid.hpp:29-36
    Context.dog_detect(pyr)      <-> initialKeypointDetection + contrast + edge      Diff_of_Gauss.cpp:254,223,79,326
    Context.dog_keypoints(imgs)  <-> ./DoG up to Diff_of_Gauss.cpp:336 (throughput path)
    Context.harris(imgs)         <-> ./Harris up to Harris_corners.cpp:179

There is no CPU fallback: if the shared library is missing or no CUDA device is present, calls raise.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HOST, DEVICE = 0, 1
MAP_BASE, MAP_GAUSS, MAP_DOG, MAP_GRADX, MAP_GRADY, MAP_MAG, MAP_ORIENT = range(7)
OK, E_INVALID, E_CUDA, E_CAPACITY, E_NOMEM = 0, -1, -2, -3, -4

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libvslam_b200.so")

KEYPOINT_DTYPE = np.dtype([("row", "<i4"), ("col", "<i4"), ("value", "<i4"), ("padding", "<i4"),
                           ("octave", "<i4"), ("level", "<i4")])
CORNER_DTYPE = np.dtype([("row", "<i4"), ("col", "<i4"), ("value", "<f4")])
CORNER8_DTYPE = np.dtype([("row", "<u2"), ("col", "<u2"), ("value", "<f4")])      # vslam_corner8 (VSLAM_COMPACT)
COMPACT = 0x100
DEFER = 0x200


class PyramidParams(C.Structure):
    _fields_ = [("num_octaves", C.c_int32), ("scale_samples", C.c_int32), ("sigma0", C.c_double),
                ("materialise", C.c_uint32)]


class DogParams(C.Structure):
    _fields_ = [("window", C.c_int32), ("contrast_thr", C.c_float), ("edge_r", C.c_float)]


class HarrisParams(C.Structure):
    _fields_ = [("k", C.c_float), ("nms2_window", C.c_int32)]


class VslamError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"vslam error {code}: {msg}")
        self.code = code


# every symbol include/vslam.h declares (tests check that the library exports all of them)
EXPORTS = [
    "vslam_ctx_create", "vslam_ctx_create_on_stream", "vslam_ctx_destroy", "vslam_last_error", "vslam_sync",
    "vslam_version", "vslam_launch_count", "vslam_calculate_sigma", "vslam_calculate_num_octaves",
    "vslam_gauss_ksize", "vslam_gauss_taps", "vslam_pyramid_build", "vslam_pyramid_rebuild", "vslam_pyramid_destroy",
    "vslam_pyramid_num_octaves", "vslam_pyramid_num_frames", "vslam_pyramid_octave_size", "vslam_pyramid_sigma",
    "vslam_pyramid_get", "vslam_pyramid_device_map", "vslam_dog_detect", "vslam_dog_keypoints", "vslam_harris", "vslam_dog_harris", "vslam_finish", "vslam_jpeg_info", "vslam_decode_jpeg_gray",
    "vslam_profile", "vslam_profile_read", "vslam_upload_frames",
    # per-function entry points used by the C++17 host layer (visualslam_b200/host)
    "vslam_bgr_to_gray", "vslam_rotate_mat_ccw", "vslam_rotated_window", "vslam_get_rotation_matrix_2d", "vslam_warp_affine", "vslam_gaussian_blur3", "vslam_sobel1", "vslam_harris_response", "vslam_convert_scale_abs", "vslam_nms3", "vslam_nms2",
    "vslam_dog_detect_octave", "vslam_edge_responses", "vslam_localize", "vslam_edge_from_windows",
    # the stage after the hot path: orientation assignment, SIFT descriptors, featureDescriptors.dat
    "vslam_gauss_ksize_f32", "vslam_gauss_taps_f32", "vslam_assign_orientations", "vslam_sift_descriptors",
    "vslam_write_descriptor_file", "vslam_describe_keypoints",
    # allocation-free variant (static device scratch); no counterpart in the reference
    "vslam_harris_response_static",
]

_lib = None


def load():
    """Load the shared library (raises if it has not been built: no fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise VslamError(E_CUDA, f"{LIB_PATH} is missing -- run `python -m visualslam_b200.build`; there is no CPU fallback")
    L = C.CDLL(LIB_PATH)
    vp, i, sz = C.c_void_p, C.c_int, C.c_size_t
    L.vslam_ctx_create.argtypes = [i, C.POINTER(vp)]
    L.vslam_ctx_create_on_stream.argtypes = [i, vp, C.POINTER(vp)]
    L.vslam_ctx_destroy.argtypes = [vp]; L.vslam_ctx_destroy.restype = None
    L.vslam_last_error.argtypes = [vp]; L.vslam_last_error.restype = C.c_char_p
    L.vslam_sync.argtypes = [vp]
    L.vslam_version.restype = C.c_char_p
    L.vslam_launch_count.argtypes = [vp]; L.vslam_launch_count.restype = C.c_longlong
    L.vslam_profile.argtypes = [vp, i]
    L.vslam_profile_read.argtypes = [vp, i, C.POINTER(C.c_double), C.POINTER(C.c_longlong)]
    L.vslam_calculate_sigma.argtypes = [i, i, C.c_double, i]; L.vslam_calculate_sigma.restype = C.c_double
    L.vslam_calculate_num_octaves.argtypes = [i, i]
    L.vslam_gauss_ksize.argtypes = [C.c_double]
    L.vslam_gauss_taps.argtypes = [C.c_double, i, C.POINTER(C.c_int32)]
    L.vslam_bgr_to_gray.argtypes = [vp, vp, i, i, i, sz, sz, vp, sz, sz]
    L.vslam_rotate_mat_ccw.argtypes = [vp, vp, i, i, sz, i, i, C.c_float, C.c_float, vp, sz]
    L.vslam_rotated_window.argtypes = [vp, vp, i, i, sz, i, i, i, C.c_float, C.c_float, vp, sz]
    L.vslam_get_rotation_matrix_2d.argtypes = [C.c_float, C.c_float, C.c_double, C.c_double, C.POINTER(C.c_double)]
    L.vslam_warp_affine.argtypes = [vp, vp, i, i, sz, C.POINTER(C.c_double), vp, i, i, sz]
    L.vslam_pyramid_build.argtypes = [vp, vp, i, i, i, sz, sz, i, C.POINTER(PyramidParams), C.POINTER(vp)]
    L.vslam_pyramid_rebuild.argtypes = [vp, vp, vp, sz, sz, i]
    L.vslam_pyramid_destroy.argtypes = [vp]; L.vslam_pyramid_destroy.restype = None
    L.vslam_pyramid_num_octaves.argtypes = [vp]
    L.vslam_pyramid_num_frames.argtypes = [vp]
    L.vslam_pyramid_octave_size.argtypes = [vp, i, C.POINTER(i), C.POINTER(i)]
    L.vslam_pyramid_sigma.argtypes = [vp, i, i, C.POINTER(C.c_double)]
    L.vslam_pyramid_get.argtypes = [vp, vp, i, i, i, i, vp, sz, i]
    L.vslam_pyramid_device_map.argtypes = [vp, i, i, i, i, C.POINTER(sz)]; L.vslam_pyramid_device_map.restype = vp
    L.vslam_dog_detect.argtypes = [vp, vp, C.POINTER(DogParams), vp, i, i, C.POINTER(i), C.POINTER(C.c_longlong)]
    L.vslam_dog_keypoints.argtypes = [vp, vp, i, i, i, sz, sz, i, C.POINTER(PyramidParams), C.POINTER(DogParams),
                                      vp, i, i, C.POINTER(i)]
    L.vslam_harris.argtypes = [vp, vp, i, i, i, sz, sz, i, C.POINTER(HarrisParams), vp, vp, vp, vp, i, i, C.POINTER(i)]
    L.vslam_dog_harris.argtypes = [vp, vp, i, i, i, sz, sz, i, C.POINTER(PyramidParams), C.POINTER(DogParams), C.POINTER(HarrisParams),
                                   vp, i, C.POINTER(i), vp, i, C.POINTER(i), i]
    L.vslam_finish.argtypes = [vp]
    L.vslam_jpeg_info.argtypes = [vp, vp, sz, C.POINTER(i), C.POINTER(i), C.POINTER(i)]
    L.vslam_decode_jpeg_gray.argtypes = [vp, C.POINTER(vp), C.POINTER(sz), i, i, i, vp, sz, sz, i]
    L.vslam_upload_frames.argtypes = [vp, vp, i, i, i, sz, sz, C.POINTER(vp), C.POINTER(sz), C.POINTER(sz)]
    L.vslam_gauss_ksize_f32.argtypes = [C.c_double]
    L.vslam_gauss_taps_f32.argtypes = [C.c_double, i, C.POINTER(C.c_float)]
    L.vslam_assign_orientations.argtypes = [vp, vp, i, i, vp, i, vp, i, C.POINTER(i), vp]
    L.vslam_sift_descriptors.argtypes = [vp, vp, i, i, vp, i, vp, vp]
    L.vslam_write_descriptor_file.argtypes = [C.c_char_p, vp, i]
    L.vslam_describe_keypoints.argtypes = [vp, vp, vp, C.POINTER(i), vp, i, C.POINTER(i), vp, vp]
    L.vslam_harris_response_static.argtypes = [vp, vp, i, i, sz, i, C.c_float, vp, vp]
    _lib = L
    return L


def get_rotation_matrix_2d(center, angle_deg, scale=1.0):
    """cv::getRotationMatrix2D(Point2f(center), angle, scale) -> 2 x 3 float64 (host arithmetic of the library)"""
    m = np.zeros(6, np.float64)
    load().vslam_get_rotation_matrix_2d(float(center[0]), float(center[1]), float(angle_deg), float(scale), m.ctypes.data_as(C.POINTER(C.c_double)))
    return m.reshape(2, 3)


def expand_corners(c8):
    """vslam_corner8 list (CORNER8_DTYPE) -> vslam_corner list (CORNER_DTYPE); the 0xffff overflow marker becomes row = -1."""
    c8 = np.asarray(c8).view(CORNER8_DTYPE).reshape(-1)
    out = np.zeros(len(c8), CORNER_DTYPE)
    out["row"] = c8["row"]; out["col"] = c8["col"]; out["value"] = c8["value"]
    lost = (c8["row"] == 0xffff) & (c8["col"] == 0xffff)
    out["row"][lost] = -1; out["col"][lost] = 0; out["value"][lost] = 0
    return out


def calculate_sigma(octave, level, sigma0=1.6, scale_samples=3):
    return load().vslam_calculate_sigma(octave, level, sigma0, scale_samples)


def gauss_taps_f32(sigma, ksize=None):
    L = load()
    if ksize is None:
        ksize = L.vslam_gauss_ksize_f32(sigma)
    t = np.zeros(ksize, np.float32)
    rc = L.vslam_gauss_taps_f32(sigma, ksize, t.ctypes.data_as(C.POINTER(C.c_float)))
    if rc:
        raise VslamError(rc, "vslam_gauss_taps_f32")
    return t


def write_descriptor_file(path, descriptors):
    d = np.ascontiguousarray(descriptors, np.float32).reshape(-1, 128)
    rc = load().vslam_write_descriptor_file(os.fsencode(path), d.ctypes.data, d.shape[0])
    if rc:
        raise VslamError(rc, "vslam_write_descriptor_file")


def gauss_taps(sigma, ksize=None):
    L = load()
    if ksize is None:
        ksize = L.vslam_gauss_ksize(sigma)
    t = (C.c_int32 * ksize)()
    rc = L.vslam_gauss_taps(sigma, ksize, t)
    if rc:
        raise VslamError(rc, "vslam_gauss_taps")
    return np.array(t[:], np.int32)


def _frames_arg(frames):
    """numpy uint8 [n,h,w] / [h,w] (host) or a torch cuda uint8 tensor (device) -> (ptr, n, h, w, pitch, stride, space, keepalive)."""
    if isinstance(frames, UploadedFrames):
        return frames.ptr, frames.n, frames.h, frames.w, frames.pitch, frames.stride, DEVICE, frames
    if isinstance(frames, np.ndarray):
        a = np.ascontiguousarray(frames, np.uint8)
        if a.ndim == 2:
            a = a[None]
        n, h, w = a.shape
        return a.ctypes.data, n, h, w, w, h * w, HOST, a
    # torch tensor on a CUDA device
    t = frames
    if t.dim() == 2:
        t = t[None]
    t = t.contiguous()
    n, h, w = t.shape
    return t.data_ptr(), n, h, w, w, h * w, DEVICE, t


class UploadedFrames:
    """A batch whose upload was started by Context.upload (vslam_upload_frames): pass it wherever frames are expected."""

    def __init__(self, ptr, n, h, w, pitch, stride, keep):
        self.ptr, self.n, self.h, self.w, self.pitch, self.stride, self.keep = ptr, n, h, w, pitch, stride, keep


class Pyramid:
    """Device-resident GaussPyramid of a batch (GaussPyramid.hpp:14-71)."""

    def __init__(self, ctx, handle):
        self.ctx, self.h = ctx, handle
        L = load()
        self.num_octaves = L.vslam_pyramid_num_octaves(handle)
        self.n_frames = L.vslam_pyramid_num_frames(handle)
        self.num_levels = 6
        self.num_scale_samples = 3

    def close(self):
        if self.h:
            load().vslam_pyramid_destroy(self.h)
            self.h = None

    __del__ = close

    def octave_size(self, octave):
        r, c = C.c_int(), C.c_int()
        self.ctx._chk(load().vslam_pyramid_octave_size(self.h, octave, C.byref(r), C.byref(c)))
        return r.value, c.value

    def sigma_at(self, octave, level):                       # getSigmaAt, GaussPyramid.cpp:187
        s = C.c_double()
        rc = load().vslam_pyramid_sigma(self.h, octave, level, C.byref(s))
        if rc:
            raise IndexError(f"octave {octave} level {level}")  # std::out_of_range in the reference (map::at)
        return s.value

    def get(self, kind, octave, level=0, frame=0):
        r, c = self.octave_size(octave)
        dt = np.uint8 if kind <= MAP_DOG else np.float32
        out = np.empty((r, c), dt)
        self.ctx._chk(load().vslam_pyramid_get(self.ctx.h, self.h, frame, kind, octave, level, out.ctypes.data,
                                               out.strides[0], HOST))
        return out

    def octave_image(self, o, frame=0): return self.get(MAP_BASE, o, 0, frame)
    def octave_blur(self, o, frame=0): return [self.get(MAP_GAUSS, o, l, frame) for l in range(6)]
    def octave_diff(self, o, frame=0): return [self.get(MAP_DOG, o, l, frame) for l in range(5)]
    def octave_grad_x(self, o, frame=0): return [self.get(MAP_GRADX, o, l, frame) for l in range(6)]
    def octave_grad_y(self, o, frame=0): return [self.get(MAP_GRADY, o, l, frame) for l in range(6)]
    def octave_grad_mag(self, o, frame=0): return [self.get(MAP_MAG, o, l, frame) for l in range(6)]
    def octave_grad_orient(self, o, frame=0): return [self.get(MAP_ORIENT, o, l, frame) for l in range(6)]
    def octave_sigma(self, o): return [self.sigma_at(o, l) for l in range(6)]


class Context:
    def __init__(self, device=0, stream=None):
        L = load()
        h = C.c_void_p()
        rc = L.vslam_ctx_create_on_stream(device, stream, C.byref(h)) if stream else L.vslam_ctx_create(device, C.byref(h))
        if rc:
            raise VslamError(rc, L.vslam_last_error(None).decode())
        self.h = h
        self.device = device
        self._own_stream = not stream      # the library's own non-blocking stream: torch's streams are not ordered with it

    def _torch_in(self, *objs):
        """Stream contract (include/vslam.h): work is enqueued on the context's stream.  A context created on the caller's
        stream is ordered with the caller's work; one with its own stream is not, so torch tensors handed to it are made
        ready first (host wait on the tensor's current stream) and device results are complete on return (_torch_out)."""
        if not self._own_stream:
            return False
        dev = False
        for o in objs:
            if o is not None and not isinstance(o, (np.ndarray, UploadedFrames)) and hasattr(o, "data_ptr"):
                import torch
                torch.cuda.current_stream(o.device).synchronize()
                dev = True
        return dev

    def _torch_out(self, dev):
        if dev:
            self.sync()

    def close(self):
        if getattr(self, "h", None):
            load().vslam_ctx_destroy(self.h)
            self.h = None

    __del__ = close

    def _chk(self, rc):
        if rc:
            raise VslamError(rc, load().vslam_last_error(self.h).decode())

    def sync(self):
        self._chk(load().vslam_sync(self.h))

    # names of the kernels behind vslam_kernel_id (include/vslam.h), as they appear in an ncu launch list
    KERNELS = ["upsample2x_k", "reflect_pad_k", "(unused2)", "dog_detect_k", "seg_reduce_k + seg_spine_k + seg_apply_k + seg_gather_k", "harris_fast_k", "grad_k",
               "octave_fused_k", "orient_hist_k + sift_k", "octave_mma_k", "octave_umma_k"]

    def profile(self, enable=True):
        self._chk(load().vslam_profile(self.h, 1 if enable else 0))

    def profile_read(self):
        out = {}
        for kid, name in enumerate(self.KERNELS):
            ms, n = C.c_double(), C.c_longlong()
            self._chk(load().vslam_profile_read(self.h, kid, C.byref(ms), C.byref(n)))
            out[name] = {"ms": ms.value, "launches": n.value}
        return out

    @property
    def launches(self):
        return load().vslam_launch_count(self.h)

    def upload(self, frames):
        """Start the chunked asynchronous upload of host frames (numpy uint8 [n,h,w]; pinned memory for full PCIe speed).
        The returned batch can be passed to dog_keypoints / harris / build_pyramid, which overlap their kernels with
        the rest of the upload; it stays valid until the next upload on this context."""
        a = np.ascontiguousarray(frames, np.uint8)
        if a.ndim == 2:
            a = a[None]
        n, h, w = a.shape
        dptr, dp, ds = C.c_void_p(), C.c_size_t(), C.c_size_t()
        self._chk(load().vslam_upload_frames(self.h, a.ctypes.data, n, h, w, w, h * w, C.byref(dptr), C.byref(dp), C.byref(ds)))
        return UploadedFrames(dptr.value, n, h, w, dp.value, ds.value, a)

    def build_pyramid(self, frames, num_octaves=4, sigma0=1.6, materialise=0):
        self._torch_in(frames)
        ptr, n, h, w, pitch, stride, space, keep = _frames_arg(frames)
        pp = PyramidParams(num_octaves, 3, sigma0, materialise)
        ph = C.c_void_p()
        self._chk(load().vslam_pyramid_build(self.h, ptr, n, h, w, pitch, stride, space, C.byref(pp), C.byref(ph)))
        self.sync()
        return Pyramid(self, ph)

    def rebuild_pyramid(self, pyr, frames):
        """vslam_pyramid_rebuild: the same pyramid object on a new batch of the same shape, in place, no allocation."""
        dev = self._torch_in(frames)
        ptr, n, h, w, pitch, stride, space, keep = _frames_arg(frames)
        assert n == pyr.n_frames
        self._chk(load().vslam_pyramid_rebuild(self.h, pyr.h, ptr, pitch, stride, space))
        self._torch_out(dev)
        return pyr

    def dog_detect(self, pyr, window=3, contrast_thr=0.03, edge_r=10.0, capacity=1 << 20, stage_counts=True):
        dp = DogParams(window, contrast_thr, edge_r)
        out = np.zeros(capacity, KEYPOINT_DTYPE)
        counts = (C.c_int * pyr.n_frames)()
        sc = (C.c_longlong * (pyr.n_frames * pyr.num_octaves * 4))() if stage_counts else None
        rc = load().vslam_dog_detect(self.h, pyr.h, C.byref(dp), out.ctypes.data, capacity, HOST, counts, sc)
        if rc and rc != E_CAPACITY:
            self._chk(rc)
        cl = list(counts)
        res = {"keypoints": out[:min(sum(cl), capacity)], "counts": cl, "overflow": rc == E_CAPACITY}
        if stage_counts:
            res["stage_counts"] = np.array(sc[:], np.int64).reshape(pyr.n_frames, pyr.num_octaves, 4)
        return res

    def dog_keypoints(self, frames, num_octaves=4, sigma0=1.6, window=3, contrast_thr=0.03, edge_r=10.0,
                      capacity=1 << 20, out=None):
        """Throughput path. `out` may be a torch cuda tensor viewed as int32 [capacity,6] to keep results on device."""
        dev = self._torch_in(frames, out)
        ptr, n, h, w, pitch, stride, space, keep = _frames_arg(frames)
        pp = PyramidParams(num_octaves, 3, sigma0, 0)
        dp = DogParams(window, contrast_thr, edge_r)
        counts = (C.c_int * n)()
        if out is None or isinstance(out, np.ndarray):
            # a caller-provided host list (e.g. a pinned buffer viewed as KEYPOINT_DTYPE) avoids a pageable copy
            host = np.zeros(capacity, KEYPOINT_DTYPE) if out is None else out.view(KEYPOINT_DTYPE).reshape(-1)
            capacity = host.shape[0]
            optr, ospace = host.ctypes.data, HOST
        else:
            host = None
            optr, ospace = out.data_ptr(), DEVICE
            capacity = out.shape[0]
        rc = load().vslam_dog_keypoints(self.h, ptr, n, h, w, pitch, stride, space, C.byref(pp), C.byref(dp), optr,
                                        capacity, ospace, counts)
        if rc and rc != E_CAPACITY:
            self._chk(rc)
        self._torch_out(dev)
        cl = list(counts)
        return {"keypoints": host[:min(sum(cl), capacity)] if host is not None else out, "counts": cl,
                "overflow": rc == E_CAPACITY}

    def dog_harris(self, frames, num_octaves=4, sigma0=1.6, contrast_thr=0.03, edge_r=10.0, k=0.04, nms2_window=5, kp_capacity=1 << 20,
                   co_capacity=1 << 22, kp_out=None, co_out=None, compact=False, defer=False):
        """Both detectors over one batch (vslam_dog_harris): the lists of dog_keypoints and harris(want_maps=False), with the corner
        list draining under the DoG kernels.  kp_out / co_out: caller-owned host buffers (numpy, e.g. pinned) or torch cuda tensors
        (both on the device then); compact: the corner list as vslam_corner8.  defer=True (host kp_out and co_out required): the call
        only enqueues; finish() completes the oldest pending call and returns its result dict (at most two calls pending)."""
        dev = self._torch_in(frames, kp_out, co_out)
        ptr, n, h, w, pitch, stride, space, keep = _frames_arg(frames)
        pp = PyramidParams(num_octaves, 3, sigma0, 0); dp = DogParams(3, contrast_thr, edge_r); hp = HarrisParams(k, nms2_window)
        kc, cc = (C.c_int * n)(), (C.c_int * n)()
        cdt = CORNER8_DTYPE if compact else CORNER_DTYPE
        on_dev = kp_out is not None and not isinstance(kp_out, np.ndarray)
        if on_dev:
            assert co_out is not None and not isinstance(co_out, np.ndarray)
            kptr, cptr, kcap, ccap, ospace = kp_out.data_ptr(), co_out.data_ptr(), kp_out.shape[0], co_out.shape[0], DEVICE
            kh = ch = None
        else:
            kh = np.zeros(kp_capacity, KEYPOINT_DTYPE) if kp_out is None else kp_out.view(KEYPOINT_DTYPE).reshape(-1)
            ch = np.zeros(co_capacity, cdt) if co_out is None else co_out.view(cdt).reshape(-1)
            kptr, cptr, kcap, ccap, ospace = kh.ctypes.data, ch.ctypes.data, kh.shape[0], ch.shape[0], HOST
        if defer:
            assert not on_dev and kp_out is not None and co_out is not None
        rc = load().vslam_dog_harris(self.h, ptr, n, h, w, pitch, stride, space, C.byref(pp), C.byref(dp), C.byref(hp), kptr, kcap, kc, cptr, ccap,
                                     cc, ospace | (COMPACT if compact else 0) | (DEFER if defer else 0))
        if rc and rc != E_CAPACITY:
            self._chk(rc)
        if defer:
            if not hasattr(self, "_pending"):
                self._pending = []
            self._pending.append((kh, ch, kc, cc, kcap, ccap, keep))       # keeps the buffers and count arrays alive until finish()
            return None
        self._torch_out(dev)
        kl, cl = list(kc), list(cc)
        return {"keypoints": kh[:min(sum(kl), kcap)] if kh is not None else kp_out, "corners": ch[:min(sum(cl), ccap)] if ch is not None else co_out,
                "kp_counts": kl, "co_counts": cl, "overflow": rc == E_CAPACITY}

    def finish(self):
        """vslam_finish: complete the oldest deferred dog_harris call; returns its result dict (None if nothing is pending)."""
        if not getattr(self, "_pending", None):
            return None
        kh, ch, kc, cc, kcap, ccap, _ = self._pending.pop(0)
        rc = load().vslam_finish(self.h)
        if rc and rc != E_CAPACITY:
            self._chk(rc)
        kl, cl = list(kc), list(cc)
        return {"keypoints": kh[:min(sum(kl), kcap)], "corners": ch[:min(sum(cl), ccap)], "kp_counts": kl, "co_counts": cl, "overflow": rc == E_CAPACITY}

    def assign_orientations(self, pyr, octave, keypoints, frame=0, histograms=False):
        """filterKeypoints' orientation stage (Diff_of_Gauss.cpp:340-369) for one octave's keypoints."""
        kp = np.ascontiguousarray(keypoints).view(KEYPOINT_DTYPE).reshape(-1)
        n = kp.shape[0]
        out = np.zeros(36 * max(n, 1), KEYPOINT_DTYPE)
        hist = np.zeros((n, 36), np.float32) if histograms else None
        cnt = C.c_int(0)
        self._chk(load().vslam_assign_orientations(self.h, pyr.h, frame, octave, kp.ctypes.data, n, out.ctypes.data, out.shape[0],
                                                   C.byref(cnt), hist.ctypes.data if histograms else None))
        return (out[:cnt.value], hist) if histograms else out[:cnt.value]

    def sift_descriptors(self, pyr, octave, oriented, frame=0):
        """SIFT() (Diff_of_Gauss.cpp:561-693) for one octave's oriented keypoints -> (n x 128 float32, n uint8 valid)."""
        kp = np.ascontiguousarray(oriented).view(KEYPOINT_DTYPE).reshape(-1)
        n = kp.shape[0]
        desc = np.zeros((n, 128), np.float32)
        valid = np.zeros(n, np.uint8)
        self._chk(load().vslam_sift_descriptors(self.h, pyr.h, frame, octave, kp.ctypes.data, n, desc.ctypes.data, valid.ctypes.data))
        return desc, valid

    def describe_keypoints(self, pyr, keypoints, counts, capacity=None, out=None):
        """Orientation assignment + SIFT for every frame of a pyramid in five launches (vslam_describe_keypoints).
        out = (oriented [cap] KEYPOINT_DTYPE or int32 [cap, 6], descriptors float32 [cap, 128], valid uint8 [cap]): caller-owned
        result buffers, e.g. views of pinned memory (a 100 MB descriptor block lands in pageable memory at a fraction of the PCIe rate)."""
        kp = np.ascontiguousarray(keypoints).view(KEYPOINT_DTYPE).reshape(-1)
        assert len(counts) == pyr.n_frames and sum(counts) == kp.shape[0]
        if out is not None:
            o, desc, valid = out
            o = o.view(KEYPOINT_DTYPE).reshape(-1)
            capacity = min(o.shape[0], desc.shape[0], valid.shape[0])
            assert desc.dtype == np.float32 and desc.shape[1] == 128 and valid.dtype == np.uint8 and desc.flags.c_contiguous
        else:
            capacity = 36 * max(kp.shape[0], 1) if capacity is None else capacity
            o = np.zeros(capacity, KEYPOINT_DTYPE)
            desc = np.zeros((capacity, 128), np.float32)
            valid = np.zeros(capacity, np.uint8)
        cin = (C.c_int * pyr.n_frames)(*[int(c) for c in counts])
        cout = (C.c_int * pyr.n_frames)()
        rc = load().vslam_describe_keypoints(self.h, pyr.h, kp.ctypes.data, cin, o.ctypes.data, capacity, cout, desc.ctypes.data,
                                             valid.ctypes.data)
        if rc and rc != E_CAPACITY:
            self._chk(rc)
        m = min(sum(cout), capacity)
        return {"oriented": o[:m], "descriptors": desc[:m], "valid": valid[:m], "counts": list(cout), "overflow": rc == E_CAPACITY}

    def bgr_to_gray(self, bgr):
        """cvtColor(COLOR_BGR2GRAY) of (rows, cols, 3) or (n, rows, cols, 3) uint8 frames -> uint8 gray frames."""
        a = np.ascontiguousarray(bgr, np.uint8)
        single = a.ndim == 3
        if single:
            a = a[None]
        if a.ndim != 4 or a.shape[3] != 3:
            raise ValueError("bgr frames must be (rows, cols, 3) or (n, rows, cols, 3)")
        n, r, c, _ = a.shape
        out = np.empty((n, r, c), np.uint8)
        self._chk(load().vslam_bgr_to_gray(self.h, a.ctypes.data, n, r, c, 3 * c, 3 * c * r, out.ctypes.data, c, c * r))
        return out[0] if single else out

    def rotate_mat_ccw(self, img, center, cos_sin):
        """Rotation::rotate_mat_CCW(I, center=(x, y), angles=(cos, sin))."""
        a = np.ascontiguousarray(img, np.uint8)
        out = np.empty_like(a)
        self._chk(load().vslam_rotate_mat_ccw(self.h, a.ctypes.data, a.shape[0], a.shape[1], a.shape[1], int(center[0]), int(center[1]),
                                              float(cos_sin[0]), float(cos_sin[1]), out.ctypes.data, a.shape[1]))
        return out

    def rotated_window(self, img, center, window, cos_sin):
        """Rotation::getRotatedWindow(I, center=(x, y), windowSize, theta) with (cos, sin) of theta."""
        a = np.ascontiguousarray(img, np.uint8)
        out = np.empty((window, window), np.uint8)
        self._chk(load().vslam_rotated_window(self.h, a.ctypes.data, a.shape[0], a.shape[1], a.shape[1], int(center[0]), int(center[1]),
                                              int(window), float(cos_sin[0]), float(cos_sin[1]), out.ctypes.data, window))
        return out

    def decode_jpeg_gray(self, streams, out=None):
        """JPEG byte strings (equal sizes) -> grey frames: numpy uint8 [n, h, w], or into `out`, a torch cuda uint8 tensor [n, h, w]
        that the detectors take directly (vslam_decode_jpeg_gray: nvJPEG + the bit-exact BGR2GRAY kernel)."""
        bufs = [np.frombuffer(bytes(s), np.uint8) for s in streams]
        n = len(bufs)
        r, c, ch = C.c_int(), C.c_int(), C.c_int()
        self._chk(load().vslam_jpeg_info(self.h, bufs[0].ctypes.data, bufs[0].size, C.byref(r), C.byref(c), C.byref(ch)))
        h, w = r.value, c.value
        ptrs = (C.c_void_p * n)(*[b.ctypes.data for b in bufs])
        lens = (C.c_size_t * n)(*[b.size for b in bufs])
        if out is None:
            res = np.empty((n, h, w), np.uint8)
            self._chk(load().vslam_decode_jpeg_gray(self.h, ptrs, lens, n, h, w, res.ctypes.data, w, h * w, HOST))
            return res
        assert tuple(out.shape) == (n, h, w) and out.is_contiguous()
        dev = self._torch_in(out)
        self._chk(load().vslam_decode_jpeg_gray(self.h, ptrs, lens, n, h, w, out.data_ptr(), w, h * w, DEVICE))
        self._torch_out(dev)
        return out

    def warp_affine(self, img, m, dsize):
        """cv::warpAffine(img, M, dsize = (width, height)): INTER_LINEAR, BORDER_CONSTANT 0, 8UC1 (rotation.cpp:198)"""
        a = np.ascontiguousarray(img, np.uint8)
        mm = np.ascontiguousarray(m, np.float64).reshape(6)
        out = np.empty((int(dsize[1]), int(dsize[0])), np.uint8)
        self._chk(load().vslam_warp_affine(self.h, a.ctypes.data, a.shape[0], a.shape[1], a.shape[1], mm.ctypes.data_as(C.POINTER(C.c_double)),
                                           out.ctypes.data, out.shape[0], out.shape[1], out.shape[1]))
        return out

    def harris(self, frames, k=0.04, nms2_window=5, want_maps=True, capacity=1 << 20, device_out=None, corners_out=None, compact=False):
        """compact=True: the corner list comes back as vslam_corner8 (CORNER8_DTYPE, 8 bytes per corner; expand_corners() gives
        the 12-byte layout); device_out["corners"] / corners_out must then hold 8 bytes per entry."""
        dev = self._torch_in(frames, *(device_out.values() if device_out else ()))
        ptr, n, h, w, pitch, stride, space, keep = _frames_arg(frames)
        hp = HarrisParams(k, nms2_window)
        counts = (C.c_int * n)()
        flag = COMPACT if compact else 0
        dt = CORNER8_DTYPE if compact else CORNER_DTYPE
        if device_out is not None:      # dict of torch cuda tensors: response/r8/mask3 (optional), corners (required)
            g = lambda name: device_out[name].data_ptr() if device_out.get(name) is not None else None
            capacity = device_out["corners"].shape[0]
            rc = load().vslam_harris(self.h, ptr, n, h, w, pitch, stride, space, C.byref(hp), g("response"), g("r8"),
                                     g("mask3"), g("corners"), capacity, DEVICE | flag, counts)
            if rc and rc != E_CAPACITY:
                self._chk(rc)
            self._torch_out(dev)
            return {"counts": list(counts), "overflow": rc == E_CAPACITY}
        resp = np.empty((n, h, w), np.float32) if want_maps else None
        r8 = np.empty((n, h, w), np.uint8) if want_maps else None
        m3 = np.empty((n, h, w), np.uint8) if want_maps else None
        corners = np.zeros(capacity, dt) if corners_out is None else corners_out.view(dt).reshape(-1)
        capacity = corners.shape[0]
        p = lambda a: a.ctypes.data if a is not None else None
        rc = load().vslam_harris(self.h, ptr, n, h, w, pitch, stride, space, C.byref(hp), p(resp), p(r8), p(m3),
                                 corners.ctypes.data, capacity, HOST | flag, counts)
        if rc and rc != E_CAPACITY:
            self._chk(rc)
        cl = list(counts)
        return {"response": resp, "r8": r8, "mask3": m3, "corners": corners[:min(sum(cl), capacity)], "counts": cl,
                "overflow": rc == E_CAPACITY}
