"""CPU tests of the host logic and the C-ABI surface: the library loads, exports every symbol include/vslam.h declares,
and its GPU-free entry points (sigma table, tap generator, error path) behave like the reference's."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import visualslam_b200 as vs
from visualslam_b200 import api, build

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    build.build()
    return api.load()


def test_exports_every_declared_symbol(lib):
    hdr = open(os.path.join(ROOT, "include", "vslam.h")).read()
    declared = set(re.findall(r"VSLAM_API\s+[^;(]*?\b(vslam_\w+)\s*\(", hdr))
    assert declared == set(api.EXPORTS), declared ^ set(api.EXPORTS)
    for name in declared:
        assert hasattr(lib, name), name


def test_struct_layouts():
    assert api.KEYPOINT_DTYPE.itemsize == 24 and api.CORNER_DTYPE.itemsize == 12
    assert C.sizeof(api.PyramidParams) == 24 and C.sizeof(api.DogParams) == 12 and C.sizeof(api.HarrisParams) == 8


def test_sigma_table_matches_reference_prints(lib):
    # tests/GaussPyramid_Test.cpp:99-106
    assert vs.calculate_sigma(0, 0) == 1.6
    assert abs(vs.calculate_sigma(0, 5) - 5.0796833) < 1e-6
    assert abs(vs.calculate_sigma(0, 3) - 3.2) < 1e-12
    assert lib.vslam_calculate_num_octaves(600, 868) == 5


def test_tap_generator_matches_cv2_tables(lib, golden):
    for key, e in golden["taps"].items():
        assert lib.vslam_gauss_ksize(e["sigma"]) == e["n"], key
        assert vs.gauss_taps(e["sigma"]).tolist() == e["taps"], key
    for e in golden["taps_sweep"]:
        assert vs.gauss_taps(e["sigma"]).tolist() == e["taps"], e["sigma"]
    assert vs.gauss_taps(0.0, 3).tolist() == [64, 128, 64]


def test_no_cpu_fallback(lib):
    """Without a CUDA device the context must fail loudly (and with a device it must succeed)."""
    import torch
    h = C.c_void_p()
    rc = lib.vslam_ctx_create(0, C.byref(h))
    if torch.cuda.is_available():
        assert rc == 0
        lib.vslam_ctx_destroy(h)
    else:
        assert rc == api.E_CUDA
        assert b"no CPU fallback" in lib.vslam_last_error(None)
        with pytest.raises(vs.VslamError):
            vs.Context(0)


def test_product_never_imports_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "visualslam_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".hpp", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle|#include\s+[\"<].*oracle", src, re.M), f
                assert "vslam_oracle" not in src and "c_oracle" not in src, f


def test_synthetic_frames_are_deterministic():
    from visualslam_b200 import synth
    a, b = synth.frame(1, 96, 128), synth.frame(1, 96, 128)
    assert np.array_equal(a, b) and a.dtype == np.uint8 and a.shape == (96, 128)
    assert synth.batch(10, 32, 48).shape == (10, 32, 48)


def test_fp32_tap_generator_and_descriptor_file_need_no_gpu(lib, tmp_path):
    """Host-only entry points of the descriptor stage: the fp32 Gaussian taps (cv::getGaussianKernel(n, sigma, CV_32F)) and
    the featureDescriptors.dat writer (Diff_of_Gauss.cpp:837-863)."""
    from oracle import c_oracle as O
    for o in range(4):
        for l in range(6):
            s = 1.5 * api.calculate_sigma(o, l)
            n = lib.vslam_gauss_ksize_f32(s)
            assert n == O.lib().vo_gauss_ksize_32f(s) and n % 2 == 1
            t = api.gauss_taps_f32(s)
            assert np.array_equal(t, O.gauss_taps_32f(s)) and abs(float(t.astype(np.float64).sum()) - 1.0) < 1e-5
            assert np.array_equal(t, t[::-1])
    assert lib.vslam_gauss_ksize_f32(1.5 * api.calculate_sigma(3, 3)) == 309
    assert lib.vslam_gauss_ksize_f32(0.0) < 0
    with pytest.raises(api.VslamError):
        api.gauss_taps_f32(2.0, 8)                                  # even kernel size
    d = np.arange(3 * 128, dtype=np.float32).reshape(3, 128) / 7
    p = tmp_path / "featureDescriptors.dat"
    api.write_descriptor_file(str(p), d)
    raw = p.read_bytes()
    assert len(raw) == 12 + 3 * 512 and np.frombuffer(raw[:12], "<i4").tolist() == [3, 128, 24]
    assert np.array_equal(np.frombuffer(raw[12:], "<f4").reshape(3, 128), d)
    api.write_descriptor_file(str(p), np.zeros((0, 128), np.float32))
    assert np.frombuffer(p.read_bytes(), "<i4").tolist() == [0, 128, 24]
    with pytest.raises(api.VslamError):
        api.write_descriptor_file(str(tmp_path / "no_such_dir" / "x.dat"), d)


def test_header_compiles_as_plain_c(tmp_path):
    """include/vslam.h is the boundary a C (cgo / ctypes / JNI) binding includes: it must parse as C99 on its own."""
    import subprocess
    src = tmp_path / "t.c"
    src.write_text('#include "vslam.h"\nint main(void) { vslam_corner8 c; c.row = 0; (void)c; return sizeof(vslam_keypoint) == 24 ? 0 : 1; }\n')
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Werror", "-fsyntax-only", "-I", os.path.join(ROOT, "include"), str(src)])
