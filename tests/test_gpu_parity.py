"""GPU parity tests (run on the B200 box): the CUDA path through the C ABI against the plain-C oracle on the same
inputs, and against the committed golden fixtures.  Integer / byte / index work is compared bit-exactly; the two fp32
maps that are not hot-path decisions (magnitude, phase) carry a stated tolerance."""
import hashlib
from ctypes import c_size_t as C_size_t

import numpy as np
import pytest

import visualslam_b200 as vs
from conftest import image_for
from oracle import c_oracle as O
from visualslam_b200 import synth

pytestmark = pytest.mark.gpu


def md5(a):
    return hashlib.md5(np.ascontiguousarray(a).tobytes()).hexdigest()


def kp_array(k):
    return np.stack([k[n] for n in ("row", "col", "value", "padding", "octave", "level")], 1).astype(np.int32)


def check_pyramid_against_oracle(ctx, img, num_octaves=4):
    pyr = ctx.build_pyramid(img, num_octaves, 1.6)
    ref = O.build_pyramid(img, num_octaves, 1.6)
    for o in range(num_octaves):
        assert pyr.octave_size(o) == ref["base"][o].shape
        assert np.array_equal(pyr.octave_image(o), ref["base"][o]), ("base", o)
        for l, (a, b) in enumerate(zip(pyr.octave_blur(o), ref["gauss"][o])):
            assert np.array_equal(a, b), ("gauss", o, l, int((a != b).sum()))
        for l, (a, b) in enumerate(zip(pyr.octave_diff(o), ref["dog"][o])):
            assert np.array_equal(a, b), ("dog", o, l)
    return pyr, ref


@pytest.mark.parametrize("stem", ["home", "building", "blox"])
def test_pyramid_golden_md5(ctx, stem, golden, images):
    g = golden["pyramid"][stem]
    pyr = ctx.build_pyramid(image_for(images, stem, "cvt"), 4, 1.6)
    assert pyr.num_octaves == 4 and pyr.num_levels == 6
    assert [list(pyr.octave_size(o)) for o in range(4)] == g["sizes"]
    assert [md5(pyr.octave_image(o)) for o in range(4)] == g["base"]
    assert [[md5(m) for m in pyr.octave_blur(o)] for o in range(4)] == g["gauss"]
    assert [[md5(m) for m in pyr.octave_diff(o)] for o in range(4)] == g["dog"]
    assert [[md5(m) for m in pyr.octave_grad_x(o)] for o in range(4)] == g["gx"]
    assert [[md5(m) for m in pyr.octave_grad_y(o)] for o in range(4)] == g["gy"]
    assert [pyr.octave_sigma(o) for o in range(4)] == g["sigma"]
    with pytest.raises(IndexError):
        pyr.sigma_at(4, 0)


def test_mag_orient_tolerance(ctx, images, golden_lists):
    pyr = ctx.build_pyramid(image_for(images, "home", "cvt"), 4, 1.6)
    mag = pyr.get(vs.MAP_MAG, 3, 1)
    ori = pyr.get(vs.MAP_ORIENT, 3, 1)
    assert np.allclose(mag, golden_lists["mag_o3l1_home"], rtol=2e-7, atol=0)         # <= 1 ulp
    assert np.abs(ori - golden_lists["orient_o3l1_home"]).max() < 1e-3                 # degrees


@pytest.mark.parametrize("shape", [(48, 64), (37, 51), (3, 5)])
def test_materialised_gradient_maps(ctx, shape):
    """vslam_pyramid_build with VSLAM_MATERIALISE_ALL keeps the four fp32 maps of processGradients (GaussPyramid.cpp:65-104)
    resident for every level and frame (grad_all_k); they equal the maps formed on demand (grad_k) bit for bit, and gx / gy
    equal the oracle's Sobel(ksize=1)."""
    rng = np.random.default_rng(shape[0] + shape[1])
    frames = rng.integers(0, 256, (3,) + shape, dtype=np.uint8)
    no = 2 if min(shape) >= 8 else 1
    lazy = ctx.build_pyramid(frames, no, 1.6)
    res = ctx.build_pyramid(frames, no, 1.6, materialise=0x78)
    ref = O.build_pyramid(frames[2], no, 1.6, gradients=True)
    for o in range(no):
        for l in range(6):
            for kind in (vs.MAP_GRADX, vs.MAP_GRADY, vs.MAP_MAG, vs.MAP_ORIENT):
                for f in (0, 2):
                    assert np.array_equal(res.get(kind, o, l, f), lazy.get(kind, o, l, f)), (o, l, kind, f)
            assert np.array_equal(res.get(vs.MAP_GRADX, o, l, 2), ref["gx"][o][l])
            assert np.array_equal(res.get(vs.MAP_GRADY, o, l, 2), ref["gy"][o][l])
    lazy.close(); res.close()


@pytest.mark.parametrize("shape", [(1, 37), (37, 1), (2, 2), (5, 7), (33, 64), (48, 31), (97, 131), (130, 258)])
def test_pyramid_awkward_sizes(ctx, shape):
    rng = np.random.default_rng(shape[0] * 1000 + shape[1])
    img = rng.integers(0, 256, shape, dtype=np.uint8)
    no = 4 if min(shape) >= 8 else 1
    check_pyramid_against_oracle(ctx, img, no)


@pytest.mark.parametrize("stem", ["home", "building", "blox"])
def test_dog_keypoints_golden(ctx, stem, golden, images, golden_lists):
    g = golden["pyramid"][stem]
    img = image_for(images, stem, "cvt")
    pyr = ctx.build_pyramid(img, 4, 1.6)
    res = ctx.dog_detect(pyr)
    want = golden_lists["dog_" + stem]
    assert res["counts"] == [len(want)]
    assert np.array_equal(kp_array(res["keypoints"]), want)            # same points, same emission order
    sc = res["stage_counts"][0]
    for o in range(4):
        c = g["counts"][o]
        assert sc[o].tolist() == [c["tested"], c["candidates"], c["contrast"], c["edge"]], o
    # the throughput path gives the same list
    res2 = ctx.dog_keypoints(img)
    assert np.array_equal(kp_array(res2["keypoints"]), want)


def test_dog_keypoints_synthetic_vs_oracle(ctx):
    frames = np.stack([synth.frame(i, 270, 480) for i in (0, 4, 7)], 0)
    res = ctx.dog_keypoints(frames)
    start = 0
    for f in range(len(frames)):
        kc, _, _ = O.dog_keypoints(frames[f])
        n = res["counts"][f]
        assert n == len(kc), f
        assert np.array_equal(kp_array(res["keypoints"][start:start + n]), kc), f
        start += n


@pytest.mark.parametrize("sigma0,octaves", [(1.2, 4), (1.9, 3), (0.9, 4)])
def test_dog_keypoints_other_sigma_vs_oracle(ctx, sigma0, octaves):
    """sigma0 != 1.6: the tensor-pipe kernel's run-time geometry (the default pyramid runs compile-time instantiations)."""
    frames = np.stack([synth.frame(i, 200, 333) for i in (1, 5)], 0)
    res = ctx.dog_keypoints(frames, num_octaves=octaves, sigma0=sigma0)
    start = 0
    for f in range(len(frames)):
        kc, _, _ = O.dog_keypoints(frames[f], num_octaves=octaves, sigma0=sigma0)
        n = res["counts"][f]
        assert n == len(kc), f
        assert np.array_equal(kp_array(res["keypoints"][start:start + n]), kc), f
        start += n


def test_dog_keypoints_kernel_variants_agree(tmp_path):
    """VSLAM_UMMA=1 (tcgen05 + TMA kernel for octaves 0 and 1 instead of the mma.sync one), VSLAM_MMA_STATIC=0 (run-time geometry),
    VSLAM_MMA_FW=32 (32-column strips for every octave) and VSLAM_FUSED_MMA=0 (dp4a/dp2a kernel) give the default kernels' lists."""
    import subprocess, sys, os
    code = ("import numpy as np, sys; import visualslam_b200 as vs; from visualslam_b200 import synth;"
            "f = np.stack([synth.frame(i, 540, 960) for i in range(8)], 0); c = vs.Context(0);"
            "r = c.dog_keypoints(f); np.save(sys.argv[1], np.asarray(r['keypoints']))")
    outs = []
    for i, env in enumerate(({}, {"VSLAM_UMMA": "1"}, {"VSLAM_MMA_STATIC": "0"}, {"VSLAM_FUSED_MMA": "0"}, {"VSLAM_MMA_FW": "32"},
                             {"VSLAM_MMA_FW": "32", "VSLAM_MMA_STATIC": "0"})):
        out = str(tmp_path / f"k{i}.npy")
        root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
        subprocess.check_call([sys.executable, "-c", code, out], env={**os.environ, **env, "PYTHONPATH": root}, timeout=600)
        outs.append(np.load(out))
    assert len(outs[0]) > 1000
    for o in outs[1:]:
        assert np.array_equal(outs[0], o)


def test_harris_kernel_variants_agree(tmp_path):
    """harris_fast_k (4-pixel-word shapes, the default) and harris_march_k (VSLAM_HARRIS_FAST=0) give identical maps and corner
    lists: strips on both image borders, a single strip, cols % 4 == 0 shapes smaller than a strip, both NMS2 windows."""
    import subprocess, sys, os
    code = ("import numpy as np, sys, hashlib; import visualslam_b200 as vs; from visualslam_b200 import synth;"
            "c = vs.Context(0); out = []\n"
            "for (n, h, w, win) in ((8, 540, 960, 5), (2, 270, 480, 3), (1, 33, 240, 5), (1, 1, 124, 5), (1, 2, 8, 3), (3, 131, 116, 5), (1, 400, 1124, 5)):\n"
            "    f = np.stack([synth.frame(i + h, h, w) for i in range(n)], 0); r = c.harris(f, nms2_window=win)\n"
            "    out += [np.frombuffer(hashlib.md5(r[k].tobytes()).digest(), np.uint8) for k in ('response', 'r8', 'mask3')]\n"
            "    out += [np.frombuffer(hashlib.md5(np.asarray(r['corners']).tobytes()).digest(), np.uint8), np.asarray(r['counts'], np.int64).view(np.uint8)]\n"
            "np.save(sys.argv[1], np.concatenate(out))")
    outs = []
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for i, env in enumerate(({}, {"VSLAM_HARRIS_FAST": "0"})):
        out = str(tmp_path / f"h{i}.npy")
        subprocess.check_call([sys.executable, "-c", code, out], env={**os.environ, **env, "PYTHONPATH": root}, timeout=600)
        outs.append(np.load(out))
    assert len(outs[0]) > 100 and np.array_equal(outs[0], outs[1])


def test_tcgen05_kernel_matches_oracle(tmp_path):
    """The tcgen05 / TMA octave kernel (fused_umma.cu, VSLAM_UMMA=1) against the plain-C oracle on awkward sizes: strips that
    end inside the image, a last step of one row, an octave smaller than one step, sizes smaller than the kernel radius."""
    import subprocess, sys, os
    code = ("import numpy as np, sys; import visualslam_b200 as vs; from visualslam_b200 import synth; from oracle import c_oracle as O;"
            "c = vs.Context(0)\n"
            "for (h, w) in ((192, 256), (300, 217), (70, 9), (9, 70), (64, 49), (129, 97), (5, 7), (257, 96)):\n"
            "    img = synth.frame(h + w, h, w); r = c.dog_keypoints(img); k = r['keypoints']\n"
            "    got = np.stack([k[n] for n in ('row', 'col', 'value', 'padding', 'octave', 'level')], 1)\n"
            "    want = O.dog_keypoints(img)[0]\n"
            "    assert np.array_equal(got, want), (h, w, len(got), len(want))\n"
            "print('ok')")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.run([sys.executable, "-c", code], env={**os.environ, "VSLAM_UMMA": "1", "PYTHONPATH": root}, timeout=900,
                         capture_output=True, text=True)
    assert out.returncode == 0 and "ok" in out.stdout, out.stdout[-2000:] + out.stderr[-2000:]


def test_dog_capacity_overflow(ctx, images):
    img = image_for(images, "building", "cvt")
    res = ctx.dog_keypoints(img, capacity=100)
    assert res["overflow"] and res["counts"] == [735] and len(res["keypoints"]) == 100
    full = ctx.dog_keypoints(img)
    assert np.array_equal(res["keypoints"], full["keypoints"][:100])


def test_harris_staging_overflow_is_marked(ctx):
    """A batch that emits more corners than the staging list holds (max(capacity, 2^20) entries) with a small capacity:
    VSLAM_E_CAPACITY, true counts, and every returned entry is either the entry of the full ordered list at that position
    or a row = -1 marker (include/vslam.h) -- never uninitialised memory."""
    frames = np.stack([synth.frame(i % 8, 1080, 1920) for i in range(24)], 0)
    full = ctx.harris(frames, want_maps=False, capacity=1 << 22)
    assert not full["overflow"] and sum(full["counts"]) > (1 << 20)
    small = ctx.harris(frames, want_maps=False, capacity=5000)
    assert small["overflow"] and small["counts"] == full["counts"] and len(small["corners"]) == 5000
    got, want = small["corners"], full["corners"][:5000]
    ok = (got["row"] == -1) | ((got["row"] == want["row"]) & (got["col"] == want["col"]) & (got["value"] == want["value"]))
    assert ok.all()


def test_harris_compact_corner_list(ctx):
    """VSLAM_COMPACT (vslam_corner8: u16 row, u16 col, f32 value) carries the same ordered list in two thirds of the bytes: host
    and device output, several launch chunks, the overflow markers."""
    import torch
    frames = np.stack([synth.frame(i % 8, 540, 960) for i in range(12)], 0)
    full = ctx.harris(frames, want_maps=False, capacity=1 << 21)
    c8 = ctx.harris(frames, want_maps=False, capacity=1 << 21, compact=True)
    assert c8["counts"] == full["counts"] and c8["corners"].dtype.itemsize == 8
    assert np.array_equal(vs.expand_corners(c8["corners"]), full["corners"])
    dev = torch.zeros((len(full["corners"]) + 7, 2), dtype=torch.int32, device="cuda")
    r = ctx.harris(torch.from_numpy(frames).cuda(), device_out={"corners": dev}, compact=True)
    assert r["counts"] == full["counts"]
    assert np.array_equal(vs.expand_corners(dev.cpu().numpy()[:len(full["corners"])].view(vs.CORNER8_DTYPE)), full["corners"])
    # staging overflow: positions of the full list or 0xffff markers (-> row = -1)
    big = np.stack([synth.frame(i % 8, 1080, 1920) for i in range(24)], 0)
    want = ctx.harris(big, want_maps=False, capacity=1 << 22)
    small = ctx.harris(big, want_maps=False, capacity=5000, compact=True)
    assert small["overflow"] and small["counts"] == want["counts"]
    got, w5 = vs.expand_corners(small["corners"]), want["corners"][:5000]
    ok = (got["row"] == -1) | ((got["row"] == w5["row"]) & (got["col"] == w5["col"]) & (got["value"] == w5["value"]))
    assert ok.all()


@pytest.mark.parametrize("shape", [(96, 160), (49, 84)])
def test_dog_window_sizes(ctx, shape):
    """Other grid strides.  (49, 84): at window 5 the last sample row / column of octave 0 (98 x 168) and octave 1 sits on
    rows - 1, so the edge test's window leaves the map (defined as repeating the border sample; the reference is undefined there)."""
    img = synth.frame(1, *shape)
    pyr = ctx.build_pyramid(img, 3, 1.6)
    ref = O.build_pyramid(img, 3, 1.6)
    for window in (3, 5):
        res = ctx.dog_detect(pyr, window=window)
        want = np.concatenate([O.detect_octave(ref["dog"][o], ref["gauss"][o], o, window)[0] for o in range(3)], 0)
        assert np.array_equal(kp_array(res["keypoints"]), want), window


@pytest.mark.parametrize("stem", ["blox", "chessboard", "building", "home"])
def test_harris_golden(ctx, stem, golden, images, golden_lists):
    g = golden["harris"][stem]
    h = ctx.harris(image_for(images, stem, "gray"))
    assert md5(h["response"][0]) == g["response"]
    assert md5(h["r8"][0]) == g["r8"]
    assert md5(h["mask3"][0]) == g["mask3"]
    assert h["counts"] == [g["n_corners"]]
    c = h["corners"]
    assert np.array_equal(np.stack([c["row"], c["col"]], 1), golden_lists["harris_" + stem])    # row-major order
    assert np.array_equal(c["value"], golden_lists["harris_val_" + stem])


@pytest.mark.parametrize("shape", [(1, 1), (1, 9), (9, 1), (2, 3), (4, 4), (5, 5), (31, 65), (64, 64), (100, 257)])
def test_harris_awkward_sizes(ctx, shape):
    rng = np.random.default_rng(shape[0] * 7 + shape[1])
    img = rng.integers(0, 256, shape, dtype=np.uint8)
    h = ctx.harris(img)
    ref = O.harris(img)
    assert np.array_equal(h["response"][0], ref["response"])
    assert np.array_equal(h["r8"][0], ref["r8"])
    assert np.array_equal(h["mask3"][0], ref["mask3"])
    c = h["corners"]
    assert np.array_equal(np.stack([c["row"], c["col"]], 1).reshape(-1, 2), ref["corners"].reshape(-1, 2))
    assert np.array_equal(c["value"], ref["corner_values"])


def test_harris_batch_and_nms3_window(ctx):
    frames = synth.batch(5, 120, 200)
    h = ctx.harris(frames, nms2_window=3)
    start = 0
    for f in range(5):
        ref = O.harris(frames[f], nms_window=3)
        n = h["counts"][f]
        assert np.array_equal(h["response"][f], ref["response"])
        c = h["corners"][start:start + n]
        assert n == len(ref["corners"]) and np.array_equal(np.stack([c["row"], c["col"]], 1), ref["corners"])
        start += n


def test_full_size_properties(ctx):
    """BASELINE.json sizes (1080p): size-independent properties instead of a full oracle run.
    - determinism / batch independence: a frame gives the same keypoints alone and inside a batch
    - the DoG of a constant frame is empty; Harris of a constant frame has no corner
    - a 1080p crop-consistent check: the oracle on one frame (seconds) equals the GPU list."""
    frames = np.stack([synth.frame(i, 1080, 1920) for i in (0, 5)], 0)
    both = ctx.dog_keypoints(frames)
    one = ctx.dog_keypoints(frames[1])
    n0, n1 = both["counts"]
    assert one["counts"] == [n1]
    assert np.array_equal(one["keypoints"], both["keypoints"][n0:n0 + n1])
    kc, _, _ = O.dog_keypoints(frames[0])
    assert n0 == len(kc) and np.array_equal(kp_array(both["keypoints"][:n0]), kc)
    flat = np.full((1080, 1920), 77, np.uint8)
    assert ctx.dog_keypoints(flat)["counts"] == [0]
    hf = ctx.harris(flat)
    assert hf["counts"] == [0] and not hf["response"].any()
    hr = ctx.harris(frames[0])
    ref = O.harris(frames[0])
    assert np.array_equal(hr["response"][0], ref["response"]) and hr["counts"] == [len(ref["corners"])]


def test_4k_frame_vs_oracle(ctx):
    """BASELINE.json's 4K size: one natural-tile frame, DoG keypoints and the Harris response / corners against the C oracle."""
    f = synth.frame(2, 2160, 3840)
    res = ctx.dog_keypoints(f)
    kc, _, _ = O.dog_keypoints(f)
    assert res["counts"] == [len(kc)] and np.array_equal(kp_array(res["keypoints"]), kc)
    hr = ctx.harris(f)
    ref = O.harris(f)
    assert np.array_equal(hr["response"][0], ref["response"])
    c = hr["corners"]
    assert hr["counts"] == [len(ref["corners"])] and np.array_equal(np.stack([c["row"], c["col"]], 1), ref["corners"])


def test_1080p_all_unique_frames_vs_oracle(ctx):
    """parity on all 8 unique frames of the 1080p workload -- natural-tile, procedural and uniform-noise frames
    drive the queue / compaction tail differently.  One batched GPU call; DoG lists and Harris corner lists per frame."""
    frames = np.stack([synth.frame(i, 1080, 1920) for i in range(8)], 0)
    res = ctx.dog_keypoints(frames, capacity=1 << 21)
    hr = ctx.harris(frames, want_maps=False, capacity=1 << 23)
    assert not res["overflow"] and not hr["overflow"]
    k0 = c0 = 0
    for f in range(8):
        kc, _, _ = O.dog_keypoints(frames[f])
        n = res["counts"][f]
        assert n == len(kc) and np.array_equal(kp_array(res["keypoints"][k0:k0 + n]), kc), f
        k0 += n
        ref = O.harris(frames[f])
        m = hr["counts"][f]
        c = hr["corners"][c0:c0 + m]
        assert m == len(ref["corners"]) and np.array_equal(np.stack([c["row"], c["col"]], 1), ref["corners"]), f
        assert np.array_equal(c["value"], ref["corner_values"]), f
        c0 += m


def test_8k_frame_vs_oracle(ctx):
    """BASELINE.json's 8K size against the C oracle itself: one natural-tile frame, DoG keypoints and Harris corners
    (tens of seconds of single-threaded CPU work)."""
    f = synth.frame(3, 4320, 7680)
    res = ctx.dog_keypoints(f, capacity=1 << 21)
    kc, _, _ = O.dog_keypoints(f)
    assert not res["overflow"] and res["counts"] == [len(kc)] and np.array_equal(kp_array(res["keypoints"]), kc)
    hr = ctx.harris(f, want_maps=False, capacity=1 << 23)
    ref = O.harris(f)
    c = hr["corners"]
    assert not hr["overflow"] and hr["counts"] == [len(ref["corners"])]
    assert np.array_equal(np.stack([c["row"], c["col"]], 1), ref["corners"]) and np.array_equal(c["value"], ref["corner_values"])


def test_8k_properties(ctx):
    """BASELINE.json's 8K size (the oracle would take minutes): batch independence, agreement of the two octave kernels'
    strip decompositions on the top-left 4K crop's interior, constant frames."""
    f = np.stack([synth.frame(i, 4320, 7680) for i in (1, 6)], 0)
    both = ctx.dog_keypoints(f)
    one = ctx.dog_keypoints(f[1])
    n0, n1 = both["counts"]
    assert n0 > 1000 and one["counts"] == [n1]
    assert np.array_equal(one["keypoints"], both["keypoints"][n0:n0 + n1])
    up = ctx.upload(f)                                            # chunked upload path at this size
    again = ctx.dog_keypoints(up)
    assert again["counts"] == [n0, n1] and np.array_equal(again["keypoints"], both["keypoints"])
    h2 = ctx.harris(f, want_maps=False, capacity=1 << 22)
    h1 = ctx.harris(f[0], want_maps=False, capacity=1 << 22)
    assert not h2["overflow"] and h1["counts"] == [h2["counts"][0]] and np.array_equal(h1["corners"], h2["corners"][:h2["counts"][0]])
    flat = np.full((4320, 7680), 200, np.uint8)
    assert ctx.dog_keypoints(flat)["counts"] == [0] and ctx.harris(flat, want_maps=False)["counts"] == [0]


def test_bgr_to_gray(ctx, images):
    """vslam_bgr_to_gray (the ingest step in front of the pyramid, Diff_of_Gauss.cpp:739) against the golden grey image and
    the oracle on awkward shapes (odd widths: the byte path; batches)."""
    import os
    with np.load(os.path.join(os.path.dirname(__file__), "golden", "home_bgr.npz")) as z:
        bgr = z["home_bgr"]
    assert np.array_equal(ctx.bgr_to_gray(bgr), images["home_cvt"])
    rng = np.random.default_rng(5)
    for shape in [(1, 1, 1), (2, 3, 5), (3, 17, 33), (1, 64, 128), (4, 31, 250)]:
        x = rng.integers(0, 256, shape + (3,), dtype=np.uint8)
        got = ctx.bgr_to_gray(x)
        for f in range(shape[0]):
            assert np.array_equal(got[f], O.bgr_to_gray(x[f])), shape


@pytest.mark.parametrize("theta", [0.0, 33.0, 45.0, 90.0, 180.0, 271.5])
def test_rotation_gathers_vs_oracle(ctx, images, theta):
    """vslam_rotate_mat_ccw / vslam_rotated_window (SLAM::Rotation, rotation.cpp:66-108) against the oracle's literal arithmetic."""
    img = images["blox"]
    cs = O.cos_sin_of_angle(theta)
    center = (img.shape[1] // 2, img.shape[0] // 2)
    assert np.array_equal(ctx.rotate_mat_ccw(img, center, cs), O.rotate_mat_ccw(img, center, cs))
    for window, c in ((16, center), (100, center), (16, (3, 250))):          # the last one reads outside the image
        assert np.array_equal(ctx.rotated_window(img, c, window, cs), O.rotated_window(img, c, window, cs))
    odd = synth.frame(3, 37, 91)
    assert np.array_equal(ctx.rotate_mat_ccw(odd, (10, 30), cs), O.rotate_mat_ccw(odd, (10, 30), cs))


def test_two_devices_in_one_process(ctx):
    """Contexts on two devices of one process (function attributes and table images are per device)."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("one GPU")
    import visualslam_b200 as vs
    f = np.stack([synth.frame(i, 300, 500) for i in (0, 7)], 0)
    a = ctx.dog_keypoints(f)
    c1 = vs.Context(1)
    b = c1.dog_keypoints(f)
    assert a["counts"] == b["counts"] and np.array_equal(a["keypoints"], b["keypoints"])
    ha, hb = ctx.harris(f), c1.harris(f)
    assert np.array_equal(ha["response"], hb["response"]) and np.array_equal(ha["corners"], hb["corners"])


def test_shared_chunked_upload_gives_the_same_lists(ctx):
    """vslam_upload_frames: one asynchronous chunked upload (several chunks at this size) serves both detectors; the
    lists equal those of the plain host-frame calls, frame by frame, and a second upload invalidates nothing it should not."""
    frames = np.stack([synth.frame(i % 8, 720, 1280) for i in range(160)], 0)        # 147 MB -> 3 upload chunks
    d0 = ctx.dog_keypoints(frames, capacity=1 << 21)
    h0 = ctx.harris(frames, want_maps=False, capacity=1 << 23)
    up = ctx.upload(frames)
    kp_buf = np.zeros((1 << 21, 6), np.int32)
    d1 = ctx.dog_keypoints(up, out=kp_buf)
    h1 = ctx.harris(up, want_maps=False, capacity=1 << 23)
    assert d1["counts"] == d0["counts"] and np.array_equal(d1["keypoints"], d0["keypoints"])
    assert h1["counts"] == h0["counts"] and np.array_equal(h1["corners"], h0["corners"])
    # frames are periodic with period 8: so are the per-frame counts
    assert d0["counts"][:8] == d0["counts"][8:16] and h0["counts"][:8] == h0["counts"][152:160]


# ---- the stage after the hot path: orientation assignment + SIFT descriptors (descriptor.cu) -----------------------------
FP32_TOL = 1e-4          # BASELINE.json north_star tolerance; against the C oracle the kernels are bit-exact


def same_floats(a, b):
    """Bit-identical apart from the payload of NaNs (0 / 0 yields a different quiet NaN on the two machines)."""
    na, nb = np.isnan(a), np.isnan(b)
    return np.array_equal(na, nb) and np.array_equal(a[~na].view(np.uint32), b[~nb].view(np.uint32))


def _features_gpu(ctx, img, kp6, num_octaves=4):
    """./DoG's per-octave loop (Diff_of_Gauss.cpp:781-790) after the hot path, through the C ABI."""
    pyr = ctx.build_pyramid(img, num_octaves, 1.6)
    hist, ori, desc, valid = [], [], [], []
    for o in range(num_octaves):
        k = np.ascontiguousarray(kp6[kp6[:, 4] == o], np.int32)
        r, h = ctx.assign_orientations(pyr, o, k, histograms=True)
        d, v = ctx.sift_descriptors(pyr, o, r)
        hist.append(h); ori.append(kp_array(r).reshape(-1, 6)); desc.append(d); valid.append(v)
    pyr.close()
    return np.concatenate(hist, 0), np.concatenate(ori, 0), np.concatenate(desc, 0), np.concatenate(valid, 0)


@pytest.mark.parametrize("stem", ["home", "building"])
def test_orientations_and_descriptors_golden(ctx, stem, images, golden_lists):
    import os
    with np.load(os.path.join(os.path.dirname(__file__), "golden", "descriptors.npz")) as z:
        gd = {k: z[k] for k in z.files}
    img = image_for(images, stem, "cvt")
    kp = golden_lists[f"dog_{stem}"]
    hist, ori, desc, valid = _features_gpu(ctx, img, kp)
    # bit-exact against the C oracle
    opyr = O.build_pyramid(img)
    want_h = np.stack([O.orientation_histogram(opyr["gauss"][k[4]][k[5]], int(k[0]), int(k[1]), 1.5 * opyr["sigma"][k[4]][k[5]])
                       for k in kp], 0)
    assert np.array_equal(hist.view(np.uint32), want_h.view(np.uint32))
    want_ori = np.concatenate([O.assign_orientations(opyr["gauss"][o], opyr["sigma"][o], kp[kp[:, 4] == o]) for o in range(4)], 0)
    assert np.array_equal(ori, want_ori)
    wd, wv = zip(*[O.sift(opyr["gauss"][o], opyr["sigma"][o], want_ori[want_ori[:, 4] == o]) for o in range(4)])
    wd, wv = np.concatenate(wd, 0), np.concatenate(wv, 0)
    assert np.array_equal(valid, wv)
    assert same_floats(desc, wd)       # incl. the flagged ones: same stand-in samples
    # within the fp32 tolerance of the cv2 fixtures (the reference's own OpenCV calls)
    assert np.abs(hist - gd[f"{stem}_hist"]).max() / gd[f"{stem}_hist"].max() <= FP32_TOL
    assert np.array_equal(ori, gd[f"{stem}_oriented"])
    assert np.array_equal(valid, gd[f"{stem}_valid"])
    ok = valid.astype(bool)
    assert np.nanmax(np.abs(desc[ok] - gd[f"{stem}_desc"][ok])) <= FP32_TOL


def test_descriptors_synthetic_all_octaves_vs_oracle(ctx):
    """A synthetic frame large enough to have keypoints in the coarse octaves (blur radii up to 154 at octave 3 level 3)."""
    img = synth.frame(3, 540, 720)
    ori_o, desc_o, valid_o, _ = O.dog_features(img)
    kc, _, _ = O.dog_keypoints(img)
    assert len(kc[kc[:, 4] >= 2]) > 0
    hist, ori, desc, valid = _features_gpu(ctx, img, kc)
    assert np.array_equal(ori, ori_o)
    assert np.array_equal(valid, valid_o)
    assert same_floats(desc, desc_o)


def test_descriptor_edge_cases(ctx, tmp_path):
    img = synth.frame(1, 96, 128)
    pyr = ctx.build_pyramid(img, 3, 1.6)
    opyr = O.build_pyramid(img, 3, 1.6)
    # empty lists
    assert len(ctx.assign_orientations(pyr, 0, np.zeros((0, 6), np.int32))) == 0
    d, v = ctx.sift_descriptors(pyr, 0, np.zeros((0, 6), np.int32))
    assert d.shape == (0, 128) and v.shape == (0,)
    # keypoints in the corners / on the borders, every level (incl. the widest kernel of the octave, wider than the map)
    rows, cols = pyr.octave_size(2)
    pts = [(0, 0), (0, cols - 1), (rows - 1, 0), (rows - 1, cols - 1), (rows // 2, cols // 2), (1, cols - 2)]
    kp = np.array([[y, x, 5, 1, 2, l] for l in range(6) for (y, x) in pts], np.int32)
    r, h = ctx.assign_orientations(pyr, 2, kp, histograms=True)
    want = O.assign_orientations(opyr["gauss"][2], opyr["sigma"][2], kp)
    assert np.array_equal(kp_array(r).reshape(-1, 6), want)
    d, v = ctx.sift_descriptors(pyr, 2, want)
    wd, wv = O.sift(opyr["gauss"][2], opyr["sigma"][2], want)
    assert np.array_equal(v, wv) and same_floats(d, wd)
    # a point outside the map is refused quietly: no orientation, NaN descriptor flagged invalid
    bad = np.array([[rows, 0, 0, 0, 2, 1]], np.int32)
    assert len(ctx.assign_orientations(pyr, 2, bad)) == 0
    d, v = ctx.sift_descriptors(pyr, 2, bad)
    assert np.isnan(d).all() and v[0] == 0
    with pytest.raises(vs.VslamError):
        ctx.assign_orientations(pyr, 7, kp)
    # featureDescriptors.dat (Diff_of_Gauss.cpp:837-863)
    path = tmp_path / "featureDescriptors.dat"
    vs.api.write_descriptor_file(str(path), wd)
    raw = path.read_bytes()
    assert np.frombuffer(raw[:12], "<i4").tolist() == [len(wd), 128, 24]
    assert np.array_equal(np.frombuffer(raw[12:], "<f4").reshape(-1, 128).view(np.uint32), wd.view(np.uint32))
    pyr.close()


def test_describe_keypoints_batch_vs_oracle(ctx):
    """All frames and octaves of a batch in two launches: per frame the oracle's oriented list and descriptors."""
    frames = np.stack([synth.frame(i, 270, 480) for i in (2, 5, 6)], 0)
    pyr = ctx.build_pyramid(frames, 4, 1.6)
    det = ctx.dog_detect(pyr, stage_counts=False)
    res = ctx.describe_keypoints(pyr, det["keypoints"], det["counts"])
    start = 0
    for f in range(len(frames)):
        ori_o, desc_o, valid_o, _ = O.dog_features(frames[f])
        n = res["counts"][f]
        assert n == len(ori_o), f
        assert np.array_equal(kp_array(res["oriented"][start:start + n]).reshape(-1, 6), ori_o), f
        assert np.array_equal(res["valid"][start:start + n], valid_o), f
        assert same_floats(res["descriptors"][start:start + n], desc_o), f
        start += n
    # a list that does not fit: totals are still reported, the stored prefix is right
    small = ctx.describe_keypoints(pyr, det["keypoints"], det["counts"], capacity=10)
    assert small["overflow"] and small["counts"] == res["counts"]
    assert np.array_equal(small["oriented"], res["oriented"][:10]) and same_floats(small["descriptors"], res["descriptors"][:10])
    # no keypoints at all
    none = ctx.describe_keypoints(pyr, np.zeros((0, 6), np.int32), [0, 0, 0])
    assert none["counts"] == [0, 0, 0] and len(none["oriented"]) == 0
    pyr.close()


def test_describe_keypoints_1080p_frame_vs_oracle(ctx):
    """BASELINE.json's frame size: every keypoint of one 1080p frame (blur radii 12 ... 154) against the oracle, bit for bit."""
    img = synth.frame(11, 1080, 1920)
    pyr = ctx.build_pyramid(img, 4, 1.6)
    det = ctx.dog_detect(pyr, stage_counts=False)
    res = ctx.describe_keypoints(pyr, det["keypoints"], det["counts"])
    pyr.close()
    ori_o, desc_o, valid_o, _ = O.dog_features(img)
    assert len(ori_o) > 2000 and set(ori_o[:, 4]) == {0, 1, 2, 3}
    assert np.array_equal(kp_array(res["oriented"]).reshape(-1, 6), ori_o)
    assert np.array_equal(res["valid"], valid_o)
    assert same_floats(res["descriptors"], desc_o)
    d = res["descriptors"][res["valid"].astype(bool)]
    d = d[~np.isnan(d).any(1)]
    assert d.min() >= 0.0 and np.all(d.max(1) == 1.0)            # size-independent: renormalised to a unit peak


def test_descriptors_five_octaves_other_sigma(ctx):
    """A deeper pyramid with another sigma: up to 489-tap fp32 kernels at octave 4, several times wider than the 68 x 90 map they blur."""
    img = synth.frame(9, 540, 720)
    pyr = ctx.build_pyramid(img, 5, 0.8)
    opyr = O.build_pyramid(img, 5, 0.8)
    rows, cols = pyr.octave_size(4)
    kp = np.array([[y, x, 3, 1, 4, l] for l in (1, 2, 3, 5) for (y, x) in ((0, 0), (rows // 2, cols // 2), (rows - 1, cols - 2))], np.int32)
    r, h = ctx.assign_orientations(pyr, 4, kp, histograms=True)
    want = O.assign_orientations(opyr["gauss"][4], opyr["sigma"][4], kp)
    assert np.array_equal(kp_array(r).reshape(-1, 6), want) and len(want) > 0
    d, v = ctx.sift_descriptors(pyr, 4, want)
    wd, wv = O.sift(opyr["gauss"][4], opyr["sigma"][4], want)
    assert np.array_equal(v, wv) and same_floats(d, wd)
    pyr.close()


def test_warp_affine_and_rotation_matrix_vs_cv2_and_oracle(ctx):
    """vslam_get_rotation_matrix_2d / vslam_warp_affine (Rotation::doubleCrop's OpenCV calls, rotation.cpp:152,185,198): the cv2
    4.13 vectors bit for bit (matrices and images), and the oracle on sizes the vectors do not hold (1-pixel images, large shifts)."""
    import os
    from conftest import GOLDEN
    with np.load(os.path.join(GOLDEN, "warp_affine.npz")) as z:
        g = {k: z[k] for k in z.files}
    for i, (h, w, ang, cx, cy, dh, dw, scale, seed) in enumerate(g["warp_cases"]):
        img = np.random.default_rng(int(seed)).integers(0, 256, (int(h), int(w)), dtype=np.uint8)
        m = vs.api.get_rotation_matrix_2d((cx, cy), ang, scale)
        assert m.tobytes() == g[f"warp_m_{i}"].tobytes(), i
        assert np.array_equal(ctx.warp_affine(img, m, (int(dw), int(dh))), g[f"warp_out_{i}"]), i
    rng = np.random.default_rng(5)
    for (h, w, dh, dw) in ((1, 1, 5, 7), (1, 37, 9, 40), (37, 1, 40, 9), (300, 217, 64, 500), (64, 64, 1, 1)):
        img = rng.integers(0, 256, (h, w), dtype=np.uint8)
        m = O.get_rotation_matrix_2d((w / 2.0, h / 2.0), float(rng.uniform(-180, 180)), float(rng.uniform(0.3, 3.0)))
        m[:, 2] += rng.uniform(-40, 40, 2)
        assert np.array_equal(ctx.warp_affine(img, m, (dw, dh)), O.warp_affine(img, m, (dw, dh))), (h, w, dh, dw)


def test_dog_harris_combined_call_matches_the_two_calls(ctx):
    """vslam_dog_harris (Harris launch in front of every chunk's DoG launches, both lists draining chunk by chunk) returns exactly
    the lists, order and counts of vslam_dog_keypoints + vslam_harris: host lists from an uploaded batch (several chunks), device
    lists, the compact corner list, a capacity overflow on either list."""
    import torch
    frames = np.stack([synth.frame(i % 8, 540, 960) for i in range(40)], 0)
    kp = ctx.dog_keypoints(frames, capacity=1 << 20)
    co = ctx.harris(frames, want_maps=False, capacity=1 << 22)
    pinned = torch.from_numpy(frames).pin_memory().numpy()
    both = ctx.dog_harris(ctx.upload(pinned), kp_capacity=1 << 20, co_capacity=1 << 22)
    assert both["kp_counts"] == kp["counts"] and both["co_counts"] == co["counts"] and not both["overflow"]
    assert np.array_equal(both["keypoints"], kp["keypoints"]) and np.array_equal(both["corners"], co["corners"])
    c8 = ctx.dog_harris(frames, kp_capacity=1 << 20, co_capacity=1 << 22, compact=True)
    assert np.array_equal(c8["keypoints"], kp["keypoints"]) and np.array_equal(vs.expand_corners(c8["corners"]), co["corners"])
    kd = torch.zeros((len(kp["keypoints"]) + 3, 6), dtype=torch.int32, device="cuda")
    cd = torch.zeros((len(co["corners"]) + 3, 3), dtype=torch.int32, device="cuda")
    r = ctx.dog_harris(torch.from_numpy(frames).cuda(), kp_out=kd, co_out=cd)
    assert r["kp_counts"] == kp["counts"] and r["co_counts"] == co["counts"]
    assert np.array_equal(kd.cpu().numpy()[:len(kp["keypoints"])].view(vs.KEYPOINT_DTYPE).reshape(-1), kp["keypoints"])
    assert np.array_equal(cd.cpu().numpy()[:len(co["corners"])].view(vs.CORNER_DTYPE).reshape(-1), co["corners"])
    small = ctx.dog_harris(frames, kp_capacity=100, co_capacity=1 << 22)
    assert small["overflow"] and small["kp_counts"] == kp["counts"] and np.array_equal(small["keypoints"], kp["keypoints"][:100])
    assert np.array_equal(small["corners"], co["corners"])
    small = ctx.dog_harris(frames, kp_capacity=1 << 20, co_capacity=1000)
    assert small["overflow"] and small["co_counts"] == co["counts"] and np.array_equal(small["keypoints"], kp["keypoints"])
    # and the separate calls still work after the combined one (the two list workspaces are back in place)
    assert np.array_equal(ctx.harris(frames[:3], want_maps=False)["corners"], co["corners"][:sum(co["counts"][:3])])


def test_pipelined_uploads_two_ingest_buffers(ctx):
    """vslam_upload_frames alternates between two ingest buffers: batch i+1 may be uploaded before the detectors run on batch i
    (include/vslam.h); lists equal those of the direct calls, in either order of use, several rounds."""
    import torch
    A = np.stack([synth.frame(i % 8, 270, 480) for i in range(24)], 0)
    B = np.stack([synth.frame((i + 3) % 8, 270, 480)[::-1].copy() for i in range(24)], 0)
    want = {}
    for name, f in (("A", A), ("B", B)):
        want[name] = (ctx.dog_keypoints(f)["keypoints"].copy(), ctx.harris(f, want_maps=False)["corners"].copy())
    pa, pb = torch.from_numpy(A).pin_memory().numpy(), torch.from_numpy(B).pin_memory().numpy()
    up = ctx.upload(pa)
    for rnd in range(4):
        name, nxt = ("A", pb) if rnd % 2 == 0 else ("B", pa)
        up_next = ctx.upload(nxt)                      # next batch on its way before this one is processed
        r = ctx.dog_harris(up)
        assert np.array_equal(r["keypoints"], want[name][0]) and np.array_equal(r["corners"], want[name][1]), rnd
        if rnd == 1:                                   # separate calls on an uploaded batch too
            assert np.array_equal(ctx.harris(up, want_maps=False)["corners"], want[name][1])
        up = up_next


def test_deferred_dog_harris_pipeline(ctx):
    """VSLAM_DEFER + vslam_finish: batch i's call only enqueues, its lists are completed by finish() after batch i+1 has been
    enqueued (two buffer sets); results equal the synchronous call's, a third pending call is refused, overflow is reported by finish."""
    import torch
    A = np.stack([synth.frame(i % 8, 270, 480) for i in range(20)], 0)
    B = np.stack([synth.frame((i + 5) % 8, 270, 480)[:, ::-1].copy() for i in range(20)], 0)
    want = {n: ctx.dog_harris(f, compact=True) for n, f in (("A", A), ("B", B))}
    pin = {"A": torch.from_numpy(A).pin_memory().numpy(), "B": torch.from_numpy(B).pin_memory().numpy()}
    bufs = [(torch.empty((1 << 18, 6), dtype=torch.int32, pin_memory=True).numpy(), torch.empty((1 << 20, 2), dtype=torch.int32, pin_memory=True).numpy())
            for _ in range(2)]
    order = ["A", "B", "B", "A", "A"]
    up = ctx.upload(pin[order[0]])
    done = []
    for i, name in enumerate(order):
        nxt = ctx.upload(pin[order[i + 1]]) if i + 1 < len(order) else None
        assert ctx.dog_harris(up, kp_out=bufs[i & 1][0], co_out=bufs[i & 1][1], compact=True, defer=True) is None
        if i:
            done.append({k: (v.copy() if isinstance(v, np.ndarray) else v) for k, v in ctx.finish().items()})   # views of bufs: reused two steps on
        up = nxt
    done.append(ctx.finish())
    assert ctx.finish() is None
    for name, r in zip(order, done):
        w = want[name]
        assert r["kp_counts"] == w["kp_counts"] and r["co_counts"] == w["co_counts"] and not r["overflow"], name
        assert np.array_equal(r["keypoints"], w["keypoints"]) and np.array_equal(r["corners"], w["corners"]), name
    # a third pending call is refused; the two pending ones still complete
    small_k = torch.empty((50, 6), dtype=torch.int32, pin_memory=True).numpy()
    ctx.dog_harris(A, kp_out=small_k, co_out=bufs[0][1], compact=True, defer=True)
    ctx.dog_harris(B, kp_out=bufs[1][0], co_out=bufs[1][1], compact=True, defer=True)
    with pytest.raises(vs.VslamError):
        ctx.dog_harris(A, kp_out=bufs[0][0], co_out=bufs[0][1], compact=True, defer=True)
    r = ctx.finish()
    assert r["overflow"] and r["kp_counts"] == want["A"]["kp_counts"] and np.array_equal(r["keypoints"], want["A"]["keypoints"][:50])
    r = ctx.finish()
    assert not r["overflow"] and np.array_equal(r["corners"], want["B"]["corners"])


def test_jpeg_ingest_decodes_to_grey_frames(ctx, images):
    """vslam_decode_jpeg_gray (nvJPEG + the bit-exact BGR2GRAY kernel).  Not a parity target -- a decoder's
    output is a fixture -- so the check is a tolerance against cv2's decode of the same stream (libjpeg-turbo): mean |diff| < 1 grey
    level, 99.9 % of the pixels within 4; a grey-scale JPEG decodes exactly; the frames feed the detectors from device memory."""
    import cv2
    import torch
    from conftest import GOLDEN
    import os
    with np.load(os.path.join(GOLDEN, "home_bgr.npz")) as z:
        bgr = z[z.files[0]] if "bgr" not in z.files else z["bgr"]
    assert bgr.ndim == 3 and bgr.shape[2] == 3
    streams, want = [], []
    for q in (95, 80):
        ok, enc = cv2.imencode(".jpg", bgr, [cv2.IMWRITE_JPEG_QUALITY, q])
        assert ok
        streams.append(enc.tobytes())
        want.append(cv2.cvtColor(cv2.imdecode(enc, cv2.IMREAD_COLOR), cv2.COLOR_BGR2GRAY))
    got = ctx.decode_jpeg_gray(streams)
    assert got.shape == (2,) + bgr.shape[:2]
    for g, w in zip(got, want):
        d = np.abs(g.astype(np.int32) - w.astype(np.int32))
        assert d.mean() < 1.0 and (d <= 4).mean() > 0.999, (d.mean(), d.max(), (d <= 4).mean())
    grey = images["home_gray"]
    ok, enc = cv2.imencode(".jpg", grey, [cv2.IMWRITE_JPEG_QUALITY, 90])
    ref = cv2.imdecode(enc, cv2.IMREAD_GRAYSCALE)
    one = ctx.decode_jpeg_gray([enc.tobytes()])[0]
    assert np.abs(one.astype(np.int32) - ref.astype(np.int32)).max() <= 2          # one component: only the IDCT rounding differs
    dev = torch.empty(got.shape, dtype=torch.uint8, device="cuda")
    ctx.decode_jpeg_gray(streams, out=dev)
    assert np.array_equal(dev.cpu().numpy(), got)
    a, b = ctx.dog_keypoints(dev), ctx.dog_keypoints(got)
    assert a["counts"] == b["counts"] and np.array_equal(a["keypoints"], b["keypoints"]) and sum(a["counts"]) > 100


def test_wide_kernel_octaves_match_oracle(ctx, images):
    """Octaves whose Gaussian kernels exceed 247 taps (octave 4 and beyond at sigma0 = 1.6: the reference's two-argument
    GaussPyramid(img, sigma) with calculateNumOctaves, GaussPyramid.cpp:150-152) run on the exact global-memory path: every map and
    the keypoint lists equal the oracle's, through the pyramid API and through the throughput entry points."""
    img = image_for(images, "building", "cvt")                       # 600 x 868: floor(log2(600)) - 4 = 5 octaves, kernels up to 489 taps
    no = vs.load().vslam_calculate_num_octaves(*img.shape)
    assert no == 5
    pyr = ctx.build_pyramid(img, 0, 1.6)                            # num_octaves <= 0: the reference's own count
    assert pyr.num_octaves == 5
    opyr = O.build_pyramid(img, 5, 1.6)
    for o in range(5):
        for l in range(6):
            assert np.array_equal(pyr.get(vs.MAP_GAUSS, o, l), opyr["gauss"][o][l]), (o, l)
        for l in range(5):
            assert np.array_equal(pyr.get(vs.MAP_DOG, o, l), opyr["dog"][o][l]), (o, l)
    want = O.dog_keypoints(img, num_octaves=5)[0]
    got = kp_array(ctx.dog_detect(pyr)["keypoints"])
    assert np.array_equal(got, want) and (want[:, 4] == 4).sum() >= 0
    assert np.array_equal(kp_array(ctx.dog_keypoints(img, num_octaves=5)["keypoints"]), want)          # throughput entry point
    both = ctx.dog_harris(np.stack([img, img[::-1].copy()], 0), num_octaves=5)
    assert np.array_equal(kp_array(both["keypoints"][:both["kp_counts"][0]]), want)
    big = synth.frame(2, 1080, 1920)                                 # six octaves, kernels up to 977 taps on a 68 x 120 octave
    p6 = ctx.build_pyramid(big, 0, 1.6)
    assert p6.num_octaves == 6
    o6 = O.build_pyramid(big, 6, 1.6)
    for o in (4, 5):
        for l in range(6):
            assert np.array_equal(p6.get(vs.MAP_GAUSS, o, l), o6["gauss"][o][l]), (o, l)


def test_pyramid_rebuild_in_place(ctx):
    """vslam_pyramid_rebuild: the same pyramid object on new frames of the same shape gives the maps, gradient maps and keypoints of
    a fresh build, at the same device addresses."""
    a = np.stack([synth.frame(i, 135, 240) for i in (0, 5)], 0)
    b = np.stack([synth.frame(i, 135, 240) for i in (6, 3)], 0)
    pyr = ctx.build_pyramid(a, 4, 1.6, materialise=0x78)
    sz = C_size_t()
    addr = vs.load().vslam_pyramid_device_map(pyr.h, 1, vs.MAP_GAUSS, 2, 3, sz)
    ctx.rebuild_pyramid(pyr, b)
    assert vs.load().vslam_pyramid_device_map(pyr.h, 1, vs.MAP_GAUSS, 2, 3, sz) == addr
    fresh = ctx.build_pyramid(b, 4, 1.6, materialise=0x78)
    for f in (0, 1):
        for o in range(4):
            for kind, levels in ((vs.MAP_GAUSS, 6), (vs.MAP_DOG, 5), (vs.MAP_GRADX, 6), (vs.MAP_ORIENT, 6)):
                for l in range(levels):
                    assert np.array_equal(pyr.get(kind, o, l, f), fresh.get(kind, o, l, f)), (f, o, kind, l)
    assert np.array_equal(ctx.dog_detect(pyr)["keypoints"], ctx.dog_detect(fresh)["keypoints"])


def test_dog_detect_batch_matches_per_frame(ctx):
    """vslam_dog_detect on a batch (one detector launch per octave for all frames, one compaction pass) gives, frame by frame, the
    lists, counts and stage counters of single-frame calls -- and the list of the fused throughput path."""
    ids = (0, 5, 6, 3, 2)
    frames = np.stack([synth.frame(i, 135, 240) for i in ids], 0)
    pyr = ctx.build_pyramid(frames, 4, 1.6)
    batch = ctx.dog_detect(pyr)
    assert len(batch["counts"]) == len(ids)
    lists, at = [], 0
    for f, i in enumerate(ids):
        one = ctx.dog_detect(ctx.build_pyramid(frames[f], 4, 1.6))
        n = batch["counts"][f]
        assert one["counts"] == [n], f
        assert np.array_equal(kp_array(batch["keypoints"][at:at + n]), kp_array(one["keypoints"])), f
        assert np.array_equal(np.asarray(batch["stage_counts"][f]), np.asarray(one["stage_counts"][0])), f
        lists.append(kp_array(one["keypoints"]))
        at += n
    assert at == len(batch["keypoints"])
    fused = ctx.dog_keypoints(frames)
    assert fused["counts"] == batch["counts"]
    assert np.array_equal(kp_array(fused["keypoints"]), np.concatenate(lists, 0))
