"""GPU test of vslam_harris_response_static (include/vslam.h): the allocation-free variant of the Harris response chain whose
intermediates live in a file-scope static __device__ array (csrc/noalloc.cu).  It has no counterpart in the reference; it is
checked, through the C ABI, against the existing entry point vslam_harris (`response` map) -- bit-exact, the bar of every
other Harris test here -- on calls in a row that share the array (different sizes, so stale contents would show).

Collected last on purpose (file name).
"""

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_PIXELS = 2097152          # VSLAM_STATIC_MAX_PIXELS
E_INVALID = -1


def _image(seed, h, w):
    rng = np.random.default_rng(seed)
    img = (rng.integers(0, 256, (h, w)) // 32 * 32).astype(np.uint8)
    img[h // 4: h // 2, w // 3: 2 * w // 3] = 220          # a block with real corners
    img[h // 2:, : w // 5] = 15
    return img


def _static(ctx, lib, torch, img, pitch=None, window=3, k=0.04):
    """one call of the new entry point: device image (optionally with a row pitch) -> device response + counter"""
    h, w = img.shape
    pitch = pitch or w
    host = np.zeros((h, pitch), np.uint8)
    host[:, :w] = img
    d_img = torch.from_numpy(host).cuda()
    d_resp = torch.full((h, w), -7.0, dtype=torch.float32, device="cuda")
    d_cnt = torch.full((1,), -1, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()                                # the context's own stream is not ordered with torch's
    rc = lib.vslam_harris_response_static(ctx.h, d_img.data_ptr(), h, w, pitch, window, k, d_resp.data_ptr(), d_cnt.data_ptr())
    assert rc == 0, lib.vslam_last_error(ctx.h).decode()
    ctx.sync()
    return d_resp.cpu().numpy(), int(d_cnt.item())


def test_static_scratch_variant_matches_vslam_harris():
    import torch
    import visualslam_b200 as vs
    from visualslam_b200 import api
    lib = api.load()
    ctx = vs.Context(0)
    try:
        # calls in a row on the shared array: large, then smaller (stale planes / counter would show), then the first again
        cases = [(_image(1, 240, 320), None), (_image(2, 135, 201), 256), (_image(1, 240, 320), None), (_image(3, 1080, 1920), None)]
        first = None
        for n, (img, pitch) in enumerate(cases):
            got, positive = _static(ctx, lib, torch, img, pitch)
            want = ctx.harris(img, want_maps=True)["response"][0]
            assert got.shape == want.shape and np.array_equal(got, want), (n, img.shape, int((got != want).sum()))
            assert positive == int((want > 0).sum()), (n, positive)
            assert positive > 0
            if n == 0:
                first = got
            if n == 2:
                assert np.array_equal(got, first)
        # a different window goes through the same path as HarrisCorner with that window: spot-check it runs and differs
        got5, _ = _static(ctx, lib, torch, cases[1][0], 256, window=5)
        assert got5.shape == (135, 201) and not np.array_equal(got5, _static(ctx, lib, torch, cases[1][0], 256)[0])
    finally:
        ctx.close()


def test_static_scratch_variant_rejects_oversize_and_bad_arguments():
    import torch
    import visualslam_b200 as vs
    from visualslam_b200 import api
    lib = api.load()
    ctx = vs.Context(0)
    try:
        d = torch.zeros(4096, dtype=torch.uint8, device="cuda")
        r = torch.zeros(4096, dtype=torch.float32, device="cuda")
        torch.cuda.synchronize()
        rows, cols = 2048, 1025                                  # 2 099 200 pixels > VSLAM_STATIC_MAX_PIXELS
        assert rows * cols > MAX_PIXELS
        rc = lib.vslam_harris_response_static(ctx.h, d.data_ptr(), rows, cols, cols, 3, 0.04, r.data_ptr(), None)
        assert rc == E_INVALID
        msg = lib.vslam_last_error(ctx.h).decode()
        assert "VSLAM_STATIC_MAX_PIXELS" in msg and str(MAX_PIXELS) in msg, msg
        assert lib.vslam_harris_response_static(ctx.h, None, 8, 8, 8, 3, 0.04, r.data_ptr(), None) == E_INVALID
        assert lib.vslam_harris_response_static(ctx.h, d.data_ptr(), 8, 8, 4, 3, 0.04, r.data_ptr(), None) == E_INVALID   # pitch < cols
        assert lib.vslam_harris_response_static(ctx.h, d.data_ptr(), 8, 8, 8, 4, 0.04, r.data_ptr(), None) == E_INVALID   # even window
        # the largest accepted shape still runs after the rejections
        img = _image(4, 1024, 2048)
        assert img.size == MAX_PIXELS
        got, positive = _static(ctx, lib, torch, img)
        want = ctx.harris(img, want_maps=True)["response"][0]
        assert np.array_equal(got, want) and positive == int((want > 0).sum())
    finally:
        ctx.close()
