// jpeg.cu -- the data format in front of the path: JPEG streams in host memory -> grey frames in HBM.
// The reference does imread(path, IMREAD_COLOR) + cvtColor(COLOR_BGR2GRAY) per image (Diff_of_Gauss.cpp:730-739,
// Harris_corners.cpp:148-149).  Here nvJPEG decodes each stream to interleaved BGR on the device and bgr2gray_k (stages.cu, the
// bit-exact cvtColor) turns it grey; the frames land where vslam_dog_keypoints / vslam_harris / vslam_dog_harris take them with
// VSLAM_DEVICE, so a stream of camera JPEGs never visits the host as pixels.
//
// NOT a parity target: a decoder's output is a fixture, not something the reference computes (nvJPEG's inverse DCT and chroma
// upsampling differ from libjpeg-turbo's by a few grey levels).  The parity tests keep using pre-decoded images; the test of this
// file holds the decode to cv2's within a tolerance.
//
// libnvjpeg is opened at run time (dlopen): the library loads and every other entry point works on a machine without it.
#include <dlfcn.h>
#include <nvjpeg.h>

#include <mutex>

#include "common.cuh"
#include "stages.cuh"

namespace vslam {
namespace {

struct NvJpeg {
    void *lib = nullptr;
    nvjpegStatus_t (*create)(nvjpegHandle_t *) = nullptr;
    nvjpegStatus_t (*destroy)(nvjpegHandle_t) = nullptr;
    nvjpegStatus_t (*state_create)(nvjpegHandle_t, nvjpegJpegState_t *) = nullptr;
    nvjpegStatus_t (*state_destroy)(nvjpegJpegState_t) = nullptr;
    nvjpegStatus_t (*info)(nvjpegHandle_t, const unsigned char *, size_t, int *, nvjpegChromaSubsampling_t *, int *, int *) = nullptr;
    nvjpegStatus_t (*decode)(nvjpegHandle_t, nvjpegJpegState_t, const unsigned char *, size_t, nvjpegOutputFormat_t, nvjpegImage_t *,
                             cudaStream_t) = nullptr;
    bool ok = false;
};

NvJpeg &nvjpeg_api() {
    static NvJpeg api;
    static std::once_flag once;
    std::call_once(once, [] {
        for (const char *name : {"libnvjpeg.so.12", "libnvjpeg.so", "/usr/local/cuda/lib64/libnvjpeg.so.12"}) {
            api.lib = dlopen(name, RTLD_NOW | RTLD_LOCAL);
            if (api.lib) break;
        }
        if (!api.lib) return;
        api.create = (decltype(api.create))dlsym(api.lib, "nvjpegCreateSimple");
        api.destroy = (decltype(api.destroy))dlsym(api.lib, "nvjpegDestroy");
        api.state_create = (decltype(api.state_create))dlsym(api.lib, "nvjpegJpegStateCreate");
        api.state_destroy = (decltype(api.state_destroy))dlsym(api.lib, "nvjpegJpegStateDestroy");
        api.info = (decltype(api.info))dlsym(api.lib, "nvjpegGetImageInfo");
        api.decode = (decltype(api.decode))dlsym(api.lib, "nvjpegDecode");
        api.ok = api.create && api.destroy && api.state_create && api.state_destroy && api.info && api.decode;
    });
    return api;
}

}  // namespace

// called by vslam_ctx_destroy
void jpeg_release(vslam_ctx *ctx) {
    NvJpeg &J = nvjpeg_api();
    if (ctx->jpeg_state && J.ok) J.state_destroy((nvjpegJpegState_t)ctx->jpeg_state);
    if (ctx->jpeg_handle && J.ok) J.destroy((nvjpegHandle_t)ctx->jpeg_handle);
    ctx->jpeg_state = ctx->jpeg_handle = nullptr;
    ctx->jpeg_bgr.release();
}

}  // namespace vslam

using namespace vslam;

extern "C" int vslam_jpeg_info(vslam_ctx *ctx, const uint8_t *stream, size_t length, int *rows, int *cols, int *channels) {
    if (!ctx || !stream || !length || !rows || !cols) return VSLAM_E_INVALID;
    NvJpeg &J = nvjpeg_api();
    if (!J.ok) VS_FAIL(ctx, VSLAM_E_CUDA, "libnvjpeg.so.12 not found (dlopen): JPEG ingest is unavailable on this machine");
    VS_CUDA(ctx, cudaSetDevice(ctx->device));
    if (!ctx->jpeg_handle && J.create((nvjpegHandle_t *)&ctx->jpeg_handle) != NVJPEG_STATUS_SUCCESS) VS_FAIL(ctx, VSLAM_E_CUDA, "nvjpegCreateSimple failed");
    int nc = 0, w[NVJPEG_MAX_COMPONENT] = {0}, h[NVJPEG_MAX_COMPONENT] = {0};
    nvjpegChromaSubsampling_t ss;
    const nvjpegStatus_t st = J.info((nvjpegHandle_t)ctx->jpeg_handle, stream, length, &nc, &ss, w, h);
    if (st != NVJPEG_STATUS_SUCCESS) VS_FAIL(ctx, VSLAM_E_INVALID, "not a JPEG stream nvJPEG can read (status %d)", (int)st);
    *rows = h[0]; *cols = w[0];
    if (channels) *channels = nc;
    return VSLAM_OK;
}

extern "C" int vslam_decode_jpeg_gray(vslam_ctx *ctx, const uint8_t *const *streams, const size_t *lengths, int n, int rows, int cols, uint8_t *gray,
                                      size_t gray_pitch, size_t gray_stride, int space) {
    if (!ctx || !streams || !lengths || !gray || n <= 0) return VSLAM_E_INVALID;
    if (rows <= 0 || cols <= 0 || rows > 65535 || cols > (1 << 15)) VS_FAIL(ctx, VSLAM_E_INVALID, "decode_jpeg: image size %dx%d", rows, cols);
    if (gray_pitch < (size_t)cols || (n > 1 && gray_stride < gray_pitch * (size_t)rows)) VS_FAIL(ctx, VSLAM_E_INVALID, "decode_jpeg: pitch / stride too small");
    if (space != VSLAM_HOST && space != VSLAM_DEVICE) VS_FAIL(ctx, VSLAM_E_INVALID, "decode_jpeg: space %d", space);
    NvJpeg &J = nvjpeg_api();
    if (!J.ok) VS_FAIL(ctx, VSLAM_E_CUDA, "libnvjpeg.so.12 not found (dlopen): JPEG ingest is unavailable on this machine");
    VS_CUDA(ctx, cudaSetDevice(ctx->device));
    if (!ctx->jpeg_handle && J.create((nvjpegHandle_t *)&ctx->jpeg_handle) != NVJPEG_STATUS_SUCCESS) VS_FAIL(ctx, VSLAM_E_CUDA, "nvjpegCreateSimple failed");
    if (!ctx->jpeg_state && J.state_create((nvjpegHandle_t)ctx->jpeg_handle, (nvjpegJpegState_t *)&ctx->jpeg_state) != NVJPEG_STATUS_SUCCESS)
        VS_FAIL(ctx, VSLAM_E_CUDA, "nvjpegJpegStateCreate failed");
    // device staging: n interleaved BGR frames (+ the grey frames when the caller wants them on the host)
    const size_t bp = round_up(3 * (size_t)cols, 256), bs = bp * (size_t)rows;
    const size_t gp = space == VSLAM_DEVICE ? gray_pitch : round_up((size_t)cols, 256), gs = space == VSLAM_DEVICE ? gray_stride : gp * (size_t)rows;
    if (ctx->jpeg_bgr.reserve(bs * (size_t)n + (space == VSLAM_DEVICE ? 0 : gs * (size_t)n))) VS_FAIL(ctx, VSLAM_E_NOMEM, "JPEG staging (%zu bytes)", bs * (size_t)n);
    u8 *d_bgr = ctx->jpeg_bgr.as<u8>();
    u8 *d_gray = space == VSLAM_DEVICE ? gray : d_bgr + bs * (size_t)n;
    for (int f = 0; f < n; ++f) {
        if (!streams[f] || !lengths[f]) VS_FAIL(ctx, VSLAM_E_INVALID, "decode_jpeg: stream %d is empty", f);
        int nc = 0, w[NVJPEG_MAX_COMPONENT] = {0}, h[NVJPEG_MAX_COMPONENT] = {0};
        nvjpegChromaSubsampling_t ss;
        if (J.info((nvjpegHandle_t)ctx->jpeg_handle, streams[f], lengths[f], &nc, &ss, w, h) != NVJPEG_STATUS_SUCCESS)
            VS_FAIL(ctx, VSLAM_E_INVALID, "decode_jpeg: stream %d is not a JPEG nvJPEG can read", f);
        if (h[0] != rows || w[0] != cols) VS_FAIL(ctx, VSLAM_E_INVALID, "decode_jpeg: stream %d is %dx%d, the batch is %dx%d", f, h[0], w[0], rows, cols);
        nvjpegImage_t out;
        memset(&out, 0, sizeof out);
        out.channel[0] = d_bgr + bs * (size_t)f;
        out.pitch[0] = bp;
        const nvjpegStatus_t st = J.decode((nvjpegHandle_t)ctx->jpeg_handle, (nvjpegJpegState_t)ctx->jpeg_state, streams[f], lengths[f], NVJPEG_OUTPUT_BGRI,
                                           &out, ctx->stream);
        if (st != NVJPEG_STATUS_SUCCESS) VS_FAIL(ctx, VSLAM_E_CUDA, "nvjpegDecode failed on stream %d (status %d)", f, (int)st);
    }
    VS_TRY(launch_bgr2gray(ctx, d_bgr, rows, cols, bp, bs, d_gray, gp, gs, n));      // cvtColor(COLOR_BGR2GRAY), bit-exact (stages.cu)
    if (space == VSLAM_HOST) {
        for (int f = 0; f < n; ++f)
            VS_CUDA(ctx, cudaMemcpy2DAsync(gray + gray_stride * f, gray_pitch, d_gray + gs * f, gp, (size_t)cols, (size_t)rows, cudaMemcpyDeviceToHost, ctx->stream));
        VS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    }
    return VSLAM_OK;
}
