// noalloc.cu -- vslam_harris_response_static: the pre-blur + Sobel + HarrisCorner chain of ./Harris
// (Harris_corners.cpp:158-164, 10-68) for callers that must not allocate: code under CUDA graph capture, tight serving
// loops, harnesses that forbid cudaMalloc.  No counterpart in the reference.
//
// The existing route to the same map (vslam_gaussian_blur3 -> vslam_sobel1 -> vslam_harris_response, or vslam_harris
// with `response` set) stages its intermediates in grow-only cudaMalloc'd context buffers.  Here the intermediates -- the
// blurred image, Ix, Iy, and one counter -- live in a file-scope `static __device__` array of a fixed maximum size that
// every GPU which loads the library holds until the process exits.  The entry point calls no cudaMalloc / cudaFree and no
// allocating library routine; it enqueues four small operations on the context's stream and returns.
//
//   * the array's name is a HOST-SIDE STAND-IN in host code, not a device address: the address every kernel receives
//     comes from cudaGetSymbolAddress, asked once per device with that device current (scratch_for below);
//   * every call shares the array, and it keeps the previous call's contents: the counter is cleared at the start of
//     each call (the three planes are fully overwritten before they are read);
//   * two calls that overlap on the GPU (different streams / contexts on one device) race on it.  Calls on one stream
//     are ordered and safe.
#include <mutex>

#include "common.cuh"

namespace vslam {
namespace {

constexpr size_t kMaxPixels = (size_t)VSLAM_STATIC_MAX_PIXELS;
constexpr size_t kHeadBytes = 256;                         // [0] = counter of positive responses (int), rest unused
constexpr size_t kBlurOff = kHeadBytes;                    // u8  [kMaxPixels]
constexpr size_t kIxOff = kBlurOff + kMaxPixels;           // f32 [kMaxPixels]
constexpr size_t kIyOff = kIxOff + 4 * kMaxPixels;         // f32 [kMaxPixels]
constexpr size_t kScratchBytes = kIyOff + 4 * kMaxPixels;  // 256 + 9 bytes per pixel
static_assert(kMaxPixels % 256 == 0, "planes stay 256-byte aligned");

static __device__ __align__(256) unsigned char g_static_scratch[kScratchBytes];

constexpr int kMaxDevices = 64;
std::mutex g_mu;
unsigned char *g_scratch_addr[kMaxDevices] = {nullptr};

// device address of g_static_scratch on `device` (which must be current): asked once per device
int scratch_for(vslam_ctx *ctx, int device, unsigned char **out) {
    if (device < 0 || device >= kMaxDevices) VS_FAIL(ctx, VSLAM_E_INVALID, "static scratch: device %d out of range", device);
    std::lock_guard<std::mutex> lock(g_mu);
    if (!g_scratch_addr[device]) {
        void *p = nullptr;
        VS_CUDA(ctx, cudaGetSymbolAddress(&p, g_static_scratch));
        if (!p) VS_FAIL(ctx, VSLAM_E_CUDA, "static scratch: cudaGetSymbolAddress returned a null address");
        g_scratch_addr[device] = (unsigned char *)p;
    }
    *out = g_scratch_addr[device];
    return VSLAM_OK;
}

// GaussianBlur(Size(3,3), 0) on 8U (Harris_corners.cpp:158): (S + 8) >> 4, S = [1 2 1; 2 4 2; 1 2 1], reflect-101
__global__ void __launch_bounds__(128) st_blur3_k(const u8 *__restrict__ src, int rows, int cols, size_t pitch, u8 *__restrict__ dst) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x >= cols) return;
    const int xl = refl101(x - 1, cols), xr = refl101(x + 1, cols);
    int s = 0;
#pragma unroll
    for (int u = -1; u <= 1; ++u) {
        const u8 *r = src + (size_t)refl101(y + u, rows) * pitch;
        s += (u == 0 ? 2 : 1) * ((int)r[xl] + 2 * (int)r[x] + (int)r[xr]);
    }
    dst[(size_t)y * cols + x] = (u8)((s + 8) >> 4);
}

// Sobel(CV_32F, 1,0 / 0,1, ksize = 1) (Harris_corners.cpp:161-164): central differences, reflect-101
__global__ void __launch_bounds__(128) st_sobel1_k(const u8 *__restrict__ b, int rows, int cols, float *__restrict__ gx, float *__restrict__ gy) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x >= cols) return;
    const u8 *r = b + (size_t)y * cols;
    const size_t o = (size_t)y * cols + x;
    gx[o] = (float)((int)r[refl101(x + 1, cols)] - (int)r[refl101(x - 1, cols)]);
    gy[o] = (float)((int)b[(size_t)refl101(y + 1, rows) * cols + x] - (int)b[(size_t)refl101(y - 1, rows) * cols + x]);
}

// StructureMatrix + HarrisCorner (Harris_corners.cpp:10-68): closed (2 pad + 1)^2 window on the replicate-padded maps, float
// accumulation in the reference's order, det through double, response = det - k tr^2, negatives to 0.  `positive` counts
// the responses above zero (one atomic per warp).
__global__ void __launch_bounds__(128) st_harris_k(const float *__restrict__ ix, const float *__restrict__ iy, int rows, int cols, int pad,
                                                   float k, float *__restrict__ out, int *__restrict__ positive) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    float resp = 0.f;
    if (x < cols) {
        float a = 0.f, b = 0.f, c = 0.f;
        for (int u = y - pad; u <= y + pad; ++u)
            for (int v = x - pad; v <= x + pad; ++v) {
                const size_t o = (size_t)clampi(u, 0, rows - 1) * cols + clampi(v, 0, cols - 1);
                const float p = ix[o], q = iy[o];
                a = __fadd_rn(a, __fmul_rn(p, p));
                b = __fadd_rn(b, __fmul_rn(q, q));
                c = __fadd_rn(c, __fmul_rn(p, q));
            }
        const float det = __double2float_rn(__dsub_rn(__dmul_rn((double)a, (double)b), __dmul_rn((double)c, (double)c)));
        const float tr = __fadd_rn(a, b);
        resp = __fsub_rn(det, __fmul_rn(k, __fmul_rn(tr, tr)));
        resp = resp > 0.f ? resp : 0.f;
        out[(size_t)y * cols + x] = resp;
    }
    const unsigned m = __ballot_sync(0xffffffffu, resp > 0.f);
    if ((threadIdx.x & 31) == 0 && m) atomicAdd(positive, __popc(m));
}

}  // namespace
}  // namespace vslam

using namespace vslam;

extern "C" int vslam_harris_response_static(vslam_ctx *ctx, const uint8_t *d_image, int rows, int cols, size_t pitch, int window, float k,
                                            float *d_response, int *d_positive) {
    if (!ctx) return VSLAM_E_INVALID;
    if (!d_image || !d_response) VS_FAIL(ctx, VSLAM_E_INVALID, "harris_response_static: null buffer");
    if (rows <= 0 || cols <= 0 || rows > 65535) VS_FAIL(ctx, VSLAM_E_INVALID, "harris_response_static: image size %dx%d", rows, cols);
    if ((size_t)rows * (size_t)cols > kMaxPixels)
        VS_FAIL(ctx, VSLAM_E_INVALID, "harris_response_static: %d x %d = %zu pixels exceeds the static scratch (VSLAM_STATIC_MAX_PIXELS = %zu); "
                "use vslam_harris", rows, cols, (size_t)rows * (size_t)cols, kMaxPixels);
    if (pitch < (size_t)cols) VS_FAIL(ctx, VSLAM_E_INVALID, "harris_response_static: pitch smaller than cols");
    if (window < 1 || !(window & 1) || window > 31) VS_FAIL(ctx, VSLAM_E_INVALID, "harris_response_static: window %d", window);
    VS_CUDA(ctx, cudaSetDevice(ctx->device));
    unsigned char *s = nullptr;
    VS_TRY(scratch_for(ctx, ctx->device, &s));
    int *cnt = (int *)s;
    u8 *blur = s + kBlurOff;
    float *ix = (float *)(s + kIxOff), *iy = (float *)(s + kIyOff);
    const dim3 grid((cols + 127) / 128, rows);
    // the array keeps the previous call's contents: the counter must start at zero
    VS_CUDA(ctx, cudaMemsetAsync(cnt, 0, kHeadBytes, ctx->stream));
    st_blur3_k<<<grid, 128, 0, ctx->stream>>>(d_image, rows, cols, pitch, blur);
    VS_LAUNCH_CHECK(ctx);
    st_sobel1_k<<<grid, 128, 0, ctx->stream>>>(blur, rows, cols, ix, iy);
    VS_LAUNCH_CHECK(ctx);
    st_harris_k<<<grid, 128, 0, ctx->stream>>>(ix, iy, rows, cols, (window - 1) / 2, k, d_response, cnt);
    VS_LAUNCH_CHECK(ctx);
    if (d_positive) VS_CUDA(ctx, cudaMemcpyAsync(d_positive, cnt, sizeof(int), cudaMemcpyDeviceToDevice, ctx->stream));
    return VSLAM_OK;
}
