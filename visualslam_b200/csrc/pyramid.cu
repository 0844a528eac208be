// pyramid.cu -- sm_100a kernels for the Gaussian scale-space pyramid.
//
//   upsample2x_k   row A   GaussPyramid.cpp:110      cv::resize x2 INTER_LINEAR, 8U, 11-bit coefficients
//   grad_k         row E   GaussPyramid.cpp:86-97    Sobel(ksize=1) central differences, magnitude, phase
//
// The Gaussian stack itself (rows B, C, D) lives in fused.cu.
#include "pyramid.cuh"

namespace vslam {

// ------------------------------------------------------------------------------------------------------------
// row A: 2x bilinear upsample.  One thread -> 4 consecutive output pixels of one output row.
// ------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void up_coef(int d, int &s, int &c0, int &c1) {
    // f = (d+0.5)*0.5-0.5; s = floor(f); frac in {0.75, 0.25}; coefficients rint(frac*2048)
    s = (d >> 1) - ((d & 1) ? 0 : 1);
    c1 = (d & 1) ? 512 : 1536;
    c0 = 2048 - c1;
}

__global__ void __launch_bounds__(256) upsample2x_k(const u8 *__restrict__ src, int rows, int cols, size_t sp, size_t sstride,
                                                    u8 *__restrict__ dst, size_t dp, size_t dstride) {
    src += (size_t)blockIdx.z * sstride;
    dst += (size_t)blockIdx.z * dstride;
    const int drows = 2 * rows, dcols = 2 * cols;
    const int x4 = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
    const int y = blockIdx.y * blockDim.y + threadIdx.y;
    if (y >= drows || x4 >= dcols) return;
    int sy, b0, b1;
    up_coef(y, sy, b0, b1);                                  // vertical: coefficients are NOT clamped, rows are
    const u8 *r0 = src + (size_t)clampi(sy, 0, rows - 1) * sp;
    const u8 *r1 = src + (size_t)clampi(sy + 1, 0, rows - 1) * sp;
    u32 packed = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int d = x4 + i;
        int s, a0, a1;
        up_coef(d, s, a0, a1);
        if (s < 0) { s = 0; a0 = 2048; a1 = 0; }             // horizontal: coefficient clamp at both ends
        if (s >= cols - 1) { s = cols - 1; a0 = 2048; a1 = 0; }
        const int s1 = min(s + 1, cols - 1);
        const int h0 = r0[s] * a0 + r0[s1] * a1;
        const int h1 = r1[s] * a0 + r1[s1] * a1;
        const int v = (((b0 * (h0 >> 4)) >> 16) + ((b1 * (h1 >> 4)) >> 16) + 2) >> 2;
        packed |= (u32)(v & 255) << (8 * i);
    }
    u8 *o = dst + (size_t)y * dp + x4;
    if (x4 + 3 < dcols) *(u32 *)o = packed;                  // dp and x4 are multiples of 4
    else for (int i = 0; x4 + i < dcols; ++i) o[i] = (u8)(packed >> (8 * i));
}

int launch_upsample2x(vslam_ctx *ctx, const u8 *src, int rows, int cols, size_t sp, size_t sstride, u8 *dst, size_t dp,
                      size_t dstride, int n_frames) {
    dim3 b(64, 4), g((2 * cols + 255) / 256, (2 * rows + 3) / 4, n_frames);
    KernelSpan span(ctx, VSLAM_K_UPSAMPLE);
    upsample2x_k<<<g, b, 0, ctx->stream>>>(src, rows, cols, sp, sstride, dst, dp, dstride);
    VS_LAUNCH_CHECK(ctx);
    return VSLAM_OK;
}

// ------------------------------------------------------------------------------------------------------------
// row E: Sobel(CV_32F, ksize=1) = central differences with reflect-101 borders (zero on the first/last
// column / row), cv::magnitude, cv::phase(degrees) (OpenCV's fastAtan polynomial, evaluated without FMA).
// ------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ float fast_atan2_deg(float y, float x) {
    const float scale = (float)(180.0 / 3.14159265358979323846);
    const float p1 = 0.9997878412794807f * scale, p3 = -0.3258083974640975f * scale,
                p5 = 0.1555786518463281f * scale, p7 = -0.04432655554792128f * scale;
    const float ax = fabsf(x), ay = fabsf(y);
    const float eps = (float)2.2204460492503131e-16;
    float a, c, c2;
    if (ax >= ay) {
        c = __fdiv_rn(ay, __fadd_rn(ax, eps));
        c2 = __fmul_rn(c, c);
        a = __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c);
    } else {
        c = __fdiv_rn(ax, __fadd_rn(ay, eps));
        c2 = __fmul_rn(c, c);
        a = __fsub_rn(90.f, __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c));
    }
    if (x < 0) a = __fsub_rn(180.f, a);
    if (y < 0) a = __fsub_rn(360.f, a);
    return a;
}

__global__ void __launch_bounds__(256) grad_k(const u8 *__restrict__ g, int rows, int cols, size_t pitch,
                                              float *__restrict__ gx, float *__restrict__ gy,
                                              float *__restrict__ mag, float *__restrict__ orient, size_t op) {
    const int x = blockIdx.x * 64 + (threadIdx.x & 63), y = blockIdx.y * 4 + (threadIdx.x >> 6);
    if (x >= cols || y >= rows) return;
    const u8 *row = g + (size_t)y * pitch;
    const float dx = (float)((int)row[refl101(x + 1, cols)] - (int)row[refl101(x - 1, cols)]);
    const float dy = (float)((int)g[(size_t)refl101(y + 1, rows) * pitch + x] - (int)g[(size_t)refl101(y - 1, rows) * pitch + x]);
    const size_t o = (size_t)y * op + x;
    if (gx) gx[o] = dx;
    if (gy) gy[o] = dy;
    if (mag) mag[o] = __fsqrt_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)));
    if (orient) orient[o] = fast_atan2_deg(dy, dx);
}

int launch_gradients(vslam_ctx *ctx, const u8 *g, int rows, int cols, size_t pitch, float *gx, float *gy, float *mag,
                     float *orient, size_t op) {
    dim3 grid((cols + 63) / 64, (rows + 3) / 4);
    KernelSpan span(ctx, VSLAM_K_GRAD);
    grad_k<<<grid, 256, 0, ctx->stream>>>(g, rows, cols, pitch, gx, gy, mag, orient, op);
    VS_LAUNCH_CHECK(ctx);
    return VSLAM_OK;
}

}  // namespace vslam
