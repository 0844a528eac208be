// pyramid.cu -- sm_100a kernels for the Gaussian scale-space pyramid.
//
//   upsample2x_k   row A   GaussPyramid.cpp:110      cv::resize x2 INTER_LINEAR, 8U, 11-bit coefficients
//   grad_k         row E   GaussPyramid.cpp:86-97    Sobel(ksize=1) central differences, magnitude, phase
//
// The Gaussian stack itself (rows B, C, D) lives in fused.cu.
#include "pyramid.cuh"
#include "grad_math.cuh"

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

__global__ void __launch_bounds__(256) upsample2x_generic_k(const u8 *__restrict__ src, int rows, int cols, size_t sp, size_t sstride,
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

// Fast path (cols % 4 == 0, 4-byte aligned rows): a thread owns 4 source columns = 8 output columns and marches down
// UP_ROWS source rows.  With q = 3*near + far (the horizontal pass, 512*q == the 11-bit fixed-point row sum, so
// (row sum >> 4) == 32*q exactly), an output pixel is ((3*q_near_row >> 2) + (q_far_row >> 2) + 2) >> 2: the two
// separately truncated vertical products of OpenCV.  Everything is evaluated on 16x2 packed halves; border clamps of
// the coefficients are index clamps here (3*p + p == 4*p).
constexpr int UP_ROWS = 16;

__device__ __forceinline__ void up_q(const u8 *__restrict__ row, int x4, int cols, u32 (&f1)[4], u32 (&f3)[4]) {
    const u32 w = *(const u32 *)(row + x4);
    const u32 pm = row[max(x4 - 1, 0)], pp = row[min(x4 + 4, cols - 1)];
    const u32 q0 = 3 * __byte_perm(w, pm, 0x5050) + __byte_perm(w, pm, 0x5154);    // (pm + 3 p0, 3 p0 + p1)
    const u32 q1 = 3 * __byte_perm(w, pm, 0x5151) + __byte_perm(w, pm, 0x5250);    // (p0 + 3 p1, 3 p1 + p2)
    const u32 q2 = 3 * __byte_perm(w, pm, 0x5252) + __byte_perm(w, pm, 0x5351);    // (p1 + 3 p2, 3 p2 + p3)
    const u32 q3 = 3 * __byte_perm(w, pm, 0x5353) + __byte_perm(w, pp, 0x5452);    // (p2 + 3 p3, 3 p3 + pp)
    const u32 q[4] = {q0, q1, q2, q3};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        f1[i] = (q[i] >> 2) & 0x3fff3fffu;              // floor(q / 4)      <- coefficient 512
        f3[i] = ((3 * q[i]) >> 2) & 0x3fff3fffu;        // floor(3 q / 4)    <- coefficient 1536
    }
}

__device__ __forceinline__ void up_store(u8 *__restrict__ o, const u32 (&near3)[4], const u32 (&far1)[4]) {
    u32 v[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) v[i] = ((near3[i] + far1[i] + 0x00020002u) >> 2) & 0x00ff00ffu;
    uint2 r;
    r.x = __byte_perm(v[0], v[1], 0x6420);
    r.y = __byte_perm(v[2], v[3], 0x6420);
    *(uint2 *)o = r;
}

__global__ void __launch_bounds__(128) upsample2x_k(const u8 *__restrict__ src, int rows, int cols, size_t sp, size_t sstride,
                                                    u8 *__restrict__ dst, size_t dp, size_t dstride) {
    src += (size_t)blockIdx.z * sstride;
    dst += (size_t)blockIdx.z * dstride;
    const int x4 = (blockIdx.x * 128 + threadIdx.x) * 4;
    if (x4 >= cols) return;
    const int k0 = blockIdx.y * UP_ROWS, k1 = min(k0 + UP_ROWS, rows);
    u32 p1[4], p3[4], c1[4], c3[4], n1[4], n3[4];       // rows k-1, k, k+1
    up_q(src + (size_t)max(k0 - 1, 0) * sp, x4, cols, p1, p3);
    up_q(src + (size_t)k0 * sp, x4, cols, c1, c3);
    for (int k = k0; k < k1; ++k) {
        up_q(src + (size_t)min(k + 1, rows - 1) * sp, x4, cols, n1, n3);
        u8 *o = dst + (size_t)(2 * k) * dp + 2 * x4;
        up_store(o, c3, p1);                             // output row 2k:   rows (k-1, k) with coefficients (512, 1536)
        up_store(o + dp, c3, n1);                        // output row 2k+1: rows (k, k+1) with coefficients (1536, 512)
#pragma unroll
        for (int i = 0; i < 4; ++i) { p1[i] = c1[i]; p3[i] = c3[i]; c1[i] = n1[i]; c3[i] = n3[i]; }
    }
}

int launch_upsample2x(vslam_ctx *ctx, const u8 *src, int rows, int cols, size_t sp, size_t sstride, u8 *dst, size_t dp,
                      size_t dstride, int n_frames) {
    KernelSpan span(ctx, VSLAM_K_UPSAMPLE);
    if (!(cols & 3) && !(sp & 3) && !(sstride & 3) && !((uintptr_t)src & 3) && !(dp & 7) && !(dstride & 7) && !((uintptr_t)dst & 7)) {
        dim3 g((cols / 4 + 127) / 128, (rows + UP_ROWS - 1) / UP_ROWS, n_frames);
        upsample2x_k<<<g, 128, 0, ctx->stream>>>(src, rows, cols, sp, sstride, dst, dp, dstride);
    } else {
        dim3 b(64, 4), g((2 * cols + 255) / 256, (2 * rows + 3) / 4, n_frames);
        upsample2x_generic_k<<<g, b, 0, ctx->stream>>>(src, rows, cols, sp, sstride, dst, dp, dstride);
    }
    VS_LAUNCH_CHECK(ctx);
    return VSLAM_OK;
}

// ------------------------------------------------------------------------------------------------------------
// row E: Sobel(CV_32F, ksize=1) = central differences with reflect-101 borders (zero on the first/last
// column / row), cv::magnitude, cv::phase(degrees) (OpenCV's fastAtan polynomial, evaluated without FMA).
// ------------------------------------------------------------------------------------------------------------
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

// magnitude and first-quadrant angle of every (|dx|, |dy|) in 0..255 x 0..255, filled by the functions grad_k evaluates per pixel
// (correctly rounded square root, OpenCV's fastAtan2 polynomial with its IEEE division): 512 KB, L2 resident.  grad_all_k is
// instruction bound without it (113 thread instructions per pixel, 90 % issue utilisation, two thirds of them this arithmetic).
__global__ void __launch_bounds__(256) grad_lut_fill_k(float2 *__restrict__ lut) {
    const int ax = threadIdx.x, ay = blockIdx.x;
    const float fx = (float)ax, fy = (float)ay;
    lut[ay * 256 + ax] = make_float2(__fsqrt_rn(__fadd_rn(__fmul_rn(fx, fx), __fmul_rn(fy, fy))), fast_atan2_deg(fy, fx));
}

// The same for every level of one octave of a batch of frames, into resident fp32 planes (vslam_pyramid_build with a
// materialise mask: what processGradients leaves behind, GaussPyramid.cpp:65-104).  Thread = 4 adjacent pixels of one
// level: 16-byte stores per map, 64 B written per 4 B read -- an HBM write stream.
// grid = (column quads / 64, rows / 4, frames * 6); Gaussian level l of frame f at g + f * gstride + l * gplane.
__global__ void __launch_bounds__(256) grad_all_k(const u8 *__restrict__ g, int rows, int cols, size_t pitch, size_t gplane, size_t gstride,
                                                  float *__restrict__ out, size_t oplane, size_t ostride, unsigned mask,
                                                  const float2 *__restrict__ lut) {
    const int xq = blockIdx.x * 64 + (threadIdx.x & 63), y = blockIdx.y * 4 + (threadIdx.x >> 6);
    const int x = 4 * xq;
    if (x >= cols || y >= rows) return;
    const int f = blockIdx.z / 6, l = blockIdx.z - 6 * f;
    const u8 *img = g + (size_t)f * gstride + (size_t)l * gplane;
    const u8 *row = img + (size_t)y * pitch, *up = img + (size_t)refl101(y - 1, rows) * pitch, *dn = img + (size_t)refl101(y + 1, rows) * pitch;
    int ix[4], iy[4];
    if (x + 3 < cols) {                                 // planes are 128-byte pitched: aligned words
        const unsigned c = *(const unsigned *)(row + x), u = *(const unsigned *)(up + x), d = *(const unsigned *)(dn + x);
        const int left = row[refl101(x - 1, cols)], right = row[refl101(x + 4, cols)];
        const int p[6] = {left, (int)(c & 255), (int)(c >> 8 & 255), (int)(c >> 16 & 255), (int)(c >> 24), right};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            ix[j] = p[j + 2] - p[j];
            iy[j] = (int)(d >> (8 * j) & 255) - (int)(u >> (8 * j) & 255);
        }
    } else {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int xx = min(x + j, cols - 1);
            ix[j] = (int)row[refl101(xx + 1, cols)] - (int)row[refl101(xx - 1, cols)];
            iy[j] = (int)dn[xx] - (int)up[xx];
        }
    }
    // maps of a level: [gx, gy, mag, orient] planes of oplane floats each, dense rows of `cols` floats
    float *o = out + (size_t)f * ostride + (size_t)l * 4 * oplane + (size_t)y * cols + x;
    float v[4][4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        v[0][j] = (float)ix[j]; v[1][j] = (float)iy[j];
        const float2 e = __ldg(lut + (abs(iy[j]) << 8 | abs(ix[j])));       // magnitude, angle of (|dx|, |dy|)
        float a = e.y;                                                     // fastAtan2's quadrant steps
        if (ix[j] < 0) a = __fsub_rn(180.f, a);
        if (iy[j] < 0) a = __fsub_rn(360.f, a);
        v[2][j] = e.x; v[3][j] = a;
    }
    const bool vec = x + 3 < cols && !(cols & 3);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        if (!(mask >> k & 1)) continue;
        float *ok = o + (size_t)k * oplane;
        if (vec) __stcs((float4 *)ok, make_float4(v[k][0], v[k][1], v[k][2], v[k][3]));      // written once, read by later kernels: streaming
        else for (int j = 0; j < 4 && x + j < cols; ++j) ok[j] = v[k][j];
    }
}

int launch_gradients_all(vslam_ctx *ctx, const u8 *g, int rows, int cols, size_t pitch, size_t gplane, size_t gstride, float *out,
                         size_t oplane, size_t ostride, unsigned mask, int n_frames) {
    if (!ctx->grad_lut.p) {                             // once per context
        if (ctx->grad_lut.reserve(65536 * sizeof(float2))) VS_FAIL(ctx, VSLAM_E_NOMEM, "gradient table");
        grad_lut_fill_k<<<256, 256, 0, ctx->stream>>>(ctx->grad_lut.as<float2>());
        VS_LAUNCH_CHECK(ctx);
    }
    for (int f0 = 0; f0 < n_frames; f0 += 10000) {      // gridDim.z <= 65535
        const int nf = n_frames - f0 < 10000 ? n_frames - f0 : 10000;
        dim3 grid(((cols + 3) / 4 + 63) / 64, (rows + 3) / 4, nf * 6);
        KernelSpan span(ctx, VSLAM_K_GRAD);
        grad_all_k<<<grid, 256, 0, ctx->stream>>>(g + (size_t)f0 * gstride, rows, cols, pitch, gplane, gstride, out + (size_t)f0 * ostride, oplane, ostride, mask,
                                                  ctx->grad_lut.as<float2>());
        VS_LAUNCH_CHECK(ctx);
    }
    return VSLAM_OK;
}

int launch_gradients(vslam_ctx *ctx, const u8 *g, int rows, int cols, size_t pitch, float *gx, float *gy, float *mag,
                     float *orient, size_t op) {
    dim3 grid((cols + 63) / 64, (rows + 3) / 4);
    KernelSpan span(ctx, VSLAM_K_GRAD);
    grad_k<<<grid, 256, 0, ctx->stream>>>(g, rows, cols, pitch, gx, gy, mag, orient, op);
    VS_LAUNCH_CHECK(ctx);
    return VSLAM_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// Exact slow path for the octaves whose 8U Gaussian kernels the fused kernels do not take (more than 247 taps: octave 4 and
// beyond at sigma0 = 1.6, i.e. the reference's two-argument GaussPyramid(img, sigma) on anything larger than 511 pixels on its
// short side).  Such octaves are tiny (at most 1/256 of the input's pixels), so this is OpenCV's arithmetic in its plainest form:
// a horizontal pass u8 x 8-bit taps -> 8.8 sums (u16), a vertical pass 8.8 x taps -> 16.16 and one rounding, reflect-101 with
// repeated bouncing (the kernels are several times wider than the image), saturating DoG, nearest decimation of level 3.
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) wide_h_k(const u8 *__restrict__ base, int rows, int cols, size_t pitch, size_t stride,
                                                const int *__restrict__ taps, int n, u16 *__restrict__ out) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y, f = blockIdx.z;
    if (x >= cols) return;
    const u8 *r = base + (size_t)f * stride + (size_t)y * pitch;
    const int h = n / 2;
    int acc = 0;
    for (int i = 0; i < n; ++i) acc += taps[i] * (int)r[refl101(x + i - h, cols)];
    out[((size_t)f * rows + y) * cols + x] = (u16)acc;                     // <= 255 * 256
}
__global__ void __launch_bounds__(128) wide_v_k(const u16 *__restrict__ hsum, int rows, int cols, const int *__restrict__ taps, int n,
                                                u8 *__restrict__ dst, size_t dpitch, size_t dstride) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y, f = blockIdx.z;
    if (x >= cols) return;
    const u16 *p = hsum + (size_t)f * rows * cols + x;
    const int h = n / 2;
    unsigned acc = 0;
    for (int i = 0; i < n; ++i) acc += (unsigned)taps[i] * (unsigned)p[(size_t)refl101(y + i - h, rows) * cols];
    dst[(size_t)f * dstride + (size_t)y * dpitch + x] = (u8)((acc + 32768u) >> 16);
}
// out = saturate(cur - prev)  (GaussPyramid.cpp:197)
__global__ void __launch_bounds__(128) wide_dog_k(const u8 *__restrict__ cur, const u8 *__restrict__ prev, int rows, int cols, size_t pitch, size_t stride,
                                                  u8 *__restrict__ out) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x >= cols) return;
    const size_t o = (size_t)blockIdx.z * stride + (size_t)y * pitch + x;
    const int d = (int)cur[o] - (int)prev[o];
    out[o] = (u8)(d > 0 ? d : 0);
}
// resize(level 3, 0.5, INTER_NEAREST)  (GaussPyramid.cpp:123-126)
__global__ void __launch_bounds__(128) wide_decimate_k(const u8 *__restrict__ src, int srows, int scols, size_t spitch, size_t sstride, u8 *__restrict__ dst,
                                                       int drows, int dcols, size_t dpitch, size_t dstride) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x >= dcols || y >= drows) return;
    const int sx = min(2 * x, scols - 1), sy = min(2 * y, srows - 1);      // the rounded-up half size reaches one past the last even index
    dst[(size_t)blockIdx.z * dstride + (size_t)y * dpitch + x] = src[(size_t)blockIdx.z * sstride + (size_t)sy * spitch + sx];
}

int launch_octave_wide(vslam_ctx *ctx, const OctavePlanes &op, size_t frame_stride, const int *d_taps, const int *tap_off, const int *ksize,
                       u16 *d_hsum, int n_frames) {
    if (n_frames <= 0 || n_frames > 65535 || op.rows > 65535) VS_FAIL(ctx, VSLAM_E_INVALID, "wide octave: %d frames of %d rows", n_frames, op.rows);
    const dim3 grid((op.cols + 127) / 128, op.rows, n_frames);
    KernelSpan span(ctx, VSLAM_K_FUSED);
    for (int l = 0; l < 6; ++l) {
        wide_h_k<<<grid, 128, 0, ctx->stream>>>(op.base, op.rows, op.cols, op.pitch, frame_stride, d_taps + tap_off[l], ksize[l], d_hsum);
        VS_LAUNCH_CHECK(ctx);
        wide_v_k<<<grid, 128, 0, ctx->stream>>>(d_hsum, op.rows, op.cols, d_taps + tap_off[l], ksize[l], op.gauss[l], op.pitch, frame_stride);
        VS_LAUNCH_CHECK(ctx);
        if (l >= 1) {
            wide_dog_k<<<grid, 128, 0, ctx->stream>>>(op.gauss[l], op.gauss[l - 1], op.rows, op.cols, op.pitch, frame_stride, op.dog[l - 1]);
            VS_LAUNCH_CHECK(ctx);
        }
    }
    if (op.next_base) {
        const dim3 g2((op.next_cols + 127) / 128, op.next_rows, n_frames);
        wide_decimate_k<<<g2, 128, 0, ctx->stream>>>(op.gauss[3], op.rows, op.cols, op.pitch, frame_stride, op.next_base, op.next_rows, op.next_cols, op.next_pitch,
                                                     frame_stride);
        VS_LAUNCH_CHECK(ctx);
    }
    return VSLAM_OK;
}

}  // namespace vslam
