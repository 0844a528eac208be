// pyramid.cu -- sm_100a kernels for the Gaussian scale-space pyramid.
//
//   upsample2x_k   row A   GaussPyramid.cpp:110      cv::resize x2 INTER_LINEAR, 8U, 11-bit coefficients
//   hblur6_k       row B   GaussPyramid.cpp:172-177  horizontal 8.8 fixed-point pass of all 6 levels (dp4a)
//   vblur6_k       rows B, C, D  :177, :197, :126    vertical 16.16 pass + rounding (dp2a), saturating DoG,
//                                                    nearest 1/2 decimation of level 3
//   grad_k         row E   GaussPyramid.cpp:86-97    Sobel(ksize=1) central differences, magnitude, phase
//
// The blur is integer-ALU bound (1738 MAC per input pixel); both passes keep every product on
// the packed-integer dot-product path: u8 pixels x u8 taps with dp4a (4 MAC/instr), u16 8.8 sums x u8 taps with
// dp2a (2 MAC/instr).  All sums are exact (<= 65 280 and <= 16 711 680), so the order of evaluation is free and
// the result is bit-identical to OpenCV's fixedSmoothInvoker.
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

__global__ void __launch_bounds__(256) upsample2x_k(const u8 *__restrict__ src, int rows, int cols, size_t sp,
                                                    u8 *__restrict__ dst, size_t dp) {
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

int launch_upsample2x(vslam_ctx *ctx, const u8 *src, int rows, int cols, size_t sp, u8 *dst, size_t dp) {
    dim3 b(64, 4), g((2 * cols + 255) / 256, (2 * rows + 3) / 4);
    KernelSpan span(ctx, VSLAM_K_UPSAMPLE);
    upsample2x_k<<<g, b, 0, ctx->stream>>>(src, rows, cols, sp, dst, dp);
    VS_LAUNCH_CHECK(ctx);
    return VSLAM_OK;
}

// ------------------------------------------------------------------------------------------------------------
// row B, horizontal pass.  Block = 128 threads = 32 rows (lane) x 4 column blocks (warp) -> tile 32 x 128 px.
// The tile's rows are staged once (reflect-101 resolved while staging) and reused by all 6 levels.
// Thread: 32 consecutive output pixels of one row = 8 groups x 4 byte alignments; the four byte-shifted views
// of the staged words come from PRMT, so every alignment needs only ceil((n+delta)/4) dp4a.
// ------------------------------------------------------------------------------------------------------------
constexpr int HB_ROWS = 32, HB_COLS = 128, HB_G = 8, HB_U = 4;

__global__ void __launch_bounds__(128) hblur6_k(const u8 *__restrict__ src, int rows, int cols, size_t sp,
                                                u16 *__restrict__ hbuf, size_t hpitch, size_t hplane,
                                                const OctaveTaps *__restrict__ taps, int sw, int stride) {
    extern __shared__ u32 smem[];
    u32 *s_pix = smem;                                   // HB_ROWS x stride words
    __shared__ u32 s_tap[kHWords];
    const int tid = threadIdx.x, lane = tid & 31, xb = tid >> 5;
    const int X0 = blockIdx.x * HB_COLS, Y0 = blockIdx.y * HB_ROWS;
    const int hp = taps->hp;
    const int xs = X0 - hp;                              // pixel of staged word 0 (multiple of 4)

    // ---- stage: s_pix[row][w] = pixels xs+4w .. xs+4w+3 of row Y0+row, borders reflected ----
    const bool aligned = ((sp & 3) == 0) && (((uintptr_t)src & 3) == 0);
    for (int idx = tid; idx < HB_ROWS * sw; idx += 128) {
        const int row = idx / sw, w = idx - row * sw;
        const int y = min(Y0 + row, rows - 1);
        const u8 *p = src + (size_t)y * sp;
        const int x = xs + 4 * w;
        u32 v;
        if (aligned && x >= 0 && x + 3 < cols) {
            v = *(const u32 *)(p + x);
        } else {
            v = (u32)p[refl101(x, cols)] | (u32)p[refl101(x + 1, cols)] << 8 | (u32)p[refl101(x + 2, cols)] << 16 |
                (u32)p[refl101(x + 3, cols)] << 24;
        }
        s_pix[row * stride + w] = v;
    }

    const int y = Y0 + lane;
    for (int l = 0; l < 6; ++l) {
        const LevelTaps &lv = taps->lv[l];
        __syncthreads();                                 // staging done / previous level's taps consumed
        for (int i = tid; i < lv.kq; i += 128) s_tap[i] = lv.hq[i];
        __syncthreads();
        const int kq = lv.kq;
        const u32 *row = s_pix + lane * stride + xb * HB_G + lv.wofs;
        u32 acc[HB_G][4];
#pragma unroll
        for (int g = 0; g < HB_G; ++g) acc[g][0] = acc[g][1] = acc[g][2] = acc[g][3] = 0;
        for (int kb = 0; kb < kq; kb += HB_U) {
            u32 d[HB_G + HB_U];
#pragma unroll
            for (int j = 0; j < HB_G + HB_U; ++j) d[j] = row[kb + j];
            u32 t[HB_U];
#pragma unroll
            for (int u = 0; u < HB_U; ++u) t[u] = s_tap[kb + u];
#pragma unroll
            for (int j = 0; j < HB_G + HB_U - 1; ++j) {
                const u32 w0 = d[j];
                const u32 w1 = __byte_perm(d[j], d[j + 1], 0x4321);
                const u32 w2 = __byte_perm(d[j], d[j + 1], 0x5432);
                const u32 w3 = __byte_perm(d[j], d[j + 1], 0x6543);
#pragma unroll
                for (int u = 0; u < HB_U; ++u) {
                    const int g = j - u;
                    if (g >= 0 && g < HB_G) {
                        acc[g][0] = __dp4a(w0, t[u], acc[g][0]);
                        acc[g][1] = __dp4a(w1, t[u], acc[g][1]);
                        acc[g][2] = __dp4a(w2, t[u], acc[g][2]);
                        acc[g][3] = __dp4a(w3, t[u], acc[g][3]);
                    }
                }
            }
        }
        if (y < rows) {
            // 32 u16 = 64 B per thread; hpitch is a multiple of 128 elements so the stores are aligned
            uint4 *o = (uint4 *)(hbuf + (size_t)l * hplane + (size_t)y * hpitch + X0 + xb * 32);
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                uint4 v;
                v.x = acc[2 * q][0] | acc[2 * q][1] << 16;
                v.y = acc[2 * q][2] | acc[2 * q][3] << 16;
                v.z = acc[2 * q + 1][0] | acc[2 * q + 1][1] << 16;
                v.w = acc[2 * q + 1][2] | acc[2 * q + 1][3] << 16;
                o[q] = v;
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------------------
// rows B (vertical pass), C, D.  Block = 128 threads = 128 columns, 32 output rows; loops over the 6 levels.
// Each thread owns one column: it stages the 8.8 sums of rows Y0-r .. as row PAIRS (2p | 2p+1 << 16) in a
// column-private strip of shared memory, then accumulates 32 outputs with dp2a: even outputs use the E tap
// pair (low half of eo), odd outputs the O pair (high half).  Rounding (v + 32768) >> 16 follows OpenCV's
// fixedSmoothInvoker; the previous level stays in registers for the saturating difference (row C); level 3 is
// also written decimated by two as the next octave's base (row D).
// ------------------------------------------------------------------------------------------------------------
constexpr int VB_ROWS = 32, VB_COLS = 128, VB_Q = VB_ROWS / 2, VB_U = 8;

__global__ void __launch_bounds__(128) vblur6_k(const u16 *__restrict__ hbuf, size_t hpitch, size_t hplane,
                                                OctavePlanes op, const OctaveTaps *__restrict__ taps) {
    extern __shared__ u32 smem[];
    __shared__ u32 s_tap[kVPairs];
    const int tx = threadIdx.x;
    const int x = blockIdx.x * VB_COLS + tx, Y0 = blockIdx.y * VB_ROWS;
    const int rows = op.rows, cols = op.cols;
    u32 prev[VB_ROWS / 4];                               // previous level, 4 rows per register
#pragma unroll
    for (int i = 0; i < VB_ROWS / 4; ++i) prev[i] = 0;

    for (int l = 0; l < 6; ++l) {
        const LevelTaps &lv = taps->lv[l];
        __syncthreads();
        for (int i = tx; i < lv.jp; i += 128) s_tap[i] = lv.eo[i];
        const int r = lv.r, jp = lv.jp;
        const int npairs = VB_Q + jp - 1;
        const u16 *hcol = hbuf + (size_t)l * hplane + x;   // x < hpitch always (hpitch multiple of 128)
        for (int p = 0; p < npairs; ++p) {
            const int ya = refl101(Y0 - r + 2 * p, rows), yb = refl101(Y0 - r + 2 * p + 1, rows);
            const u32 lo = hcol[(size_t)ya * hpitch], hi = hcol[(size_t)yb * hpitch];
            smem[p * VB_COLS + tx] = lo | hi << 16;
        }
        __syncthreads();                                 // taps visible (the strip itself is column-private)
        u32 acc[VB_ROWS];
#pragma unroll
        for (int m = 0; m < VB_ROWS; ++m) acc[m] = 32768u;  // rounding term of (v + (1<<15)) >> 16
        for (int jb = 0; jb < jp; jb += VB_U) {
            u32 d[VB_Q + VB_U - 1];
#pragma unroll
            for (int i = 0; i < VB_Q + VB_U - 1; ++i) d[i] = smem[(jb + i) * VB_COLS + tx];
#pragma unroll
            for (int u = 0; u < VB_U; ++u) {
                const u32 t = s_tap[jb + u];
#pragma unroll
                for (int q = 0; q < VB_Q; ++q) {
                    acc[2 * q] = __dp2a_lo(d[q + u], t, acc[2 * q]);
                    acc[2 * q + 1] = __dp2a_hi(d[q + u], t, acc[2 * q + 1]);
                }
            }
        }
        const bool xin = x < cols;
#pragma unroll
        for (int m4 = 0; m4 < VB_ROWS / 4; ++m4) {
            u32 cur = 0;
#pragma unroll
            for (int b = 0; b < 4; ++b) cur |= (acc[4 * m4 + b] >> 16) << (8 * b);
            const u32 pv = prev[m4];
            prev[m4] = cur;
#pragma unroll
            for (int b = 0; b < 4; ++b) {
                const int y = Y0 + 4 * m4 + b;
                if (xin && y < rows) {
                    const int g = (cur >> (8 * b)) & 255;
                    op.gauss[l][(size_t)y * op.pitch + x] = (u8)g;
                    if (l > 0) {
                        const int dgv = g - (int)((pv >> (8 * b)) & 255);     // GaussPyramid.cpp:197 saturating
                        op.dog[l - 1][(size_t)y * op.pitch + x] = (u8)max(dgv, 0);
                    }
                    if (l == 3 && op.next_base && !(y & 1) && !(x & 1) && (y >> 1) < op.next_rows &&
                        (x >> 1) < op.next_cols)
                        op.next_base[(size_t)(y >> 1) * op.next_pitch + (x >> 1)] = (u8)g;   // :126 INTER_NEAREST
                }
            }
        }
    }
}

int launch_octave_blur(vslam_ctx *ctx, const OctavePlanes &op, const OctaveTaps *d_taps, const OctaveTaps &h_taps,
                       u16 *hbuf, size_t hpitch) {
    const size_t hplane = hpitch * (size_t)op.rows;
    {
        const int sw = h_taps.hp / 2 + 40, stride = sw | 1;
        const size_t sm = (size_t)HB_ROWS * stride * sizeof(u32);
        dim3 g((op.cols + HB_COLS - 1) / HB_COLS, (op.rows + HB_ROWS - 1) / HB_ROWS);
        KernelSpan span(ctx, VSLAM_K_HBLUR);
        hblur6_k<<<g, 128, sm, ctx->stream>>>(op.base, op.rows, op.cols, op.pitch, hbuf, hpitch, hplane, d_taps, sw,
                                              stride);
        VS_LAUNCH_CHECK(ctx);
    }
    {
        int jpmax = 0;
        for (int l = 0; l < 6; ++l) jpmax = h_taps.lv[l].jp > jpmax ? h_taps.lv[l].jp : jpmax;
        const size_t sm = (size_t)(VB_Q + jpmax - 1) * VB_COLS * sizeof(u32);
        if (sm > 48 * 1024)
            VS_CUDA(ctx, cudaFuncSetAttribute(vblur6_k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
        dim3 g((op.cols + VB_COLS - 1) / VB_COLS, (op.rows + VB_ROWS - 1) / VB_ROWS);
        KernelSpan span(ctx, VSLAM_K_VBLUR);
        vblur6_k<<<g, 128, sm, ctx->stream>>>(hbuf, hpitch, hplane, op, d_taps);
        VS_LAUNCH_CHECK(ctx);
    }
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
