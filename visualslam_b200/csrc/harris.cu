// harris.cu -- fused Harris corner path: one kernel from the 8-bit frame to the response
// map, the 8-bit response, the 3x3 NMS mask and the ordered NMS2 corner list.
//
//   row K  Harris_corners.cpp:158      GaussianBlur(3x3, sigma 0) 8U  == (S121x121 + 8) >> 4, reflect-101
//   row L  :163-164                    Sobel(CV_32F, ksize=1)        == central differences, reflect-101
//   row M  :10-29, :42-43              3x3 box sums of Ix^2, Iy^2, IxIy on replicate-padded gradients
//   row N  :54-62                      det (double det2 -> float) - 0.04f * (tr*tr), negatives zeroed
//   row O  :176, :70-81                convertScaleAbs, centre-excluded 3x3 dilate, strict '>'
//   row P  :83-129                     NMS2: 4x4 half-open window, '>=', interior only, row-major emission
//
// Tile = 32 x 64 pixels per 256-thread block.  Halo budget: NMS2 reads R at offsets -2..+1, R needs gradients at
// +-1, gradients need the blurred image at +-1, the blur needs the frame at +-1 -> the frame tile is (32+9)x(64+9).
// All sums are small integers, so they are accumulated in int32 / int64 and converted once; the float steps use
// round-to-nearest intrinsics (no FMA contraction) to reproduce the reference's three separately rounded ops.
#include "harris.cuh"

namespace vslam {

constexpr int TH = 32, TW = 64;
constexpr int RAW_H = TH + 9, RAW_W = TW + 9, RAW_S = 76;
constexpr int BL_H = TH + 7, BL_W = TW + 7, BL_S = 72;
constexpr int GR_H = TH + 5, GR_W = TW + 5, GR_S = 70;
constexpr int RS_H = TH + 3, RS_W = TW + 3, RS_S = 68;

__device__ __forceinline__ int r8_of(float r) {
    // saturate_cast<uchar>(cvRound(|r|)); cvtss2si returns INT_MIN for |r| >= 2^31, which saturates to 0
    const float a = fabsf(r);
    if (a >= 2147483648.f) return 0;
    return min(__float2int_rn(a), 255);
}

__global__ void __launch_bounds__(256) harris_fused_k(HarrisArgs a) {
    __shared__ u8 s_raw[RAW_H * RAW_S];
    __shared__ u8 s_bl[BL_H * BL_S];
    __shared__ short s_ix[GR_H * GR_S], s_iy[GR_H * GR_S];
    __shared__ float s_r[RS_H * RS_S];
    __shared__ u32 s_flag[TH][2];
    __shared__ int s_rowoff[TH];
    __shared__ int s_base;
    const int tid = threadIdx.x;
    const int X0 = blockIdx.x * TW, Y0 = blockIdx.y * TH;
    const int rows = a.rows, cols = a.cols;
    const u8 *__restrict__ img = a.frame;

    // frame tile, reflect-101 extension: origin (Y0-5, X0-5)
    for (int idx = tid; idx < RAW_H * RAW_W; idx += 256) {
        const int ly = idx / RAW_W, lx = idx - ly * RAW_W;
        const int gy = refl101(Y0 - 5 + ly, rows), gx = refl101(X0 - 5 + lx, cols);
        s_raw[ly * RAW_S + lx] = img[(size_t)gy * a.pitch + gx];
    }
    __syncthreads();
    // row K: blurred, origin (Y0-4, X0-4)
    for (int idx = tid; idx < BL_H * BL_W; idx += 256) {
        const int ly = idx / BL_W, lx = idx - ly * BL_W;
        const u8 *p = s_raw + ly * RAW_S + lx;       // top-left of the 3x3 window
        const int s = p[0] + 2 * p[1] + p[2] + 2 * (p[RAW_S] + 2 * p[RAW_S + 1] + p[RAW_S + 2]) + p[2 * RAW_S] +
                      2 * p[2 * RAW_S + 1] + p[2 * RAW_S + 2];
        s_bl[ly * BL_S + lx] = (u8)((s + 8) >> 4);
    }
    __syncthreads();
    // row L on replicate-clamped coordinates (row M pads the GRADIENT maps by replication), origin (Y0-3, X0-3)
    for (int idx = tid; idx < GR_H * GR_W; idx += 256) {
        const int ly = idx / GR_W, lx = idx - ly * GR_W;
        const int cy = clampi(Y0 - 3 + ly, 0, rows - 1) - (Y0 - 4), cx = clampi(X0 - 3 + lx, 0, cols - 1) - (X0 - 4);
        const u8 *p = s_bl + cy * BL_S + cx;
        s_ix[ly * GR_S + lx] = (short)((int)p[1] - (int)p[-1]);
        s_iy[ly * GR_S + lx] = (short)((int)p[BL_S] - (int)p[-BL_S]);
    }
    __syncthreads();
    // rows M, N: response, origin (Y0-2, X0-2); zero outside the image
    for (int idx = tid; idx < RS_H * RS_W; idx += 256) {
        const int ly = idx / RS_W, lx = idx - ly * RS_W;
        const int gy = Y0 - 2 + ly, gx = X0 - 2 + lx;
        float resp = 0.f;
        if (gy >= 0 && gy < rows && gx >= 0 && gx < cols) {
            int ix2 = 0, iy2 = 0, ixy = 0;
#pragma unroll
            for (int u = 0; u < 3; ++u)
#pragma unroll
                for (int v = 0; v < 3; ++v) {
                    const int x = s_ix[(ly + u) * GR_S + lx + v], y = s_iy[(ly + u) * GR_S + lx + v];
                    ix2 += x * x; iy2 += y * y; ixy += x * y;
                }
            const long long det_i = (long long)ix2 * iy2 - (long long)ixy * ixy;   // exact; == double det2
            const float det = __ll2float_rn(det_i);
            const float tr = (float)(ix2 + iy2);                                     // exact (< 2^24)
            const float r = __fsub_rn(det, __fmul_rn(a.k, __fmul_rn(tr, tr)));       // Harris_corners.cpp:57
            resp = r > 0.f ? r : 0.f;
        }
        s_r[ly * RS_S + lx] = resp;
    }
    __syncthreads();
    // outputs + NMS, 4 rows per pass; a warp covers half a tile row
    const int p2 = a.nms2_pad;
    const int lane = tid & 31;
    for (int pass = 0; pass < TH / 4; ++pass) {
        const int ty = pass * 4 + (tid >> 6), tx = tid & 63;
        const int gy = Y0 + ty, gx = X0 + tx;
        const bool in = gy < rows && gx < cols;
        const float *c = s_r + (ty + 2) * RS_S + tx + 2;
        const float r = *c;
        bool corner = false;
        if (in) {
            const size_t o = (size_t)gy * cols + gx;
            if (a.response) a.response[o] = r;
            const int v8 = r8_of(r);
            if (a.r8) a.r8[o] = (u8)v8;
            if (a.mask3) {
                int m = 0;
#pragma unroll
                for (int u = -1; u <= 1; ++u)
#pragma unroll
                    for (int v = -1; v <= 1; ++v)
                        if (u || v) m = max(m, r8_of(c[u * RS_S + v]));
                a.mask3[o] = v8 > m ? 255 : 0;
            }
            if (r > 0.f && gy >= p2 && gy < rows - p2 && gx >= p2 && gx < cols - p2) {
                float m = 0.f;
                for (int u = -p2; u < p2; ++u)
                    for (int v = -p2; v < p2; ++v) m = fmaxf(m, c[u * RS_S + v]);
                corner = r >= m;
            }
        }
        const unsigned bal = __ballot_sync(0xffffffffu, corner);
        if (lane == 0) s_flag[ty][tx >> 5] = bal;
    }
    __syncthreads();
    if (tid < 32) {
        const int cnt = __popc(s_flag[tid][0]) + __popc(s_flag[tid][1]);
        int incl = cnt;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, incl, d);
            if (tid >= d) incl += t;
        }
        const int total = __shfl_sync(0xffffffffu, incl, 31);
        int base = 0;
        if (tid == 0 && total) base = atomicAdd(a.cursor, total);
        base = __shfl_sync(0xffffffffu, base, 0);
        s_rowoff[tid] = base + incl - cnt;
        if (Y0 + tid < rows) {
            const int seg = a.seg_base + (Y0 + tid) * gridDim.x + blockIdx.x;
            a.seg_off[seg] = base + incl - cnt;
            a.seg_cnt[seg] = cnt;
        }
    }
    __syncthreads();
    for (int pass = 0; pass < TH / 4; ++pass) {
        const int ty = pass * 4 + (tid >> 6), tx = tid & 63;
        const u32 f = s_flag[ty][tx >> 5];
        if (f >> (tx & 31) & 1u) {
            int pos = s_rowoff[ty] + __popc(f & ((1u << (tx & 31)) - 1u));
            if (tx >= 32) pos += __popc(s_flag[ty][0]);
            if (pos < a.staging_cap) {
                vslam_corner k;
                k.row = Y0 + ty; k.col = X0 + tx; k.value = s_r[(ty + 2) * RS_S + tx + 2];
                a.staging[pos] = k;
            }
        }
    }
}

void harris_grid(int rows, int cols, int *ntx, int *nty) { *ntx = (cols + TW - 1) / TW; *nty = (rows + TH - 1) / TH; }

int launch_harris(vslam_ctx *ctx, const HarrisArgs &a) {
    int ntx, nty;
    harris_grid(a.rows, a.cols, &ntx, &nty);
    KernelSpan span(ctx, VSLAM_K_HARRIS);
    harris_fused_k<<<dim3(ntx, nty), 256, 0, ctx->stream>>>(a);
    VS_LAUNCH_CHECK(ctx);
    return VSLAM_OK;
}

}  // namespace vslam
