// harris.cu -- fused Harris corner path: one kernel from the 8-bit frames to the response
// map, the 8-bit response, the 3x3 NMS mask and the ordered NMS2 corner list.
//
//   row K  Harris_corners.cpp:158      GaussianBlur(3x3, sigma 0) 8U  == (S121x121 + 8) >> 4, reflect-101
//   row L  :163-164                    Sobel(CV_32F, ksize=1)        == central differences, reflect-101
//   row M  :10-29, :42-43              3x3 box sums of Ix^2, Iy^2, IxIy on replicate-padded gradients
//   row N  :54-62                      det (double det2 -> float) - 0.04f * (tr*tr), negatives zeroed
//   row O  :176, :70-81                convertScaleAbs, centre-excluded 3x3 dilate, strict '>'
//   row P  :83-129                     NMS2: 4x4 half-open window, '>=', interior only, row-major emission
//
// Design: a streaming register pipeline.  A warp owns a strip of 112 columns (lane = 4 adjacent pixels; two halo
// lanes on each side) and marches down a band of rows.  Every stage keeps its sliding window of rows in registers
// (packed 16x2 where the values allow it), horizontal neighbours come from the adjacent lanes with SHFL, so each frame
// byte is loaded once (plus band/strip halos) and nothing but the requested outputs is written: the kernel moves the
// algorithmic 6 B per pixel in pipeline mode (1 read + 4 response + 1 mask).  All sums are exact integers; det is
// formed in int64 and rounded once (== OpenCV's double det2 -> float); the three float steps of the response use
// round-to-nearest intrinsics (no FMA contraction).  Corners are staged per warp in shared memory and flushed with
// one atomic per 8 rows; compaction.cu orders the (frame, row, strip) segments.
#include <stdlib.h>

#include <type_traits>

#include "harris.cuh"

namespace vslam {

constexpr int HS = kHarrisStrip, HB = kHarrisBand;
constexpr int H_WARPS = 4;                 // warps per CTA, each an independent (strip, band)
#ifndef HARRIS_MIN_CTAS
#define HARRIS_MIN_CTAS 3        // <= 128 registers: 16 warps per SM (4.68 -> 4.06 ms per 1080p x 256; 5 or 6 CTAs spill and lose)
#endif
constexpr int H_CAP = 256;                 // staged corners per warp before a flush
constexpr int H_ROWS = 8;                  // rows per flush

__device__ __forceinline__ u32 pk(u32 lo, u32 hi, u32 sel) { return __byte_perm(lo, hi, sel); }

// saturate_cast<uchar>(cvRound(|r|)) for r >= 0; cvRound returns INT_MIN for r >= 2^31, which saturates to 0
__device__ __forceinline__ int r8_of(float r) {
    const float m = __fadd_rn(fminf(r, 255.0f), 12582912.0f);          // round half to even in the low mantissa bits
    return r >= 2147483648.f ? 0 : (__float_as_int(m) & 0xff);
}

template <int P2>
__global__ void __launch_bounds__(32 * H_WARPS, HARRIS_MIN_CTAS) harris_march_k(const HarrisArgs a) {
    __shared__ vslam_corner s_buf[H_WARPS][H_CAP];
    __shared__ int s_rowoff[H_WARPS][H_ROWS], s_rowcnt[H_WARPS][H_ROWS];
    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int unit = blockIdx.x * H_WARPS + warp;
    if (unit >= a.n_strips * a.n_bands) return;                 // warps are independent: no block-wide barrier below
    const int strip = unit % a.n_strips, band = unit / a.n_strips, frame = blockIdx.z;
    const int rows = a.rows, cols = a.cols;
    const int X0 = strip * HS, c = X0 - 8 + 4 * lane;           // this lane's 4 pixels: columns c .. c+3
    const int Y0 = band * HB, Y1 = min(Y0 + HB, rows);
    const u8 *__restrict__ img = a.frames + (size_t)frame * a.frame_stride;
    const bool fast_ld = c >= 0 && c + 3 < cols && !(a.pitch & 3) && !((uintptr_t)img & 3);
    const bool own = lane >= 2 && lane < 30;
    bool inx[4], nmsx[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        inx[j] = (unsigned)(c + j) < (unsigned)cols;
        nmsx[j] = own && c + j >= P2 && c + j < cols - P2;      // NMS2 interior columns (Harris_corners.cpp:95)
    }
    // replicate padding of the GRADIENT maps (row M): column -1 := column 0, column cols := column cols-1
    const bool repL = c == 0, rep0 = c + 1 == cols, rep1 = c + 2 == cols, rep2 = c + 3 == cols, repR = c + 4 == cols;
    const bool vec_out = !(cols & 3) && !((uintptr_t)a.response & 15) && !((uintptr_t)a.r8 & 3) && !((uintptr_t)a.mask3 & 3);
    const size_t fpx = (size_t)frame * rows * cols;

    u32 ra0 = 0, ra1 = 0, rb0 = 0, rb1 = 0;                     // raw rows yr-2, yr-1 (16x2)
    u32 bA0 = 0, bA1 = 0, bB0 = 0, bB1 = 0;                     // blurred rows yr-3, yr-2 (16x2)
    int hs[4][3][4];                                            // horizontal 3-sums of xx, yy, xy; ring by row & 3
    float hmax[4][4];                                           // horizontal window maxima of R; ring by row & 3
    int m3[4][4];                                               // 3-wide maxima of the 8-bit response; ring by row & 3
    float Rp[4];                                                // R of row yr-4 (the output row)
    int q8p[4], m2p[4];                                         // 8-bit response of row yr-4 and its left/right maximum
#pragma unroll
    for (int s = 0; s < 4; ++s)
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            hs[s][0][j] = hs[s][1][j] = hs[s][2][j] = 0; hmax[s][j] = 0.f; m3[s][j] = 0;
            Rp[j] = 0.f; q8p[j] = 0; m2p[j] = 0;
        }
    int cur = 0, ri = 0, batch_y0 = Y0;                         // staged corners / rows of the current flush batch

    auto flush = [&]() {
        __syncwarp();
        if (cur > 0) {
            int base = 0;
            if (lane == 0) base = atomicAdd(a.cursor, cur);
            base = __shfl_sync(FULL, base, 0);
            for (int i = lane; i < cur; i += 32)
                if (base + i < a.staging_cap) a.staging[base + i] = s_buf[warp][i];
            if (lane < ri && s_rowcnt[warp][lane] > 0) {
                const size_t sg = ((size_t)frame * rows + (batch_y0 + lane)) * a.n_strips + strip;
                a.seg_off[sg] = base + s_rowoff[warp][lane];
                a.seg_cnt[sg] = s_rowcnt[warp][lane];
            }
        }
        __syncwarp();
        cur = 0; ri = 0;
    };

    // raw row yr, reflect-101
    auto load_row = [&](int yr) -> u32 {
        const u8 *__restrict__ src = img + (size_t)refl101(yr, rows) * a.pitch;
        if (fast_ld) return *(const u32 *)(src + c);
        return (u32)src[refl101(c, cols)] | (u32)src[refl101(c + 1, cols)] << 8 |
               (u32)src[refl101(c + 2, cols)] << 16 | (u32)src[refl101(c + 3, cols)] << 24;
    };
    const int yr_begin = (Y0 - 5) & ~3;                         // ring slots are compile-time: start on a multiple of 4
    u32 wn[4];                                                  // the four raw rows of the NEXT group: loads stay one group ahead
#pragma unroll
    for (int ph = 0; ph < 4; ++ph) wn[ph] = load_row(yr_begin + ph);
    for (int yr4 = yr_begin; yr4 <= Y1 + 3; yr4 += 4) {
        u32 wc[4];
#pragma unroll
        for (int ph = 0; ph < 4; ++ph) wc[ph] = wn[ph];
        if (yr4 + 4 <= Y1 + 3) {
#pragma unroll
            for (int ph = 0; ph < 4; ++ph) wn[ph] = load_row(yr4 + 4 + ph);
        }
#pragma unroll
        for (int ph = 0; ph < 4; ++ph) {
            const int yr = yr4 + ph;                            // (yr & 3) == ph
            const u32 w = wc[ph];
            const u32 n0 = pk(w, 0, 0x4140), n1 = pk(w, 0, 0x4342);
            // ---- row K: blurred row yr-1 = (121 x 121 + 8) >> 4 -----------------------------------------------
            const u32 V0 = ra0 + 2 * rb0 + n0, V1 = ra1 + 2 * rb1 + n1;
            ra0 = rb0; ra1 = rb1; rb0 = n0; rb1 = n1;
            u32 Bn0, Bn1;
            {
                const u32 L = __shfl_up_sync(FULL, V1, 1), R = __shfl_down_sync(FULL, V0, 1);
                const u32 Vm = pk(L, V0, 0x5432), Vc = pk(V0, V1, 0x5432), Vp = pk(V1, R, 0x5432);
                Bn0 = ((Vm + 2 * V0 + Vc + 0x00080008u) >> 4) & 0x00ff00ffu;
                Bn1 = ((Vc + 2 * V1 + Vp + 0x00080008u) >> 4) & 0x00ff00ffu;
            }
            // ---- row L: gradients of row yr-2 (biased by 256, 16x2) --------------------------------------------
            u32 e[6];                                           // per pixel ix | iy << 16 for columns c-1 .. c+4
            {
                const u32 L = __shfl_up_sync(FULL, bB1, 1), R = __shfl_down_sync(FULL, bB0, 1);
                const u32 Bm = pk(L, bB0, 0x5432), Bc = pk(bB0, bB1, 0x5432), Bp = pk(bB1, R, 0x5432);
                const u32 IX01 = Bc + 0x01000100u - Bm, IX23 = Bp + 0x01000100u - Bc;
                const u32 IY01 = Bn0 + 0x01000100u - bA0, IY23 = Bn1 + 0x01000100u - bA1;
                e[1] = pk(IX01, IY01, 0x5410); e[2] = pk(IX01, IY01, 0x7632);
                e[3] = pk(IX23, IY23, 0x5410); e[4] = pk(IX23, IY23, 0x7632);
                e[0] = __shfl_up_sync(FULL, e[4], 1); e[5] = __shfl_down_sync(FULL, e[1], 1);
                if (repL) e[0] = e[1];
                if (rep0) e[2] = e[1];
                if (rep1) e[3] = e[2];
                if (rep2) e[4] = e[3];
                if (repR) e[5] = e[4];
            }
            bA0 = bB0; bA1 = bB1; bB0 = Bn0; bB1 = Bn1;
            // ---- row M, horizontal part: 3-sums of the products, row yr-2 -> ring slot (ph + 2) & 3 ---------------
            {
                int xx[6], yy[6], xy[6];
#pragma unroll
                for (int i = 0; i < 6; ++i) {
                    const int ix = (int)(e[i] & 0xffffu) - 256, iy = (int)(e[i] >> 16) - 256;
                    xx[i] = ix * ix; yy[i] = iy * iy; xy[i] = ix * iy;
                }
                const int S = (ph + 2) & 3;
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    hs[S][0][j] = xx[j] + xx[j + 1] + xx[j + 2];
                    hs[S][1][j] = yy[j] + yy[j + 1] + yy[j + 2];
                    hs[S][2][j] = xy[j] + xy[j + 1] + xy[j + 2];
                }
            }
            // ---- rows M (vertical part), N: response of row rb = yr-3 ----------------------------------------------
            const int rb = yr - 3;
            float Rn[4];
            if (rb >= 0 && rb < rows) {
                // replicate padding in y: rows -1 and `rows` of the gradient maps repeat rows 0 and rows-1
                const int SM = (ph + 1) & 3, ST = ph & 3, SB = (ph + 2) & 3;
                int sx[4], sy[4], sxy[4];
                if (rb == 0 && rb == rows - 1) {
#pragma unroll
                    for (int j = 0; j < 4; ++j) { sx[j] = 3 * hs[SM][0][j]; sy[j] = 3 * hs[SM][1][j]; sxy[j] = 3 * hs[SM][2][j]; }
                } else if (rb == 0) {
#pragma unroll
                    for (int j = 0; j < 4; ++j) { sx[j] = 2 * hs[SM][0][j] + hs[SB][0][j]; sy[j] = 2 * hs[SM][1][j] + hs[SB][1][j]; sxy[j] = 2 * hs[SM][2][j] + hs[SB][2][j]; }
                } else if (rb == rows - 1) {
#pragma unroll
                    for (int j = 0; j < 4; ++j) { sx[j] = 2 * hs[SM][0][j] + hs[ST][0][j]; sy[j] = 2 * hs[SM][1][j] + hs[ST][1][j]; sxy[j] = 2 * hs[SM][2][j] + hs[ST][2][j]; }
                } else {
#pragma unroll
                    for (int j = 0; j < 4; ++j) { sx[j] = hs[ST][0][j] + hs[SM][0][j] + hs[SB][0][j]; sy[j] = hs[ST][1][j] + hs[SM][1][j] + hs[SB][1][j]; sxy[j] = hs[ST][2][j] + hs[SM][2][j] + hs[SB][2][j]; }
                }
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const long long det_i = (long long)sx[j] * sy[j] - (long long)sxy[j] * sxy[j];   // exact; == double det2
                    const float det = __ll2float_rn(det_i);
                    const float tr = __fsub_rn(__int_as_float(0x4B000000 | (sx[j] + sy[j])), 8388608.0f);   // exact int -> float
                    const float r = __fsub_rn(det, __fmul_rn(a.k, __fmul_rn(tr, tr)));                 // Harris_corners.cpp:57
                    Rn[j] = inx[j] ? fmaxf(r, 0.f) : 0.f;                                               // :60-62
                }
            } else {
#pragma unroll
                for (int j = 0; j < 4; ++j) Rn[j] = 0.f;
            }
            // ---- row P, horizontal part: window maxima of row rb -> ring slot (ph + 1) & 3 ---------------------------
            {
                const int S = (ph + 1) & 3;
                const float l3 = __shfl_up_sync(FULL, Rn[3], 1);
                if (P2 == 2) {
                    const float l2 = __shfl_up_sync(FULL, Rn[2], 1), r0 = __shfl_down_sync(FULL, Rn[0], 1);
                    const float q0 = fmaxf(l2, l3), q1 = fmaxf(l3, Rn[0]), q2 = fmaxf(Rn[0], Rn[1]), q3 = fmaxf(Rn[1], Rn[2]),
                                q4 = fmaxf(Rn[2], Rn[3]), q5 = fmaxf(Rn[3], r0);
                    hmax[S][0] = fmaxf(q0, q2); hmax[S][1] = fmaxf(q1, q3); hmax[S][2] = fmaxf(q2, q4); hmax[S][3] = fmaxf(q3, q5);
                } else {
                    hmax[S][0] = fmaxf(l3, Rn[0]); hmax[S][1] = fmaxf(Rn[0], Rn[1]); hmax[S][2] = fmaxf(Rn[1], Rn[2]);
                    hmax[S][3] = fmaxf(Rn[2], Rn[3]);
                }
            }
            // ---- row O, horizontal part: 8-bit response of row rb, its 3-wide maxima -> ring slot (ph + 1) & 3 ----------
            int q8n[4], m2n[4];
            const bool want8 = a.r8 || a.mask3;
            if (want8) {
#pragma unroll
                for (int j = 0; j < 4; ++j) q8n[j] = r8_of(Rn[j]);
                const int l = __shfl_up_sync(FULL, q8n[3], 1), r = __shfl_down_sync(FULL, q8n[0], 1);
                m2n[0] = max(l, q8n[1]); m2n[1] = max(q8n[0], q8n[2]); m2n[2] = max(q8n[1], q8n[3]); m2n[3] = max(q8n[2], r);
                const int S = (ph + 1) & 3;
#pragma unroll
                for (int j = 0; j < 4; ++j) m3[S][j] = max(m2n[j], q8n[j]);
            }
            // ---- outputs of row yo = yr-4 ---------------------------------------------------------------------------
            const int yo = yr - 4;
            if (yo >= Y0 && yo < Y1) {
                const size_t o = fpx + (size_t)yo * cols + c;
                if (own && inx[0]) {
                    if (a.response) {
                        if (vec_out && inx[3]) *(float4 *)(a.response + o) = make_float4(Rp[0], Rp[1], Rp[2], Rp[3]);
                        else {
#pragma unroll
                            for (int j = 0; j < 4; ++j) if (inx[j]) a.response[o + j] = Rp[j];
                        }
                    }
                    if (want8) {
                        u32 w8 = 0, wm = 0;
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            // centre-excluded 3x3 dilate of the 8-bit response, strict '>' (Harris_corners.cpp:76-79)
                            const int d = max(max(m3[(ph + 3) & 3][j], m3[(ph + 1) & 3][j]), m2p[j]);
                            w8 |= (u32)q8p[j] << (8 * j);
                            wm |= (q8p[j] > d ? 255u : 0u) << (8 * j);
                        }
                        if (vec_out && inx[3]) {
                            if (a.r8) *(u32 *)(a.r8 + o) = w8;
                            if (a.mask3) *(u32 *)(a.mask3 + o) = wm;
                        } else {
#pragma unroll
                            for (int j = 0; j < 4; ++j)
                                if (inx[j]) {
                                    if (a.r8) a.r8[o + j] = (u8)(w8 >> (8 * j));
                                    if (a.mask3) a.mask3[o + j] = (u8)(wm >> (8 * j));
                                }
                        }
                    }
                }
                // NMS2 (Harris_corners.cpp:94-118): R >= max of the half-open window, interior pixels, positive responses
                if (a.want_corners) {
                    bool cr[4];
                    const bool rowok = yo >= P2 && yo < rows - P2;
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        float m = fmaxf(hmax[(ph + 3) & 3][j], hmax[ph & 3][j]);                 // rows yo-1, yo
                        if (P2 == 2) m = fmaxf(m, fmaxf(hmax[(ph + 2) & 3][j], hmax[(ph + 1) & 3][j]));   // rows yo-2, yo+1
                        cr[j] = rowok && nmsx[j] && Rp[j] > 0.f && Rp[j] >= m;
                    }
                    if (cur + HS > H_CAP || ri == H_ROWS) flush();
                    if (ri == 0) batch_y0 = yo;
                    const unsigned b0 = __ballot_sync(FULL, cr[0]), b1 = __ballot_sync(FULL, cr[1]),
                                   b2 = __ballot_sync(FULL, cr[2]), b3 = __ballot_sync(FULL, cr[3]);
                    const int total = __popc(b0) + __popc(b1) + __popc(b2) + __popc(b3);
                    if (total) {
                        const unsigned lt = (1u << lane) - 1u;
                        int pos = cur + __popc(b0 & lt) + __popc(b1 & lt) + __popc(b2 & lt) + __popc(b3 & lt);
#pragma unroll
                        for (int j = 0; j < 4; ++j)
                            if (cr[j]) {
                                vslam_corner k;
                                k.row = yo; k.col = c + j; k.value = Rp[j];
                                s_buf[warp][pos++] = k;
                            }
                    }
                    if (lane == 0) { s_rowoff[warp][ri] = cur; s_rowcnt[warp][ri] = total; }
                    cur += total;
                    ++ri;
                }
            }
            // the new row becomes the output row of the next step
#pragma unroll
            for (int j = 0; j < 4; ++j) { Rp[j] = Rn[j]; if (want8) { q8p[j] = q8n[j]; m2p[j] = m2n[j]; } }
        }
    }
    if (a.want_corners) flush();
}

// ---------------------------------------------------------------------------------------------------------------
// harris_fast_k: the same pipeline for the shapes the throughput path sees (cols, pitch, frame stride multiples of 4,
// aligned pointers, sizes below 65536), reorganised around the instruction count.  Round 2 measured harris_march_k at 435
// warp instructions per warp row (139 thread instructions per pixel), a quarter of them control: predicates re-evaluated per
// row, BSSY / BSYNC / BRA around conditionally updated state, register moves of three-generation chains in a loop unrolled
// by four.  Here:
//   * a group of four rows runs one of two instantiations of the row body: STEADY (no image border in y within reach,
//     every output on: no row predicate at all; ~84 % of the groups of a band) or general (the warm-up and the groups at
//     the top / bottom border).  Image borders in x sit behind warp-uniform branches that interior strips never take; y
//     borders copy a ring slot (replicate padding of the gradient rows) instead of choosing among formulas;
//   * every state with three generations lives in a ring indexed by row & 3 (compile time), so nothing moves;
//   * the blurred row travels to the neighbour lanes once, packed; gradients of the halo columns are formed locally from
//     unpacked ints (no packed ix|iy words, no second exchange, no bias to remove before the products);
//   * the 8-bit response, its dilate and the strict '>' run packed 16x2 (VIMNMX.U16x2, one biased subtraction);
//   * NMS2's vertical window is two pair-maxima; corners are staged as 8-byte entries (row << 16 | col, value), row
//     bookkeeping stays in registers (lane r holds row r of the batch) and a band is 216 rows (warm-up 4 %).
// Arithmetic, emission order and outputs are those of harris_march_k bit for bit (tests: ...harris_kernel_variants_agree).
// ---------------------------------------------------------------------------------------------------------------
#ifndef HARRIS_FAST_MIN_CTAS
#define HARRIS_FAST_MIN_CTAS 3   // 168 registers, no spills: 3.17 ms per 1080p x 256; 4 CTAs (128 registers, spills) 3.63; 5 CTAs 4.87
#endif
constexpr int HBF = 216;                   // rows per band of the fast kernel (divides 1080, 2160, 4320)
constexpr int HF_CAP = 512;                // staged corners per warp before a flush
constexpr int HF_ROWS = 8;                 // rows per flush

template <int N> using IC = std::integral_constant<int, N>;

template <int P2, int CTAS>
__global__ void __launch_bounds__(32 * H_WARPS, CTAS) harris_fast_k(const HarrisArgs a) {
    __shared__ uint2 s_buf[H_WARPS][HF_CAP];
    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int unit = blockIdx.x * H_WARPS + warp;
    if (unit >= a.n_strips * a.n_bands) return;                 // warps are independent: no block-wide barrier below
    const int strip = unit % a.n_strips, band = unit / a.n_strips, frame = blockIdx.z;
    const int rows = a.rows, cols = a.cols;
    const int X0 = strip * HS, c = X0 - 8 + 4 * lane;           // this lane's 4 pixels: columns c .. c+3 (cols % 4 == 0: all in or all out)
    const int Y0 = band * HBF, Y1 = min(Y0 + HBF, rows);
    const u8 *__restrict__ img = a.frames + (size_t)frame * a.frame_stride;
    const bool xedge = X0 < 8 || X0 + 120 >= cols;              // warp-uniform: some lane sits on or beyond an image border
    const bool lane_in = c >= 0 && c < cols;
    const bool own = lane >= 2 && lane < 30 && lane_in;
    const bool repL = c == 0, repR = c + 4 == cols;             // replicate padding of the gradient maps (row M)
    const size_t fpx = (size_t)frame * rows * cols;
    const bool want8 = a.r8 || a.mask3;
    const float kk = a.k;

    u32 raw[4][2];                                              // raw rows (16x2), ring by row & 3
    int U[4][6];                                                // blurred rows at columns c-1 .. c+4 as ints, ring by row & 3
    int ixm[6];                                                 // ix of row yr-2 at columns c-1 .. c+4
    int hs[4][3][4];                                            // horizontal 3-sums of xx, yy, xy; ring by row & 3
    float PY[4][4];                                             // NMS2: PY[r & 3] = max of the window maxima of rows r, r+1
    float hB[4], Rp[4];                                         // window maxima and R of the output row yr-4
    u32 Q[2] = {0, 0}, m2[2] = {0, 0}, M3[4][2];                // 8-bit response (16x2) of row yo, its left/right maxima; 3-wide maxima ring
#pragma unroll
    for (int s = 0; s < 4; ++s) {
        raw[s][0] = raw[s][1] = 0; M3[s][0] = M3[s][1] = 0;
#pragma unroll
        for (int i = 0; i < 6; ++i) U[s][i] = 0;
#pragma unroll
        for (int j = 0; j < 4; ++j) { hs[s][0][j] = hs[s][1][j] = hs[s][2][j] = 0; PY[s][j] = 0.f; }
    }
#pragma unroll
    for (int i = 0; i < 6; ++i) ixm[i] = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) { hB[j] = 0.f; Rp[j] = 0.f; }
    int cur = 0, ri = 0, batch_y0 = Y0, my_off = 0, my_cnt = 0; // staged corners / rows of the current flush batch; lane r: row r's range

    auto flush = [&]() {
        __syncwarp();
        if (cur > 0) {
            int base = 0;
            if (lane == 0) base = atomicAdd(a.cursor, cur);
            base = __shfl_sync(FULL, base, 0);
            for (int i = lane; i < cur; i += 32)
                if (base + i < a.staging_cap) {
                    const uint2 e = s_buf[warp][i];
                    vslam_corner k;
                    k.row = (int)(e.x >> 16); k.col = (int)(e.x & 0xffffu); k.value = __uint_as_float(e.y);
                    a.staging[base + i] = k;
                }
            if (lane < ri && my_cnt > 0) {
                const size_t sg = ((size_t)frame * rows + (batch_y0 + lane)) * a.n_strips + strip;
                a.seg_off[sg] = base + my_off;
                a.seg_cnt[sg] = my_cnt;
            }
        }
        __syncwarp();
        cur = 0; ri = 0;
    };

    // raw row yr; steady groups touch neither border
    auto load_row = [&](int yr, bool plain) -> u32 {
        if (plain) return *(const u32 *)(img + (size_t)yr * a.pitch + c);
        int ys = yr;
        if ((unsigned)ys >= (unsigned)rows) ys = refl101(ys, rows);
        const u8 *__restrict__ src = img + (size_t)ys * a.pitch;
        if (lane_in) return *(const u32 *)(src + c);
        return (u32)src[refl101(c, cols)] | (u32)src[refl101(c + 1, cols)] << 8 |
               (u32)src[refl101(c + 2, cols)] << 16 | (u32)src[refl101(c + 3, cols)] << 24;
    };
    u32 wc[4], wn[4];                                           // the raw rows of this group and of the next: loads stay one group ahead

    // one row of the pipeline; ph = yr & 3 and ST (steady group) are compile-time
    auto row = [&](auto ph_, auto st_, const int yr) {
        constexpr int ph = decltype(ph_)::value;
        constexpr bool ST = decltype(st_)::value != 0;
        constexpr int s0 = ph, s1 = (ph + 1) & 3, s2 = (ph + 2) & 3, s3 = (ph + 3) & 3;   // ring slots of rows yr, yr-3, yr-2, yr-1
        const u32 w = wc[ph];
        raw[s0][0] = pk(w, 0, 0x4140); raw[s0][1] = pk(w, 0, 0x4342);
        // ---- row K: blurred row yr-1 = (121 x 121 + 8) >> 4 ----------------------------------------------------------
        const u32 V0 = raw[s2][0] + 2 * raw[s3][0] + raw[s0][0], V1 = raw[s2][1] + 2 * raw[s3][1] + raw[s0][1];
        u32 Bn0, Bn1;
        {
            const u32 L = __shfl_up_sync(FULL, V1, 1), R = __shfl_down_sync(FULL, V0, 1);
            const u32 Vm = pk(L, V0, 0x5432), Vc = pk(V0, V1, 0x5432), Vp = pk(V1, R, 0x5432);
            Bn0 = ((Vm + 2 * V0 + Vc + 0x00080008u) >> 4) & 0x00ff00ffu;
            Bn1 = ((Vc + 2 * V1 + Vp + 0x00080008u) >> 4) & 0x00ff00ffu;
        }
        // ---- row L: the new blurred row at columns c-2 .. c+5 as ints; ix of row yr-1, iy of row yr-2 -------------------
        int ixn[6], iy[6];
        {
            int(&un)[6] = U[s3];
            const u32 BL = __shfl_up_sync(FULL, Bn1, 1), BR = __shfl_down_sync(FULL, Bn0, 1);
            const int um2 = (int)(BL & 0xffffu), up5 = (int)(BR >> 16);
            un[0] = (int)(BL >> 16); un[1] = (int)(Bn0 & 0xffffu); un[2] = (int)(Bn0 >> 16);
            un[3] = (int)(Bn1 & 0xffffu); un[4] = (int)(Bn1 >> 16); un[5] = (int)(BR & 0xffffu);
            ixn[0] = un[1] - um2; ixn[5] = up5 - un[4];
#pragma unroll
            for (int i = 1; i < 5; ++i) ixn[i] = un[i + 1] - un[i - 1];
#pragma unroll
            for (int i = 0; i < 6; ++i) iy[i] = un[i] - U[s1][i];
            if (xedge) {                                        // replicate: column -1 := column 0, column cols := column cols-1
                if (repL) { ixn[0] = ixn[1]; iy[0] = iy[1]; }
                if (repR) { ixn[5] = ixn[4]; iy[5] = iy[4]; }
            }
        }
        // ---- row M, horizontal part: 3-sums of the products of row g = yr-2 -> ring slot s2 ------------------------------
        {
            int xx[6], yy[6], xy[6];
#pragma unroll
            for (int i = 0; i < 6; ++i) { xx[i] = ixm[i] * ixm[i]; yy[i] = iy[i] * iy[i]; xy[i] = ixm[i] * iy[i]; }
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                hs[s2][0][j] = xx[j] + xx[j + 1] + xx[j + 2];
                hs[s2][1][j] = yy[j] + yy[j + 1] + yy[j + 2];
                hs[s2][2][j] = xy[j] + xy[j + 1] + xy[j + 2];
            }
            if (!ST) {                                          // replicate padding in y: gradient row -1 := row 0, row `rows` := row rows-1
                const int g = yr - 2;
                if (g == 0) {
#pragma unroll
                    for (int j = 0; j < 4; ++j) { hs[s1][0][j] = hs[s2][0][j]; hs[s1][1][j] = hs[s2][1][j]; hs[s1][2][j] = hs[s2][2][j]; }
                }
                if (g == rows) {
#pragma unroll
                    for (int j = 0; j < 4; ++j) { hs[s2][0][j] = hs[s1][0][j]; hs[s2][1][j] = hs[s1][1][j]; hs[s2][2][j] = hs[s1][2][j]; }
                }
            }
        }
#pragma unroll
        for (int i = 0; i < 6; ++i) ixm[i] = ixn[i];
        // ---- rows M (vertical part), N: response of row rb = yr-3 (ring slots: rb-1 -> s0, rb -> s1, rb+1 -> s2) ---------------
        const int rb = yr - 3;
        float Rn[4];
        if (ST || (rb >= 0 && rb < rows)) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int sx = hs[s0][0][j] + hs[s1][0][j] + hs[s2][0][j], sy = hs[s0][1][j] + hs[s1][1][j] + hs[s2][1][j],
                          sxy = hs[s0][2][j] + hs[s1][2][j] + hs[s2][2][j];
                const long long det_i = (long long)sx * sy - (long long)sxy * sxy;                  // exact; == double det2
                const float det = __ll2float_rn(det_i);
                const float tr = __fsub_rn(__int_as_float(0x4B000000 | (sx + sy)), 8388608.0f);   // exact int -> float
                const float r = __fsub_rn(det, __fmul_rn(kk, __fmul_rn(tr, tr)));                  // Harris_corners.cpp:57
                Rn[j] = fmaxf(r, 0.f);                                                             // :60-62
            }
            if (xedge && !lane_in) {
#pragma unroll
                for (int j = 0; j < 4; ++j) Rn[j] = 0.f;
            }
        } else {
#pragma unroll
            for (int j = 0; j < 4; ++j) Rn[j] = 0.f;
        }
        // ---- row P, horizontal part: window maxima of row rb ------------------------------------------------------------------
        float hn[4];
        {
            const float l3 = __shfl_up_sync(FULL, Rn[3], 1);
            if (P2 == 2) {
                const float l2 = __shfl_up_sync(FULL, Rn[2], 1), r0 = __shfl_down_sync(FULL, Rn[0], 1);
                const float q0 = fmaxf(l2, l3), q1 = fmaxf(l3, Rn[0]), q2 = fmaxf(Rn[0], Rn[1]), q3 = fmaxf(Rn[1], Rn[2]),
                            q4 = fmaxf(Rn[2], Rn[3]), q5 = fmaxf(Rn[3], r0);
                hn[0] = fmaxf(q0, q2); hn[1] = fmaxf(q1, q3); hn[2] = fmaxf(q2, q4); hn[3] = fmaxf(q3, q5);
            } else {
                hn[0] = fmaxf(l3, Rn[0]); hn[1] = fmaxf(Rn[0], Rn[1]); hn[2] = fmaxf(Rn[1], Rn[2]); hn[3] = fmaxf(Rn[2], Rn[3]);
            }
        }
        // ---- row O, horizontal part: 8-bit response of row rb (16x2), its 3-wide maxima -> ring slot s1 -----------------------
        u32 Qn0 = 0, Qn1 = 0, m2n0 = 0, m2n1 = 0;
        if (want8) {
            u32 mb[4];
#pragma unroll
            for (int j = 0; j < 4; ++j)     // saturate_cast<uchar>(cvRound(r)): round half to even in the low mantissa bits; cvRound(r >= 2^31) = INT_MIN -> 0
                mb[j] = __float_as_uint(__fadd_rn(fminf(Rn[j] >= 2147483648.f ? 0.f : Rn[j], 255.0f), 12582912.0f));
            Qn0 = pk(mb[0], mb[1], 0x5410); Qn1 = pk(mb[2], mb[3], 0x5410);     // bytes 0, 1 of 0x4B4000vv are vv, 00
            const u32 QL = __shfl_up_sync(FULL, Qn1, 1), QR = __shfl_down_sync(FULL, Qn0, 1);
            const u32 Qm = pk(QL, Qn0, 0x5432), Qc = pk(Qn0, Qn1, 0x5432), Qq = pk(Qn1, QR, 0x5432);
            m2n0 = __vmaxu2(Qm, Qc); m2n1 = __vmaxu2(Qc, Qq);
            M3[s1][0] = __vmaxu2(m2n0, Qn0); M3[s1][1] = __vmaxu2(m2n1, Qn1);
        }
        // ---- outputs of row yo = yr-4 ((yo & 3) == ph) -------------------------------------------------------------------------
        const int yo = yr - 4;
        float pn[4];                                            // pair maxima of rows (yo, yo+1)
#pragma unroll
        for (int j = 0; j < 4; ++j) pn[j] = fmaxf(hB[j], hn[j]);
        if (ST || (yo >= Y0 && yo < Y1)) {
            const size_t o = fpx + (size_t)yo * cols + c;
            if (own) {
                if (a.response) *(float4 *)(a.response + o) = make_float4(Rp[0], Rp[1], Rp[2], Rp[3]);
                if (a.r8) *(u32 *)(a.r8 + o) = pk(Q[0], Q[1], 0x6420);
                if (a.mask3) {
                    // centre-excluded 3x3 dilate of the 8-bit response, strict '>' (Harris_corners.cpp:76-79): d + 256 - q
                    // has bit 8 set exactly when d >= q
                    const u32 d0 = __vmaxu2(__vmaxu2(M3[s3][0], M3[s1][0]), m2[0]), d1 = __vmaxu2(__vmaxu2(M3[s3][1], M3[s1][1]), m2[1]);
                    const u32 t0 = (~(d0 + 0x01000100u - Q[0]) >> 8) & 0x00010001u, t1 = (~(d1 + 0x01000100u - Q[1]) >> 8) & 0x00010001u;
                    *(u32 *)(a.mask3 + o) = pk(t0 * 255u, t1 * 255u, 0x6420);
                }
            }
            // NMS2 (Harris_corners.cpp:94-118): R >= max of the half-open window, interior pixels, positive responses
            if (a.want_corners) {
                bool cr[4];
                const bool lok = ST ? own : (own && yo >= P2 && yo < rows - P2);
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float m = P2 == 2 ? fmaxf(PY[s2][j], pn[j]) : PY[s3][j];       // rows yo-2 .. yo+1 / yo-1, yo
                    cr[j] = lok && Rp[j] > 0.f && Rp[j] >= m;
                }
                if (xedge) {                                    // NMS2 interior columns (Harris_corners.cpp:95)
#pragma unroll
                    for (int j = 0; j < 4; ++j) cr[j] = cr[j] && c + j >= P2 && c + j < cols - P2;
                }
                if (!ST && (cur + HS > HF_CAP || ri == HF_ROWS)) flush();
                if (ri == 0) batch_y0 = yo;
                const unsigned b0 = __ballot_sync(FULL, cr[0]), b1 = __ballot_sync(FULL, cr[1]),
                               b2 = __ballot_sync(FULL, cr[2]), b3 = __ballot_sync(FULL, cr[3]);
                const int total = __popc(b0) + __popc(b1) + __popc(b2) + __popc(b3);
                if (total) {
                    const unsigned lt = (1u << lane) - 1u;
                    int pos = cur + __popc(b0 & lt) + __popc(b1 & lt) + __popc(b2 & lt) + __popc(b3 & lt);
                    const u32 rc = (u32)yo << 16 | (u32)c;
#pragma unroll
                    for (int j = 0; j < 4; ++j)
                        if (cr[j]) s_buf[warp][pos++] = make_uint2(rc + j, __float_as_uint(Rp[j]));
                }
                if (lane == ri) { my_off = cur; my_cnt = total; }
                cur += total;
                ++ri;
            }
        }
        // ---- the new row becomes the output row of the next step ------------------------------------------------------------------
#pragma unroll
        for (int j = 0; j < 4; ++j) { Rp[j] = Rn[j]; PY[s0][j] = pn[j]; hB[j] = hn[j]; }
        if (want8) { Q[0] = Qn0; Q[1] = Qn1; m2[0] = m2n0; m2[1] = m2n1; }
    };

    const int yr_begin = (Y0 - 5) & ~3;                         // ring slots are compile-time: start on a multiple of 4
#pragma unroll
    for (int ph = 0; ph < 4; ++ph) wn[ph] = load_row(yr_begin + ph, false);
    for (int yr4 = yr_begin; yr4 <= Y1 + 3; yr4 += 4) {
#pragma unroll
        for (int ph = 0; ph < 4; ++ph) wc[ph] = wn[ph];
        // steady group: rows yr4 .. yr4+3 and the prefetched yr4+4 .. yr4+7 lie inside the image, gradient rows yr-2 are neither
        // 0 nor `rows`, response rows yr-3 exist, output rows yr-4 lie in [Y0, Y1) and are NMS2-interior
        const bool steady = yr4 >= max(Y0, P2) + 4 && yr4 + 7 < rows && yr4 + 1 <= Y1;
        if (steady) {
#pragma unroll
            for (int ph = 0; ph < 4; ++ph) wn[ph] = load_row(yr4 + 4 + ph, !xedge);
            if (a.want_corners && (cur + 4 * HS > HF_CAP || ri + 4 > HF_ROWS)) flush();
            row(IC<0>(), IC<1>(), yr4); row(IC<1>(), IC<1>(), yr4 + 1); row(IC<2>(), IC<1>(), yr4 + 2); row(IC<3>(), IC<1>(), yr4 + 3);
        } else {
            if (yr4 + 4 <= Y1 + 3) {
#pragma unroll
                for (int ph = 0; ph < 4; ++ph) wn[ph] = load_row(yr4 + 4 + ph, false);
            }
            row(IC<0>(), IC<0>(), yr4); row(IC<1>(), IC<0>(), yr4 + 1); row(IC<2>(), IC<0>(), yr4 + 2); row(IC<3>(), IC<0>(), yr4 + 3);
        }
    }
    if (a.want_corners) flush();
}

int harris_strips(int cols) { return (cols + HS - 1) / HS; }

int launch_harris(vslam_ctx *ctx, HarrisArgs &a, int n_frames) {
    if (n_frames <= 0 || n_frames > 65535) VS_FAIL(ctx, VSLAM_E_INVALID, "harris: %d frames per launch", n_frames);
    a.n_strips = harris_strips(a.cols);
    // harris_fast_k wants whole 4-pixel words everywhere: every other shape (and VSLAM_HARRIS_FAST=0) runs harris_march_k
    static const bool fast_env = [] { const char *e = getenv("VSLAM_HARRIS_FAST"); return !e || atoi(e) != 0; }();
    const bool fast = fast_env && !(a.cols & 3) && !(a.pitch & 3) && !(a.frame_stride & 3) && !((uintptr_t)a.frames & 3) &&
                      !((uintptr_t)a.response & 15) && !((uintptr_t)a.r8 & 3) && !((uintptr_t)a.mask3 & 3) && a.rows < 65536 && a.cols < 65536;
    a.n_bands = fast ? (a.rows + HBF - 1) / HBF : (a.rows + HB - 1) / HB;
    const dim3 grid((a.n_strips * a.n_bands + H_WARPS - 1) / H_WARPS, 1, n_frames);
    KernelSpan span(ctx, VSLAM_K_HARRIS);
    if (fast) {
        if (a.nms2_pad == 2) harris_fast_k<2, HARRIS_FAST_MIN_CTAS><<<grid, 32 * H_WARPS, 0, ctx->stream>>>(a);
        else harris_fast_k<1, HARRIS_FAST_MIN_CTAS><<<grid, 32 * H_WARPS, 0, ctx->stream>>>(a);
    } else if (a.nms2_pad == 2) harris_march_k<2><<<grid, 32 * H_WARPS, 0, ctx->stream>>>(a);
    else harris_march_k<1><<<grid, 32 * H_WARPS, 0, ctx->stream>>>(a);
    VS_LAUNCH_CHECK(ctx);
    return VSLAM_OK;
}

}  // namespace vslam
