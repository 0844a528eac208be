// fused_umma.cu -- the fused octave kernel on Blackwell's own instruction forms: TMA tile loads (cp.async.bulk.tensor),
// tcgen05.mma.kind::i8 issued by one thread with the accumulators in tensor memory, tcgen05.ld epilogues.
//
// Same rows of SURVEY.md 8(a) as fused.cu / fused_mma.cu (B, C, D, G, H, I, J) and the same results bit for bit: OpenCV's
// 8U GaussianBlur is exact integer arithmetic (GaussPyramid.cpp:166-185 blurs every level from the octave base), so each
// pass is a product with a banded Toeplitz matrix of taps.  Because both passes are exact until the single rounding at the
// end, their order is free; this kernel runs the VERTICAL pass first:
//
//   ring     the octave base (reflect-101 margins on all four sides in HBM) arrives in chunks of 32 rows x 128 bytes by TMA,
//            128B-swizzled.  That layout is at once the MN-major operand of the V pass (K = rows) -- no transposition, no
//            staging code, no border path.
//   V pass   D_v[128 out rows x NV cols] = T_v[128 x 32 KV] * ring chunks; all six levels read the same ring.
//   X        warps read D_v from TMEM (thread = row), split the 16-bit sums into two byte planes and store them K-major for
//            the H pass (the rounding bias of that pass is one more tcgen05.mma into the low-byte accumulator).
//   H pass   D_lo / D_hi[128 rows x 64 out cols] = X_lo / X_hi[128 x 32 KH] * T_h[64 x 32 KH]^T.
//   epilogue out = (256 D_hi + D_lo) >> 16, saturating DoG against the previous level (registers), byte tiles in shared
//            memory for the detector.
//   T_v, T_h banded Toeplitz matrices are Hankel once their rows are reversed: 8 x 16-byte core matrix j of a level holds
//            tap[8 j + row + byte - shift], and the no-swizzle K-major descriptor (SBO 128, LBO 256) walks OVERLAPPING core
//            matrices -- 4.4 KB per level instead of 20 KB.  Reversed rows only permute TMEM lanes / columns.
//
// One CTA per SM, thirty warps (960 threads, 64 registers): 24 tile warps that turn D_lo / D_hi into Gaussian / DoG tile rows
// and then run the detector (8 columns each), 4 X warps that turn D_v into the byte planes, a control warp (TMA producer +
// the single MMA-issuing thread) and a service warp (queued localisations, ordered emission); TMEM lane quadrant =
// warp % 4.  D_v and D_h are double buffered in TMEM, so the tensor pipe, the X warps and the tile warps work on three
// different levels at once.  The detector, localisation, edge test, ordered emission and decimation are those of
// fused_mma.cu on a row-major byte tile.  Measured against fused_mma.cu in DESIGN.md 4.1b (slower: opt-in, VSLAM_UMMA=1).
#include <math.h>
#include <string.h>

#include <mutex>
#include <type_traits>
#include <vector>

#include "detect_math.cuh"
#include "fused_umma.cuh"
#include "umma.cuh"

namespace vslam {

using namespace umma;

namespace {
constexpr int FR = kUmmaRows, FWO = kUmmaFWO, NS = FWO / 3;
constexpr int kThreads = 960;                   // 28 epilogue warps (7 per TMEM lane quadrant), the control warp, the service warp
constexpr int NE = 896;                         // threads of the epilogue warps
constexpr int W_CTL = 28, W_SVC = 29;
constexpr int TROWS = FR + 4;                   // tile rows: 4 rows of history + the step
constexpr int TILE_B = TROWS * 64;              // bytes per tile: 64 columns (49 used), 16-byte chunks XOR-swizzled by the row pair
constexpr int N_G = 3, N_D = 5;                 // G tiles: Gaussian levels 1..3; D tiles: DoG levels 0..4
constexpr int MAX_SROW = FR / 3 + 2;            // sample rows of one step
constexpr int kTmemCols = 512;
}  // namespace

// ---------------------------------------------------------------------------------------------------------------
// host: geometry and tables
// ---------------------------------------------------------------------------------------------------------------
int umma_geometry(double sigma0, int scale_samples, int octave, UmmaGeom *g) {
    memset(g, 0, sizeof *g);
    const double k = pow(2.0, 1.0 / (double)scale_samples);          // GaussPyramid.hpp:69
    int32_t t[kMaxKsize];
    for (int l = 0; l < 6; ++l) {
        const double sig = pow(2, octave) * sigma0 * pow(k, l);      // GaussPyramid.cpp:157
        const int n = host_gauss_ksize(sig);
        if (n > 63) return 1;
        if (host_gauss_taps(sig, n, t)) return 1;
        g->n[l] = n; g->r[l] = n / 2;
        for (int i = 0; i < n; ++i) {
            if (t[i] < 0 || t[i] > 255) return 1;
            g->taps[l][i] = (u8)t[i];
        }
        if (g->r[l] > g->rmax) g->rmax = g->r[l];
    }
    g->P = g->rmax <= 15 ? 16 : 32;
    g->KV = (FR + 2 * g->P) / 32;
    g->NV = (g->P + FWO + g->rmax + 15) / 16 * 16;
    g->KH = (g->NV + 31) / 32;
    g->JV = 14 + 4 * g->KV; g->JH = 6 + 4 * g->KH;
    g->tab_bytes = 6 * (g->JV + g->JH) * 128;
    return 0;
}

struct UmmaArgs : OctaveIO {
    const uint4 *tab_image;         // compact Toeplitz tables of the six levels: V tables, then H tables
    int plane_x0, plane_y0;         // coordinates of image pixel (0, 0) inside the TMA tensor (the padded plane)
    long long *dbg;                 // VSLAM_UMMA_DBG=1: cycle counters of CTA (0, 0, 0) per role (diagnosis only)
};

namespace {

__device__ __forceinline__ void named_bar(int id, int n) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory"); }

__device__ __forceinline__ void tmem_ld8(uint32_t taddr, uint32_t (&v)[8]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]) : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld4(uint32_t taddr, uint32_t (&v)[4]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]) : "r"(taddr));
}

// 16 consecutive columns as 8 registers: the low 16 bits of two adjacent columns per register (the 8.8 sums fit 16 bits)
__device__ __forceinline__ void tmem_ld16_pack(uint32_t taddr, uint32_t (&v)[8]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.pack::16b.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]) : "r"(taddr));
}

// byte offset of pixel (tile row tr, column n) inside a tile
__device__ __forceinline__ int toff(int tr, int n) { return tr * 64 + ((((n >> 4) ^ (tr >> 1)) & 3) << 4) + (n & 15); }

// computeEdgeResponse + threshold (Diff_of_Gauss.cpp:79-109, :336) on the G tile `gt` of a level around sample (xc, yc).
// rb = ring row of row y0 - 4 (the tiles are rings of TROWS rows); tile column of image column x is x - X0 + 1.
// One copy in the kernel (instruction cache): the detector calls it from a single site per level.
__device__ __noinline__ bool edge_pass_tile(const u8 *gt, int xc, int yc, int y0, int rb, int X0, int rows, int cols, float edge_thr) {
    int ix2 = 0, iy2 = 0, ixy = 0;
#pragma unroll
    for (int u = 0; u <= 1; ++u) {
        const int yy = yc + u;
        const int yd = (yy + 1 < rows ? yy + 1 : max(rows - 2, 0)), yu = (yy > 0 ? yy - 1 : min(1, rows - 1));   // reflect-101, single bounce
        int ty = rb + yy - y0 + 4, tyd = rb + yd - y0 + 4, tyu = rb + yu - y0 + 4;
        ty -= ty >= TROWS ? TROWS : 0; tyd -= tyd >= TROWS ? TROWS : 0; tyu -= tyu >= TROWS ? TROWS : 0;
#pragma unroll
        for (int v = 0; v <= 1; ++v) {
            const int xx = xc + v;
            const int xr = (xx + 1 < cols ? xx + 1 : max(cols - 2, 0)), xl = (xx > 0 ? xx - 1 : min(1, cols - 1));
            const int tx = xx - X0 + 1;
            const int gxv = (int)gt[toff(ty, xr - X0 + 1)] - (int)gt[toff(ty, xl - X0 + 1)];
            const int gyv = (int)gt[toff(tyd, tx)] - (int)gt[toff(tyu, tx)];
            ix2 += gxv * gxv; iy2 += gyv * gyv; ixy += gxv * gyv;
        }
    }
    return edge_from_sums((float)ix2, (float)iy2, (float)ixy) < edge_thr;
}

}  // namespace

// KV row chunks / KH column chunks / NV staged columns: (5, 3, 80) for kernels up to 31 taps, (6, 4, 112) up to 63 taps
// XBUF: buffers of the byte planes X (2 when shared memory allows)
template <int KV, int KH, int NV, int XBUF>
__global__ void __launch_bounds__(kThreads, 1) octave_umma_k(const __grid_constant__ CUtensorMap tmap, const __grid_constant__ UmmaArgs a) {
    constexpr int P = 16 * (KV - 4);                // rows above / columns left of the strip in the staged window
    constexpr int JV = 14 + 4 * KV, JH = 6 + 4 * KH, TVB = JV * 128, THB = JH * 128;
    constexpr int RC = KV + 4;                      // ring chunks: one step's window + the next step's four new chunks
    constexpr int XPB = 2 * KH * 2048;              // bytes of one byte plane: 2 KH K-units of [16 row groups][8 rows][16 bytes]
    constexpr int NP = NV / 16, PMID = (NP + 1) / 2;   // 16-column pieces of D_v: the first PMID go to column half 0
    constexpr int NFLAG = (3 * MAX_SROW + 31) / 32, NRES = 3 * MAX_SROW * NS;
    constexpr int QS = 64;                          // queued samples of a step the service warp finishes; the detector threads take the rest
    extern __shared__ __align__(1024) u8 smem_raw[];
    u8 *smem = (u8 *)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    u8 *ring = smem;
    u8 *xbuf = ring + RC * 4096;
    u8 *tab = xbuf + XBUF * 2 * XPB;
    u8 *tiles = tab + 6 * (TVB + THB);
    u8 *biasop = tiles + (N_G + N_D) * TILE_B;     // operands of the rounding-bias MMA: 512 bytes of 0x80, then 8 x [row pattern | zeros]
    __shared__ __align__(8) unsigned long long s_bar[17];
    __shared__ uint32_t s_tmem;
    // detector results of a step, double buffered: the service warp finishes and emits step s while the others are in s + 1
    __shared__ short s_res[2][NRES];                          // keypoint value per (sample row, level, sample column), or -1
    __shared__ uint2 s_queue[NRES];                           // samples that need the full localisation, with their derivatives and edge test
    __shared__ unsigned int s_flag[2][NFLAG];                 // segments (sample row, level) of the step with survivors
    __shared__ int s_nq[2];                                   // queue lengths: front (full localisation), back (zero derivative)
    __shared__ float s_div[511];                              // float(i - 255) / 255f, the reference's divisions

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int strip = blockIdx.x, frame = blockIdx.z;
    const int rows = a.rows, cols = a.cols;
    const int X0 = strip * FWO;
    const int ys = blockIdx.y * a.seg_rows, ye = min(ys + a.seg_rows, rows);
    const int n_steps = (ye - ys + FR - 1) / FR;

    // barriers: 0,1 ring full; 2,3 D_v full; 4,5 D_v free; 6,7 X full; 8,9 D_h full; 10,11 D_h free; 12,13 detector results
    // full; 14,15 detector results free; 16 queue free
    const uint32_t bar0 = smem_u32(&s_bar[0]);
    auto bar = [&](int i) { return bar0 + 8u * (uint32_t)i; };
    if (tid == 0) {
        mbar_init(bar(0), 1); mbar_init(bar(1), 1);
        mbar_init(bar(2), 1); mbar_init(bar(3), 1);
        mbar_init(bar(4), 4); mbar_init(bar(5), 4);
        mbar_init(bar(6), 4); mbar_init(bar(7), 4);
        mbar_init(bar(8), 1); mbar_init(bar(9), 1);
        mbar_init(bar(10), 24); mbar_init(bar(11), 24);
        mbar_init(bar(12), 28); mbar_init(bar(13), 28);
        mbar_init(bar(14), 1); mbar_init(bar(15), 1); mbar_init(bar(16), 1);
        fence_barrier_init();
    }
    if (warp == 0) tmem_alloc(smem_u32(&s_tmem), kTmemCols);
    for (int i = tid; i < 6 * (TVB + THB) / 16; i += kThreads) ((uint4 *)tab)[i] = a.tab_image[i];
    for (int i = tid; i < 511; i += kThreads) s_div[i] = __fdiv_rn((float)(i - 255), 255.0f);
    // rounding bias of the second pass, (sum + 32768) >> 16, as one more tcgen05.mma per level into the low-byte accumulator:
    // A = 128 x 32 bytes of 0x80 (every core matrix aliases the same 512-byte block), B[n][0] = B[n][1] = 0x80, else 0:
    // D_lo += 2 * 128 * 128.  Costs ~50 tensor-pipe cycles per level and removes an add per 8.8 sum from the X warps.
    for (int i = tid; i < (512 + 8 * 256) / 4; i += kThreads) {
        const int o = 4 * i;
        ((uint32_t *)biasop)[i] = o < 512 ? 0x80808080u : ((((o - 512) & 255) < 128 && ((o - 512) & 15) == 0) ? 0x00008080u : 0u);
    }
    if (tid < 2 * NFLAG) (&s_flag[0][0])[tid] = 0;
    if (tid < 2) s_nq[tid] = 0;
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tm = s_tmem;
    // TMEM columns: D_v buffers at 0 / 128, D_h buffers at 256 / 384 (low-byte plane, +64 high-byte plane)
    // completion count of a per-level barrier pair (index l & 1) at level l of step s: 3 s + (l >> 1)
#ifdef VSLAM_UMMA_DBG_BUILD                                     // cycle counters per warp in shared memory (lane 0 only): a separate build
    __shared__ long long s_dbg[32][8];                          // slots: 0 wait, 1 work, 2 det, 3 tests, 4 decim, 5 edge, 6 tail, 7 last stamp
    if (tid < 256) (&s_dbg[0][0])[tid] = 0;
    __syncthreads();
    if (lane == 0) s_dbg[warp][7] = clock64();
#define c_wait 0
#define c_work 1
#define c_det 2
#define c_d1 3
#define c_d2 4
#define c_d3 5
#define c_d4 6
#define c_xfull 0
#define c_free 1
#define DBG_T(slot) do { if (lane == 0) { const long long t_ = clock64(); s_dbg[warp][slot] += t_ - s_dbg[warp][7]; s_dbg[warp][7] = t_; } } while (0)
#define DBG_T2(slot, slot2) do { if (lane == 0) { const long long t_ = clock64(); s_dbg[warp][slot] += t_ - s_dbg[warp][7]; s_dbg[warp][slot2] += t_ - s_dbg[warp][7]; s_dbg[warp][7] = t_; } } while (0)
#define DBG_WAIT(slot, stmt) do { const long long t_ = clock64(); stmt; s_dbg[warp][slot] += clock64() - t_; } while (0)
#define DBG_ONLY(stmt) stmt
#else
#define DBG_T(acc) do { } while (0)
#define DBG_T2(acc, acc2) do { } while (0)
#define DBG_WAIT(acc, stmt) do { stmt; } while (0)
#define DBG_ONLY(stmt)
#endif

    if (warp == W_CTL) {
        // =============================== control: TMA producer + MMA issuer (one thread) ===============================
        if (lane == 0) {
            const int tx = X0 - P + a.plane_x0;                       // first staged column inside the padded plane: multiple of 16
            const int ty = ys - P + a.plane_y0;
            const uint32_t ring_sa = smem_u32(ring), x_sa = smem_u32(xbuf), tab_sa = smem_u32(tab);
            constexpr uint32_t iv = idesc_u8(NV, 0, 1), ih = idesc_u8(64, 0, 0);
            // descriptor halves that never change
            constexpr uint64_t tv_d = smem_desc(0, 256, 128, kLayoutNone), img_d = smem_desc(0, 4096, 1024, kLayoutSw128),
                               x_d = smem_desc(0, 2048, 128, kLayoutNone);
            const uint64_t ba_d = smem_desc(smem_u32(biasop), 16, 16, kLayoutNone), bb_d = smem_desc(smem_u32(biasop) + 512, 128, 256, kLayoutNone);
            mbar_expect_tx(bar(0), KV * 4096);
            for (int k = 0; k < KV; ++k) tma_load_3d(ring_sa + 4096 * k, &tmap, bar(0), tx, ty + 32 * k, frame);
            // The issuing thread is the pacemaker of the CTA: every instruction it spends between two tcgen05.mma delays all
            // 29 other warps.  Levels are compile-time here (descriptor offsets are immediates), the ring slots of a step are
            // computed once, barrier parities come from the step's parity.
            const uint32_t tab_lo = tab_sa >> 4, x_lo = x_sa >> 4, ring_lo = ring_sa >> 4;
            constexpr uint32_t tv_hi = (uint32_t)(tv_d >> 32), img_hi = (uint32_t)(img_d >> 32), x_hi = (uint32_t)(x_d >> 32);
            constexpr uint32_t tv_l = (uint32_t)tv_d, img_l = (uint32_t)img_d, x_l = (uint32_t)x_d;
            const uint32_t ba_lo = (uint32_t)ba_d, ba_hi = (uint32_t)(ba_d >> 32), bb_lo = (uint32_t)bb_d, bb_hi = (uint32_t)(bb_d >> 32);
            auto mma2 = [&](uint32_t d, uint32_t alo, uint32_t ahi, uint32_t blo, uint32_t bhi, uint32_t idesc, uint32_t acc) {
                asm volatile("{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\tmov.b64 da, {%1, %2};\n\tmov.b64 db, {%3, %4};\n\tsetp.ne.b32 p, %6, 0;\n\t"
                             "tcgen05.mma.cta_group::1.kind::i8 [%0], da, db, %5, {%7, %7, %7, %7}, p;\n\t}"
                             ::"r"(d), "r"(alo), "r"(ahi), "r"(blo), "r"(bhi), "r"(idesc), "r"(acc), "r"(0u) : "memory");
            };
            int slot0 = 0;                                            // ring slot of the step's first chunk: (4 s) % RC
            for (int s = 0; s < n_steps; ++s) {
                const uint32_t sp = (uint32_t)s & 1u;
                if (s + 1 < n_steps) {
                    // the next step's four new chunks replace the four oldest of step s - 1: its last V pass must have completed
                    if (s > 0) mbar_wait(bar(3), sp ^ 1u);            // completion 3 (s - 1) + 2 of D_v-full[1]
                    const uint32_t b = bar((s + 1) & 1);
                    mbar_expect_tx(b, 4 * 4096);
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        int sl = slot0 + KV + j;
                        sl -= sl >= RC ? RC : 0; sl -= sl >= RC ? RC : 0;
                        tma_load_3d(ring_sa + 4096 * sl, &tmap, b, tx, ty + 32 * (KV + 4 * s + j), frame);
                    }
                }
                uint32_t rs[KV];                                      // descriptor address fields of the step's ring chunks
#pragma unroll
                for (int k = 0; k < KV; ++k) {
                    int sl = slot0 + k;
                    sl -= sl >= RC ? RC : 0;
                    rs[k] = img_l + ring_lo + 256u * (uint32_t)sl;
                }
                mbar_wait(bar(s & 1), (s >> 1) & 1);
                auto v_pass = [&](auto l_) {
                    constexpr int l = decltype(l_)::value;
                    constexpr uint32_t lp = (l >> 1) & 1;             // completion 3 s + (l >> 1) has parity sp ^ lp
                    if (s > 0 || l >= 2) DBG_WAIT(c_free, mbar_wait(bar(4 + (l & 1)), sp ^ lp ^ 1u));
                    tc_fence_after();
#pragma unroll
                    for (int k = 0; k < KV; ++k)
                        mma2(tm + 128 * (l & 1), tv_l + tab_lo + (uint32_t)((l * TVB + 512 * k) >> 4), tv_hi, rs[k], img_hi, iv, k > 0);
                    mma_commit(bar(2 + (l & 1)));
                };
                auto h_pass = [&](auto l_) {
                    constexpr int l = decltype(l_)::value;
                    constexpr uint32_t lp = (l >> 1) & 1;
                    constexpr int xb = XBUF == 2 ? (l & 1) : 0;
                    DBG_WAIT(c_xfull, mbar_wait(bar(6 + xb), XBUF == 2 ? (sp ^ lp) : (uint32_t)(l & 1)));
                    if (s > 0 || l >= 2) DBG_WAIT(c_free, mbar_wait(bar(10 + (l & 1)), sp ^ lp ^ 1u));
                    tc_fence_after();
                    const uint32_t d = tm + 256 + 128 * (l & 1);
#pragma unroll
                    for (int k = 0; k < KH; ++k) {
                        const uint32_t blo = tv_l + tab_lo + (uint32_t)((6 * TVB + l * THB + 512 * k) >> 4);
                        mma2(d, x_l + x_lo + (uint32_t)((xb * 2 * XPB + 4096 * k) >> 4), x_hi, blo, tv_hi, ih, k > 0);
                        mma2(d + 64, x_l + x_lo + (uint32_t)((xb * 2 * XPB + XPB + 4096 * k) >> 4), x_hi, blo, tv_hi, ih, k > 0);
                    }
                    mma2(d, ba_lo, ba_hi, bb_lo, bb_hi, ih, 1);      // + 32768
                    mma_commit(bar(8 + (l & 1)));
                };
                using std::integral_constant;
                // the V pass stays two levels ahead of the H pass
                v_pass(integral_constant<int, 0>()); v_pass(integral_constant<int, 1>()); h_pass(integral_constant<int, 0>());
                v_pass(integral_constant<int, 2>()); h_pass(integral_constant<int, 1>()); v_pass(integral_constant<int, 3>());
                h_pass(integral_constant<int, 2>()); v_pass(integral_constant<int, 4>()); h_pass(integral_constant<int, 3>());
                v_pass(integral_constant<int, 5>()); h_pass(integral_constant<int, 4>()); h_pass(integral_constant<int, 5>());
                slot0 += 4;
                slot0 -= slot0 >= RC ? RC : 0;
            }
        }
    } else if (warp == W_SVC) {
        // =============================== service warp: queued localisations + ordered emission of a finished step ===============================
        for (int s = 0; s < n_steps; ++s) {
            const int b = s & 1;
            const int y0 = ys + FR * s;
            const int si_lo = ((y0 == 0 ? 0 : y0 - 2) + 2) / 3;
            mbar_wait(bar(12 + b), (s >> 1) & 1);
            DBG_T(c_wait);
            // FeaturePointLocalization's full sequence for the samples whose three derivatives are all non-zero, one lane per
            // sample; their derivatives and edge tests were taken from the tiles by the detector threads
            // (the first QS entries; the detector threads took the rest themselves)
            const int nq_all = s_nq[0], nq = min(nq_all, QS);
            DBG_ONLY(if (lane == 0) s_dbg[warp][3] += nq_all;)
            for (int e = lane; e < nq; e += 32) {
                const uint2 qe = s_queue[e];
                if (!((qe.x >> 16) & 1)) continue;                                           // fails the edge test (:336)
                const int k = qe.x & 31, level = (qe.x >> 5) & 3, sr = (qe.x >> 8) & 255, c = (qe.x >> 17) & 255;
                const int dx = (int)(qe.y & 511) - 255, dy = (int)((qe.y >> 9) & 511) - 255, ds = (int)((qe.y >> 18) & 511) - 255;
                const float zhat = localize_f(s_div[dx + 255], s_div[dy + 255], s_div[ds + 255], s_div[c + 255]);
                if (zhat > a.contrast_thr) {                                                 // Diff_of_Gauss.cpp:245
                    const int task = sr * 3 + level - 1;
                    s_res[b][task * NS + k] = (short)(int)__fmul_rn(zhat, 255.0f);           // :246
                    atomicOr(&s_flag[b][task >> 5], 1u << (task & 31));
                }
            }
            __syncwarp();
            if (lane == 0) { s_nq[0] = 0; s_nq[1] = 0; mbar_arrive(bar(16)); }   // the queue may be refilled
            // ordered emission: flagged segments (level, sample row, strip) get consecutive staging ranges behind ONE atomic per step
            if (a.detect) {
                int total = 0;
                for (int wd = 0; wd < NFLAG; ++wd)
                    for (unsigned mk = s_flag[b][wd]; mk; mk &= mk - 1) {
                        const int task = 32 * wd + __ffs((int)mk) - 1;
                        const int res = lane < NS ? s_res[b][task * NS + lane] : -1;
                        total += __popc(__ballot_sync(0xffffffffu, res >= 0));
                    }
                int off = 0;
                if (lane == 0 && total) off = atomicAdd(a.cursor, total);
                off = __shfl_sync(0xffffffffu, off, 0);
                for (int wd = 0; wd < NFLAG; ++wd)
                    for (unsigned mk = s_flag[b][wd]; mk; mk &= mk - 1) {
                        const int task = 32 * wd + __ffs((int)mk) - 1;
                        const int res = lane < NS ? s_res[b][task * NS + lane] : -1;
                        const unsigned bal = __ballot_sync(0xffffffffu, res >= 0);
                        const int level = 1 + task % 3, si = si_lo + task / 3;
                        const int cnt = __popc(bal);
                        if (lane == 0) {
                            const int sg = frame * a.segs_per_frame + a.seg_base + ((level - 1) * a.ni + si) * a.n_strips + strip;
                            a.seg_off[sg] = off;
                            a.seg_cnt[sg] = cnt;
                        }
                        if (res >= 0) {
                            const int pos = off + __popc(bal & ((1u << lane) - 1u));
                            if (pos < a.staging_cap) {
                                vslam_keypoint kp;
                                kp.row = 3 * si + 1; kp.col = X0 + 3 * lane + 1; kp.value = res; kp.padding = 1; kp.octave = a.octave;
                                kp.level = level;
                                a.staging[pos] = kp;
                            }
                        }
                        off += cnt;
                    }
            }
            __syncwarp();
            if (lane < NFLAG) s_flag[b][lane] = 0;
            __syncwarp();
            if (lane == 0) mbar_arrive(bar(14 + b));
            DBG_T(c_work);
        }
    } else {
        // =============================== epilogue warps ===============================
        // Seven warps per TMEM lane quadrant (q = warp % 4), role r = warp / 4: r < 6 turns the 8 tile columns 8 r .. 8 r + 7 of
        // D_lo / D_hi into Gaussian / DoG tile bytes (r == 5 also columns 48..51), r == 6 turns D_v into the byte planes X.
        // Many light warps instead of a few heavy ones: every stage is a dependent chain (TMEM load -> wait -> pack -> store),
        // so the issue slots are filled by warp count, not by instruction-level parallelism.
        const int q = warp & 3, r = warp >> 2;
        const bool has_ev = r == 6, has_eh = r < 6;
        const int mp = 32 * q + lane;                                 // TMEM lane = reversed row of the step
        const int m = 127 - mp;
        const uint32_t lane_off = (uint32_t)(32 * q) << 16;
        constexpr int DT0 = N_G * TILE_B;                             // byte offset of the first D tile

        // ---- EV(s, l): D_v -> byte planes X (+ the rounding bias of the H pass) ----
        auto ev = [&](int s, int l) {
            const int n = 3 * s + (l >> 1);
            if (XBUF == 2) {
                if (n > 0) mbar_wait(bar(8 + (l & 1)), (n - 1) & 1);  // H(l - 2) has read this X buffer
            } else if (6 * s + l > 0) {
                const int lp = l == 0 ? 5 : l - 1, sp = l == 0 ? s - 1 : s;
                mbar_wait(bar(8 + (lp & 1)), (3 * sp + (lp >> 1)) & 1);   // H(l - 1) has read X
            }
            mbar_wait(bar(2 + (l & 1)), n & 1);
            DBG_T(c_wait);
            tc_fence_after();
            const uint32_t tb = tm + 128 * (l & 1) + lane_off;
            u8 *xl = xbuf + (XBUF == 2 ? (l & 1) : 0) * 2 * XPB + mp * 16;   // (mp >> 3) * 128 + (mp & 7) * 16
            // up to four 16-column pieces in flight (two 16-bit sums per register)
#pragma unroll
            for (int pb = 0; pb < NP; pb += 4) {
                uint32_t v[4][8];
#pragma unroll
                for (int p = 0; p < 4; ++p)
                    if (pb + p < NP) tmem_ld16_pack(tb + 16 * (pb + p), v[p]);
                tmem_ld_wait();
                if (pb + 4 >= NP) {                                   // last loads of the level: the D_v buffer is free again
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(bar(4 + (l & 1)));
                }
#pragma unroll
                for (int p = 0; p < 4; ++p)
                    if (pb + p < NP) {
                        *(uint4 *)(xl + (pb + p) * 2048) = make_uint4(__byte_perm(v[p][0], v[p][1], 0x6420), __byte_perm(v[p][2], v[p][3], 0x6420),
                                                                       __byte_perm(v[p][4], v[p][5], 0x6420), __byte_perm(v[p][6], v[p][7], 0x6420));
                        *(uint4 *)(xl + XPB + (pb + p) * 2048) = make_uint4(__byte_perm(v[p][0], v[p][1], 0x7531), __byte_perm(v[p][2], v[p][3], 0x7531),
                                                                             __byte_perm(v[p][4], v[p][5], 0x7531), __byte_perm(v[p][6], v[p][7], 0x7531));
                    }
            }
            fence_proxy_async();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar(6 + (XBUF == 2 ? (l & 1) : 0)));
            DBG_T(c_work);
        };

        // ---- EH(s, l): D_lo, D_hi -> Gaussian level l, DoG level l - 1 -> tiles ----
        // The tiles are rings of TROWS = 132 rows: row y of step s sits at ring row (128 s + (y - y0) + 4) mod 132, so the
        // last four rows of step s - 1 are the four rows in front of step s without a copy
        u8 *trow = tiles;
        int tswz = 0;
        uint32_t prev[6];                                             // previous level, 16-bit halves: 2 pixels per register
        auto eh = [&](int s, int l) {
            mbar_wait(bar(8 + (l & 1)), (3 * s + (l >> 1)) & 1);
            DBG_T(c_wait);
            tc_fence_after();
            const uint32_t dlo = tm + 256 + 128 * (l & 1) + lane_off;
            {   // accumulator columns n' = 56 - 8 r .. 63 - 8 r (column-reversed) = image columns n = 8 r .. 8 r + 7
                uint32_t lo[8], hi[8];
                tmem_ld8(dlo + 56 - 8 * r, lo);
                tmem_ld8(dlo + 64 + 56 - 8 * r, hi);
                tmem_ld_wait();
                uint32_t g2[4], d2[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const uint32_t va = (hi[7 - 2 * j] << 8) + lo[7 - 2 * j], vb = (hi[6 - 2 * j] << 8) + lo[6 - 2 * j];
                    g2[j] = __byte_perm(va, vb, 0x7632);              // byte 2 of each sum: (256 hi + lo) >> 16, bias included
                    d2[j] = __vmaxu2(g2[j], prev[j]) - prev[j];       // max(cur - prev, 0), GaussPyramid.cpp:197
                    prev[j] = g2[j];
                }
                const int ch = (((r >> 1) ^ tswz) << 4) + ((r & 1) << 3);
                if (l >= 1 && l <= 3) *(uint2 *)(trow + (l - 1) * TILE_B + ch) = make_uint2(__byte_perm(g2[0], g2[1], 0x6420), __byte_perm(g2[2], g2[3], 0x6420));
                if (l >= 1) *(uint2 *)(trow + DT0 + (l - 1) * TILE_B + ch) = make_uint2(__byte_perm(d2[0], d2[1], 0x6420), __byte_perm(d2[2], d2[3], 0x6420));
            }
            if (r == 5) {                                             // columns n = 48 .. 51 (49 .. 51 are never read)
                uint32_t lo[4], hi[4];
                tmem_ld4(dlo + 12, lo);
                tmem_ld4(dlo + 64 + 12, hi);
                tmem_ld_wait();
                uint32_t g2[2], d2[2];
#pragma unroll
                for (int j = 0; j < 2; ++j) {
                    const uint32_t va = (hi[3 - 2 * j] << 8) + lo[3 - 2 * j], vb = (hi[2 - 2 * j] << 8) + lo[2 - 2 * j];
                    g2[j] = __byte_perm(va, vb, 0x7632);
                    d2[j] = __vmaxu2(g2[j], prev[4 + j]) - prev[4 + j];
                    prev[4 + j] = g2[j];
                }
                const int ch = (3 ^ tswz) * 16;
                if (l >= 1 && l <= 3) *(uint32_t *)(trow + (l - 1) * TILE_B + ch) = __byte_perm(g2[0], g2[1], 0x6420);
                if (l >= 1) *(uint32_t *)(trow + DT0 + (l - 1) * TILE_B + ch) = __byte_perm(d2[0], d2[1], 0x6420);
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar(10 + (l & 1)));            // D_h buffer is free again
            DBG_T(c_work);
        };

        // ================= detector tests of a finished step (rows G, H, I), next octave's base: all 16 warps =================
        // Thread = one grid position (sample column k, sample row sr) for the three levels.  Result per sample in s_res (value
        // or -1); samples whose three derivatives are all non-zero need the full 3x3 inverse in double: they are queued with
        // their derivatives and edge test for the service warp, which also emits the step's survivors.
        auto det_phase = [&](int s) {
            const int y0 = ys + FR * s, b = s & 1;
            const int rb = (128 * s) % TROWS;                         // ring row of row y0 - 4
            auto rr = [&](int ty) { const int r = rb + ty; return r >= TROWS ? r - TROWS : r; };   // ring row of tile row ty = y - y0 + 4
            if (s >= 2) mbar_wait(bar(14 + b), ((s >> 1) - 1) & 1);   // the service warp is done with this result buffer
            if (s >= 1) mbar_wait(bar(16), (s - 1) & 1);              // ... and with the queue
            named_bar(2, NE);                                         // every tile warp's rows of the step are in place
            DBG_T(c_wait);
            const int yc_lo = y0 == 0 ? 0 : y0 - 2;
            const int yc_hi = (y0 + FR >= rows) ? rows - 2 : y0 + FR - 3;         // centre row yc needs rows up to yc + 2
            const int si_lo = (yc_lo + 2) / 3;
            if (a.detect) {
                int n_srow = yc_hi >= 0 ? yc_hi / 3 - si_lo + 1 : 0;
                if (n_srow < 0) n_srow = 0;
                const int k = tid & (NS - 1), xc = X0 + 3 * k, tx = 3 * k + 1;
                const int txm = xc > 0 ? tx - 1 : tx, txp = xc + 1 < cols ? tx + 1 : tx;
                // column parts of the tile offsets: 16-byte chunk (XOR-swizzled by the row pair) and byte inside it
                const int gm = txm >> 4, bm = txm & 15, g0 = tx >> 4, b0 = tx & 15, gp = txp >> 4, bp = txp & 15;
                const u8 *d = tiles + DT0;
                const int sr = tid >> 4;                              // 56 sample rows of threads, at most 44 live
                if (sr < n_srow) {
                    const int yc = 3 * (si_lo + sr);
                    if (xc <= cols - 2) {
                        // D tiles at (xc, yc); neighbours clamped at the image border (padOctave's replicate)
                        const int t0 = yc - y0 + 4;
                        const int ty = rr(t0), tym = yc > 0 ? rr(t0 - 1) : ty, typ = yc + 1 < rows ? rr(t0 + 1) : ty;
                        const int sw = (ty >> 1) & 3, swm = (tym >> 1) & 3, swp = (typ >> 1) & 3;
                        const int o00 = tym * 64 + ((gm ^ swm) << 4) + bm, o01 = tym * 64 + ((g0 ^ swm) << 4) + b0;
                        const int o10 = ty * 64 + ((gm ^ sw) << 4) + bm, o11 = ty * 64 + ((g0 ^ sw) << 4) + b0;
                        int w01[5], w10[5], w11[5], mn[5], mx[5];
#pragma unroll
                        for (int l = 0; l < 5; ++l) {
                            const int w00 = d[l * TILE_B + o00];
                            w01[l] = d[l * TILE_B + o01]; w10[l] = d[l * TILE_B + o10]; w11[l] = d[l * TILE_B + o11];
                            mn[l] = min(min(w00, w01[l]), min(w10[l], w11[l]));
                            mx[l] = max(max(w00, w01[l]), max(w10[l], w11[l]));
                        }
#pragma unroll
                        for (int level = 1; level <= 3; ++level) {
                            const int c = w11[level];
                            const int lo3 = min(min(mn[level - 1], mn[level]), mn[level + 1]);
                            const int hi3 = max(max(mx[level - 1], mx[level]), mx[level + 1]);
                            if (c == lo3 || c == hi3) {                                          // Diff_of_Gauss.cpp:287
                                const int dx = w10[level] - (int)d[level * TILE_B + ty * 64 + ((gp ^ sw) << 4) + bp];    // :226-228
                                const int dy = w01[level] - (int)d[level * TILE_B + typ * 64 + ((g0 ^ swp) << 4) + b0];
                                const int ds = w11[level - 1] - w11[level + 1];
                                const unsigned code = (unsigned)(k | level << 5 | sr << 8) | (unsigned)c << 17;
                                // all three derivatives non-zero: the full localisation (front of the queue); a zero derivative:
                                // cv::invert of the singular A*A^T gives zeros, dog_zhat = value/255 > 0.03 <=> c >= cmin (back of the queue)
                                if (dx != 0 && dy != 0 && ds != 0)
                                    s_queue[atomicAdd(&s_nq[0], 1)] = make_uint2(code, (unsigned)(dx + 255) | (unsigned)(dy + 255) << 9 | (unsigned)(ds + 255) << 18);
                                else if (c >= a.cmin)
                                    s_queue[NRES - 1 - atomicAdd(&s_nq[1], 1)].x = code;
                            }
                        }
                    }
#pragma unroll
                    for (int level = 1; level <= 3; ++level) s_res[b][(sr * 3 + level - 1) * NS + k] = (short)-1;
                }
            }
            DBG_T2(c_d1, c_det);
            // ---- next octave's base: level 3 at even rows and columns (GaussPyramid.cpp:123-126); thread = (row, four output columns) ----
            if (a.next_base && tid < (FR / 2) * (FWO / 8)) {
                const int g = tid % (FWO / 8), w = tid / (FWO / 8);
                const int oy = (y0 >> 1) + w, ox = X0 / 2 + 4 * g;
                if (oy < a.next_rows && ox < a.next_cols) {
                    const u8 *t3 = tiles + 2 * TILE_B;
                    const int tr = rr(2 * w + 4);
                    u8 bb[4];
#pragma unroll
                    for (int i = 0; i < 4; ++i) bb[i] = t3[toff(tr, 8 * g + 2 * i + 1)];
                    u8 *dst = a.next_base + (size_t)frame * a.next_stride + (size_t)oy * a.next_pitch + ox;
                    if (ox + 3 < a.next_cols) *(uint32_t *)dst = (uint32_t)bb[0] | (uint32_t)bb[1] << 8 | (uint32_t)bb[2] << 16 | (uint32_t)bb[3] << 24;
                    else for (int i = 0; i < 4 && ox + i < a.next_cols; ++i) dst[i] = bb[i];
                }
            }
            DBG_T2(c_d2, c_det);
            named_bar(3, NE);                                         // the queue is complete
            // ---- the edge test of the queued samples, one thread per sample (a dense pass: the survivors of the extrema test
            // are few, a per-sample branch would run the test in nearly every warp) ----
            const int nqb = s_nq[0], nqa = s_nq[1];
            for (int e = tid; e < nqb + nqa; e += NE) {
                const int qi = e < nqb ? e : NRES - 1 - (e - nqb);
                const unsigned code = s_queue[qi].x;
                const int k = code & 31, level = (code >> 5) & 3, sr = (code >> 8) & 255;
                const bool ok = edge_pass_tile(tiles + (level - 1) * TILE_B, X0 + 3 * k, 3 * (si_lo + sr), y0, rb, X0, rows, cols, a.edge_thr);
                if (e < nqb) {
                    if (ok) s_queue[qi].x = code | 1u << 16;
                } else if (ok) {
                    const int task = sr * 3 + level - 1;
                    s_res[b][task * NS + k] = (short)a.val_lut[(code >> 17) & 255];
                    atomicOr(&s_flag[b][task >> 5], 1u << (task & 31));
                }
            }
            DBG_T2(c_d3, c_det);
            named_bar(1, NE);                                         // the tiles may be overwritten
            // a long queue (textured frames): everything beyond the service warp's share is finished here, one thread per sample
            for (int e = QS + tid; e < nqb; e += NE) {
                const uint2 qe = s_queue[e];
                if (!((qe.x >> 16) & 1)) continue;                                           // fails the edge test (:336)
                const int k = qe.x & 31, level = (qe.x >> 5) & 3, sr = (qe.x >> 8) & 255, c = (qe.x >> 17) & 255;
                const int dx = (int)(qe.y & 511) - 255, dy = (int)((qe.y >> 9) & 511) - 255, ds = (int)((qe.y >> 18) & 511) - 255;
                const float zhat = localize_f(s_div[dx + 255], s_div[dy + 255], s_div[ds + 255], s_div[c + 255]);
                if (zhat > a.contrast_thr) {                                                 // Diff_of_Gauss.cpp:245
                    const int task = sr * 3 + level - 1;
                    s_res[b][task * NS + k] = (short)(int)__fmul_rn(zhat, 255.0f);           // :246
                    atomicOr(&s_flag[b][task >> 5], 1u << (task & 31));
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(bar(12 + b));                  // this warp's results are complete
            DBG_T2(c_d4, c_det);
        };

        // One loop over (step, level) for all sixteen warps (one copy of every stage in the instruction stream).  The X stage
        // runs two levels ahead of the tile stage, across the step boundary: levels 0 and 1 of step s + 1 come before the
        // detector of step s.
        const int n_lv = 6 * n_steps;
        if (has_ev) { ev(0, 0); ev(0, 1); }
        int s = 0, l = 0, s2 = 0, l2 = 2;                             // (s, l): tile stage; (s2, l2): X stage
#pragma unroll 1
        for (int gl = 0; gl < n_lv; ++gl) {
            if (has_eh) {
                if (l == 0) {
                    const int tr_own = (128 * s + m + 4) % TROWS;     // ring row of this thread's output row
                    trow = tiles + tr_own * 64;
                    tswz = (tr_own >> 1) & 3;
                }
                eh(s, l);
            }
            if (has_ev && gl + 2 < n_lv) ev(s2, l2);
            if (l == 5) det_phase(s);
            s += l == 5; l = l == 5 ? 0 : l + 1;
            s2 += l2 == 5; l2 = l2 == 5 ? 0 : l2 + 1;
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tm, kTmemCols);
#ifdef VSLAM_UMMA_DBG_BUILD
    if (a.dbg && blockIdx.x == 1 && blockIdx.z == 1 && tid < 256) a.dbg[tid] = (&s_dbg[0][0])[tid];
#endif
}

// ---------------------------------------------------------------------------------------------------------------
// host: launch
// ---------------------------------------------------------------------------------------------------------------
namespace {
typedef void (*umma_kern_t)(const CUtensorMap, const UmmaArgs);
typedef CUresult (*encode_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *, const cuuint32_t *,
                              const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

struct UmmaPlan { umma_kern_t k; size_t smem; UmmaGeom g; };

template <int KV, int KH, int NV, int XBUF> constexpr size_t umma_smem() {
    return 1024 + (size_t)(KV + 4) * 4096 + (size_t)XBUF * 2 * (2 * KH * 2048) + 6 * (size_t)((14 + 4 * KV) + (6 + 4 * KH)) * 128 + (size_t)(N_G + N_D) * TILE_B + 512 + 8 * 256;
}

bool umma_plan(vslam_ctx *ctx, double sigma0, int scale_samples, int octave, UmmaPlan *p) {
    // opt-in: bit-exact and covered by the GPU tests, but measured slower than the mma.sync kernel of fused_mma.cu on this
    // workload (31.9 against 25.6 ms per 1080p x 256 step for the four octave launches)
    static const int mode = getenv("VSLAM_UMMA") ? atoi(getenv("VSLAM_UMMA")) : 0;
    if (!mode) return false;
    if (umma_geometry(sigma0, scale_samples, octave, &p->g)) return false;
    if (p->g.KV == 5) { p->k = octave_umma_k<5, 3, 80, 2>; p->smem = umma_smem<5, 3, 80, 2>(); }
    else { p->k = octave_umma_k<6, 4, 112, 1>; p->smem = umma_smem<6, 4, 112, 1>(); }
    static std::mutex mu;
    static bool done[64];
    std::lock_guard<std::mutex> lock(mu);
    const int di = ctx->device >= 0 && ctx->device < 64 ? ctx->device : 0;
    if (!done[di]) {
        if (cudaFuncSetAttribute(octave_umma_k<5, 3, 80, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)umma_smem<5, 3, 80, 2>()) != cudaSuccess ||
            cudaFuncSetAttribute(octave_umma_k<6, 4, 112, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)umma_smem<6, 4, 112, 1>()) != cudaSuccess) {
            cudaGetLastError();
            return false;
        }
        done[di] = true;
    }
    return true;
}

// device image of an octave's compact Toeplitz tables: V tables [level][core matrix j][row i][byte kk] = tap[8 j + i + kk -
// 127 + r - P] (rows of the 128-row output block reversed), H tables likewise with 63 + (P - 1): the staged window starts P
// columns left of the strip and output column n is image column X0 - 1 + n
namespace {
struct UmmaTabEntry { int dev; double sigma0; int ss, octave; void *d; };
std::vector<UmmaTabEntry> g_umma_tabs;
std::mutex g_umma_tabs_mu;
}  // namespace

const uint4 *umma_tab_image(vslam_ctx *ctx, double sigma0, int scale_samples, int octave, const UmmaGeom &g) {
    typedef UmmaTabEntry Entry;
    std::vector<Entry> &cache = g_umma_tabs;
    std::lock_guard<std::mutex> lock(g_umma_tabs_mu);
    for (const Entry &e : cache)
        if (e.dev == ctx->device && e.sigma0 == sigma0 && e.ss == scale_samples && e.octave == octave) return (const uint4 *)e.d;
    std::vector<u8> img((size_t)g.tab_bytes, 0);
    for (int l = 0; l < 6; ++l) {
        for (int e = 0; e < g.JV * 128; ++e) {
            const int j = e >> 7, i = (e >> 4) & 7, kk = e & 15, idx = 8 * j + i + kk - 127 + g.r[l] - g.P;
            img[(size_t)l * g.JV * 128 + e] = (idx >= 0 && idx < g.n[l]) ? g.taps[l][idx] : (u8)0;
        }
        for (int e = 0; e < g.JH * 128; ++e) {
            const int j = e >> 7, i = (e >> 4) & 7, kk = e & 15, idx = 8 * j + i + kk - 63 - (g.P - 1) + g.r[l];
            img[(size_t)6 * g.JV * 128 + (size_t)l * g.JH * 128 + e] = (idx >= 0 && idx < g.n[l]) ? g.taps[l][idx] : (u8)0;
        }
    }
    void *d = nullptr;
    if (cudaMalloc(&d, img.size()) != cudaSuccess) return nullptr;
    if (cudaMemcpy(d, img.data(), img.size(), cudaMemcpyHostToDevice) != cudaSuccess) { cudaFree(d); return nullptr; }
    cache.push_back({ctx->device, sigma0, scale_samples, octave, d});
    return (const uint4 *)d;
}
}  // namespace

// called when the last context of `device` is destroyed (capi.cu)
void umma_tab_release(int device) {
    std::lock_guard<std::mutex> lock(g_umma_tabs_mu);
    for (size_t i = 0; i < g_umma_tabs.size();)
        if (g_umma_tabs[i].dev == device) { cudaFree(g_umma_tabs[i].d); g_umma_tabs.erase(g_umma_tabs.begin() + i); }
        else ++i;
}


bool octave_uses_umma(vslam_ctx *ctx, double sigma0, int scale_samples, int octave) {
    UmmaPlan p;
    return umma_plan(ctx, sigma0, scale_samples, octave, &p);
}

// reflect-101 margins the kernel wants around an octave base: rows above / below and columns left / right (multiples of 16)
int umma_base_pad(vslam_ctx *ctx, double sigma0, int scale_samples, int octave) {
    UmmaPlan p;
    return umma_plan(ctx, sigma0, scale_samples, octave, &p) ? p.g.P : 0;
}

int launch_octave_umma(vslam_ctx *ctx, const OctaveIO &io, int n_frames, double sigma0, int scale_samples) {
    if (io.gauss[0]) return 1;                                       // materialising launches stay on fused.cu
    UmmaPlan p;
    if (!umma_plan(ctx, sigma0, scale_samples, io.octave, &p)) return 1;
    if (io.base_pad < p.g.P || io.base_row_pad < p.g.P || (io.base_pad & 15)) return 1;
    if (n_frames <= 0 || n_frames > 65535) VS_FAIL(ctx, VSLAM_E_INVALID, "fused octave: %d frames per launch", n_frames);
    if ((io.base_pitch & 15) || ((uintptr_t)io.base & 15) || (io.base_stride & 15))
        VS_FAIL(ctx, VSLAM_E_INVALID, "fused octave: base plane must be 16-byte aligned (pitch %zu)", io.base_pitch);
    static encode_fn encode = nullptr;
    if (!encode) {
        cudaDriverEntryPointQueryResult qres;
        void *fn = nullptr;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) != cudaSuccess || !fn || qres != cudaDriverEntryPointSuccess) {
            cudaGetLastError();
            VS_FAIL(ctx, VSLAM_E_CUDA, "cuTensorMapEncodeTiled is not available from this driver");
        }
        encode = (encode_fn)fn;
    }
    UmmaArgs a;
    memset(&a, 0, sizeof a);
    static_cast<OctaveIO &>(a) = io;
    a.tab_image = umma_tab_image(ctx, sigma0, scale_samples, io.octave, p.g);
    if (!a.tab_image) VS_FAIL(ctx, VSLAM_E_NOMEM, "tap table image of octave %d", io.octave);
    a.n_strips = umma_strips(a.cols);
    a.n_segs = 1;
    a.seg_rows = (a.rows + FR - 1) / FR * FR;
    a.plane_x0 = io.base_pad; a.plane_y0 = io.base_row_pad;
    if (getenv("VSLAM_UMMA_NODETECT")) a.detect = 0;               // timing experiments only: the blur pipeline without the detector
    // the padded plane as a 3-D tensor (byte column, row, frame); boxes of 128 bytes x 32 rows, 128B-swizzled
    CUtensorMap tmap;
    u8 *origin = const_cast<u8 *>(io.base) - (size_t)io.base_row_pad * io.base_pitch - io.base_pad;
    const cuuint64_t gdim[3] = {(cuuint64_t)io.base_pitch, (cuuint64_t)(io.rows + 2 * io.base_row_pad), (cuuint64_t)n_frames};
    const cuuint64_t gstr[2] = {(cuuint64_t)io.base_pitch, (cuuint64_t)io.base_stride};
    const cuuint32_t box[3] = {128, 32, 1}, estr[3] = {1, 1, 1};
    const CUresult cr = encode(&tmap, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, origin, gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                               CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (cr != CUDA_SUCCESS) VS_FAIL(ctx, VSLAM_E_CUDA, "cuTensorMapEncodeTiled -> %d (octave %d)", (int)cr, io.octave);
    const dim3 grid(a.n_strips, a.n_segs, n_frames);
    static const bool dbg = getenv("VSLAM_UMMA_DBG") && atoi(getenv("VSLAM_UMMA_DBG"));
    static long long *d_dbg = nullptr;
    if (dbg && !d_dbg) cudaMalloc(&d_dbg, 256 * sizeof(long long));
    a.dbg = dbg ? d_dbg : nullptr;
    {
        KernelSpan span(ctx, VSLAM_K_OCTAVE_UMMA);
        p.k<<<grid, kThreads, p.smem, ctx->stream>>>(tmap, a);
    }
    VS_LAUNCH_CHECK(ctx);
    if (dbg) {
        long long h[256];
        cudaStreamSynchronize(ctx->stream);
        cudaMemcpy(h, d_dbg, sizeof h, cudaMemcpyDeviceToHost);
        fprintf(stderr, "umma dbg octave %d (CTA 1,0,1), cycles per warp: wait work det | tests decim edge tail\n", io.octave);
        for (int w = 0; w < 30; w += (w < 24 ? 4 : 1))
            fprintf(stderr, "  warp %2d (role %d): %8lld %8lld %8lld | %8lld %8lld %8lld %8lld\n", w, w >> 2, h[8 * w], h[8 * w + 1], h[8 * w + 2], h[8 * w + 3],
                    h[8 * w + 4], h[8 * w + 5], h[8 * w + 6]);
    }
    return VSLAM_OK;
}

}  // namespace vslam
