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
//   X        warps read D_v from TMEM (thread = row), add the rounding bias of the second pass (128 per 8.8 sum: the taps
//            sum to 256), split the 16-bit sums into two byte planes and store them K-major for the H pass.
//   H pass   D_lo / D_hi[128 rows x 64 out cols] = X_lo / X_hi[128 x 32 KH] * T_h[64 x 32 KH]^T.
//   epilogue out = (256 D_hi + D_lo) >> 16, saturating DoG against the previous level (registers), byte tiles in shared
//            memory for the detector.
//   T_v, T_h banded Toeplitz matrices are Hankel once their rows are reversed: 8 x 16-byte core matrix j of a level holds
//            tap[8 j + row + byte - shift], and the no-swizzle K-major descriptor (SBO 128, LBO 256) walks OVERLAPPING core
//            matrices -- 4.4 KB per level instead of 20 KB.  Reversed rows only permute TMEM lanes / columns.
//
// One CTA per SM: a control warp (TMA producer + the single MMA-issuing thread) and eight epilogue warps (TMEM lane
// quadrant = warp % 4, column half = warp / 4); D_v and D_h are double buffered in TMEM (512 columns), so the tensor pipe
// works on level l + 1 while the warps drain level l.  The detector, localisation, edge test, ordered emission and
// decimation are those of fused_mma.cu on a row-major byte tile.
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
constexpr int kThreads = 288;                   // eight epilogue warps + the control warp
constexpr int NE = 256;                         // epilogue threads
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
};

namespace {

__device__ __forceinline__ void named_bar(int id, int n) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory"); }

__device__ __forceinline__ void tmem_ld4(uint32_t taddr, uint32_t (&v)[4]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]) : "r"(taddr));
}

// byte offset of pixel (tile row tr, column n) inside a tile
__device__ __forceinline__ int toff(int tr, int n) { return tr * 64 + ((((n >> 4) ^ (tr >> 1)) & 3) << 4) + (n & 15); }

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
    extern __shared__ __align__(1024) u8 smem_raw[];
    u8 *smem = (u8 *)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    u8 *ring = smem;
    u8 *xbuf = ring + RC * 4096;
    u8 *tab = xbuf + XBUF * 2 * XPB;
    u8 *tiles = tab + 6 * (TVB + THB);
    __shared__ __align__(8) unsigned long long s_bar[12];
    __shared__ uint32_t s_tmem;
    __shared__ int s_res[3 * MAX_SROW * NS];                  // detector results of one step
    __shared__ unsigned short s_queue[3 * MAX_SROW * NS];     // samples that need the full localisation
    __shared__ unsigned int s_flag[(3 * MAX_SROW + 31) / 32]; // segments (sample row, level) of the step with survivors
    __shared__ int s_nq;
    __shared__ float s_div[511];                              // float(i - 255) / 255f, the reference's divisions

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int strip = blockIdx.x, frame = blockIdx.z;
    const int rows = a.rows, cols = a.cols;
    const int X0 = strip * FWO;
    const int ys = blockIdx.y * a.seg_rows, ye = min(ys + a.seg_rows, rows);
    const int n_steps = (ye - ys + FR - 1) / FR;

    // barriers: [0,1] ring full, [2,3] D_v full, [4,5] D_v free, [6,7] X full, [8,9] D_h full, [10,11] D_h free
    uint32_t bar[12];
#pragma unroll
    for (int i = 0; i < 12; ++i) bar[i] = smem_u32(&s_bar[i]);
    if (tid == 0) {
        mbar_init(bar[0], 1); mbar_init(bar[1], 1);
        mbar_init(bar[2], 1); mbar_init(bar[3], 1);
        mbar_init(bar[4], 8); mbar_init(bar[5], 8);
        mbar_init(bar[6], 8); mbar_init(bar[7], 8);
        mbar_init(bar[8], 1); mbar_init(bar[9], 1);
        mbar_init(bar[10], 8); mbar_init(bar[11], 8);
        fence_barrier_init();
    }
    if (warp == 0) tmem_alloc(smem_u32(&s_tmem), kTmemCols);
    for (int i = tid; i < 6 * (TVB + THB) / 16; i += kThreads) ((uint4 *)tab)[i] = a.tab_image[i];
    for (int i = tid; i < 511; i += kThreads) s_div[i] = __fdiv_rn((float)(i - 255), 255.0f);
    if (tid < (3 * MAX_SROW + 31) / 32) s_flag[tid] = 0;
    if (tid == 0) s_nq = 0;
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tm = s_tmem;
    // TMEM columns: D_v buffers at 0 / 128, D_h buffers at 256 / 384 (low-byte plane, +64 high-byte plane)

    if (warp == 8) {
        // =============================== control: TMA producer + MMA issuer (one thread) ===============================
        if (lane == 0) {
            const int tx = X0 - P + a.plane_x0;                       // first staged column inside the padded plane: multiple of 16
            const int ty = ys - P + a.plane_y0;
            const uint32_t ring_sa = smem_u32(ring), x_sa = smem_u32(xbuf), tab_sa = smem_u32(tab);
            constexpr uint32_t iv = idesc_u8(NV, 0, 1), ih = idesc_u8(64, 0, 0);
            // descriptor halves that never change
            constexpr uint64_t tv_d = smem_desc(0, 256, 128, kLayoutNone), img_d = smem_desc(0, 4096, 1024, kLayoutSw128),
                               x_d = smem_desc(0, 2048, 128, kLayoutNone);
            mbar_expect_tx(bar[0], KV * 4096);
            for (int k = 0; k < KV; ++k) tma_load_3d(ring_sa + 4096 * k, &tmap, bar[0], tx, ty + 32 * k, frame);
            for (int s = 0; s < n_steps; ++s) {
                if (s + 1 < n_steps) {
                    // the next step's four new chunks replace the four oldest of step s - 1: its last V pass must have completed
                    if (s > 0) mbar_wait(bar[3], (3 * (s - 1) + 2) & 1);
                    const uint32_t b = bar[(s + 1) & 1];
                    mbar_expect_tx(b, 4 * 4096);
                    for (int j = 0; j < 4; ++j) {
                        const int c = KV + 4 * s + j;
                        tma_load_3d(ring_sa + 4096 * (c % RC), &tmap, b, tx, ty + 32 * c, frame);
                    }
                }
                mbar_wait(bar[s & 1], (s >> 1) & 1);
                auto v_pass = [&](int l) {
                    const int n = 3 * s + (l >> 1);                   // use count of this D_v buffer
                    if (n > 0) mbar_wait(bar[4 + (l & 1)], (n - 1) & 1);
                    tc_fence_after();
#pragma unroll
                    for (int k = 0; k < KV; ++k) {
                        const uint32_t slot = (uint32_t)((4 * s + k) % RC);
                        mma_i8_ss(tm + 128 * (l & 1), tv_d + ((tab_sa + l * TVB + 512 * k) >> 4), img_d + ((ring_sa + 4096 * slot) >> 4), iv, k > 0);
                    }
                    mma_commit(bar[2 + (l & 1)]);
                };
                auto h_pass = [&](int l) {
                    const int xb = XBUF == 2 ? (l & 1) : 0;
                    const int nx = XBUF == 2 ? 3 * s + (l >> 1) : 6 * s + l;
                    mbar_wait(bar[6 + xb], nx & 1);
                    const int n = 3 * s + (l >> 1);
                    if (n > 0) mbar_wait(bar[10 + (l & 1)], (n - 1) & 1);
                    tc_fence_after();
                    const uint32_t xl = x_sa + xb * 2 * XPB, th = tab_sa + 6 * TVB + l * THB, d = tm + 256 + 128 * (l & 1);
#pragma unroll
                    for (int k = 0; k < KH; ++k) {
                        const uint64_t bd = tv_d + ((th + 512 * k) >> 4);
                        mma_i8_ss(d, x_d + ((xl + 4096 * k) >> 4), bd, ih, k > 0);
                        mma_i8_ss(d + 64, x_d + ((xl + XPB + 4096 * k) >> 4), bd, ih, k > 0);
                    }
                    mma_commit(bar[8 + (l & 1)]);
                };
                v_pass(0); v_pass(1); h_pass(0); v_pass(2); h_pass(1); v_pass(3); h_pass(2); v_pass(4); h_pass(3); v_pass(5); h_pass(4); h_pass(5);
            }
        }
    } else {
        // =============================== epilogue warps ===============================
        const int q = warp & 3, h = warp >> 2;
        const int mp = 32 * q + lane;                                 // TMEM lane = reversed row of the step
        const int m = 127 - mp;
        const uint32_t lane_off = (uint32_t)(32 * q) << 16;
        const int xoff = mp * 16;                                     // (mp >> 3) * 128 + (mp & 7) * 16
        const int tr_own = m + 4;                                     // tile row of this thread's output row
        const int tswz = (tr_own >> 1) & 3;
        constexpr int DT0 = N_G * TILE_B;                             // byte offset of the first D tile
        uint32_t prev[18];                                            // previous level, 16-bit halves: 2 pixels per register

        // ---- computeEdgeResponse + threshold on the G tile of `level` around sample (xc, yc); y0 = first row of the step ----
        auto edge_pass = [&](int level, int xc, int yc, int y0) -> bool {
            const u8 *gt = tiles + (level - 1) * TILE_B;
            int ix2 = 0, iy2 = 0, ixy = 0;
#pragma unroll
            for (int u = 0; u <= 1; ++u)
#pragma unroll
                for (int v = 0; v <= 1; ++v) {
                    // reflect-101 of an index at most one outside [0, len): a single bounce (len == 1 -> 0)
                    const int xx = xc + v, yy = yc + u;
                    const int xr = (xx + 1 < cols ? xx + 1 : max(cols - 2, 0)), xl = (xx > 0 ? xx - 1 : min(1, cols - 1));
                    const int yd = (yy + 1 < rows ? yy + 1 : max(rows - 2, 0)), yu = (yy > 0 ? yy - 1 : min(1, rows - 1));
                    const int ty = yy - y0 + 4, tx = xx - X0 + 1;
                    const int gxv = (int)gt[toff(ty, xr - X0 + 1)] - (int)gt[toff(ty, xl - X0 + 1)];
                    const int gyv = (int)gt[toff(yd - y0 + 4, tx)] - (int)gt[toff(yu - y0 + 4, tx)];
                    ix2 += gxv * gxv; iy2 += gyv * gyv; ixy += gxv * gyv;
                }
            return edge_from_sums((float)ix2, (float)iy2, (float)ixy) < a.edge_thr;   // Diff_of_Gauss.cpp:336
        };

        for (int s = 0; s < n_steps; ++s) {
            const int y0 = ys + FR * s;
            // ---- EV(l): D_v -> byte planes X (+ the rounding bias of the H pass) ----
            auto ev = [&](int l) {
                if (XBUF == 1 && l > 0) mbar_wait(bar[8 + ((l - 1) & 1)], (3 * s + ((l - 1) >> 1)) & 1);   // H(l-1) has read X
                mbar_wait(bar[2 + (l & 1)], (3 * s + (l >> 1)) & 1);
                tc_fence_after();
                const uint32_t tb = tm + 128 * (l & 1) + lane_off;
                u8 *xl = xbuf + (XBUF == 2 ? (l & 1) : 0) * 2 * XPB + xoff;
                constexpr int NPH = PMID;                             // pieces of the larger half
                uint32_t v[NPH][16];
                const int p0 = h ? PMID : 0, np = h ? NP - PMID : PMID;
#pragma unroll
                for (int p = 0; p < NPH; ++p)
                    if (p < np) tmem_ld16(tb + 16 * (p0 + p), v[p]);
                tmem_ld_wait();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(bar[4 + (l & 1)]);         // D_v buffer is free again
#pragma unroll
                for (int p = 0; p < NPH; ++p)
                    if (p < np) {
                        uint32_t lo[4], hi[4];
#pragma unroll
                        for (int w = 0; w < 4; ++w) {
                            const uint32_t t0 = __byte_perm(v[p][4 * w], v[p][4 * w + 1], 0x5410) + 0x00800080u;
                            const uint32_t t1 = __byte_perm(v[p][4 * w + 2], v[p][4 * w + 3], 0x5410) + 0x00800080u;
                            lo[w] = __byte_perm(t0, t1, 0x6420);
                            hi[w] = __byte_perm(t0, t1, 0x7531);
                        }
                        *(uint4 *)(xl + (p0 + p) * 2048) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
                        *(uint4 *)(xl + XPB + (p0 + p) * 2048) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
                    }
                fence_proxy_async();
                __syncwarp();
                if (lane == 0) mbar_arrive(bar[6 + (XBUF == 2 ? (l & 1) : 0)]);
            };
            // ---- EH(l): D_lo, D_hi -> Gaussian level l, DoG level l - 1 -> tiles ----
            // 16 columns n' = c0 .. c0 + 15 of the (column-reversed) accumulators = image columns n = 48 - c0 .. 63 - c0;
            // pi = index of the piece's first register pair in prev[]
            auto eh_piece16 = [&](int l, uint32_t dlo, int c0, int pi) {
                uint32_t lo[16], hi[16];
                tmem_ld16(dlo + c0, lo);
                tmem_ld16(dlo + 64 + c0, hi);
                tmem_ld_wait();
                uint32_t g2[8], d2[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const uint32_t va = (hi[15 - 2 * j] << 8) + lo[15 - 2 * j], vb = (hi[14 - 2 * j] << 8) + lo[14 - 2 * j];
                    g2[j] = __byte_perm(va, vb, 0x7632);              // byte 2 of each sum: (256 hi + lo) >> 16, bias included
                    d2[j] = __vmaxu2(g2[j], prev[pi + j]) - prev[pi + j];   // max(cur - prev, 0), GaussPyramid.cpp:197
                    prev[pi + j] = g2[j];
                }
                const int ch = ((48 - c0) >> 4) ^ tswz;
                if (l >= 1 && l <= 3)
                    *(uint4 *)(tiles + (l - 1) * TILE_B + tr_own * 64 + ch * 16) =
                        make_uint4(__byte_perm(g2[0], g2[1], 0x6420), __byte_perm(g2[2], g2[3], 0x6420), __byte_perm(g2[4], g2[5], 0x6420), __byte_perm(g2[6], g2[7], 0x6420));
                if (l >= 1)
                    *(uint4 *)(tiles + DT0 + (l - 1) * TILE_B + tr_own * 64 + ch * 16) =
                        make_uint4(__byte_perm(d2[0], d2[1], 0x6420), __byte_perm(d2[2], d2[3], 0x6420), __byte_perm(d2[4], d2[5], 0x6420), __byte_perm(d2[6], d2[7], 0x6420));
            };
            auto eh = [&](int l) {
                mbar_wait(bar[8 + (l & 1)], (3 * s + (l >> 1)) & 1);
                tc_fence_after();
                const uint32_t dlo = tm + 256 + 128 * (l & 1) + lane_off;
                if (h) {                                              // columns n = 0 .. 31
                    eh_piece16(l, dlo, 48, 0);
                    eh_piece16(l, dlo, 32, 8);
                } else {                                              // columns n = 32 .. 51 (49 .. 51 are never read)
                    eh_piece16(l, dlo, 16, 0);
                    uint32_t lo[4], hi[4];
                    tmem_ld4(dlo + 12, lo);
                    tmem_ld4(dlo + 64 + 12, hi);
                    tmem_ld_wait();
                    uint32_t g2[2], d2[2];
#pragma unroll
                    for (int j = 0; j < 2; ++j) {
                        const uint32_t va = (hi[3 - 2 * j] << 8) + lo[3 - 2 * j], vb = (hi[2 - 2 * j] << 8) + lo[2 - 2 * j];
                        g2[j] = __byte_perm(va, vb, 0x7632);
                        d2[j] = __vmaxu2(g2[j], prev[8 + j]) - prev[8 + j];
                        prev[8 + j] = g2[j];
                    }
                    const int ch = 3 ^ tswz;
                    if (l >= 1 && l <= 3) *(uint32_t *)(tiles + (l - 1) * TILE_B + tr_own * 64 + ch * 16) = __byte_perm(g2[0], g2[1], 0x6420);
                    if (l >= 1) *(uint32_t *)(tiles + DT0 + (l - 1) * TILE_B + tr_own * 64 + ch * 16) = __byte_perm(d2[0], d2[1], 0x6420);
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(bar[10 + (l & 1)]);        // D_h buffer is free again
            };
            ev(0); ev(1); eh(0); ev(2); eh(1); ev(3); eh(2); ev(4); eh(3); ev(5); eh(4); eh(5);

            // ================= detector of the step (rows G, H, I, J), next octave's base, tile history =================
            named_bar(1, NE);                                         // every warp's tile rows are in place
            int si_lo = 0, n_srow = 0;
            if (a.detect) {
                const int yc_lo = y0 == 0 ? 0 : y0 - 2;
                const int yc_hi = (y0 + FR >= rows) ? rows - 2 : y0 + FR - 3;     // centre row yc needs rows up to yc + 2
                si_lo = (yc_lo + 2) / 3;
                n_srow = yc_hi >= 0 ? yc_hi / 3 - si_lo + 1 : 0;
                if (n_srow < 0) n_srow = 0;
                // ---- 4a. thread = one grid position (sample column k, sample row sr) for the three levels ----
                for (int e = tid; e < n_srow * NS; e += NE) {
                    const int k = e % NS, sr = e / NS;
                    const int yc = 3 * (si_lo + sr), xc = X0 + 3 * k;
                    int res[3] = {-1, -1, -1};
                    if (xc <= cols - 2) {
                        // D tiles at (xc, yc); neighbours clamped at the image border (padOctave's replicate)
                        const int ty = yc - y0 + 4, tx = 3 * k + 1;
                        const int tym = yc > 0 ? ty - 1 : ty, typ = yc + 1 < rows ? ty + 1 : ty;
                        const int txm = xc > 0 ? tx - 1 : tx, txp = xc + 1 < cols ? tx + 1 : tx;
                        const int o00 = toff(tym, txm), o01 = toff(tym, tx), o10 = toff(ty, txm), o11 = toff(ty, tx);
                        const u8 *d = tiles + DT0;
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
                                const int dx = w10[level] - (int)d[level * TILE_B + toff(ty, txp)];   // :226-228
                                const int dy = w01[level] - (int)d[level * TILE_B + toff(typ, tx)];
                                const int ds = w11[level - 1] - w11[level + 1];
                                if (dx == 0 || dy == 0 || ds == 0) {
                                    // cv::invert of the singular A*A^T gives zeros: dog_zhat = value/255 (a.cmin, a.val_lut: the
                                    // same float ops, tabulated on the host)
                                    if (c >= a.cmin && edge_pass(level, xc, yc, y0)) {
                                        res[level - 1] = a.val_lut[c];
                                        atomicOr(&s_flag[(sr * 3 + level - 1) >> 5], 1u << ((sr * 3 + level - 1) & 31));
                                    }
                                } else {
                                    s_queue[atomicAdd(&s_nq, 1)] = (unsigned short)(k | level << 5 | sr << 8);
                                }
                            }
                        }
                    }
#pragma unroll
                    for (int level = 1; level <= 3; ++level) s_res[(sr * 3 + level - 1) * NS + k] = res[level - 1];
                }
            }
            named_bar(1, NE);
            // ---- queued samples: FeaturePointLocalization's full sequence, one thread per sample ----
            {
                const int nq = s_nq;
                for (int e = tid; e < nq; e += NE) {
                    const int code = s_queue[e];                                             // k | level << 5 | sample row << 8
                    const int k = code & 31, level = (code >> 5) & 3, sr = code >> 8;
                    const int yc = 3 * (si_lo + sr), xc = X0 + 3 * k;
                    const u8 *d = tiles + DT0 + level * TILE_B;
                    const int ty = yc - y0 + 4, tx = 3 * k + 1;
                    const int tym = yc > 0 ? ty - 1 : ty, typ = yc + 1 < rows ? ty + 1 : ty;
                    const int txm = xc > 0 ? tx - 1 : tx, txp = xc + 1 < cols ? tx + 1 : tx;
                    const int c = d[toff(ty, tx)];
                    const int dx = (int)d[toff(ty, txm)] - (int)d[toff(ty, txp)], dy = (int)d[toff(tym, tx)] - (int)d[toff(typ, tx)];
                    const int ds = (int)d[toff(ty, tx) - TILE_B] - (int)d[toff(ty, tx) + TILE_B];
                    const float zhat = localize_f(s_div[dx + 255], s_div[dy + 255], s_div[ds + 255], s_div[c + 255]);
                    if (zhat > a.contrast_thr && edge_pass(level, xc, yc, y0)) {             // :245
                        const int task = sr * 3 + level - 1;
                        s_res[task * NS + k] = (int)__fmul_rn(zhat, 255.0f);                 // :246
                        atomicOr(&s_flag[task >> 5], 1u << (task & 31));
                    }
                }
            }
            named_bar(1, NE);
            // ---- ordered emission (warp 0): flagged segments (level, sample row, strip) get consecutive staging ranges
            // behind ONE atomic per step ----
            if (warp == 0 && a.detect) {
                int total = 0;
                for (int wd = 0; wd < (3 * MAX_SROW + 31) / 32; ++wd)
                    for (unsigned mk = s_flag[wd]; mk; mk &= mk - 1) {
                        const int task = 32 * wd + __ffs((int)mk) - 1;
                        const int res = lane < NS ? s_res[task * NS + lane] : -1;
                        total += __popc(__ballot_sync(0xffffffffu, res >= 0));
                    }
                int off = 0;
                if (lane == 0 && total) off = atomicAdd(a.cursor, total);
                off = __shfl_sync(0xffffffffu, off, 0);
                for (int wd = 0; wd < (3 * MAX_SROW + 31) / 32; ++wd)
                    for (unsigned mk = s_flag[wd]; mk; mk &= mk - 1) {
                        const int task = 32 * wd + __ffs((int)mk) - 1;
                        const int res = lane < NS ? s_res[task * NS + lane] : -1;
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
            // ---- next octave's base: level 3 at even rows and columns (GaussPyramid.cpp:123-126); thread = (row, four output columns) ----
            if (a.next_base) {
                for (int e = tid; e < (FR / 2) * (FWO / 8); e += NE) {
                    const int g = e % (FWO / 8), w = e / (FWO / 8);
                    const int oy = (y0 >> 1) + w, ox = X0 / 2 + 4 * g;
                    if (oy < a.next_rows && ox < a.next_cols && y0 + 2 * w < ye) {
                        const u8 *t3 = tiles + 2 * TILE_B;
                        const int tr = 2 * w + 4;
                        u8 b[4];
#pragma unroll
                        for (int i = 0; i < 4; ++i) b[i] = t3[toff(tr, 8 * g + 2 * i + 1)];
                        u8 *dst = a.next_base + (size_t)frame * a.next_stride + (size_t)oy * a.next_pitch + ox;
                        if (ox + 3 < a.next_cols) *(uint32_t *)dst = (uint32_t)b[0] | (uint32_t)b[1] << 8 | (uint32_t)b[2] << 16 | (uint32_t)b[3] << 24;
                        else for (int i = 0; i < 4 && ox + i < a.next_cols; ++i) dst[i] = b[i];
                    }
                }
            }
            named_bar(1, NE);                                         // emission and decimation have read s_res / the tiles
            // ---- rows y0 + 124 .. y0 + 127 become the history of the next step (same swizzle: 128 + i and i share (tr >> 1) & 3) ----
            if (tid < (N_G + N_D) * 16) {
                const int t = tid >> 4, i = tid & 15;                 // 4 rows x 4 chunks
                uint4 *tp = (uint4 *)(tiles + t * TILE_B);
                tp[i] = tp[FR * 4 + i];
            }
            if (tid < (3 * MAX_SROW + 31) / 32) s_flag[tid] = 0;
            if (tid == 0) s_nq = 0;
            named_bar(1, NE);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tm, kTmemCols);
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
    return 1024 + (size_t)(KV + 4) * 4096 + (size_t)XBUF * 2 * (2 * KH * 2048) + 6 * (size_t)((14 + 4 * KV) + (6 + 4 * KH)) * 128 + (size_t)(N_G + N_D) * TILE_B;
}

bool umma_plan(vslam_ctx *ctx, double sigma0, int scale_samples, int octave, UmmaPlan *p) {
    static const int mode = getenv("VSLAM_UMMA") ? atoi(getenv("VSLAM_UMMA")) : 1;
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
const uint4 *umma_tab_image(vslam_ctx *ctx, double sigma0, int scale_samples, int octave, const UmmaGeom &g) {
    struct Entry { int dev; double sigma0; int ss, octave; void *d; };
    static std::vector<Entry> cache;
    static std::mutex mu;
    std::lock_guard<std::mutex> lock(mu);
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
    KernelSpan span(ctx, VSLAM_K_FUSED);
    p.k<<<grid, kThreads, p.smem, ctx->stream>>>(tmap, a);
    VS_LAUNCH_CHECK(ctx);
    return VSLAM_OK;
}

}  // namespace vslam
