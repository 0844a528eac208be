// fused_mma.cu -- the fused octave kernel with both blur passes on the integer tensor pipe (IMMA.16832.U8.U8).
//
// Same rows of SURVEY.md 8(a) as fused.cu (B, C, D, G, H, I, J) and the same marching scheme; what changes is how the
// 1738 MAC per input pixel of the Gaussian stack are issued.  OpenCV's 8U GaussianBlur is exact integer arithmetic
// (u8 pixels x 8-bit taps -> 8.8 sums, 8.8 sums x 8-bit taps -> 16.16 sums), so each pass is a product with a banded
// Toeplitz matrix of taps that mma.sync.m16n8k32 (u8 x u8 -> s32) evaluates exactly -- bit-identical to the dp4a/dp2a
// kernel and to OpenCV, at 569 TMAC/s instead of 74 / 37 TMAC/s (profiles/imma_peak_r1.json vs alu_peak_r1.json).
//
//   H pass   D[16 rows x 8 cols] = Img[16 rows x 32 staged cols] * T[32 x 8],  T[k][n] = tap[col_k - (col_n - r)]
//            The image fragment of a 32-byte chunk is two LDS.64 per row pair and is shared by every 8-column block
//            whose band touches the chunk.  Two M blocks give a thread four consecutive rows of two columns, which it
//            stores as one 16-byte word {lo bytes, hi bytes} x 2 columns of the row quad (the "h plane").
//   V pass   D[16 cols x 8 rows] = H^T[16 cols x 32 rows] * T'[32 x 8], once for the low and once for the high bytes
//            of the 8.8 sums (out = (256 * hi + lo + 32768) >> 16).  The h plane is a ring of row quads per level, so
//            the A fragment is two LDS.128 per K step; the B fragment (taps) is one pair of words per level and K step.
//            Epilogue in registers: rounding, saturating DoG against the previous level, 2x2 packed tile words.
//   The detector, localisation, edge test, ordered emission and decimation are those of fused.cu on the new tile layout.
//
// Tensor cores on a "stencil": the north star expected these kernels to be memory bound; the survey showed the Gaussian
// stack is ALU bound by a factor of ~50.  The MMA form removes that bound without touching exactness.
#include <math.h>
#include <string.h>

#include <algorithm>
#include <mutex>
#include <vector>
#include <type_traits>

#include "detect_math.cuh"
#include "fused_mma.cuh"

namespace vslam {

namespace {
constexpr int FR = kFusedR;
constexpr int kMmaThreads = 288;            // eight compute warps and a service warp (queued localisations, emission)
constexpr int TRP = 20;                     // words per column pair of a tile: 2 history + 16 row pairs (+2: conflict-free STS);
                                            // every block of 8 column pairs is skewed by one more word (tcol): 20 m mod 32 repeats with
                                            // period 8, and the detector / decimation read column pairs 8, 16, 24 apart in one warp
constexpr int N_G = 3, N_D = 5;             // G tiles: Gaussian levels 1..3; D tiles: DoG levels 0..4
constexpr int MAX_NS = 20;                  // sample columns per strip at FW = 64
constexpr int MAX_SROW = FR / 3 + 2;
__host__ __device__ constexpr int tile_words(int fw) { return (fw / 2) * TRP + 4; }
}  // namespace

// ---------------------------------------------------------------------------------------------------------------
// host: tables
// ---------------------------------------------------------------------------------------------------------------
int build_mma_taps(double sigma0, int scale_samples, int octave, int fw, MmaTaps *out) {
    memset(out, 0, sizeof *out);
    out->fw = fw;
    const int PPW = mma_ppw(fw);
    int32_t t[kMaxKsize];
    const double k = pow(2.0, 1.0 / (double)scale_samples);          // GaussPyramid.hpp:69
    int rmin = 1 << 30, rmax = 0, poff = 0;
    for (int l = 0; l < 6; ++l) {
        const double sig = pow(2, octave) * sigma0 * pow(k, l);      // GaussPyramid.cpp:157
        const int n = host_gauss_ksize(sig);
        if (n > 247) return VSLAM_E_INVALID;
        int rc = host_gauss_taps(sig, n, t);
        if (rc) return rc;
        MmaLevel &lv = out->lv[l];
        lv.n = n; lv.r = n / 2;
        for (int i = 0; i < n; ++i) {
            if (t[i] < 0 || t[i] > 255) return VSLAM_E_INVALID;
            out->taps[l][i] = (u8)t[i];
        }
        if (lv.r < rmin) rmin = lv.r;
        if (lv.r > rmax) rmax = lv.r;
    }
    // rows between a step's first output row and its H chunk: rmax, or up to 3 more when that saves V-pass K steps
    auto v_ksteps = [&](int v) {
        int s = 0;
        for (int l = 0; l < 6; ++l) s += (out->lv[l].n + 7 + (4 - (v + out->lv[l].r) % 4) % 4 + 31) / 32;
        return s;
    };
    int rv = rmax;
    for (int v = rmax + 1; v < rmax + 4; ++v) if (v_ksteps(v) < v_ksteps(rv)) rv = v;
    for (int l = 0; l < 6; ++l) {
        MmaLevel &lv = out->lv[l];
        const int n = lv.n;
        // every level's H chunk of a step is the same 32 image rows y0 + rv ..., so the A fragments are shared by the
        // levels; a level's V window for output rows y0 + 8j ... starts rv + r + sv rows above the chunk's row 8j
        lv.sv = (4 - (rv + lv.r) % 4) % 4;
        lv.ksv = (n + 7 + lv.sv + 31) / 32;
        lv.cq = (rv + lv.r + lv.sv) / 4;
        lv.np = (lv.cq + 8 + 2) / 2;                                 // aligned quads Q = q + cq of a step: 8i .. 8i + 7 + cq
        if (lv.np < 8) lv.np = 8;                                    // single-step wrap of every slot computation (needs >= 6)
        lv.plane_off = poff;
        poff += 2 * lv.np * PPW;
    }
    // tap copies (fused_mma.cuh): the left pad covers the H pass's chunks in front of the band, padl + r == 0 (mod 4)
    // rounded up to an instantiated chunk count: every chunk is multiplied, so the zero padding must cover all of them
    // a block's chunks start on the 16-byte group its widest band starts in (not on a 32-byte boundary): the band begins at most 15
    // bytes into the first chunk, ceil((15 + 8 + 2 rmax) / 32) chunks cover it -- 2 / 3 / 5 / 9 for the default octaves
    out->nch = (2 * rmax + 23 + 31) / 32;
    {
        const int have64[] = {2, 3, 5}, have32[] = {5, 9};
        int pick = 0;
        for (int c : have64) if (fw == 64 && !pick && c >= out->nch) pick = c;
        for (int c : have32) if (fw == 32 && !pick && c >= out->nch) pick = c;
        if (!pick) return VSLAM_E_INVALID;                           // bands wider than the widest instantiation
        out->nch = pick;
    }
    int cp = 0;
    for (int l = 0; l < 6; ++l) {
        MmaLevel &lv = out->lv[l];
        lv.padl = rmax - lv.r + 41 + (4 - (rmax + 41) % 4) % 4;
        const int need_h = lv.padl + lv.r - rmax + 12 + 32 * (out->nch - 1) + 20;
        const int ksvs = (lv.n + 7 + lv.sv + 8 * (fw / 32 - 1) + 31) / 32;     // static kernels: the row blocks of a warp share a K step's A fragments
        const int need_v = lv.padl + 24 + 32 * (ksvs - 1) + 8;
        const int need_t = lv.padl + lv.n + 3;
        cp = std::max(cp, std::max(need_h, std::max(need_v, need_t)));
    }
    cp = (cp + 4 + 3) & ~3;                                          // 4 spare bytes: copies s > rv % 4 start one word later (below)
    // Copies 32 bytes apart modulo 128.  A V-pass fragment load touches 8 consecutive words of every copy: words -1 .. 6 of the
    // copies s <= rv % 4, words -2 .. 5 of the others; moving those one word up tiles the 32 banks exactly (round 2: the load ran
    // at two wavefronts per request, 22 M excess wavefronts per 32 frames at octaves 0 and 1).  The H pass touches 5 words per copy.
    while (cp % 128 != 32) cp += 4;
    out->cp = cp;
    for (int l = 0; l < 6; ++l) {
        MmaLevel &lv = out->lv[l];
        lv.tab_hoff = 4 * cp * l + lv.padl + lv.r;
        lv.tab_voff = 4 * cp * l + lv.padl - lv.sv - rv % 4;
    }
    out->rmin = rmin; out->rmax = rmax; out->rv = rv;
    out->stage_rows = FR;
    out->stage_pitch = (2 * rmax + 15 + fw + 127) / 128 * 128;       // left halo rmax (+ up to 15 alignment), 64 columns, right halo
    {   // row = (q * magic) >> 16 for q < stage_rows * (pitch / 16): checked exhaustively
        const int spq = out->stage_pitch >> 4, nq = out->stage_rows * spq;
        out->spq_magic = ((1 << 16) + spq - 1) / spq;
        for (int q = 0; q < nq; ++q)
            if ((int)(((unsigned)q * (unsigned)out->spq_magic) >> 16) != q / spq) return VSLAM_E_INVALID;
    }
    out->plane_words = poff;
    out->tab_bytes = (24 * cp + 4 + 15) & ~15;
    return VSLAM_OK;
}

size_t mma_smem_bytes(const MmaTaps &t) {
    return (size_t)t.plane_words * 4 + (size_t)t.stage_rows * t.stage_pitch + (size_t)(N_G + N_D) * tile_words(t.fw) * 4 + (size_t)t.tab_bytes;
}

int mma_strips(int cols, int fw) { return (cols + mma_fwo(fw) - 1) / mma_fwo(fw); }

// ---------------------------------------------------------------------------------------------------------------
// device
// ---------------------------------------------------------------------------------------------------------------
struct MmaArgs : OctaveIO {
    MmaTaps mt;
    const uint4 *tab_image;         // mt.tab_bytes of shifted tap copies, built on the host once per (device, pyramid, octave, width)
};

namespace {

__device__ __forceinline__ void imma(int (&d)[4], u32 a0, u32 a1, u32 a2, u32 a3, u32 b0, u32 b1) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.u8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+r"(d[0]), "+r"(d[1]), "+r"(d[2]), "+r"(d[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

// same with a constant accumulator input (the compiler materialises it once; RZ for 0)
__device__ __forceinline__ void imma_c(int (&d)[4], u32 a0, u32 a1, u32 a2, u32 a3, u32 b0, u32 b1, int c) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.u8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%10,%10,%10,%10};"
                 : "=r"(d[0]), "=r"(d[1]), "=r"(d[2]), "=r"(d[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1), "r"(c));
}

// split barrier (arrive early, wait late) on an mbarrier in shared memory
__device__ __forceinline__ void mbar_init(u32 sa, u32 count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(sa), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(u32 sa) {
    asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.shared::cta.b64 st, [%0];\n\t}" ::"r"(sa) : "memory");
}
__device__ __forceinline__ void mbar_wait(u32 sa, u32 parity) {
    asm volatile("{\n\t.reg .pred p;\n\tMB_WAIT:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra MB_DONE;\n\t"
                 "bra MB_WAIT;\n\tMB_DONE:\n\t}" ::"r"(sa), "r"(parity) : "memory");
}

// byte offset of pixel (x, ty) inside a tile: words of 2 x 2 pixels [x even/odd][row even/odd], column pairs TRP words apart
__device__ __forceinline__ int tcol(int x) { return (x >> 1) * (TRP * 4) + ((x >> 4) << 2) + ((x & 1) << 1); }
__device__ __forceinline__ int trow(int ty) { return ((ty >> 1) << 2) + (ty & 1); }

}  // namespace

// NCH = 32-byte chunks of a staged row that the bands of one 8-column block can touch, starting on the 16-byte group the
//       widest band starts in: ceil((15 + 8 + 2 rmax) / 32)
// GEO = the level geometry: MmaStaticGeom<n0..n5> (immediates) or MmaDynGeom (read from the argument)
// FW  = columns a CTA computes: 64, or 32 with the warps regrouped (H pass: 4 column blocks x 2 groups of three levels;
//       V pass: 2 column blocks x 4 row blocks)
// MAT = the launch materialises the octave (GaussPyramid's getters): every Gaussian level and every DoG level leaves the V pass's
//       registers for its plane in HBM as well (rows in pairs, two columns per store)
template <int NCH, class GEO, int FW, bool MAT = false>
__global__ void __launch_bounds__(kMmaThreads, NCH <= 3 ? 2 : 1) octave_mma_k(const __grid_constant__ MmaArgs a) {
    constexpr int NT = 256;
    extern __shared__ __align__(16) u32 smem[];
    constexpr int NW = NT / 32;
    constexpr int FWO = mma_fwo(FW), PPW = mma_ppw(FW), TILE_W = tile_words(FW), NS = FWO / 3;
    constexpr int HCB = FW / 8, LG = NW / HCB, LPW = 6 / LG;    // H pass: 8-column blocks, level groups, levels per warp
    constexpr int VCB = FW / 16;                // V pass: 16-column blocks
    constexpr int RBW = 4 * VCB / NW;           // 8-row blocks of one 16-column block a warp owns in the V pass
    const MmaTaps &T = a.mt;
    const GEO geo(a.mt);
    u8 *planes = (u8 *)smem;
    u8 *stage = (u8 *)(smem + geo.plane_words());
    const int SP = geo.stage_pitch();
    u32 *tiles32 = (u32 *)(stage + FR * SP);
    const u8 *tiles = (const u8 *)tiles32;
    u8 *tab = (u8 *)(tiles32 + (N_G + N_D) * TILE_W);
    __shared__ int s_res[3 * MAX_SROW * MAX_NS];             // detector results of one step
    __shared__ unsigned long long s_mask[2];                  // step parity -> segments (level, sample row) of the step with survivors
    __shared__ unsigned short s_queue[3 * MAX_SROW * MAX_NS]; // samples of the step that need the full localisation
    __shared__ int s_nq;
    __shared__ __align__(8) unsigned long long s_mbar;        // split barrier between the detector and the queue
    __shared__ float s_div[511];                              // float(i - 255) / 255f, the reference's divisions

    // special registers are read once through volatile asm: under the 128-register cap the compiler otherwise re-reads
    // them (S2R, long latency) inside the marching loop
    int tid, strip, seg, frame;
    asm volatile("mov.u32 %0, %%tid.x;" : "=r"(tid));
    asm volatile("mov.u32 %0, %%ctaid.x;" : "=r"(strip));
    asm volatile("mov.u32 %0, %%ctaid.y;" : "=r"(seg));
    asm volatile("mov.u32 %0, %%ctaid.z;" : "=r"(frame));
    const int lane = tid & 31, warp = tid >> 5;
    const bool svc = tid >= NT;                  // service warp: no blur passes, no detector
    const int g = lane >> 2, t = lane & 3;
    const int rows = a.rows, cols = a.cols;
    const int X0 = strip * FWO, X0c = X0 - 2, xs = X0c - geo.rmax();
    const int xs16 = xs & ~15;                   // staged rows start on a 16-byte boundary of the plane
    const int cboff = X0c - xs16;                // staged column of the strip's first computed column
    const u8 *__restrict__ base = a.base + (size_t)frame * a.base_stride;
    const int ys = seg * a.seg_rows;
    const int ye = min(ys + a.seg_rows, rows);
    const int yv = seg > 0 ? ys - FR : 0;        // first step with a V pass (one extra step feeds the detector's history)
    const int K = (geo.rmax() + geo.rv() + FR - 1) / FR;    // warm-up steps: H pass only

    // ---- stage base rows y0 + rv ..., columns xs16 ..., reflect-101 resolved (16-byte cp.async inside the image) ----
    const int SPQ = SP >> 4, n_stage = FR * SPQ;
    auto issue_stage = [&](int y0n) {
        for (int q = svc ? n_stage : tid; q < n_stage; q += NT) {
            const int row = (int)(((u32)q * (u32)geo.spq_magic()) >> 16), j = q - row * SPQ;
            const int y = refl101(y0n + geo.rv() + row, rows);
            const u8 *__restrict__ src = base + (size_t)y * a.base_pitch;
            u32 *dst = (u32 *)(stage + row * SP) + 4 * (j ^ ((row >> 2) & 7));   // 16-byte groups XOR-swizzled by the row quad
            const int x = xs16 + 16 * j;
            if (x >= -a.base_pad && x + 15 < cols + a.base_pad) {            // inside the image or its reflected margins
                const u32 sa = (u32)__cvta_generic_to_shared(dst);
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sa), "l"(src + x) : "memory");
            } else {
#pragma unroll
                for (int w = 0; w < 4; ++w) {
                    const int xw = x + 4 * w;
                    dst[w] = (u32)src[refl101(xw, cols)] | (u32)src[refl101(xw + 1, cols)] << 8 |
                             (u32)src[refl101(xw + 2, cols)] << 16 | (u32)src[refl101(xw + 3, cols)] << 24;
                }
            }
        }
    };
    issue_stage(yv - K * FR);
    // ---- tables: four byte-shifted, zero-padded copies of every level's taps (host-built image); the division table ----
    for (int i = tid; i < (T.tab_bytes >> 4); i += NT) ((uint4 *)tab)[i] = a.tab_image[i];
    for (int i = tid; i < 511; i += NT) s_div[i] = __fdiv_rn((float)(i - 255), 255.0f);
    if (tid < 2) s_mask[tid] = 0;
    if (tid == 2) s_nq = 0;
    const u32 mb_sa = (u32)__cvta_generic_to_shared(&s_mbar);
    if (tid == 3) mbar_init(mb_sa, NT);

    constexpr int DT0 = N_G * TILE_W * 4;                                     // byte offset of the first D tile
    constexpr int LS = TILE_W * 4;                                            // bytes between tiles
    // computeEdgeResponse + threshold on the G tile of `level` around sample (xc, yc); yb = first row of the tiles' step
    auto edge_pass = [&](int level, int xc, int yc, int yb) -> bool {
        const u8 *gt = tiles + (level - 1) * LS;
        int ix2 = 0, iy2 = 0, ixy = 0;
#pragma unroll
        for (int u = 0; u <= 1; ++u)
#pragma unroll
            for (int v = 0; v <= 1; ++v) {
                // reflect-101 of an index at most one outside [0, len): a single bounce (len == 1 -> 0)
                const int xx = xc + v, yy = yc + u;
                const int xr = (xx + 1 < cols ? xx + 1 : max(cols - 2, 0)), xl = (xx > 0 ? xx - 1 : min(1, cols - 1));
                const int yd = (yy + 1 < rows ? yy + 1 : max(rows - 2, 0)), yu = (yy > 0 ? yy - 1 : min(1, rows - 1));
                const int ty = trow(yy - yb + 4), tx = tcol(xx - X0c);
                const int gxv = (int)gt[tcol(xr - X0c) + ty] - (int)gt[tcol(xl - X0c) + ty];
                const int gyv = (int)gt[tx + trow(yd - yb + 4)] - (int)gt[tx + trow(yu - yb + 4)];
                ix2 += gxv * gxv; iy2 += gyv * gyv; ixy += gxv * gyv;
            }
        return edge_from_sums((float)ix2, (float)iy2, (float)ixy) < a.edge_thr;   // Diff_of_Gauss.cpp:336
    };
    // samples queued by the detector (all three derivatives non-zero) of the step whose tiles start at row yb:
    // FeaturePointLocalization's full sequence, one lane per queued sample; survivors flag their segment in s_mask[mp]
    // The service warp takes the first 32 samples while the compute warps are still in their H pass; the compute warps
    // take what is left after it.
    auto run_queue = [&](int yb, int si_lo, int nq, int mp) {
        for (int e = svc ? lane : 32 + tid; e < nq; e += NT + 32) {
            const int code = s_queue[e];                                             // k | level << 5 | sample row << 8
            const int k = code & 31, level = (code >> 5) & 3, sr = code >> 8;
            const int yc = 3 * (si_lo + sr), xc = X0 + 3 * k;
            const u8 *d = tiles + DT0 + level * LS;
            const int x = xc - X0c, ty = yc - yb + 4;
            const int cx = tcol(x), cy = trow(ty);
            const int cxm = xc > 0 ? tcol(x - 1) : cx, cym = yc > 0 ? trow(ty - 1) : cy;
            const int cxp = xc + 1 < cols ? tcol(x + 1) : cx, cyp = yc + 1 < rows ? trow(ty + 1) : cy;
            const int c = d[cx + cy];
            const int dx = (int)d[cxm + cy] - (int)d[cxp + cy], dy = (int)d[cx + cym] - (int)d[cx + cyp];
            const int ds = (int)d[cx + cy - LS] - (int)d[cx + cy + LS];
            const float zhat = localize_f(s_div[dx + 255], s_div[dy + 255], s_div[ds + 255], s_div[c + 255]);
            if (zhat > a.contrast_thr && edge_pass(level, xc, yc, yb)) {             // :245
                const int task = sr * 3 + level - 1;
                s_res[task * NS + k] = (int)__fmul_rn(zhat, 255.0f);                 // :246
                atomicOr((unsigned *)&s_mask[mp] + (task >> 5), 1u << (task & 31));      // 32-bit halves: a native shared atomic, not a CAS loop
            }
        }
    };
    // ordered emission (service warp): the flagged segments (level, sample row, strip) of a finished step get consecutive
    // staging ranges behind ONE atomic
    auto emit = [&](int si_lo, unsigned long long mask) {
        int total = 0;
        for (unsigned long long m = mask; m; m &= m - 1) {
            const int task = __ffsll((long long)m) - 1;
            const int res = lane < NS ? s_res[task * NS + lane] : -1;
            total += __popc(__ballot_sync(0xffffffffu, res >= 0));
        }
        int off = 0;
        if (lane == 0) off = atomicAdd(a.cursor, total);
        off = __shfl_sync(0xffffffffu, off, 0);
        for (unsigned long long m = mask; m; m &= m - 1) {
            const int task = __ffsll((long long)m) - 1;
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
    };

    // ---- lane constants of the two passes (nothing below depends on the step) ----
    // H pass, A fragment through ldmatrix.x4: lane L supplies the address of row (L & 7) of matrix (L >> 3); matrices 0/1
    // are the M rows 0-7 / 8-15 of K bytes 0-15, matrices 2/3 the same rows of K bytes 16-31.  M row m <-> image row
    // 4 pi(m & 7) + (m >> 3) (+2 for the second M block), so that a thread ends up with four consecutive rows of a quad.
    const int h_row = 4 * (((lane & 7) >> 1) + 4 * (lane & 1)) + ((lane >> 3) & 1);   // staged row this lane addresses, M block 0
    const u32 stage_sa = (u32)__cvta_generic_to_shared(stage);
    const int hw = warp % HCB, hlg = warp / HCB;                              // H pass: warp = (8 output columns, group of levels)
    const int hcb = cboff + 8 * hw;                                           // staged column of the warp's first output column
    const int c16 = (hcb - geo.rmax()) >> 4;                                      // 16-byte group of a staged row the widest band starts in: chunk 0 starts there
    const bool h_short = hcb - 16 * c16 + 7 + geo.rmax() < 32 * (NCH - 1);          // (warp-uniform) the widest band ends before the last chunk
    u32 saA[NCH], saB[NCH];                                                   // ldmatrix addresses of the chunks (XOR swizzle resolved)
#pragma unroll
    for (int ci = 0; ci < NCH; ++ci) {
        const int ch = c16 + 2 * ci + (lane >> 4);
        saA[ci] = stage_sa + h_row * SP + ((ch ^ ((h_row >> 2) & 7)) << 4);
        saB[ci] = stage_sa + (h_row + 2) * SP + ((ch ^ (((h_row + 2) >> 2) & 7)) << 4);
    }
    // B fragments: b0 = taps idx0 .. idx0 + 3, b1 = taps idx0 + 16 ..; H pass idx0 = r + 16 c16 + 32 ci - hcb + 4 t - g,
    // V pass (K rows remapped so that a thread's 8 rows are consecutive) b1 = taps idx0 + 4 .., idx0 = 32 k + 8 t - g - sv.
    // One lane pointer per pass (byte-shifted copy + word offset); the level is an offset (fused_mma.cuh)
    const int hu = 4 * t - g - hcb + 16 * c16, vu = 8 * t - g + geo.rv() % 4;
    const u8 *hbase = tab + (hu & 3) * geo.cp() + ((hu & 3) > geo.rv() % 4 ? 4 : 0) + (hu - (hu & 3));   // two's complement: fine for hu < 0
    const u8 *vbase = tab + (vu & 3) * geo.cp() + ((vu & 3) > geo.rv() % 4 ? 4 : 0) + (vu - (vu & 3));
    const int d_pi = (g >> 1) + 4 * (g & 1);                                  // quad of the H chunk this thread's D rows belong to
    const int vcb = warp % VCB, vjb = (warp / VCB) * RBW;                         // V pass: warp = (16-column block, RBW blocks of 8 rows)
    u32 hoff[6], voff[6];                                                     // ring byte offsets: H store slot, V window start (row block vjb)
#pragma unroll
    for (int l = 0; l < 6; ++l) {
        // the lane's column part (< one pair row) rides in the ring offsets: the wrap test does not see it
        hoff[l] = (u32)((d_pi + geo.cq(l)) >> 1) * (PPW * 4) + 16 * (4 * hw + t) + 8 * ((d_pi + geo.cq(l)) & 1);   // < np pairs
        voff[l] = (u32)(vjb + t) * (PPW * 4) + 16 * (8 * vcb + g);
    }
    const int wi0 = (8 * vcb + g) * TRP + vcb + 2 + 4 * vjb + t;                    // tile word of the thread's first V output
    // materialising launches: the thread's V outputs are columns mat_col, mat_col + 1 (rows 8 (vjb + rb) + 2t, + 1 of the step)
    // (cols is even: the launcher sends odd widths to fused.cu); per step: 32-bit plane offsets of the thread's row pairs and
    // which of their rows are stored -- the map pointers themselves are CTA-uniform
    const int mat_col = X0c + 16 * vcb + 2 * g;
    const bool mat_own = MAT && !svc && 16 * vcb + 2 * g >= 2 && 16 * vcb + 2 * g < 2 + FWO && mat_col < cols;
    const size_t mat_fo = (size_t)blockIdx.z * a.map_stride;                  // (blockIdx, not `frame`: the compiler keeps the plane pointers uniform)
    const u32 mat_pitch = (u32)a.map_pitch;
    u32 mat_off[RBW];
    unsigned mat_ok = 0;                                                      // bit 2 rb: row pair rb's first row is stored, bit 2 rb + 1: its second
    auto mat_put = [&](u8 *plane, int rb, u32 q) {                            // q = {first row: two columns, second row: two columns}
        u8 *p = plane + mat_off[rb];
        if (mat_ok & (1u << (2 * rb))) *(u16 *)p = (u16)q;
        if (mat_ok & (2u << (2 * rb))) *(u16 *)(p + mat_pitch) = (u16)(q >> 16);
    };
    // detector: thread = (sample column det_k, sample row det_sr of the step)
    const int det_k = tid % NS, det_sr = tid / NS;                            // 256 threads: rows 0 .. 12 (at most 12 are live)
    const int det_xc = X0 + 3 * det_k;
    const bool det_colok = det_xc <= cols - 2;
    const int det_cx = tcol(det_xc - X0c), det_cxm = det_xc > 0 ? tcol(det_xc - X0c - 1) : det_cx;
    const int det_cxp = det_xc + 1 < cols ? tcol(det_xc - X0c + 1) : det_cx;
    // ---- 4. detector of a finished step (its tiles start at row y0), decimation of its level 3 ----
    auto detect_step = [&](int y0, bool own_step, int par) -> int {
        // ---- 4a. detector, test.  Thread = one grid position (sample column, sample row) for the three levels: the five
        // DoG levels' 2 x 2 windows are read once, their minima / maxima shared by the levels.  Result per sample in s_res:
        // the keypoint's value or -1.  Samples whose three derivatives are all non-zero need the full 3x3 inverse: they are
        // queued (s_queue) and finished, one lane per sample, by run_queue behind the split barrier.  Segments with
        // survivors are flagged in s_mask: emit (next phase) touches nothing else.
        int si_lo = 0, n_srow = 0;
        if (own_step && a.detect) {
            const int yc_lo = y0 == 0 ? 0 : y0 - 2;
            const int yc_hi = (y0 + FR >= rows) ? rows - 2 : y0 + FR - 3;   // centre row yc needs rows up to yc + 2
            si_lo = (yc_lo + 2) / 3;
            n_srow = yc_hi >= 0 ? yc_hi / 3 - si_lo + 1 : 0;
            if (n_srow < 0) n_srow = 0;
            if (det_sr < n_srow) {
                const int yc = 3 * (si_lo + det_sr);
                int res[3] = {-1, -1, -1};
                if (det_colok) {
                    // D tiles at (xc, yc); neighbours clamped at the image border (padOctave's replicate)
                    const int ty = yc - y0 + 4;
                    const int cy = trow(ty), cym = yc > 0 ? trow(ty - 1) : cy, cyp = yc + 1 < rows ? trow(ty + 1) : cy;
                    const u8 *d = tiles + DT0;
                    int w00[5], w01[5], w10[5], w11[5], mn[5], mx[5];
#pragma unroll
                    for (int l = 0; l < 5; ++l) {
                        w00[l] = d[l * LS + det_cxm + cym]; w01[l] = d[l * LS + det_cx + cym];
                        w10[l] = d[l * LS + det_cxm + cy]; w11[l] = d[l * LS + det_cx + cy];
                        mn[l] = min(min(w00[l], w01[l]), min(w10[l], w11[l]));
                        mx[l] = max(max(w00[l], w01[l]), max(w10[l], w11[l]));
                    }
#pragma unroll
                    for (int level = 1; level <= 3; ++level) {
                        const int c = w11[level];
                        const int lo3 = min(min(mn[level - 1], mn[level]), mn[level + 1]);
                        const int hi3 = max(max(mx[level - 1], mx[level]), mx[level + 1]);
                        if (c == lo3 || c == hi3) {                                          // Diff_of_Gauss.cpp:287
                            const int dx = w10[level] - (int)d[level * LS + det_cxp + cy];   // :226-228
                            const int dy = w01[level] - (int)d[level * LS + det_cx + cyp];
                            const int ds = w11[level - 1] - w11[level + 1];
                            if (dx == 0 || dy == 0 || ds == 0) {
                                // cv::invert of the singular A*A^T gives zeros: dog_zhat = value/255 (tables built on the
                                // host with the same float ops: a.cmin, a.val_lut)
                                if (c >= a.cmin && edge_pass(level, det_xc, yc, y0)) {
                                    res[level - 1] = a.val_lut[c];
                                    atomicOr((unsigned *)&s_mask[par] + ((det_sr * 3 + level - 1) >> 5), 1u << ((det_sr * 3 + level - 1) & 31));
                                }
                            } else {
                                s_queue[atomicAdd(&s_nq, 1)] = (unsigned short)(det_k | level << 5 | det_sr << 8);
                            }
                        }
                    }
                }
#pragma unroll
                for (int level = 1; level <= 3; ++level) s_res[(det_sr * 3 + level - 1) * NS + det_k] = res[level - 1];
            }
        }
        // ---- 4b. next octave's base: level 3 at even rows and columns (GaussPyramid.cpp:123-126) = byte 0 of the words;
        // thread = (row, pair of output columns): the strip's 30 columns start on an even column of the next base ----
        if (own_step && a.next_base && tid < (FWO / 4) * (FR / 2)) {
            const int oc = 2 * (tid % (FWO / 4)), w = tid / (FWO / 4);
            const int ox = X0 / 2 + oc, oy = (y0 >> 1) + w;
            if (oy < a.next_rows && ox < a.next_cols) {
                const u32 v0 = tiles32[2 * TILE_W + (oc + 1) * TRP + ((oc + 1) >> 3) + 2 + w], v1 = tiles32[2 * TILE_W + (oc + 2) * TRP + ((oc + 2) >> 3) + 2 + w];
                u8 *dst = a.next_base + (size_t)frame * a.next_stride + (size_t)oy * a.next_pitch + ox;
                if (ox + 1 < a.next_cols) *(u16 *)dst = (u16)((v0 & 0xffu) | (v1 & 0xffu) << 8);
                else *dst = (u8)v0;
            }
        }
        return si_lo;
    };
    // Per step two barriers and a split one.  The detector of the PREVIOUS step shares a phase with this step's H pass
    // (they touch different buffers); its queue is finished behind an mbarrier every thread arrived at before its H
    // pass, so nobody waits there:
    //   X (after Bx)  detector tests + decimation of step c-1 | arrive | H pass of step c | wait | queued localisations of c-1
    //   Y (after By)  prefetch of step c+1's base rows, V pass of step c -> tiles, ordered emission of step c-1
    // A ninth (service) warp skips the passes: it waits for the detector, runs the first 32 queued localisations while
    // the compute warps are in their H pass, and emits the survivors (one global atomic per step) during the V pass.
    int p_y0 = 0;                                                             // the previous step (tiles hold it)
    bool p_own = false;
    int par = 0;                                                              // parity of the step the detector works on
    u32 mb_phase = 0;
    int si_lo = 0;
    for (int y0 = yv - K * FR; y0 < ye; y0 += FR) {
        const bool do_v = y0 >= yv;
        asm volatile("cp.async.wait_all;" ::: "memory");
        __syncthreads();                                                      // Bx: staged rows landed; V pass / emission of c-1 done
        si_lo = detect_step(p_y0, p_own, par);
        if (!svc) mbar_arrive(mb_sa);                                         // the service warp goes straight to the wait
        // ---- 2. H pass: image rows y0 + rv .. + 31 for all six levels -> 8 quads of every level's ring ----
        // the image fragments of the warp's chunks are loaded once and serve every level whose band touches them.
        // One basic block, software pipelined by hand: every chunk is multiplied (chunks outside a level's band meet the
        // zero padding of its tap copies: no band test, no branch), level l + 1's taps are loaded before level l's MMAs
        // and level l - 1's stores come after them (the compiler cannot hoist a shared load above a shared store it
        // cannot disambiguate).
        auto h_pass = [&](auto nc_) {
            constexpr int NC = decltype(nc_)::value;           // chunks this warp multiplies: NCH, or NCH - 1 when no band reaches the last
            u32 fa[NC][4], fb[NC][4];
#pragma unroll
            for (int ci = 0; ci < NC; ++ci) {
                asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                             : "=r"(fa[ci][0]), "=r"(fa[ci][1]), "=r"(fa[ci][2]), "=r"(fa[ci][3]) : "r"(saA[ci]));
                asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                             : "=r"(fb[ci][0]), "=r"(fb[ci][1]), "=r"(fb[ci][2]), "=r"(fb[ci][3]) : "r"(saB[ci]));
            }
            u32 tb[2][NC][2];
            int acc[2][2][4];
            auto load_taps = [&](int l, int buf) {
#pragma unroll
                for (int ci = 0; ci < NC; ++ci) {
                    tb[buf][ci][0] = *(const u32 *)(hbase + geo.tab_hoff(l) + 32 * ci);
                    tb[buf][ci][1] = *(const u32 *)(hbase + geo.tab_hoff(l) + 32 * ci + 16);
                }
            };
            // thread: rows 4 d_pi .. + 3 of columns 2t, 2t + 1 -> low bytes {c0, c1} and high bytes {c0, c1} of one quad,
            // in the pair-aligned planes (aligned quad Q = q + cq, pair Q >> 1, position Q & 1)
            auto h_store = [&](int l, const int (&ac)[2][4]) {
                const u32 ring = (u32)geo.np(l) * (PPW * 4);
                u8 *plo = planes + 4 * geo.plane_off(l) + hoff[l];
                const u32 e01 = __byte_perm(ac[0][0], ac[0][2], 0x5410), e23 = __byte_perm(ac[1][0], ac[1][2], 0x5410);
                const u32 o01 = __byte_perm(ac[0][1], ac[0][3], 0x5410), o23 = __byte_perm(ac[1][1], ac[1][3], 0x5410);
                uint2 vl, vh;
                vl.x = __byte_perm(e01, e23, 0x6420); vh.x = __byte_perm(e01, e23, 0x7531);
                vl.y = __byte_perm(o01, o23, 0x6420); vh.y = __byte_perm(o01, o23, 0x7531);
                *(uint2 *)plo = vl;
                *(uint2 *)(plo + ring) = vh;
                const u32 nx = hoff[l] + 4 * (PPW * 4);                       // the next step's chunk is four pairs further on
                hoff[l] = min(nx, nx - ring);
            };
            auto h_levels = [&](auto l0_) {
            constexpr int L0 = decltype(l0_)::value;
            load_taps(L0, 0);
#pragma unroll
            for (int l = L0; l < L0 + LPW; ++l) {
                const int b = (l - L0) & 1;
                if (l < L0 + LPW - 1) load_taps(l + 1, b ^ 1);
#pragma unroll
                for (int ci = 0; ci < NC; ++ci) {
                    if (ci == 0) {
                        imma_c(acc[b][0], fa[ci][0], fa[ci][1], fa[ci][2], fa[ci][3], tb[b][ci][0], tb[b][ci][1], 0);
                        imma_c(acc[b][1], fb[ci][0], fb[ci][1], fb[ci][2], fb[ci][3], tb[b][ci][0], tb[b][ci][1], 0);
                    } else {
                        imma(acc[b][0], fa[ci][0], fa[ci][1], fa[ci][2], fa[ci][3], tb[b][ci][0], tb[b][ci][1]);
                        imma(acc[b][1], fb[ci][0], fb[ci][1], fb[ci][2], fb[ci][3], tb[b][ci][0], tb[b][ci][1]);
                    }
                }
                if (l > L0) h_store(l - 1, acc[b ^ 1]);
            }
            h_store(L0 + LPW - 1, acc[(LPW - 1) & 1]);
            };
            if (LG == 1 || hlg == 0) h_levels(std::integral_constant<int, 0>());
            else h_levels(std::integral_constant<int, 6 - LPW>());
        };
        if (!svc) {
            if (h_short) h_pass(std::integral_constant<int, NCH - 1>());
            else h_pass(std::integral_constant<int, NCH>());
        }
        mbar_wait(mb_sa, mb_phase);                                           // every thread's detector results are visible
        mb_phase ^= 1;
        {
            const int nq = s_nq;
            if (nq) run_queue(p_y0, si_lo, nq, par);
        }
        __syncthreads();                                                      // By
        if (y0 + FR < ye) issue_stage(y0 + FR);                  // the H pass is done with the staging buffer
        // ---- 3. V pass + rounding + DoG -> tiles, levels in order ----
        if (do_v && !svc) {
            u32 prev0[RBW], prev1[RBW];                          // previous level, columns 2g and 2g + 1, rows as 16-bit halves
#pragma unroll
            for (int rb = 0; rb < RBW; ++rb) prev0[rb] = prev1[rb] = 0;
            if constexpr (MAT) {
                mat_ok = 0;
#pragma unroll
                for (int rb = 0; rb < RBW; ++rb) {
                    const int y = y0 + 8 * (vjb + rb) + 2 * t;
                    mat_off[rb] = (u32)y * mat_pitch + (u32)mat_col;             // < 2^32: checked by the launcher
                    if (mat_own && y0 >= ys) mat_ok |= (y < ye ? 1u << (2 * rb) : 0u) | (y + 1 < ye ? 2u << (2 * rb) : 0u);
                }
            }
            {   // rows y0+28 .. y0+31 of the last step become the history of the G and D tiles (before anything is overwritten)
                const int hrb = 3 - vjb;
                if (hrb >= 0 && hrb < RBW && t >= 2) {
                    const int wi = wi0 + 4 * hrb;
                    u32 hv[N_G + N_D];
#pragma unroll
                    for (int k2 = 0; k2 < N_G + N_D; ++k2) hv[k2] = tiles32[k2 * TILE_W + wi];
#pragma unroll
                    for (int k2 = 0; k2 < N_G + N_D; ++k2) tiles32[k2 * TILE_W + wi - 16] = hv[k2];
                }
            }
            // (256 hi + lo + 32768) >> 16 lands in byte 2; rows 2t, 2t+1 of a column share a register as 16-bit halves
            auto v_out = [&](int l, const int (&lo)[RBW][4], const int (&hi)[RBW][4]) {
#pragma unroll
                for (int rb = 0; rb < RBW; ++rb) {
                    const u32 v0 = (u32)(hi[rb][0] << 8) + (u32)lo[rb][0], v1 = (u32)(hi[rb][1] << 8) + (u32)lo[rb][1];
                    const u32 v2 = (u32)(hi[rb][2] << 8) + (u32)lo[rb][2], v3 = (u32)(hi[rb][3] << 8) + (u32)lo[rb][3];
                    const u32 c0 = __byte_perm(v0, v1, 0x7632), c1 = __byte_perm(v2, v3, 0x7632);
                    const int wi = wi0 + 4 * rb;
                    if (l >= 1 && l <= 3) tiles32[(l - 1) * TILE_W + wi] = __byte_perm(c0, c1, 0x6420);
                    u32 d0 = 0, d1 = 0;
                    if (l >= 1) {
                        d0 = __vmaxu2(c0, prev0[rb]) - prev0[rb]; d1 = __vmaxu2(c1, prev1[rb]) - prev1[rb];   // max(cur - prev, 0)
                        tiles32[(N_G + l - 1) * TILE_W + wi] = __byte_perm(d0, d1, 0x6420);
                    }
                    if constexpr (MAT) {                                 // GaussPyramid::octaveBlur / octaveDiff planes (own rows of the segment)
                        mat_put(a.gauss[l] + mat_fo, rb, __byte_perm(c0, c1, 0x6240));
                        if (l >= 1) mat_put(a.dog[l - 1] + mat_fo, rb, __byte_perm(d0, d1, 0x6240));
                    }
                    prev0[rb] = c0; prev1[rb] = c1;
                }
            };
            // pair of (row block j, K step k, lane t): 4i + j + 4k + t; slot = that modulo np
            if constexpr (GEO::is_static) {
                // every (level, K step) is an item of one unrolled pipeline: item i + 1's fragments are loaded before item
                // i's MMAs are issued, level l - 1's epilogue (tile stores) follows level l's first MMAs
                // Both row blocks of the warp read the SAME 32 ring rows per K step (row block rb through taps shifted by 8 rb: B_rb[k][n] =
                // B_0[k - 8 rb][n]), so a step costs one pair of LDS.128 instead of one per block -- the h-plane loads were 45 % of the
                // kernel's shared-memory wavefronts, and that pipe (68-72 % busy at octaves 0 / 1) is what bounds it.  (block, step)
                // products whose taps are all zeros are left out at compile time.
                constexpr int NI = GEO::items();
                static_assert(GEO::rbw() == RBW, "row blocks per warp");
                uint4 fl[2], fh[2];
                u32 tb[2][RBW][2], off = 0;
                int lo[2][RBW][4], hi[2][RBW][4];
                auto load_item = [&](int i, int buf) {
                    const int l = GEO::item_l(i), k = GEO::item_k(i);
                    const u32 ring = (u32)geo.np(l) * (PPW * 4);
                    const u8 *plane = planes + 4 * geo.plane_off(l);
                    if (k == 0) {
                        off = voff[l];
                    } else {
                        const u32 nx = off + 4 * (PPW * 4);
                        off = min(nx, nx - ring);                // wrap: nx - ring is huge (unsigned) unless nx >= ring
                    }
#pragma unroll
                    for (int rb = 0; rb < RBW; ++rb)
                        if (GEO::v_needed(l, k, rb)) {
                            tb[buf][rb][0] = *(const u32 *)(vbase + geo.tab_voff(l) + 32 * k - 8 * rb);
                            tb[buf][rb][1] = *(const u32 *)(vbase + geo.tab_voff(l) + 32 * k - 8 * rb + 4);
                        }
                    fl[buf] = *(const uint4 *)(plane + off);
                    fh[buf] = *(const uint4 *)(plane + off + ring);
                };
                load_item(0, 0);
#pragma unroll
                for (int i = 0; i < NI; ++i) {
                    const int l = GEO::item_l(i), k = GEO::item_k(i), b = i & 1, lb = l & 1;
                    if (i + 1 < NI) load_item(i + 1, b ^ 1);
#pragma unroll
                    for (int rb = 0; rb < RBW; ++rb) {
                        if (k == 0) {   // K step 0 seeds the accumulators: 32768 (rounding) for the low bytes, 0 for the high bytes
                            imma_c(lo[lb][rb], fl[b].x, fl[b].y, fl[b].z, fl[b].w, tb[b][rb][0], tb[b][rb][1], 32768);
                            imma_c(hi[lb][rb], fh[b].x, fh[b].y, fh[b].z, fh[b].w, tb[b][rb][0], tb[b][rb][1], 0);
                        } else if (GEO::v_needed(l, k, rb)) {
                            imma(lo[lb][rb], fl[b].x, fl[b].y, fl[b].z, fl[b].w, tb[b][rb][0], tb[b][rb][1]);
                            imma(hi[lb][rb], fh[b].x, fh[b].y, fh[b].z, fh[b].w, tb[b][rb][0], tb[b][rb][1]);
                        }
                    }
                    if (k == 0 && l > 0) v_out(l - 1, lo[lb ^ 1], hi[lb ^ 1]);
                }
                v_out(5, lo[1], hi[1]);
            } else {
#pragma unroll
                for (int l = 0; l < 6; ++l) {
                    const u32 ring = (u32)geo.np(l) * (PPW * 4);              // bytes of one plane; the high-byte plane follows
                    const u8 *plane = planes + 4 * geo.plane_off(l);
                    u32 off[RBW];
                    off[0] = voff[l];
#pragma unroll
                    for (int rb = 1; rb < RBW; ++rb) {
                        const u32 nx = off[rb - 1] + PPW * 4;
                        off[rb] = min(nx, nx - ring);
                    }
                    int lo[RBW][4], hi[RBW][4];
                    {   // K step 0 seeds the accumulators: 32768 (rounding) for the low bytes, 0 for the high bytes
                        const u32 b0 = *(const u32 *)(vbase + geo.tab_voff(l)), b1 = *(const u32 *)(vbase + geo.tab_voff(l) + 4);
#pragma unroll
                        for (int rb = 0; rb < RBW; ++rb) {
                            const uint4 vl = *(const uint4 *)(plane + off[rb]), vh = *(const uint4 *)(plane + off[rb] + ring);
                            imma_c(lo[rb], vl.x, vl.y, vl.z, vl.w, b0, b1, 32768);
                            imma_c(hi[rb], vh.x, vh.y, vh.z, vh.w, b0, b1, 0);
                        }
                    }
#pragma unroll 1
                    for (int k = 1; k < geo.ksv(l); ++k) {
                        const u32 b0 = *(const u32 *)(vbase + geo.tab_voff(l) + 32 * k), b1 = *(const u32 *)(vbase + geo.tab_voff(l) + 32 * k + 4);
#pragma unroll
                        for (int rb = 0; rb < RBW; ++rb) {
                            const u32 nx = off[rb] + 4 * (PPW * 4);
                            off[rb] = min(nx, nx - ring);        // wrap: nx - ring is huge (unsigned) unless nx >= ring
                            const uint4 vl = *(const uint4 *)(plane + off[rb]), vh = *(const uint4 *)(plane + off[rb] + ring);
                            imma(lo[rb], vl.x, vl.y, vl.z, vl.w, b0, b1);
                            imma(hi[rb], vh.x, vh.y, vh.z, vh.w, b0, b1);
                        }
                    }
                    v_out(l, lo, hi);
                }
            }
        }
        // the V windows of the next step start four pairs further on (also during the warm-up steps)
#pragma unroll
        for (int l = 0; l < 6; ++l) {
            const u32 ring = (u32)geo.np(l) * (PPW * 4), nx = voff[l] + 4 * (PPW * 4);
            voff[l] = min(nx, nx - ring);
        }
        {
            const unsigned long long p_mask = s_mask[par];
            if (tid == 0) { s_nq = 0; s_mask[par ^ 1] = 0; }                  // the next step's queue and segment flags
            if (svc && p_mask) emit(si_lo, p_mask);
        }
        p_y0 = y0; p_own = do_v && y0 >= ys; par ^= 1;
    }
    // drain: the last step's detector, queue and emission
    __syncthreads();
    si_lo = detect_step(p_y0, p_own, par);
    __syncthreads();
    {
        const int nq = s_nq;
        if (nq) run_queue(p_y0, si_lo, nq, par);
    }
    __syncthreads();
    {
        const unsigned long long p_mask = s_mask[par];
        if (svc && p_mask) emit(si_lo, p_mask);
    }
}

// ---------------------------------------------------------------------------------------------------------------
// host: launch
// ---------------------------------------------------------------------------------------------------------------
namespace {
typedef void (*mma_kern_t)(const MmaArgs);
struct MmaVariant { mma_kern_t k; int fw, nch; bool (*match)(const MmaTaps &); mma_kern_t k_mat; };

// Which instantiation runs an octave: 64-column strips when the rings fit shared memory, else 32-column strips, else
// none (the dp4a/dp2a kernel).  The default pyramid's octaves get the instantiations with compile-time geometry.
const MmaVariant *mma_plan(vslam_ctx *ctx, double sigma0, int scale_samples, int octave, MmaTaps *mt, size_t *smem) {
    using G0 = MmaStaticGeom<64, 11, 13, 17, 21, 25, 31>;
    using G1 = MmaStaticGeom<64, 21, 25, 31, 39, 49, 63>;
    using G2 = MmaStaticGeom<64, 39, 49, 63, 79, 99, 123>;
    using G3 = MmaStaticGeom<32, 79, 99, 123, 155, 195, 245>;
    static const MmaVariant variants[] = {
        {octave_mma_k<2, G0, 64>, 64, 2, G0::matches, octave_mma_k<2, G0, 64, true>},
        {octave_mma_k<3, G1, 64>, 64, 3, G1::matches, octave_mma_k<3, G1, 64, true>},
        {octave_mma_k<5, G2, 64>, 64, 5, G2::matches, octave_mma_k<5, G2, 64, true>},
        {octave_mma_k<9, G3, 32>, 32, 9, G3::matches, octave_mma_k<9, G3, 32, true>},
        {octave_mma_k<2, MmaDynGeom, 64>, 64, 2, nullptr, nullptr}, {octave_mma_k<3, MmaDynGeom, 64>, 64, 3, nullptr, nullptr},
        {octave_mma_k<5, MmaDynGeom, 64>, 64, 5, nullptr, nullptr}, {octave_mma_k<5, MmaDynGeom, 32>, 32, 5, nullptr, nullptr},
        {octave_mma_k<9, MmaDynGeom, 32>, 32, 9, nullptr, nullptr}};      // materialising launches: the default pyramid's octaves only
    constexpr int NV = sizeof variants / sizeof variants[0];
    static const int mode = getenv("VSLAM_FUSED_MMA") ? atoi(getenv("VSLAM_FUSED_MMA")) : 1;
    static const int use_static = getenv("VSLAM_MMA_STATIC") ? atoi(getenv("VSLAM_MMA_STATIC")) : 1;
    static const int force_fw = getenv("VSLAM_MMA_FW") ? atoi(getenv("VSLAM_MMA_FW")) : 0;
    if (!mode) return nullptr;
    // function attributes belong to a device: set them once per device this process launches on
    static int static_smem_dev[64];
    static bool dev_init = false;
    static std::mutex mu;
    int static_smem;
    {
        std::lock_guard<std::mutex> lock(mu);
        if (!dev_init) { for (int &v : static_smem_dev) v = -1; dev_init = true; }
        const int di = ctx->device >= 0 && ctx->device < 64 ? ctx->device : 0;
        if (static_smem_dev[di] < 0) {
            cudaFuncAttributes fa;
            int mx = 0;
            for (int i = 0; i < NV; ++i)
                for (mma_kern_t k : {variants[i].k, variants[i].k_mat}) {
                    if (!k) continue;
                    if (cudaFuncGetAttributes(&fa, k) != cudaSuccess) return nullptr;
                    if ((int)fa.sharedSizeBytes > mx) mx = (int)fa.sharedSizeBytes;
                }
            for (int i = 0; i < NV; ++i)
                for (mma_kern_t k : {variants[i].k, variants[i].k_mat})
                    if (k && cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - mx) != cudaSuccess) return nullptr;
            static_smem_dev[di] = mx;
        }
        static_smem = static_smem_dev[di];
    }
    for (int fw = (force_fw == 32 ? 32 : 64); fw >= 32; fw -= 32) {
        if (build_mma_taps(sigma0, scale_samples, octave, fw, mt)) continue;
        *smem = mma_smem_bytes(*mt);
        if (*smem + static_smem > 227 * 1024) continue;
        for (int pass = use_static ? 0 : 1; pass < 2; ++pass)
            for (int i = 0; i < NV; ++i) {
                const MmaVariant &c = variants[i];
                if (c.fw != fw || (pass == 0) != (c.match != nullptr)) continue;
                if (c.nch == mt->nch && (pass == 1 || c.match(*mt))) return &c;
            }
    }
    return nullptr;
}
}  // namespace

// device image of an octave's shifted tap copies: tab[(4 l + s) cp + j] = tap_l[j + s - padl_l] or 0
namespace {
struct MmaTabEntry { int dev; double sigma0; int ss, octave, fw; void *d; };
std::vector<MmaTabEntry> g_mma_tabs;                  // device images of the tap copies, one per (device, pyramid, octave, width)
std::mutex g_mma_tabs_mu;
}  // namespace

// called when the last context of `device` is destroyed (capi.cu): the images live as long as some context can launch with them
void mma_tab_release(int device) {
    std::lock_guard<std::mutex> lock(g_mma_tabs_mu);
    for (size_t i = 0; i < g_mma_tabs.size();)
        if (g_mma_tabs[i].dev == device) { cudaFree(g_mma_tabs[i].d); g_mma_tabs.erase(g_mma_tabs.begin() + i); }
        else ++i;
}

const uint4 *mma_tab_image(vslam_ctx *ctx, double sigma0, int scale_samples, int octave, const MmaTaps &mt) {
    typedef MmaTabEntry Entry;
    std::vector<Entry> &cache = g_mma_tabs;
    std::lock_guard<std::mutex> lock(g_mma_tabs_mu);
    for (const Entry &e : cache)
        if (e.dev == ctx->device && e.sigma0 == sigma0 && e.ss == scale_samples && e.octave == octave && e.fw == mt.fw) return (const uint4 *)e.d;
    std::vector<u8> img((size_t)mt.tab_bytes, 0);
    for (int l = 0; l < 6; ++l)
        for (int sft = 0; sft < 4; ++sft)
            for (int j = 0; j < mt.cp; ++j) {
                const int idx = j + sft - mt.lv[l].padl;
                img[(size_t)(4 * l + sft) * mt.cp + (sft > mt.rv % 4 ? 4 : 0) + j] = (idx >= 0 && idx < mt.lv[l].n) ? mt.taps[l][idx] : (u8)0;
            }
    void *d = nullptr;
    if (cudaMalloc(&d, img.size()) != cudaSuccess) return nullptr;
    if (cudaMemcpy(d, img.data(), img.size(), cudaMemcpyHostToDevice) != cudaSuccess) { cudaFree(d); return nullptr; }
    cache.push_back({ctx->device, sigma0, scale_samples, octave, mt.fw, d});
    return (const uint4 *)d;
}

// strips per frame of the kernel that will run the octave (the caller sizes the segment arrays with it)
int octave_strips(vslam_ctx *ctx, double sigma0, int scale_samples, int octave, int cols) {
    if (octave_uses_umma(ctx, sigma0, scale_samples, octave)) return (cols + 47) / 48;   // fused_umma.cuh: kUmmaFWO
    MmaTaps mt;
    size_t sm;
    const MmaVariant *v = mma_plan(ctx, sigma0, scale_samples, octave, &mt, &sm);
    return v ? mma_strips(cols, mt.fw) : fused_strips(cols);
}

int octave_base_row_pad(vslam_ctx *ctx, double sigma0, int scale_samples, int octave) {
    return umma_base_pad(ctx, sigma0, scale_samples, octave);
}

int octave_base_pad(vslam_ctx *ctx, double sigma0, int scale_samples, int octave) {
    if (const int up = umma_base_pad(ctx, sigma0, scale_samples, octave)) return up;
    MmaTaps mt;
    size_t sm;
    if (!mma_plan(ctx, sigma0, scale_samples, octave, &mt, &sm)) return 0;
    // staged columns: ((X0 - 2 - rmax) & ~15) + [0, stage_pitch), X0 < cols
    return (mt.stage_pitch - mt.rmax + 30) & ~15;
}

// columns of the LEFT margin the kernel that runs the octave really stages: the mma.sync kernel's leftmost strip starts at
// (-2 - rmax) & ~15, well inside the symmetric margin the layout reserves; the tcgen05 kernel's TMA boxes use all of theirs
int octave_base_pad_left(vslam_ctx *ctx, double sigma0, int scale_samples, int octave) {
    if (const int up = umma_base_pad(ctx, sigma0, scale_samples, octave)) return up;
    MmaTaps mt;
    size_t sm;
    if (!mma_plan(ctx, sigma0, scale_samples, octave, &mt, &sm)) return 0;
    return (mt.rmax + 2 + 15) & ~15;
}

// ... and of the RIGHT margin: the last strip's staged row ends at ((X0 - 2 - rmax) & ~15) + stage_pitch
int octave_base_pad_right(vslam_ctx *ctx, double sigma0, int scale_samples, int octave, int cols) {
    if (const int up = umma_base_pad(ctx, sigma0, scale_samples, octave)) return up;
    MmaTaps mt;
    size_t sm;
    if (!mma_plan(ctx, sigma0, scale_samples, octave, &mt, &sm)) return 0;
    const int x0 = (mma_strips(cols, mt.fw) - 1) * mma_fwo(mt.fw);
    const int need = ((x0 - 2 - mt.rmax) & ~15) + mt.stage_pitch - cols;
    const int full = (mt.stage_pitch - mt.rmax + 30) & ~15;
    return need < 4 ? 4 : need > full ? full : (need + 3) & ~3;
}

// thread = one aligned word of a margin (pad_left columns on the left, pad on the right); the right margin starts at the last
// aligned word that still holds image columns (those bytes are rewritten with their own values)
__global__ void reflect_pad_k(u8 *base, int rows, int cols, size_t pitch, size_t stride, int pad, int pad_left) {
    const int wl = pad_left >> 2, x_r = cols & ~3, wr = (cols + pad - x_r + 3) >> 2, wpr = wl + wr;
    u8 *plane = base + (size_t)blockIdx.y * stride;
    const int n = rows * wpr;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const int y = i / wpr, w = i - y * wpr;
        const int x = w < wl ? 4 * w - pad_left : x_r + 4 * (w - wl);
        u8 *row = plane + (size_t)y * pitch;
        const u32 v = (u32)row[refl101(x, cols)] | (u32)row[refl101(x + 1, cols)] << 8 | (u32)row[refl101(x + 2, cols)] << 16 |
                      (u32)row[refl101(x + 3, cols)] << 24;
        *(u32 *)(row + x) = v;
    }
}

// thread = one 16-byte word of a margin row: row -k copies row refl101(-k), row rows - 1 + k copies row refl101(rows - 1 + k)
// (whole padded rows; runs after reflect_pad_k)
__global__ void reflect_pad_rows_k(u8 *base, int rows, size_t pitch, size_t stride, int pad, int row_pad) {
    u8 *plane = base + (size_t)blockIdx.y * stride;
    const int wpr = (int)(pitch >> 4), n = 2 * row_pad * wpr;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const int j = i / wpr, w = i - j * wpr;
        const int y = j < row_pad ? j - row_pad : rows + (j - row_pad);
        const uint4 *src = (const uint4 *)(plane + (ptrdiff_t)refl101(y, rows) * (ptrdiff_t)pitch - pad);
        uint4 *dst = (uint4 *)(plane + (ptrdiff_t)y * (ptrdiff_t)pitch - pad);
        dst[w] = src[w];
    }
}

int launch_reflect_pad(vslam_ctx *ctx, u8 *base, int rows, int cols, size_t pitch, size_t stride, int pad, int row_pad, int n_frames, int pad_left,
                       int pad_right) {
    if (pad <= 0) return VSLAM_OK;
    if (pad_left <= 0 || pad_left > pad || row_pad > 0) pad_left = pad;      // (whole padded rows are copied when there are row margins)
    if (pad_right <= 0 || pad_right > pad || row_pad > 0) pad_right = pad;
    const int full_pad = pad;
    pad = pad_right;                                                         // the kernel's `pad` is the right margin it writes
    const int wpr = (pad_left >> 2) + ((cols + pad - (cols & ~3) + 3) >> 2);
    int bx = (rows * wpr + 255) / 256;
    if (bx > 64) bx = 64;
    const dim3 grid(bx, n_frames);
    KernelSpan span(ctx, VSLAM_K_PAD);
    reflect_pad_k<<<grid, 256, 0, ctx->stream>>>(base, rows, cols, pitch, stride, pad, pad_left);
    VS_LAUNCH_CHECK(ctx);
    if (row_pad > 0) {
        int by = (int)((2 * row_pad * (pitch >> 4) + 255) / 256);
        if (by > 64) by = 64;
        reflect_pad_rows_k<<<dim3(by, n_frames), 256, 0, ctx->stream>>>(base, rows, pitch, stride, full_pad, row_pad);
        VS_LAUNCH_CHECK(ctx);
    }
    return VSLAM_OK;
}

int launch_octave_mma(vslam_ctx *ctx, const OctaveIO &io, int n_frames, double sigma0, int scale_samples) {
    if (n_frames <= 0 || n_frames > 65535) VS_FAIL(ctx, VSLAM_E_INVALID, "fused octave: %d frames per launch", n_frames);
    // materialising launches write two columns per store at 32-bit plane offsets: other shapes stay on fused.cu
    if (io.gauss[0] && ((io.map_pitch & 1) || (io.map_stride & 1) || (io.cols & 1) || (unsigned long long)io.rows * io.map_pitch >= (1ull << 32))) return 1;
    if ((io.base_pitch & 15) || ((uintptr_t)io.base & 15) || (io.base_stride & 15))
        VS_FAIL(ctx, VSLAM_E_INVALID, "fused octave: base plane must be 16-byte aligned (pitch %zu)", io.base_pitch);
    MmaArgs a;
    memset(&a, 0, sizeof a);
    size_t sm = 0;
    const MmaVariant *v = mma_plan(ctx, sigma0, scale_samples, io.octave, &a.mt, &sm);
    if (!v) return 1;
    if (io.gauss[0] && !v->k_mat) return 1;                          // materialising launches of other pyramids stay on fused.cu
    static_cast<OctaveIO &>(a) = io;
    a.tab_image = mma_tab_image(ctx, sigma0, scale_samples, io.octave, a.mt);
    if (!a.tab_image) VS_FAIL(ctx, VSLAM_E_NOMEM, "tap table image of octave %d", io.octave);
    a.n_strips = mma_strips(a.cols, a.mt.fw);
    // vertical split only when the batch alone cannot fill the machine (every extra segment repeats the warm-up)
    const long long want = (long long)ctx->sm_count * 4;
    const long long have = (long long)a.n_strips * n_frames;
    int nseg = (int)((want + have - 1) / have);
    const int max_seg = a.rows / (4 * FR) > 0 ? a.rows / (4 * FR) : 1;
    if (nseg > max_seg) nseg = max_seg;
    if (nseg < 1) nseg = 1;
    const int seg_rows = ((a.rows + nseg - 1) / nseg + FR - 1) / FR * FR;
    a.seg_rows = seg_rows;
    a.n_segs = (a.rows + seg_rows - 1) / seg_rows;
    const dim3 grid(a.n_strips, a.n_segs, n_frames);
    KernelSpan span(ctx, VSLAM_K_OCTAVE_MMA);
    (io.gauss[0] ? v->k_mat : v->k)<<<grid, kMmaThreads, sm, ctx->stream>>>(a);
    VS_LAUNCH_CHECK(ctx);
    return VSLAM_OK;
}

}  // namespace vslam
