// fused.cu -- one sm_100a kernel per octave: Gaussian stack + DoG + decimation + scale-space extrema with the
// contrast and edge tests, nothing but keypoints and the next octave's
// base leaving the SM unless the caller asks for the maps.
//
//   GaussVector              GaussPyramid.cpp:166-185   six 8U GaussianBlur(Size(0,0), sigma) of the octave base
//   Diff_of_Gauss            GaussPyramid.cpp:191-200   saturating differences
//   createPyramid            GaussPyramid.cpp:123-126   level 3 decimated by two = next octave's base
//   initialKeypointDetection Diff_of_Gauss.cpp:254-297  stride-3 grid, half-open 2x2x3 window, ties kept
//   FeaturePointLocalization Diff_of_Gauss.cpp:223-251  contrast test (exact IEEE sequence, detect_math.cuh)
//   computeEdgeResponse      Diff_of_Gauss.cpp:79-109, 326-338
//
// Design.  The blur is integer-ALU bound (1738 MAC per input pixel), so the kernel is built
// around the packed dot-product pipe: u8 pixels x u8 taps with IDP.4A in the horizontal pass, 8.8 sums x u8 taps
// with IDP.2A in the vertical pass, all sums exact, hence bit-identical to OpenCV's fixed-point filter.
// A CTA owns a strip of 60 columns (computes 64: two halo columns each side for the detector) of one frame and
// MARCHES down it in steps of 32 rows.  Per step:
//   1. stage   base rows [y0 + rmin, y0 + rmax + 32) with reflect-101 resolved            global/L2 -> smem
//   2. H pass  the 32 NEW 8.8 rows of each level (rows y0 + r_l ...), as packed row pairs  smem -> smem (IDP.4A)
//   3. V pass  32 output rows of each level from the level's r_l + 16 resident row pairs   smem -> regs (IDP.2A)
//              + rounding, DoG against the previous level (registers), column-major u8 tiles in smem
//   4. detect  extrema / contrast / edge on the tiles (one warp per (level, sample row): ballot-ordered emission),
//              decimated level 3 -> next base, shift the h buffers up by 16 pairs
// so every 8.8 row is computed once (no vertical halo recomputation) and never leaves shared memory.
#include "detect_math.cuh"
#include "fused.cuh"

namespace vslam {

constexpr int FW = kFusedW, FWO = kFusedWO, FR = kFusedR;
constexpr int FTR = FR + 4;                 // tile rows: 4 rows of history + the step's 32
constexpr int FTW = FTR / 4;                // words per tile column (9: odd -> conflict-free column-major access)
constexpr int F_NG = 3, F_ND = 5;           // G tiles: Gaussian levels 1..3; D tiles: DoG levels 0..4

// ---------------------------------------------------------------------------------------------------------------
// host: taps
// ---------------------------------------------------------------------------------------------------------------
int build_fused_taps(double sigma0, int scale_samples, int octave, FusedTaps *out) {
    memset(out, 0, sizeof *out);
    int32_t t[kMaxKsize];
    const double k = pow(2.0, 1.0 / (double)scale_samples);          // GaussPyramid.hpp:69
    int rmin = 1 << 30, rmax = 0, hoff = 0;
    for (int l = 0; l < 6; ++l) {
        const double sig = pow(2, octave) * sigma0 * pow(k, l);      // GaussPyramid.cpp:157
        const int n = host_gauss_ksize(sig);
        if (n > kMaxKsize - 8) return VSLAM_E_INVALID;
        int rc = host_gauss_taps(sig, n, t);
        if (rc) return rc;
        FusedLevel &lv = out->lv[l];
        lv.n = n; lv.r = n / 2;
        lv.nw = ((n + 3) / 4 + 1) & ~1;
        lv.jp = (lv.r + 1 + 1) & ~1;
        if (lv.nw > kFusedHWords || lv.jp > kFusedVPairs) return VSLAM_E_INVALID;
        for (int i = 0; i < n; ++i) {
            if (t[i] < 0 || t[i] > 255) return VSLAM_E_INVALID;
            lv.hq[i >> 2] |= (u32)t[i] << (8 * (i & 3));
        }
        auto tap = [&](int i) -> u32 { return (i >= 0 && i < n) ? (u32)t[i] : 0u; };
        for (int j = 0; j <= lv.r; ++j)
            lv.eo[j] = tap(2 * j) | tap(2 * j + 1) << 8 | tap(2 * j - 1) << 16 | tap(2 * j) << 24;
        lv.np = lv.r + FR / 2;
        lv.hoff = hoff;
        hoff += lv.np * FW;
        if (lv.r < rmin) rmin = lv.r;
        if (lv.r > rmax) rmax = lv.r;
    }
    out->hbuf_words = hoff + 2 * FW;                                 // spare pair rows: padded (zero) taps read them
    out->rmin = rmin; out->rmax = rmax;
    int rpad = rmax;
    while ((rpad & 3) != 2) ++rpad;
    out->rpad = rpad;
    out->stage_rows = FR + rmax - rmin;
    // widest read of the H pass: word w0 + nw + 4 with w0 = (48 + rpad - r) >> 2
    int need = 0;
    for (int l = 0; l < 6; ++l) {
        const int w = ((FW - 16 + rpad - out->lv[l].r) >> 2) + out->lv[l].nw + 6;
        if (w > need) need = w;
    }
    const int wmin = (FW + 2 * rpad + 3) / 4 + 1;
    if (need < wmin) need = wmin;
    need += 3;                                                       // the 16-byte alignment shift of the origin
    need += 4;                                                       // LDS.128 reads whole 16-byte blocks
    out->stage_pitch = (need * 4 + 15) & ~15;
    while ((out->stage_pitch & 63) != 32) out->stage_pitch += 16;   // conflict-free LDS.128 across the row pairs of a warp
    {   // row = (q * magic) >> 20 for q < stage_rows * (pitch / 16): checked exhaustively
        const int spq = out->stage_pitch >> 4, nq = out->stage_rows * spq;
        out->spq_magic = ((1 << 20) + spq - 1) / spq;
        for (int q = 0; q < nq; ++q)
            if ((int)(((unsigned)q * (unsigned)out->spq_magic) >> 20) != q / spq) return VSLAM_E_INVALID;
    }
    return VSLAM_OK;
}

size_t fused_smem_bytes(const FusedTaps &t) {
    return (size_t)t.hbuf_words * 4 + (size_t)t.stage_rows * t.stage_pitch + (size_t)(F_NG + F_ND) * FW * FTR;   // + 2880 static
}

int fused_strips(int cols) { return (cols + FWO - 1) / FWO; }

// ---------------------------------------------------------------------------------------------------------------
// device: the two blur passes
// ---------------------------------------------------------------------------------------------------------------
// H pass body: UH tap words (4 taps each) for 2 rows x 4 groups x 4 byte alignments.  The thread's window of staged
// words starts WS words into a 16-byte aligned block, so the loads are three conflict-free LDS.128 per row (the staged
// pitch is 32 mod 64 bytes).  The words are first re-aligned by the level's byte offset (sel), then the three shifted
// views of every word come from PRMT (ALU pipe), so the dot-product pipe sees exactly one IDP.4A per 4 taps per output.
template <int UH, int WS>
__device__ __forceinline__ void hbody(const uint4 *__restrict__ rowA, const uint4 *__restrict__ rowB,
                                      const u32 *__restrict__ hq, int it, u32 sel, u32 (&acc)[2][4][4]) {
    u32 t[UH];
#pragma unroll
    for (int u = 0; u < UH; ++u) t[u] = hq[it + u];
#pragma unroll
    for (int row = 0; row < 2; ++row) {
        const uint4 *p = (row ? rowB : rowA) + (it >> 2);
        constexpr int NQ = (WS + UH + 5 + 3) / 4;
        u32 raw[4 * NQ], e[UH + 4];
#pragma unroll
        for (int q = 0; q < NQ; ++q) {
            const uint4 v = p[q];
            raw[4 * q] = v.x; raw[4 * q + 1] = v.y; raw[4 * q + 2] = v.z; raw[4 * q + 3] = v.w;
        }
#pragma unroll
        for (int j = 0; j < UH + 4; ++j) e[j] = __byte_perm(raw[WS + j], raw[WS + j + 1], sel);
#pragma unroll
        for (int j = 0; j < UH + 3; ++j) {
            const u32 v0 = e[j];
            const u32 v1 = __byte_perm(e[j], e[j + 1], 0x4321);
            const u32 v2 = __byte_perm(e[j], e[j + 1], 0x5432);
            const u32 v3 = __byte_perm(e[j], e[j + 1], 0x6543);
#pragma unroll
            for (int u = 0; u < UH; ++u) {
                const int g = j - u;
                if (g >= 0 && g < 4) {
                    acc[row][g][0] = __dp4a(v0, t[u], acc[row][g][0]);
                    acc[row][g][1] = __dp4a(v1, t[u], acc[row][g][1]);
                    acc[row][g][2] = __dp4a(v2, t[u], acc[row][g][2]);
                    acc[row][g][3] = __dp4a(v3, t[u], acc[row][g][3]);
                }
            }
        }
    }
}

template <int WS>
__device__ __forceinline__ void hloop(const uint4 *__restrict__ rowA, const uint4 *__restrict__ rowB, const u32 *__restrict__ hq,
                                      int nw, u32 sel, u32 (&acc)[2][4][4]) {
    int it = 0;
    for (; it + 4 <= nw; it += 4) hbody<4, WS>(rowA, rowB, hq, it, sel, acc);
    if (nw & 2) hbody<2, WS>(rowA, rowB, hq, it, sel, acc);
}

// V pass body: UV tap pairs for RB output rows of one column.  d[] are packed row pairs (2p | 2p+1 << 16); even
// output rows use the low (E) tap pair of the word, odd rows the high (O) pair.
template <int UV, int RB>
__device__ __forceinline__ void vbody(const u32 *__restrict__ hb, const u32 *__restrict__ eo, int jb, u32 (&acc)[RB]) {
    constexpr int Q = RB / 2;
    u32 d[Q + UV - 1];
#pragma unroll
    for (int i = 0; i < Q + UV - 1; ++i) d[i] = hb[(jb + i) * FW];
#pragma unroll
    for (int u = 0; u < UV; ++u) {
        const u32 t = eo[jb + u];
#pragma unroll
        for (int q = 0; q < Q; ++q) {
            acc[2 * q] = __dp2a_lo(d[q + u], t, acc[2 * q]);
            acc[2 * q + 1] = __dp2a_hi(d[q + u], t, acc[2 * q + 1]);
        }
    }
}

// level order of the H-pass task list, balanced so that every group of 64 threads gets a similar number of taps
template <int NT> __device__ __forceinline__ int hperm(int i) {
    if (NT == 128) return (0x542310 >> (4 * i)) & 15;   // 0,1,3,2,4,5 -> halves {0,3,4} / {1,2,5}
    return (0x234510 >> (4 * i)) & 15;                  // 0,1,5,4,3,2 -> groups {0,3} {1,2} {5} {4}
}

// ---------------------------------------------------------------------------------------------------------------
// the kernel
// ---------------------------------------------------------------------------------------------------------------
template <int NT>
__global__ void __launch_bounds__(NT) octave_fused_k(const __grid_constant__ FusedArgs a) {
    extern __shared__ __align__(16) u32 smem[];
    constexpr int NW = NT / 32;
    constexpr int RG = NT / FW;                 // row groups of the V pass
    constexpr int RB = FR / RG;                 // output rows per thread
    const FusedTaps &T = a.taps;
    u32 *hbuf = smem;
    u8 *stage = (u8 *)(smem + T.hbuf_words);
    const int SP = T.stage_pitch;                  // a multiple of 16
    u8 *tiles = stage + T.stage_rows * SP;
    u32 *tiles32 = (u32 *)tiles;
    u8 *gt = tiles, *dt = tiles + F_NG * FW * FTR;
    __shared__ int s_res[3 * (FR / 3 + 2) * (FWO / 3)];              // detector results of one step
    __shared__ unsigned short s_queue[3 * (FR / 3 + 2) * (FWO / 3)];  // samples that need the full localisation
    __shared__ int s_nq;
    __shared__ float s_div[511];                                      // float(i - 255) / 255f, the reference's divisions

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int strip = blockIdx.x, seg = blockIdx.y, frame = blockIdx.z;
    const int rows = a.rows, cols = a.cols;
    const int X0 = strip * FWO, X0c = X0 - 2, xs = X0c - T.rpad;
    const int xs16 = xs & ~15, xshift = xs - xs16;   // staged rows start on a 16-byte boundary of the plane
    const u8 *__restrict__ base = a.base + (size_t)frame * a.base_stride;
    const int ys = seg * a.seg_rows;
    const int ye = min(ys + a.seg_rows, rows);
    const int yv = seg > 0 ? ys - FR : 0;       // first step with a V pass (one extra step feeds the detector's history)
    const int K = (2 * T.rmax + FR - 1) / FR;   // warm-up steps: H pass only

    // stage base rows y0 + rmin ..., columns xs16 ..., reflect-101 resolved.  16-byte groups inside the image go through
    // cp.async so the copy of step c+1 overlaps the V pass and the detector of step c; border groups are gathered byte
    // by byte.  One thread per (row, group) of the flattened tile.
    const int SPQ = SP >> 4, n_stage = T.stage_rows * SPQ;
    auto issue_stage = [&](int y0n) {
        for (int q = tid; q < n_stage; q += NT) {
            const int row = (int)(((u32)q * (u32)T.spq_magic) >> 20), j = q - row * SPQ;
            const int y = refl101(y0n + T.rmin + row, rows);
            const u8 *__restrict__ src = base + (size_t)y * a.base_pitch;
            u32 *dst = (u32 *)(stage + row * SP) + 4 * j;
            const int x = xs16 + 16 * j;
            if (x >= 0 && x + 15 < cols) {
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
    for (int i = threadIdx.x; i < 511; i += NT) s_div[i] = __fdiv_rn((float)(i - 255), 255.0f);
    if (threadIdx.x == 0) s_nq = 0;

    constexpr int NS = FWO / 3;                                               // 20 sample columns per strip
    constexpr int LS = FW * FTR;                                              // bytes between tiles
    // computeEdgeResponse + threshold on the G tile of `level` around sample (xc, yc); yb = first row of the tiles' step
    auto edge_pass = [&](int level, int xc, int yc, int yb) -> bool {
        const u8 *g = gt + ((level - 1) * FW + (xc - X0c)) * FTR + (yc - yb + 4);
        int ix2 = 0, iy2 = 0, ixy = 0;
#pragma unroll
        for (int u = 0; u <= 1; ++u)
#pragma unroll
            for (int v = 0; v <= 1; ++v) {
                // reflect-101 of an index at most one outside [0, len): a single bounce (len == 1 -> 0)
                const int xx = xc + v, yy = yc + u;
                const int xr = (xx + 1 < cols ? xx + 1 : max(cols - 2, 0)) - xc, xl = (xx > 0 ? xx - 1 : min(1, cols - 1)) - xc;
                const int yd = (yy + 1 < rows ? yy + 1 : max(rows - 2, 0)) - yc, yu = (yy > 0 ? yy - 1 : min(1, rows - 1)) - yc;
                const int gxv = (int)g[xr * FTR + u] - (int)g[xl * FTR + u];
                const int gyv = (int)g[v * FTR + yd] - (int)g[v * FTR + yu];
                ix2 += gxv * gxv; iy2 += gyv * gyv; ixy += gxv * gyv;
            }
        return edge_from_sums((float)ix2, (float)iy2, (float)ixy) < a.edge_thr;   // Diff_of_Gauss.cpp:336
    };
    // queued samples of the step whose tiles start at row yb: FeaturePointLocalization's full sequence, all lanes busy
    auto run_queue = [&](int yb, int si_lo) {
        const int nq = s_nq;
        for (int i = tid; i < nq; i += NT) {
            const int s = s_queue[i];
            const int k = s % NS, t = s / NS;
            const int level = 1 + t % 3, sr = t / 3;
            const int yc = 3 * (si_lo + sr), xc = X0 + 3 * k;
            const u8 *p = dt + (level * FW + (xc - X0c)) * FTR + (yc - yb + 4);
            const int oxm = xc > 0 ? -FTR : 0, oym = yc > 0 ? -1 : 0;
            const int oxp = xc + 1 < cols ? FTR : 0, oyp = yc + 1 < rows ? 1 : 0;
            const int c = p[0];
            const int dx = (int)p[oxm] - (int)p[oxp], dy = (int)p[oym] - (int)p[oyp], ds = (int)p[-LS] - (int)p[LS];
            const float zhat = localize_f(s_div[dx + 255], s_div[dy + 255], s_div[ds + 255], s_div[c + 255]);
            if (zhat > a.contrast_thr && edge_pass(level, xc, yc, yb))               // :245
                s_res[s] = (int)__fmul_rn(zhat, 255.0f);                             // :246
        }
    };
    // ordered emission: one warp per segment (level, sample row, strip) of a finished step
    auto emit = [&](int si_lo, int n_srow) {
        for (int task = warp; task < 3 * n_srow; task += NW) {
            const int res = lane < NS ? s_res[task * NS + lane] : -1;
            const unsigned bal = __ballot_sync(0xffffffffu, res >= 0);
            if (bal) {
                const int level = 1 + task % 3, si = si_lo + task / 3;
                const int cnt = __popc(bal);
                const int sg = frame * a.segs_per_frame + a.seg_base + ((level - 1) * a.ni + si) * a.n_strips + strip;
                int off = 0;
                if (lane == 0) {
                    off = atomicAdd(a.cursor, cnt);
                    a.seg_off[sg] = off;
                    a.seg_cnt[sg] = cnt;
                }
                off = __shfl_sync(0xffffffffu, off, 0);
                if (res >= 0) {
                    const int pos = off + __popc(bal & ((1u << lane) - 1u));
                    if (pos < a.staging_cap) {
                        vslam_keypoint kp;
                        kp.row = 3 * si + 1; kp.col = X0 + 3 * lane + 1; kp.value = res; kp.padding = 1; kp.octave = a.octave;
                        kp.level = level;
                        a.staging[pos] = kp;
                    }
                }
            }
        }
    };

    // Per step three barriers.  Small jobs of the PREVIOUS step ride along with the two blur passes:
    //   P1 (after B1)  H pass of step c            + queued localisations of step c-1 (its tiles are still intact)
    //   P2 (after B2)  V pass of step c -> tiles   + ordered emission of step c-1, prefetch of step c+1's base rows
    //   P3 (after B3)  detector tests of step c, decimated level 3 -> next base, slide of the h buffers
    int p_y0 = 0, p_si_lo = 0, p_nsrow = 0;                                   // the previous step's detector state
    for (int y0 = yv - K * FR; y0 < ye; y0 += FR) {
        const bool do_v = y0 >= yv;
        asm volatile("cp.async.wait_all;" ::: "memory");
        __syncthreads();                                                      // B1: staged rows landed; P3 of c-1 done
        // ---- 2. H pass: rows y0 + r_l .. y0 + r_l + 31 of every level -> pairs r_l .. r_l + 15 ----------------
        for (int t = tid; t < 6 * 64; t += NT) {
            const int l = hperm<NT>(t >> 6);
            const FusedLevel &lv = T.lv[l];
            if (y0 + FR <= yv - 2 * lv.r) continue;          // warm-up: this level's rows are not needed yet
            const int p = (t & 63) >> 2, cg = t & 3;
            const int boff = 16 * cg + T.rpad - lv.r + xshift;
            const u32 sel = 0x3210u + 0x1111u * (u32)(boff & 3);
            const int wl = boff >> 2;                            // first staged word of the thread's window
            const uint4 *rowA = (const uint4 *)(stage + (lv.r - T.rmin + 2 * p) * SP) + (wl >> 2);
            const uint4 *rowB = rowA + (SP >> 4);
            u32 acc[2][4][4];
#pragma unroll
            for (int i = 0; i < 32; ++i) (&acc[0][0][0])[i] = 0;
            switch (wl & 3) {                                    // warp-uniform (same level, 16-column groups)
            case 0: hloop<0>(rowA, rowB, lv.hq, lv.nw, sel, acc); break;
            case 1: hloop<1>(rowA, rowB, lv.hq, lv.nw, sel, acc); break;
            case 2: hloop<2>(rowA, rowB, lv.hq, lv.nw, sel, acc); break;
            default: hloop<3>(rowA, rowB, lv.hq, lv.nw, sel, acc); break;
            }
            uint4 *o = (uint4 *)(hbuf + lv.hoff + (lv.r + p) * FW + 16 * cg);
#pragma unroll
            for (int g = 0; g < 4; ++g) {
                uint4 v;
                v.x = acc[0][g][0] | acc[1][g][0] << 16;
                v.y = acc[0][g][1] | acc[1][g][1] << 16;
                v.z = acc[0][g][2] | acc[1][g][2] << 16;
                v.w = acc[0][g][3] | acc[1][g][3] << 16;
                o[g] = v;
            }
        }
        if (p_nsrow) run_queue(p_y0, p_si_lo);
        __syncthreads();                                                      // B2
        if (y0 + FR < ye) issue_stage(y0 + FR);                  // the H pass is done with the staging buffer
        // ---- 3. V pass + rounding + DoG -> tiles (column-major: tile k, column x, rows contiguous) -------------
        if (do_v) {
            const int x = tid & (FW - 1), rg = tid / FW;
            const int gx = X0c + x;
            const bool own_x = gx >= X0 && gx < X0 + FWO && gx < cols;
            u32 prev[RB / 2];
#pragma unroll
            for (int q = 0; q < RB / 2; ++q) prev[q] = 0;
            for (int l = 0; l < 6; ++l) {
                const FusedLevel &lv = T.lv[l];
                const u32 *hb = hbuf + lv.hoff + rg * (RB / 2) * FW + x;
                u32 acc[RB];
#pragma unroll
                for (int m = 0; m < RB; ++m) acc[m] = 32768u;             // (v + (1 << 15)) >> 16
                int jb = 0;
                for (; jb + 4 <= lv.jp; jb += 4) vbody<4, RB>(hb, lv.eo, jb, acc);
                if (lv.jp & 2) vbody<2, RB>(hb, lv.eo, jb, acc);
                u32 cur[RB / 2];                                          // two rows per word, 16 bits each
#pragma unroll
                for (int q = 0; q < RB / 2; ++q) cur[q] = __byte_perm(acc[2 * q], acc[2 * q + 1], 0x7632);
                const int wofs = 1 + rg * (RB / 4);
                if (l >= 1 && l <= 3) {
                    u32 *g = tiles32 + ((l - 1) * FW + x) * FTW + wofs;
                    if (rg == RG - 1) g[-wofs] = g[FTW - 1 - wofs];       // the last 4 rows become the next step's history
#pragma unroll
                    for (int k = 0; k < RB / 4; ++k) g[k] = __byte_perm(cur[2 * k], cur[2 * k + 1], 0x6420);
                }
                u32 dg[RB / 2];
#pragma unroll
                for (int q = 0; q < RB / 2; ++q) dg[q] = __vmaxu2(cur[q], prev[q]) - prev[q];   // max(cur - prev, 0)
                if (l >= 1) {
                    u32 *d = tiles32 + ((F_NG + l - 1) * FW + x) * FTW + wofs;
                    if (rg == RG - 1) d[-wofs] = d[FTW - 1 - wofs];
#pragma unroll
                    for (int k = 0; k < RB / 4; ++k) d[k] = __byte_perm(dg[2 * k], dg[2 * k + 1], 0x6420);
                }
                if (a.gauss[0] && own_x) {                                // materialise (GaussPyramid getters)
                    u8 *gp = a.gauss[l] + (size_t)frame * a.map_stride + gx;
                    u8 *dp = l ? a.dog[l - 1] + (size_t)frame * a.map_stride + gx : nullptr;
#pragma unroll
                    for (int m = 0; m < RB; ++m) {
                        const int y = y0 + rg * RB + m;
                        if (y >= ys && y < ye) {
                            gp[(size_t)y * a.map_pitch] = (u8)(cur[m >> 1] >> (16 * (m & 1)));
                            if (l) dp[(size_t)y * a.map_pitch] = (u8)(dg[m >> 1] >> (16 * (m & 1)));
                        }
                    }
                }
#pragma unroll
                for (int q = 0; q < RB / 2; ++q) prev[q] = cur[q];
            }
        }
        if (p_nsrow) emit(p_si_lo, p_nsrow);
        if (tid == 0) s_nq = 0;
        __syncthreads();                                                      // B3
        // ---- 4a. detector, test.  Thread = (sample column, level), marching over the step's sample rows.  Result per
        // sample in s_res: the keypoint's value, or -1; samples whose three derivatives are all non-zero need the full
        // 3x3 inverse (rare) and are queued for P1 of the next step.
        const bool own_step = do_v && y0 >= ys;
        int si_lo = 0, n_srow = 0;
        if (own_step && a.detect) {
            const int yc_lo = y0 == 0 ? 0 : y0 - 2;
            const int yc_hi = (y0 + FR >= rows) ? rows - 2 : y0 + FR - 3;   // centre row yc needs rows up to yc + 2
            si_lo = (yc_lo + 2) / 3;
            n_srow = yc_hi >= 0 ? yc_hi / 3 - si_lo + 1 : 0;
            if (n_srow < 0) n_srow = 0;
            constexpr int NG = NT / (3 * NS);                                 // sample-row groups
            if (tid < NG * 3 * NS) {
                const int id = tid % (3 * NS), grp = tid / (3 * NS);
                const int k = id % NS, level = 1 + id / NS;
                const int xc = X0 + 3 * k;
                const bool colok = xc <= cols - 2;
                const int oxm = xc > 0 ? -FTR : 0, oxp = xc + 1 < cols ? FTR : 0;
                const u8 *pc = dt + (level * FW + (xc - X0c)) * FTR + 4 - y0;
                for (int sr = grp; sr < n_srow; sr += NG) {
                    const int yc = 3 * (si_lo + sr);
                    const int s = (sr * 3 + level - 1) * NS + k;
                    int res = -1;
                    if (colok) {
                        // D tile of DoG `level` at (xc, yc); neighbours at constant strides, clamped at the image border
                        const u8 *p = pc + yc;
                        const int oym = yc > 0 ? -1 : 0, oyp = yc + 1 < rows ? 1 : 0;
                        const int c = p[0];
                        const int a0 = p[oxm + oym], a1 = p[oym], a2 = p[oxm];                    // level
                        const int b0 = p[-LS + oxm + oym], b1 = p[-LS + oym], b2 = p[-LS + oxm], b3 = p[-LS];   // level - 1
                        const int c0 = p[LS + oxm + oym], c1 = p[LS + oym], c2 = p[LS + oxm], c3 = p[LS];       // level + 1
                        const int mn = min(min(min(a0, a1), min(a2, c)), min(min(min(b0, b1), min(b2, b3)), min(min(c0, c1), min(c2, c3))));
                        const int mx = max(max(max(a0, a1), max(a2, c)), max(max(max(b0, b1), max(b2, b3)), max(max(c0, c1), max(c2, c3))));
                        if (c == mn || c == mx) {                                            // Diff_of_Gauss.cpp:287
                            const int dx = a2 - (int)p[oxp], dy = a1 - (int)p[oyp], ds = b3 - c3;   // :226-228
                            if (dx == 0 || dy == 0 || ds == 0) {
                                // cv::invert of the singular A*A^T gives zeros: dog_zhat = value/255 (tables built on the
                                // host with the same float ops: a.cmin, a.val_lut)
                                if (c >= a.cmin && edge_pass(level, xc, yc, y0)) res = a.val_lut[c];
                            } else {
                                s_queue[atomicAdd(&s_nq, 1)] = (unsigned short)s;
                            }
                        }
                    }
                    s_res[s] = res;
                }
            }
        }
        // ---- 4b. next octave's base: level 3 at even rows and columns (GaussPyramid.cpp:123-126) ------------------
        if (own_step && a.next_base) {
            u8 *nb = a.next_base + (size_t)frame * a.next_stride;
            for (int t = tid; t < (FWO / 2) * (FR / 4); t += NT) {
                const int oc = t % (FWO / 2), w = t / (FWO / 2);
                const int ox = X0 / 2 + oc, oy = (y0 >> 1) + 2 * w;
                const u32 v = tiles32[(2 * FW + 2 * oc + 2) * FTW + 1 + w];
                if (ox < a.next_cols) {
                    if (oy < a.next_rows) nb[(size_t)oy * a.next_pitch + ox] = (u8)v;
                    if (oy + 1 < a.next_rows) nb[(size_t)(oy + 1) * a.next_pitch + ox] = (u8)(v >> 16);
                }
            }
        }
        // ---- 4c. slide the h buffers up by 16 pairs: thread = (level, 4-column group), rows in order ---------------
        if (tid < 6 * (FW / 4)) {
            const FusedLevel &lv = T.lv[tid / (FW / 4)];
            uint4 *p = (uint4 *)(hbuf + lv.hoff) + (tid % (FW / 4));
            for (int i = 0; i < lv.r; i += 8) {
                uint4 v[8];
#pragma unroll
                for (int k = 0; k < 8; ++k)
                    if (i + k < lv.r) v[k] = p[(i + k + FR / 2) * (FW / 4)];
#pragma unroll
                for (int k = 0; k < 8; ++k)
                    if (i + k < lv.r) p[(i + k) * (FW / 4)] = v[k];
            }
        }
        p_y0 = y0; p_si_lo = si_lo; p_nsrow = n_srow;
    }
    // drain: the last step's queue and emission
    __syncthreads();
    if (p_nsrow) {
        run_queue(p_y0, p_si_lo);
        __syncthreads();
        emit(p_si_lo, p_nsrow);
    }
}

// ---------------------------------------------------------------------------------------------------------------
// host: launch
// ---------------------------------------------------------------------------------------------------------------
int launch_octave_fused(vslam_ctx *ctx, FusedArgs &a, int n_frames) {
    if (n_frames <= 0 || n_frames > 65535) VS_FAIL(ctx, VSLAM_E_INVALID, "fused octave: %d frames per launch", n_frames);
    if ((a.base_pitch & 15) || ((uintptr_t)a.base & 15) || (a.base_stride & 15))
        VS_FAIL(ctx, VSLAM_E_INVALID, "fused octave: base plane must be 16-byte aligned (pitch %zu)", a.base_pitch);
    const size_t sm = fused_smem_bytes(a.taps);
    if (sm > 223 * 1024) VS_FAIL(ctx, VSLAM_E_INVALID, "fused octave %d needs %zu bytes of shared memory", a.octave, sm);
    a.n_strips = fused_strips(a.cols);
    // vertical split only when the batch alone cannot fill the machine: every extra segment repeats the H-pass
    // warm-up and one V step
    const long long want = (long long)ctx->sm_count * 4;
    const long long have = (long long)a.n_strips * n_frames;
    int nseg = (int)((want + have - 1) / have);
    const int max_seg = a.rows / (4 * FR) > 0 ? a.rows / (4 * FR) : 1;
    if (nseg > max_seg) nseg = max_seg;
    if (nseg < 1) nseg = 1;
    int seg_rows = ((a.rows + nseg - 1) / nseg + FR - 1) / FR * FR;
    a.seg_rows = seg_rows;
    a.n_segs = (a.rows + seg_rows - 1) / seg_rows;
    const dim3 grid(a.n_strips, a.n_segs, n_frames);
    static bool attr_done_dev[64] = {false};                         // function attributes belong to a device
    bool &attr_done = attr_done_dev[ctx->device >= 0 && ctx->device < 64 ? ctx->device : 0];
    if (!attr_done) {
        cudaFuncAttributes fa;
        VS_CUDA(ctx, cudaFuncGetAttributes(&fa, octave_fused_k<128>));
        VS_CUDA(ctx, cudaFuncSetAttribute(octave_fused_k<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - (int)fa.sharedSizeBytes));
        VS_CUDA(ctx, cudaFuncGetAttributes(&fa, octave_fused_k<256>));
        VS_CUDA(ctx, cudaFuncSetAttribute(octave_fused_k<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - (int)fa.sharedSizeBytes));
        attr_done = true;
    }
    KernelSpan span(ctx, VSLAM_K_FUSED);
    static const long nt128_max = getenv("VSLAM_NT128_MAX") ? atol(getenv("VSLAM_NT128_MAX")) : 75 * 1024;
    if ((long)sm <= nt128_max) octave_fused_k<128><<<grid, 128, sm, ctx->stream>>>(a);
    else octave_fused_k<256><<<grid, 256, sm, ctx->stream>>>(a);
    VS_LAUNCH_CHECK(ctx);
    return VSLAM_OK;
}

}  // namespace vslam
