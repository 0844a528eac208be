// umma_probe.cu -- stand-alone check of the tcgen05 / TMA building blocks of fused_umma.cu on one tile:
//   TMA 128B-swizzled row chunks -> V pass (tcgen05.mma kind::i8, image as MN-major B, Toeplitz taps as K-major A)
//   -> 8.8 sums split into byte planes in shared memory -> H pass (planes as K-major A, Toeplitz as K-major B) -> u8.
// Compares every level of one 128 x 64 tile with a plain CPU loop, and times the two MMA shapes.
// Descriptor strides are command-line parameters so that one GPU call can sweep the layout assumptions.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o umma_probe tools/umma_probe.cu visualslam_b200/csrc/taps.cu
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include <vector>
#include <type_traits>

#include "../visualslam_b200/csrc/common.cuh"
#include "../visualslam_b200/csrc/umma.cuh"

using namespace vslam;
using namespace vslam::umma;

struct ProbeCfg {
    int n[6], r[6];
    int rmax, P, Pc;          // P = rows of the V window above the first output row (multiple of 16, >= rmax); Pc likewise for columns (> rmax)
    int KV, KH, NV;           // V chunks (32 rows), H chunks (32 columns), V-pass N (columns incl. both halos)
    int JV, JH;               // core matrices per level of the compact Toeplitz tables
    uint32_t va_lbo, va_sbo;  // V pass A (taps), K-major no swizzle
    uint32_t vb_lbo, vb_sbo;  // V pass B (image ring), MN-major 128B swizzle
    uint32_t xa_lbo, xa_sbo;  // H pass A (byte planes)
    uint32_t hb_lbo, hb_sbo;  // H pass B (taps)
    uint32_t idesc_v, idesc_h;
    int x0, y0;               // tile origin in the padded plane (first staged column / row)
    int dbg_level, reps, timing, stage, bmn;
    u8 taps[6][248];
};

constexpr int kThreads = 288;

__global__ void __launch_bounds__(kThreads, 1) probe_k(const __grid_constant__ CUtensorMap tmap, const __grid_constant__ ProbeCfg c,
                                                       u8 *out, int *dbg_v, long long *cyc) {
    extern __shared__ __align__(1024) u8 smem_raw[];
    u8 *smem = (u8 *)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int TVB = c.JV * 128, THB = c.JH * 128, XPB = 2 * c.KH * 2048;
    u8 *ring = smem;                                   // KV chunks of 32 rows x 128 bytes
    u8 *tv = ring + c.KV * 4096;
    u8 *th = tv + 6 * TVB;
    u8 *xlo = (u8 *)(((uintptr_t)(th + 6 * THB) + 127) & ~(uintptr_t)127);
    u8 *xhi = xlo + XPB;
    __shared__ __align__(8) unsigned long long s_bar[2];
    __shared__ uint32_t s_tmem;
    const uint32_t bar_tma = smem_u32(&s_bar[0]), bar_mma = smem_u32(&s_bar[1]);

    if (tid == 0) { mbar_init(bar_tma, 1); mbar_init(bar_mma, 1); fence_barrier_init(); }
    if (warp == 1) tmem_alloc(smem_u32(&s_tmem), 256);
    // compact Toeplitz tables: core matrix j, row i, byte kk holds tap[8 j + i + kk - shift]
    for (int l = 0; l < 6; ++l) {
        for (int e = tid; e < TVB; e += kThreads) {
            const int j = e >> 7, i = (e >> 4) & 7, kk = e & 15, idx = 8 * j + i + kk - 127 + c.r[l] - c.P;
            tv[l * TVB + e] = (idx >= 0 && idx < c.n[l]) ? c.taps[l][idx] : (u8)0;
        }
        for (int e = tid; e < THB; e += kThreads) {
            const int j = e >> 7, i = (e >> 4) & 7, kk = e & 15, idx = 8 * j + i + kk - 63 - (c.Pc - 1) + c.r[l];
            th[l * THB + e] = (idx >= 0 && idx < c.n[l]) ? c.taps[l][idx] : (u8)0;
        }
    }
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tm = s_tmem;
    const uint32_t d_v = tm, d_lo = tm + 128, d_hi = tm + 192;
    const bool ctl = warp == 8 && lane == 0;
    if (c.stage == 1) { __syncthreads(); if (warp == 1) tmem_dealloc(tm, 256); return; }
    if (ctl) {
        mbar_expect_tx(bar_tma, (uint32_t)c.KV * 4096u);
        for (int k = 0; k < c.KV; ++k) tma_load_3d(smem_u32(ring + 4096 * k), &tmap, bar_tma, c.x0, c.y0 + 32 * k, blockIdx.x % 4);
        mbar_wait(bar_tma, 0);
        tc_fence_after();
    }
    if (c.stage == 2) { __syncthreads(); if (warp == 1) tmem_dealloc(tm, 256); return; }
    if (c.stage == 3 || c.stage == 4) {
        if (ctl) {
            if (c.stage == 3)
                mma_i8_ss(d_v, smem_desc(smem_u32(tv), c.va_lbo, c.va_sbo, kLayoutNone), smem_desc(smem_u32(ring), c.vb_lbo, c.vb_sbo, kLayoutSw128), c.idesc_v, 0);
            else
                mma_i8_ss(d_lo, smem_desc(smem_u32(xlo), c.xa_lbo, c.xa_sbo, kLayoutNone), smem_desc(smem_u32(th), c.hb_lbo, c.hb_sbo, kLayoutNone), c.idesc_h, 0);
            mma_commit(bar_mma);
            mbar_wait(bar_mma, 0);
        }
        __syncthreads(); if (warp == 1) tmem_dealloc(tm, 256); return;
    }
    uint32_t ph = 0;
    const int q = warp & 3, h = warp >> 2;                 // TMEM lane quadrant, column half (epilogue warps 0..7)
    const int mp = 32 * q + lane;                          // TMEM lane = reversed output row
    const int m = 127 - mp;
    const uint32_t lane_off = (uint32_t)(32 * q) << 16;

    if (c.timing) {
        // tensor-pipe time of the two instruction shapes with the descriptors precomputed (the issuing thread executes the
        // tcgen05.mma and nothing else): reps x 6 x KV V instructions, then reps x 6 x 2 KH H instructions
        if (ctl) {
            auto body = [&](auto kv_, auto kh_) {
                constexpr int KV = decltype(kv_)::value, KH = decltype(kh_)::value;
                uint64_t va[KV], vb[KV], xa[KH], xb[KH], hb[KH];
#pragma unroll
                for (int k = 0; k < KV; ++k) {
                    va[k] = smem_desc(smem_u32(tv) + 512 * k, c.va_lbo, c.va_sbo, kLayoutNone);
                    vb[k] = smem_desc(smem_u32(ring) + 4096 * k, c.vb_lbo, c.vb_sbo, kLayoutSw128);
                }
#pragma unroll
                for (int k = 0; k < KH; ++k) {
                    hb[k] = smem_desc(smem_u32(th) + 512 * k, c.hb_lbo, c.hb_sbo, kLayoutNone);
                    xa[k] = smem_desc(smem_u32(xlo) + 4096 * k, c.xa_lbo, c.xa_sbo, kLayoutNone);
                    xb[k] = smem_desc(smem_u32(xhi) + 4096 * k, c.xa_lbo, c.xa_sbo, kLayoutNone);
                }
                const uint32_t iv = c.idesc_v, ih = c.idesc_h;
                long long t0 = clock64();
                for (int rep = 0; rep < 6 * c.reps; ++rep) {
#pragma unroll
                    for (int k = 0; k < KV; ++k) mma_i8_ss(d_v, va[k], vb[k], iv, k > 0);
                }
                mma_commit(bar_mma);
                mbar_wait(bar_mma, 0);
                long long t1 = clock64();
                for (int rep = 0; rep < 6 * c.reps; ++rep) {
#pragma unroll
                    for (int k = 0; k < KH; ++k) {
                        mma_i8_ss(d_lo, xa[k], hb[k], ih, k > 0);
                        mma_i8_ss(d_hi, xb[k], hb[k], ih, k > 0);
                    }
                }
                mma_commit(bar_mma);
                mbar_wait(bar_mma, 1);
                long long t2 = clock64();
                if (blockIdx.x == 0) { cyc[0] = t1 - t0; cyc[1] = t2 - t1; }
            };
            if (c.KV == 5) body(std::integral_constant<int, 5>(), std::integral_constant<int, 3>());
            else body(std::integral_constant<int, 6>(), std::integral_constant<int, 4>());
        }
    } else {
        for (int l = 0; l < 6; ++l) {
            // ---- V pass ----
            if (ctl) {
                for (int k = 0; k < c.KV; ++k)
                    mma_i8_ss(d_v, smem_desc(smem_u32(tv + l * TVB) + 512 * k, c.va_lbo, c.va_sbo, kLayoutNone),
                              smem_desc(smem_u32(ring) + 4096 * k, c.vb_lbo, c.vb_sbo, kLayoutSw128), c.idesc_v, k > 0);
                mma_commit(bar_mma);
            }
            if (warp < 8) {
                mbar_wait(bar_mma, ph);
                tc_fence_after();
                const int np_all = c.NV / 16, p_mid = (np_all + 1) / 2;    // 16-column pieces: the first half of them or the rest
                for (int p = h ? p_mid : 0; p < (h ? np_all : p_mid); ++p) {
                    uint32_t v[16];
                    const int col0 = 16 * p;
                    tmem_ld16(d_v + lane_off + col0, v);
                    tmem_ld_wait();
                    if (l == c.dbg_level && blockIdx.x == 0)
                        for (int i = 0; i < 16; ++i) dbg_v[m * c.NV + col0 + i] = (int)v[i];
                    uint32_t lo[4], hi[4];
#pragma unroll
                    for (int w = 0; w < 4; ++w) {
                        const uint32_t t0 = __byte_perm(v[4 * w], v[4 * w + 1], 0x5410) + 0x00800080u;   // + 128: the rounding bias of the second pass
                        const uint32_t t1 = __byte_perm(v[4 * w + 2], v[4 * w + 3], 0x5410) + 0x00800080u;
                        lo[w] = __byte_perm(t0, t1, 0x6420);
                        hi[w] = __byte_perm(t0, t1, 0x7531);
                    }
                    const int u = col0 >> 4;
                    const int off = u * 2048 + (mp >> 3) * 128 + (mp & 7) * 16;
                    *(uint4 *)(xlo + off) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
                    *(uint4 *)(xhi + off) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
                }
                fence_proxy_async();
                tc_fence_before();
            }
            ph ^= 1;
            __syncthreads();
            // ---- H pass ----
            if (ctl) {
                tc_fence_after();
                for (int k = 0; k < c.KH; ++k) {
                    const uint64_t bd = smem_desc(smem_u32(th + l * THB) + 512 * k, c.hb_lbo, c.hb_sbo, kLayoutNone);
                    mma_i8_ss(d_lo, smem_desc(smem_u32(xlo) + 4096 * k, c.xa_lbo, c.xa_sbo, kLayoutNone), bd, c.idesc_h, k > 0);
                    mma_i8_ss(d_hi, smem_desc(smem_u32(xhi) + 4096 * k, c.xa_lbo, c.xa_sbo, kLayoutNone), bd, c.idesc_h, k > 0);
                }
                mma_commit(bar_mma);
            }
            if (warp < 8) {
                mbar_wait(bar_mma, ph);
                tc_fence_after();
                for (int p = 0; p < 2; ++p) {
                    uint32_t a[16], b[16];
                    const int n0 = 32 * h + 16 * p;
                    tmem_ld16(d_lo + lane_off + n0, a);
                    tmem_ld16(d_hi + lane_off + n0, b);
                    tmem_ld_wait();
                    for (int i = 0; i < 16; ++i) {
                        const uint32_t v = (b[i] << 8) + a[i];
                        if (blockIdx.x == 0) out[(l * 128 + m) * 64 + (63 - (n0 + i))] = (u8)(v >> 16);
                    }
                }
                tc_fence_before();
            }
            ph ^= 1;
            __syncthreads();
            if (ctl) tc_fence_after();
        }
    }
    __syncthreads();
    if (warp == 1) tmem_dealloc(tm, 256);
}

typedef CUresult (*encode_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                              const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                              CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at line %d\n", cudaGetErrorString(e_), __LINE__); return 2; } } while (0)

int main(int argc, char **argv) {
    // usage: umma_probe octave [key=value ...]   keys: va_lbo va_sbo vb_lbo vb_sbo xa_lbo xa_sbo hb_lbo hb_sbo nv timing reps
    const int octave = argc > 1 ? atoi(argv[1]) : 0;
    ProbeCfg c;
    memset(&c, 0, sizeof c);
    long ux0 = 32, swz = 1;
    const double k = pow(2.0, 1.0 / 3.0);
    c.rmax = 0;
    for (int l = 0; l < 6; ++l) {
        const double sig = pow(2, octave) * 1.6 * pow(k, l);
        const int n = host_gauss_ksize(sig);
        int32_t t[kMaxKsize];
        if (host_gauss_taps(sig, n, t)) { printf("taps failed\n"); return 2; }
        c.n[l] = n; c.r[l] = n / 2;
        for (int i = 0; i < n; ++i) c.taps[l][i] = (u8)t[i];
        if (c.r[l] > c.rmax) c.rmax = c.r[l];
    }
    c.P = (c.rmax + 15) / 16 * 16;
    c.KV = (128 + 2 * c.P) / 32;
    c.Pc = (c.rmax + 1 + 15) / 16 * 16;
    c.NV = (c.Pc + 48 + c.rmax + 15) / 16 * 16;
    c.va_lbo = 256; c.va_sbo = 128;
    c.vb_lbo = 4096; c.vb_sbo = 1024;
    c.xa_lbo = 2048; c.xa_sbo = 128;
    c.hb_lbo = 256; c.hb_sbo = 128;
    c.dbg_level = 0; c.reps = 1; c.timing = 0; c.stage = 0; c.bmn = 1;
    for (int i = 2; i < argc; ++i) {
        char key[32]; long val;
        if (sscanf(argv[i], "%31[^=]=%ld", key, &val) != 2) continue;
        if (!strcmp(key, "va_lbo")) c.va_lbo = val; else if (!strcmp(key, "va_sbo")) c.va_sbo = val;
        else if (!strcmp(key, "vb_lbo")) c.vb_lbo = val; else if (!strcmp(key, "vb_sbo")) c.vb_sbo = val;
        else if (!strcmp(key, "xa_lbo")) c.xa_lbo = val; else if (!strcmp(key, "xa_sbo")) c.xa_sbo = val;
        else if (!strcmp(key, "hb_lbo")) c.hb_lbo = val; else if (!strcmp(key, "hb_sbo")) c.hb_sbo = val;
        else if (!strcmp(key, "nv")) c.NV = val; else if (!strcmp(key, "timing")) c.timing = val;
        else if (!strcmp(key, "reps")) c.reps = val; else if (!strcmp(key, "dbg")) c.dbg_level = val;
        else if (!strcmp(key, "stage")) c.stage = val; else if (!strcmp(key, "bmn")) c.bmn = val;
        else if (!strcmp(key, "x0")) ux0 = val; else if (!strcmp(key, "swz")) swz = val;
    }
    c.KH = (c.NV + 31) / 32;
    c.JV = 14 + 4 * c.KV; c.JH = 6 + 4 * c.KH;
    c.idesc_v = idesc_u8(c.NV, 0, c.bmn);
    c.idesc_h = idesc_u8(64, 0, 0);
    printf("octave %d: rmax %d P %d KV %d KH %d NV %d  n = %d %d %d %d %d %d\n", octave, c.rmax, c.P, c.KV, c.KH, c.NV, c.n[0], c.n[1], c.n[2], c.n[3], c.n[4], c.n[5]);

    // padded plane: 4 frames of 512 rows x 1024 bytes
    const int PW = 1024, PH = 512, NF = 4;
    std::vector<u8> img((size_t)PW * PH * NF);
    srand(1234);
    for (auto &v : img) v = (u8)(rand() >> 7);
    u8 *d_img; CK(cudaMalloc(&d_img, img.size()));
    CK(cudaMemcpy(d_img, img.data(), img.size(), cudaMemcpyHostToDevice));
    c.x0 = (int)ux0; c.y0 = 21;                                  // deliberately unaligned tile origin

    encode_fn encode = nullptr;
    cudaDriverEntryPointQueryResult qres;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", (void **)&encode, cudaEnableDefault, &qres));
    if (!encode || qres != cudaDriverEntryPointSuccess) { printf("no cuTensorMapEncodeTiled\n"); return 2; }
    CUtensorMap tmap;
    const cuuint64_t gdim[3] = {(cuuint64_t)PW, (cuuint64_t)PH, (cuuint64_t)NF};
    const cuuint64_t gstr[2] = {(cuuint64_t)PW, (cuuint64_t)PW * PH};
    const cuuint32_t box[3] = {128, 32, 1}, estr[3] = {1, 1, 1};
    CUresult cr = encode(&tmap, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, d_img, gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                         swz ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (cr != CUDA_SUCCESS) { printf("cuTensorMapEncodeTiled -> %d\n", (int)cr); return 2; }

    u8 *d_out; int *d_dbg; long long *d_cyc;
    CK(cudaMalloc(&d_out, 6 * 128 * 64)); CK(cudaMemset(d_out, 0xee, 6 * 128 * 64));
    CK(cudaMalloc(&d_dbg, 128 * 256 * 4)); CK(cudaMemset(d_dbg, 0xff, 128 * 256 * 4));
    CK(cudaMalloc(&d_cyc, 16)); CK(cudaMemset(d_cyc, 0, 16));
    const size_t smem = 1024 + (size_t)c.KV * 4096 + 6 * (size_t)(c.JV + c.JH) * 128 + 128 + 2 * (size_t)(2 * c.KH) * 2048;
    CK(cudaFuncSetAttribute(probe_k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int grid = c.timing ? 148 : 1;
    probe_k<<<grid, kThreads, smem>>>(tmap, c, d_out, d_dbg, d_cyc);
    CK(cudaGetLastError());
    CK(cudaDeviceSynchronize());

    if (c.stage) { printf("STAGE %d ok\n", c.stage); return 0; }
    if (c.timing) {
        long long cyc[2];
        CK(cudaMemcpy(cyc, d_cyc, 16, cudaMemcpyDeviceToHost));
        const double nv = (double)c.reps * 6 * c.KV, nh = (double)c.reps * 6 * c.KH * 2;
        printf("TIMING V: %lld cycles for %.0f MMAs (N=%d) = %.1f cyc/MMA (floor %d);  H: %lld cycles for %.0f MMAs (N=64) = %.1f cyc/MMA (floor 32)\n",
               cyc[0], nv, c.NV, cyc[0] / nv, c.NV / 2, cyc[1], nh, cyc[1] / nh);
        return 0;
    }
    std::vector<u8> out(6 * 128 * 64);
    std::vector<int> dbg(128 * 256);
    CK(cudaMemcpy(out.data(), d_out, out.size(), cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(dbg.data(), d_dbg, dbg.size() * 4, cudaMemcpyDeviceToHost));
    // CPU: window row k <-> plane row y0 + k, output row m <-> window row P + m; staged column j <-> plane column x0 + j,
    // output column n <-> staged column rmax + n
    int bad_v = 0, bad_g = 0, first = 1;
    for (int l = 0; l < 6; ++l) {
        std::vector<int> v1(128 * c.NV);
        for (int m = 0; m < 128; ++m)
            for (int j = 0; j < c.NV; ++j) {
                int s = 0;
                for (int t = 0; t < c.n[l]; ++t) s += c.taps[l][t] * img[(size_t)(c.y0 + c.P + m - c.r[l] + t) * PW + c.x0 + j];
                v1[m * c.NV + j] = s;
                if (l == c.dbg_level && dbg[m * c.NV + j] != s) {
                    if (bad_v < 5) printf("  V mismatch l=%d m=%d j=%d got %d want %d\n", l, m, j, dbg[m * c.NV + j], s);
                    ++bad_v;
                }
            }
        for (int m = 0; m < 128; ++m)
            for (int n = 0; n < 64; ++n) {
                if (c.Pc - 1 + n + c.r[l] >= c.NV) continue;       // band leaves the staged columns: not a computed column
                long long s = 0;
                for (int t = 0; t < c.n[l]; ++t) s += (long long)c.taps[l][t] * (v1[m * c.NV + c.Pc - 1 + n - c.r[l] + t] + 128);
                const int want = (int)(s >> 16), got = out[(l * 128 + m) * 64 + n];
                if (got != want) {
                    if (bad_g < 8 || first) printf("  G mismatch l=%d m=%d n=%d got %d want %d\n", l, m, n, got, want);
                    first = 0;
                    ++bad_g;
                }
            }
    }
    printf("RESULT va=%u/%u vb=%u/%u xa=%u/%u hb=%u/%u nv=%d : V mismatches %d / %d, G mismatches %d / %d  %s\n", c.va_lbo, c.va_sbo, c.vb_lbo,
           c.vb_sbo, c.xa_lbo, c.xa_sbo, c.hb_lbo, c.hb_sbo, c.NV, bad_v, 128 * c.NV, bad_g, 6 * 128 * 64, (bad_v || bad_g) ? "FAIL" : "PASS");
    return (bad_v || bad_g) ? 1 : 0;
}
