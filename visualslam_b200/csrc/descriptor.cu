// descriptor.cu -- the stage that follows the hot path in ./DoG, one CTA per keypoint:
//   orient_hist_k   filterKeypoints' second half   Diff_of_Gauss.cpp:340-369 + orientationHistogram :112-133
//   sift_k          SIFT + rotateImageSection      Diff_of_Gauss.cpp:561-693, :528-559
// Both blur a 16x16 magnitude window with OpenCV's fp32 GaussianBlur (kernel size cvRound(sigma*8+1)|1, up to 489 taps):
// every output is one sequential chain of fused multiply-adds in the order OpenCV's RowFilter (taps left to right) and
// SymmColumnFilter (centre tap, then symmetric pairs) use, so the result does not depend on how the work is split over threads.
// Magnitude and orientation are formed from the level's u8 Gaussian on the fly (grad_math.cuh): the fp32 maps of
// GaussPyramid::processGradients never exist in HBM.
#include "descriptor.cuh"
#include "grad_math.cuh"

namespace vslam {

static const int W = 16;                                           // windowSize (Diff_of_Gauss.cpp:305, :565)

__host__ __device__ __forceinline__ int stage_stride(int r) {     // floats per staged row: >= 16 + 2r, == 16 (mod 32)
    const int need = W + 2 * r;
    return ((need + 15) / 32) * 32 + 16;
}

// One output of OpenCV's RowFilter: taps left to right over n consecutive samples, one fused multiply-add each.  The tap
// table starts 16-byte aligned, so four taps arrive per shared-memory load.
__device__ __forceinline__ float row_chain(const float *__restrict__ taps, const float *__restrict__ s, int n) {
    float a = __fmul_rn(taps[0], s[0]);
    int u = 1;
    for (; u < 4 && u < n; ++u) a = __fmaf_rn(taps[u], s[u], a);
    for (; u + 3 < n; u += 4) {
        const float4 t = *reinterpret_cast<const float4 *>(taps + u);
        a = __fmaf_rn(t.x, s[u], a);
        a = __fmaf_rn(t.y, s[u + 1], a);
        a = __fmaf_rn(t.z, s[u + 2], a);
        a = __fmaf_rn(t.w, s[u + 3], a);
    }
    for (; u < n; ++u) a = __fmaf_rn(taps[u], s[u], a);
    return a;
}
// One output of SymmColumnFilter: centre tap, then the symmetric pairs; h points at the centre row, rows are `stride` apart.
__device__ __forceinline__ float col_chain(const float *__restrict__ taps, const float *__restrict__ h, int r, int stride) {
    float a = __fmul_rn(taps[r], h[0]);
    for (int v = 1; v <= r; ++v) a = __fmaf_rn(taps[r + v], __fadd_rn(h[v * stride], h[-v * stride]), a);
    return a;
}

size_t orient_smem_bytes(int r) {
    const int n = 2 * r + 1, span = W + 2 * r;
    return sizeof(float) * ((size_t)((n + 3) & ~3) + (size_t)W * stage_stride(r) + (size_t)span * W + 256 + 36) + sizeof(int) * (3 * (size_t)span + 256);
}

// The window is Rect(x, y, 16, 16) of the level's maps replicate-padded by 8 (:318-320, :341-343).  The blur of that ROI
// reads the padded parent around it and reflects (101) at the parent's edges, so source row q of the (16 + 2r)^2
// neighbourhood is padded row refl101(y + q - r, rows + 16), i.e. map row clamp(. - 8).
// Resolves a keypoint to its level's Gaussian, map size and tap table; false for points outside the pyramid.
struct DescSite { const u8 *g; const float *taps; int rows, cols, n; size_t pitch; };
__device__ __forceinline__ bool desc_site(const DescPyr &P, const vslam_keypoint &k, int frame, DescSite &s) {
    const int o = P.octave >= 0 ? P.octave : k.octave;
    if (o < 0 || o >= P.num_octaves || k.level < 0 || k.level >= 6) return false;
    s.rows = P.orows[o]; s.cols = P.ocols[o]; s.pitch = P.opitch[o];
    s.n = P.ksize[o][k.level];
    if (s.n <= 0 || k.row < 0 || k.row >= s.rows || k.col < 0 || k.col >= s.cols) return false;
    s.g = P.maps + (size_t)frame * P.frame_bytes + P.ooff[o] + P.oplane[o] * (size_t)(1 + k.level);
    s.taps = P.taps + P.tap_off[o][k.level];
    return true;
}

// order: the keypoints this launch works on (widest kernels first), or nullptr for all of them in list order.
__global__ void __launch_bounds__(256, 6) orient_hist_k(const vslam_keypoint *__restrict__ kp, const int *__restrict__ frame_of,
                                                        const int *__restrict__ order, int n_kp, const __grid_constant__ DescPyr P,
                                                        unsigned long long *__restrict__ masks, float *__restrict__ histos) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int t = threadIdx.x;
    const int kid = order ? order[blockIdx.x] : (int)blockIdx.x;
    const vslam_keypoint k = kp[kid];
    DescSite site;
    if (!desc_site(P, k, frame_of ? frame_of[kid] : P.frame0, site)) {
        if (t == 0) masks[kid] = 0;
        if (histos && t < 36) histos[(size_t)kid * 36 + t] = 0.f;
        return;
    }
    const int rows = site.rows, cols = site.cols;
    const size_t pitch = site.pitch;
    const int n = site.n, r = n / 2, span = W + 2 * r, SP = stage_stride(r), pad = 8;
    const u8 *__restrict__ g = site.g;
    float *taps = (float *)smem_raw;                               // n
    float *S = taps + ((n + 3) & ~3);                              // 16 x SP staged magnitudes
    float *H = S + W * SP;                                         // span x 16 row-filtered
    float *wv = H + span * W;                                      // 256 weighted magnitudes
    float *hs = wv + 256;                                          // 36 bins
    int *cidx = (int *)(hs + 36);                                  // span source columns
    int *cpl = cidx + span, *cmi = cpl + span;                     // their reflect-101 neighbours (Sobel ksize 1)
    int *wb = cmi + span;                                          // 256 bin indices
    for (int i = t; i < n; i += 256) taps[i] = site.taps[i];
    for (int c = t; c < span; c += 256) {
        const int sx = clampi(refl101(k.col + c - r, cols + 2 * pad) - pad, 0, cols - 1);
        cidx[c] = sx; cpl[c] = refl101(sx + 1, cols); cmi[c] = refl101(sx - 1, cols);
    }
    __syncthreads();
    const int q = t >> 4, j = t & 15;
    for (int q0 = 0; q0 < span; q0 += W) {
        // stage 16 source rows of magnitudes: thread (q, j) walks row q in steps of 16 columns, four samples in flight
        if (q0 + q < span) {
            const int sy = clampi(refl101(k.row + q0 + q - r, rows + 2 * pad) - pad, 0, rows - 1);
            const u8 *__restrict__ r0 = g + (size_t)sy * pitch;
            const u8 *__restrict__ rm = g + (size_t)refl101(sy - 1, rows) * pitch;
            const u8 *__restrict__ rp = g + (size_t)refl101(sy + 1, rows) * pitch;
            for (int c0 = j; c0 < span; c0 += 64) {
                int a[4], b[4], u[4], d[4];
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const int c = c0 + 16 * e;
                    const int cc = c < span ? c : span - 1, sx = cidx[cc];
                    a[e] = r0[cpl[cc]]; b[e] = r0[cmi[cc]];
                    d[e] = rp[sx]; u[e] = rm[sx];
                }
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const int c = c0 + 16 * e;
                    const float dx = (float)(a[e] - b[e]), dy = (float)(d[e] - u[e]);
                    if (c < span) S[q * SP + c] = __fsqrt_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)));
                }
            }
        }
        __syncthreads();
        if (q0 + q < span) {
            H[(q0 + q) * W + j] = row_chain(taps, S + q * SP + j, n);
        }
        __syncthreads();
    }
    {   // column pass for window sample (i, j) = (q, j)
        const float a = col_chain(taps, H + (q + r) * W + j, r, W);
        float dx, dy;
        grad_at(g, rows, cols, pitch, clampi(k.row + q - pad, 0, rows - 1), clampi(k.col + j - pad, 0, cols - 1), dx, dy);
        const float coeff = 36.f / 360.0f;                         // :114
        wv[t] = a;
        wb[t] = clampi((int)__fmul_rn(fast_atan2_deg(dy, dx), coeff), 0, 35);     // :125 (orientations are < 359.8: always inside)
    }
    __syncthreads();
    // :128 in the reference's order (rows, then columns): warp 0 walks the eight chunks of 32 samples in order; one match finds
    // the lanes of a chunk that share a bin, and the lowest lane of each group adds the group's values to the bin one by one
    if (t < 32) {
        if (t < 4) hs[32 + t] = 0.f;
        hs[t] = 0.f;
        __syncwarp();
        for (int c = 0; c < 256; c += 32) {
            const int b = wb[c + t];
            const float v = wv[c + t];
            const unsigned grp = __match_any_sync(0xffffffffu, b);
            const bool leader = (__ffs(grp) - 1) == t;
            float a = leader ? hs[b] : 0.f;
            unsigned m = leader ? grp : 0u;
            while (__any_sync(0xffffffffu, m != 0)) {
                const float x = __shfl_sync(0xffffffffu, v, m ? __ffs(m) - 1 : t);
                if (m) { a = __fadd_rn(a, x); m &= m - 1; }
            }
            if (leader) hs[b] = a;
            __syncwarp();
        }
    }
    __syncthreads();
    if (histos && t < 36) histos[(size_t)kid * 36 + t] = hs[t];
    if (t == 0) {
        float peak = hs[0];
        for (int b = 1; b < 36; ++b) peak = hs[b] > peak ? hs[b] : peak;
        const float thr = __fmul_rn(peak, 0.8f);                   // :359
        unsigned long long m = 0;
        for (int b = 0; b < 36; ++b)
            if (hs[b] > thr) m |= 1ull << b;
        masks[kid] = m;
    }
}

size_t sift_smem_bytes(int r) { return sizeof(float) * ((size_t)((2 * r + 4) & ~3) + (size_t)W * stage_stride(r)) + sizeof(int) * (size_t)(W + 2 * r); }

// One oriented keypoint: the 16x16 rotated section of the maps replicate-padded by 20 (:578-580), read the way
// rotateImageSection does -- list index i * 16 + j into the 17-wide point list (:545) and at<>(point.x, point.y), x as
// the row (:549-554).  A sample whose linear address leaves the padded map is undefined in the reference; it is taken
// from the nearest end of the buffer and the descriptor is flagged invalid.
__global__ void __launch_bounds__(256) sift_k(const vslam_keypoint *__restrict__ kp, const int *__restrict__ frame_of, const float2 *__restrict__ cs,
                                              int n_kp, const __grid_constant__ DescPyr P, float *__restrict__ desc, u8 *__restrict__ valid) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ float M[256], Hh[256], Wt[256], D[128], red[4];
    __shared__ int B[256], bad;
    const int t = threadIdx.x, i = t >> 4, j = t & 15;
    const vslam_keypoint k = kp[blockIdx.x];
    float *out = desc + (size_t)blockIdx.x * 128;
    DescSite site;
    if (!desc_site(P, k, frame_of ? frame_of[blockIdx.x] : P.frame0, site)) {
        if (t < 128) out[t] = __int_as_float(0x7fc00000);
        if (t == 0) valid[blockIdx.x] = 0;
        return;
    }
    const int rows = site.rows, cols = site.cols;
    const size_t pitch = site.pitch;
    const int n = site.n, r = n / 2, pad = 20;
    const u8 *__restrict__ g = site.g;
    const int span = W + 2 * r, SP = stage_stride(r);
    float *taps = (float *)smem_raw;
    float *E = taps + ((n + 3) & ~3);                              // the window extended by reflection: 16 x SP, rows then columns
    int *ridx = (int *)(E + W * SP);                               // ridx[c] = refl101(c - r, 16), c in [0, span)
    for (int c = t; c < n; c += 256) taps[c] = site.taps[c];
    for (int c = t; c < span; c += 256) ridx[c] = refl101(c - r, W);
    if (t == 0) bad = 0;
    __syncthreads();
    {
        const int PR = rows + 2 * pad, PC = cols + 2 * pad, cx = k.col + pad, cy = k.row + pad;
        const int p = i * W + j, pi = p / (W + 1), pj = p - pi * (W + 1);
        const float c = cs[blockIdx.x].x, s = cs[blockIdx.x].y;
        const int rx = pj - 8, ry = pi - 8;                        // rotate_pt_CW (rotation.cpp:19-28)
        const int xr = (int)__fsub_rn(__fmul_rn((float)rx, c), __fmul_rn((float)ry, s)) + cx;
        const int yr = (int)__fadd_rn(__fmul_rn((float)rx, s), __fmul_rn((float)ry, c)) + cy;
        long long lin = (long long)xr * PC + yr;
        const long long total = (long long)PR * PC;
        if (lin < 0 || lin >= total) { bad = 1; lin = lin < 0 ? 0 : total - 1; }
        const int sy = clampi((int)(lin / PC) - pad, 0, rows - 1), sx = clampi((int)(lin % PC) - pad, 0, cols - 1);
        float dx, dy;
        grad_at(g, rows, cols, pitch, sy, sx, dx, dy);
        M[t] = __fsqrt_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)));
        B[t] = (int)__fmul_rn(fast_atan2_deg(dy, dx), 8.f / 360.0f);
    }
    __syncthreads();
    // GaussianBlur of the stand-alone 16x16 Mat (:611): reflect-101 on the window itself.  E[i][c] = M[i][refl101(c - r)]
    for (int c = j; c < span; c += W) E[i * SP + c] = M[i * W + ridx[c]];
    __syncthreads();
    Hh[t] = row_chain(taps, E + i * SP + j, n);
    __syncthreads();
    for (int c = j; c < span; c += W) E[i * SP + c] = Hh[ridx[c] * W + i];   // E[x][c] = Hh[refl101(c - r)][x]: columns laid out as rows
    __syncthreads();
    Wt[j * W + i] = col_chain(taps, E + i * SP + j + r, r, 1);   // this thread: window row j, column i (lanes walk a column)
    __syncthreads();
    if (t < 128) {                                                 // :630-646: cell t / 8 (columns first), bin t % 8
        const int cell = t >> 3, b = t & 7, r0 = 4 * (cell >> 2), c0 = 4 * (cell & 3);
        float a = 0.f;
        for (int y = r0; y < r0 + 4; ++y)
            for (int x = c0; x < c0 + 4; ++x)
                if (B[y * W + x] == b) a = __fadd_rn(a, Wt[y * W + x]);
        D[t] = a;
    }
    for (int pass = 0; pass < 2; ++pass) {                         // :652-669: / max, min(., 0.2), / max
        __syncthreads();
        if (t < 128) {                                             // max_element: NaN never replaces the running maximum
            float v = D[t];
            for (int o = 16; o; o >>= 1) { const float w = __shfl_xor_sync(0xffffffffu, v, o); v = w > v ? w : v; }
            if ((t & 31) == 0) red[t >> 5] = v;
        }
        __syncthreads();
        if (t < 128) {
            float peak = red[0];
            for (int w = 1; w < 4; ++w) peak = red[w] > peak ? red[w] : peak;
            float v = __fdiv_rn(D[t], peak);
            if (pass == 0) v = 0.2f < v ? 0.2f : v;
            D[t] = v;
        }
    }
    if (t < 128) out[t] = D[t];
    if (t == 0) valid[blockIdx.x] = bad ? 0 : 1;
}

int launch_orient_hist(vslam_ctx *ctx, const vslam_keypoint *kp, const int *frame_of, const int *order, int n, const DescPyr &P, int max_r,
                       unsigned long long *masks, float *histos) {
    const size_t smem = orient_smem_bytes(max_r);
    if (smem > 227 * 1024) VS_FAIL(ctx, VSLAM_E_INVALID, "orientation window blur needs %zu bytes of shared memory (sigma too large)", smem);
    VS_CUDA(ctx, cudaFuncSetAttribute(orient_hist_k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    KernelSpan span(ctx, VSLAM_K_DESCRIPTOR);
    orient_hist_k<<<n, 256, smem, ctx->stream>>>(kp, frame_of, order, n, P, masks, histos);
    VS_LAUNCH_CHECK(ctx);
    return VSLAM_OK;
}

int launch_sift(vslam_ctx *ctx, const vslam_keypoint *kp, const int *frame_of, const float2 *cs, int n, const DescPyr &P, int max_r, float *desc,
                u8 *valid) {
    const size_t smem = sift_smem_bytes(max_r);
    if (smem > 200 * 1024) VS_FAIL(ctx, VSLAM_E_INVALID, "descriptor window blur needs %zu bytes of shared memory (sigma too large)", smem);
    VS_CUDA(ctx, cudaFuncSetAttribute(sift_k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    KernelSpan span(ctx, VSLAM_K_DESCRIPTOR);
    sift_k<<<n, 256, smem, ctx->stream>>>(kp, frame_of, cs, n, P, desc, valid);
    VS_LAUNCH_CHECK(ctx);
    return VSLAM_OK;
}

// ---- filterKeypoints :362-368 on the device: one point per histogram bin above 0.8 of the peak, bins in ascending order ----------
__global__ void __launch_bounds__(256) orient_count_k(const unsigned long long *__restrict__ masks, int n, int *__restrict__ cnt) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) cnt[i] = __popcll(masks[i] & ((1ull << 36) - 1));
}
__global__ void __launch_bounds__(256) orient_expand_k(const vslam_keypoint *__restrict__ kp, const int *__restrict__ frame_of,
                                                       const unsigned long long *__restrict__ masks, const int *__restrict__ pos, int n, int capacity,
                                                       const OrientLut lut, vslam_keypoint *__restrict__ out, float2 *__restrict__ cs,
                                                       int *__restrict__ oframe) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const vslam_keypoint k = kp[i];
    int m = pos[i];
    for (unsigned long long bits = masks[i] & ((1ull << 36) - 1); bits && m < capacity; bits &= bits - 1, ++m) {
        const int bin = __ffsll((long long)bits) - 1;
        vslam_keypoint o;
        o.row = k.row; o.col = k.col; o.value = bin * 10; o.padding = 0; o.octave = k.octave; o.level = k.level;
        out[m] = o;
        cs[m] = lut.cs[bin];
        oframe[m] = frame_of[i];
    }
}

int launch_orient_expand(vslam_ctx *ctx, const vslam_keypoint *kp, const int *frame_of, const unsigned long long *masks, int n, int capacity,
                         const OrientLut &lut, int *cnt, int *pos, int *partial, int *d_total, vslam_keypoint *out, float2 *cs, int *oframe,
                         int phase) {
    KernelSpan span(ctx, VSLAM_K_DESCRIPTOR);
    if (phase == 0) {                                             // counts and their exclusive scan; *d_total = number of oriented points
        orient_count_k<<<(n + 255) / 256, 256, 0, ctx->stream>>>(masks, n, cnt);
        VS_LAUNCH_CHECK(ctx);
        return device_exclusive_scan(ctx, cnt, n, partial, pos, d_total);
    }
    orient_expand_k<<<(n + 255) / 256, 256, 0, ctx->stream>>>(kp, frame_of, masks, pos, n, capacity, lut, out, cs, oframe);
    VS_LAUNCH_CHECK(ctx);
    return VSLAM_OK;
}

}  // namespace vslam
