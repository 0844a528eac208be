// compaction.cu -- segment-ordered stream compaction shared by the DoG detector and the Harris NMS2 stage.
//
// The reference emits keypoints in loop order ((level, i, j) per octave, Diff_of_Gauss.cpp:264-268; row-major
// for NMS2, Harris_corners.cpp:94-95).  Producer blocks compact their survivors in that order internally,
// reserve space in a staging list with ONE atomicAdd per block (block-aggregated atomics), and record
// (offset, count) per *segment* -- a unit whose index order equals the reference's emission order.  The two
// kernels below turn that into the final ordered list: an exclusive scan over segment counts and a gather.
// Cost is O(#segments + #entries); nothing is sorted.
#include "common.cuh"

namespace vslam {

// Three-phase exclusive scan over the segment counts (n can be tens of millions for a batch of frames):
//   seg_reduce_k  one partial sum per block of SCAN_ITEMS counts
//   seg_spine_k   single block: exclusive scan of the partials, seeded with the running total *d_total
//   seg_apply_k   every block re-scans its counts from its seed -> dst[]; per-frame counts from the frame boundaries
constexpr int SCAN_T = 256, SCAN_V = 16, SCAN_ITEMS = SCAN_T * SCAN_V;

__device__ __forceinline__ int block_exclusive_scan(int v, int *s_warp, int &total) {   // SCAN_T threads
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    int incl = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += t;
    }
    if (lane == 31) s_warp[wid] = incl;
    __syncthreads();
    if (wid == 0) {
        int w = lane < SCAN_T / 32 ? s_warp[lane] : 0;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, w, d);
            if (lane >= d) w += t;
        }
        s_warp[lane] = w;
    }
    __syncthreads();
    total = s_warp[SCAN_T / 32 - 1];
    const int r = (wid ? s_warp[wid - 1] : 0) + incl - v;
    __syncthreads();
    return r;
}

__global__ void __launch_bounds__(SCAN_T) seg_reduce_k(const int *__restrict__ cnt, long long n, int *__restrict__ partial) {
    __shared__ int s_warp[32];
    const long long base = (long long)blockIdx.x * SCAN_ITEMS + (long long)threadIdx.x * SCAN_V;
    int s = 0;
    if (base + SCAN_V <= n) {
        const int4 *p = (const int4 *)(cnt + base);
#pragma unroll
        for (int i = 0; i < SCAN_V / 4; ++i) { const int4 v = p[i]; s += v.x + v.y + v.z + v.w; }
    } else {
        for (int i = 0; i < SCAN_V; ++i) if (base + i < n) s += cnt[base + i];
    }
    int total;
    block_exclusive_scan(s, s_warp, total);
    if (threadIdx.x == 0) partial[blockIdx.x] = total;
}

__global__ void __launch_bounds__(SCAN_T) seg_spine_k(int *__restrict__ partial, int nb, int *__restrict__ d_total) {
    __shared__ int s_warp[32];
    int run = *d_total;
    for (int b0 = 0; b0 < nb; b0 += SCAN_T) {
        const int i = b0 + threadIdx.x;
        const int v = i < nb ? partial[i] : 0;
        int total;
        const int ex = block_exclusive_scan(v, s_warp, total);
        if (i < nb) partial[i] = run + ex;
        run += total;
    }
    __syncthreads();
    if (threadIdx.x == 0) *d_total = run;
}

__global__ void __launch_bounds__(SCAN_T) seg_apply_k(const int *__restrict__ cnt, long long n, const int *__restrict__ partial,
                                                      int *__restrict__ dst) {
    __shared__ int s_warp[32];
    const long long base = (long long)blockIdx.x * SCAN_ITEMS + (long long)threadIdx.x * SCAN_V;
    int v[SCAN_V];
    int s = 0;
    if (base + SCAN_V <= n) {
        const int4 *p = (const int4 *)(cnt + base);
#pragma unroll
        for (int i = 0; i < SCAN_V / 4; ++i) { const int4 q = p[i]; v[4 * i] = q.x; v[4 * i + 1] = q.y; v[4 * i + 2] = q.z; v[4 * i + 3] = q.w; }
    } else {
#pragma unroll
        for (int i = 0; i < SCAN_V; ++i) v[i] = base + i < n ? cnt[base + i] : 0;
    }
#pragma unroll
    for (int i = 0; i < SCAN_V; ++i) s += v[i];
    int total;
    int run = partial[blockIdx.x] + block_exclusive_scan(s, s_warp, total);
    if (base + SCAN_V <= n) {
        int4 *o = (int4 *)(dst + base);
#pragma unroll
        for (int i = 0; i < SCAN_V / 4; ++i) {
            int4 q;
            q.x = run; run += v[4 * i]; q.y = run; run += v[4 * i + 1]; q.z = run; run += v[4 * i + 2]; q.w = run; run += v[4 * i + 3];
            o[i] = q;
        }
    } else {
#pragma unroll
        for (int i = 0; i < SCAN_V; ++i) { if (base + i < n) dst[base + i] = run; run += v[i]; }
    }
}

// counts of frames [0, n_frames): difference of the scanned positions at the frame boundaries
__global__ void frame_counts_k(const int *__restrict__ dst, long long segs_per_frame, int n_frames, const int *__restrict__ d_total,
                               int *__restrict__ frame_count) {
    const int f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= n_frames) return;
    const int lo = dst[(long long)f * segs_per_frame];
    const int hi = f + 1 < n_frames ? dst[(long long)(f + 1) * segs_per_frame] : *d_total;
    frame_count[f] = hi - lo;
}

__global__ void __launch_bounds__(256) seg_gather_k(const int *__restrict__ off, const int *__restrict__ cnt,
                                                    const int *__restrict__ dst, long long n, int words,
                                                    const u32 *__restrict__ staging, int staging_cap,
                                                    u32 *__restrict__ out, int capacity, int compact) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int c = cnt[i];
    if (!c) return;
    const int o = off[i], d = dst[i];
    if (compact) {                                       // vslam_corner {row, col, value} -> vslam_corner8 {u16 row, u16 col, value}
        for (int e = 0; e < c; ++e) {
            if (d + e >= capacity) break;
            uint2 t = make_uint2(0xffffffffu, 0u);       // no staging slot: row = col = 0xffff
            if (o + e < staging_cap) {
                const u32 *s = staging + (size_t)(o + e) * 3;
                t = make_uint2((s[0] & 0xffffu) | s[1] << 16, s[2]);
            }
            *(uint2 *)(out + (size_t)(d + e) * 2) = t;
        }
        return;
    }
    for (int e = 0; e < c; ++e) {
        if (d + e >= capacity) break;
        u32 *t = out + (size_t)(d + e) * words;
        if (o + e >= staging_cap) {                      // the producer had no staging slot left for this entry: mark it (row = -1)
            t[0] = 0xffffffffu;
            for (int w = 1; w < words; ++w) t[w] = 0;
            continue;
        }
        const u32 *s = staging + (size_t)(o + e) * words;
        for (int w = 0; w < words; ++w) t[w] = s[w];
    }
}

int seg_prepare(vslam_ctx *ctx, size_t n_segments, size_t entry_bytes, size_t entries) {
    const size_t staging_bytes = entry_bytes * entries;
    ctx->staging_entries = (int)entries;
    if (ctx->seg_off.reserve(n_segments * 4) || ctx->seg_cnt.reserve(n_segments * 4) ||
        ctx->seg_dst.reserve(n_segments * 4) || ctx->staging.reserve(staging_bytes))
        VS_FAIL(ctx, VSLAM_E_NOMEM, "segment workspace allocation failed (%zu segments)", n_segments);
    // producers write every segment they own; zero once so that never-launched (empty) grids stay empty
    VS_CUDA(ctx, cudaMemsetAsync(ctx->seg_cnt.p, 0, n_segments * 4, ctx->stream));
    return VSLAM_OK;
}

int seg_finalise(vslam_ctx *ctx, size_t n_segments, int entry_bytes, void *dst, int capacity, int *d_total,
                 int *d_frame_count, int n_frames, int compact) {
    if (compact && entry_bytes != (int)sizeof(vslam_corner)) VS_FAIL(ctx, VSLAM_E_INVALID, "compact lists exist for corners only");
    const long long n = (long long)n_segments;
    const int nb = (int)((n + SCAN_ITEMS - 1) / SCAN_ITEMS);
    if (ctx->seg_partial.reserve((size_t)nb * 4 + 16)) VS_FAIL(ctx, VSLAM_E_NOMEM, "scan partials");
    int *partial = ctx->seg_partial.as<int>();
    KernelSpan span(ctx, VSLAM_K_COMPACT);
    seg_reduce_k<<<nb, SCAN_T, 0, ctx->stream>>>(ctx->seg_cnt.as<int>(), n, partial);
    VS_LAUNCH_CHECK(ctx);
    seg_spine_k<<<1, SCAN_T, 0, ctx->stream>>>(partial, nb, d_total);
    VS_LAUNCH_CHECK(ctx);
    seg_apply_k<<<nb, SCAN_T, 0, ctx->stream>>>(ctx->seg_cnt.as<int>(), n, partial, ctx->seg_dst.as<int>());
    VS_LAUNCH_CHECK(ctx);
    if (d_frame_count) {
        frame_counts_k<<<(n_frames + 127) / 128, 128, 0, ctx->stream>>>(ctx->seg_dst.as<int>(), n / n_frames, n_frames, d_total,
                                                                         d_frame_count);
        VS_LAUNCH_CHECK(ctx);
    }
    seg_gather_k<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(ctx->seg_off.as<int>(), ctx->seg_cnt.as<int>(),
                                                                        ctx->seg_dst.as<int>(), n, entry_bytes / 4,
                                                                        ctx->staging.as<u32>(), ctx->staging_entries, (u32 *)dst,
                                                                        capacity, compact);
    VS_LAUNCH_CHECK(ctx);
    return VSLAM_OK;
}

// exclusive scan of n ints on the device (the three kernels above): dst[i] = sum of cnt[0 .. i-1], *d_total = the sum.
// partial: at least ceil(n / 4096) + 4 ints of workspace; cnt / dst 16-byte aligned.
int device_exclusive_scan(vslam_ctx *ctx, const int *cnt, long long n, int *partial, int *dst, int *d_total) {
    const int nb = (int)((n + SCAN_ITEMS - 1) / SCAN_ITEMS);
    VS_CUDA(ctx, cudaMemsetAsync(d_total, 0, sizeof(int), ctx->stream));
    seg_reduce_k<<<nb, SCAN_T, 0, ctx->stream>>>(cnt, n, partial);
    VS_LAUNCH_CHECK(ctx);
    seg_spine_k<<<1, SCAN_T, 0, ctx->stream>>>(partial, nb, d_total);
    VS_LAUNCH_CHECK(ctx);
    seg_apply_k<<<nb, SCAN_T, 0, ctx->stream>>>(cnt, n, partial, dst);
    VS_LAUNCH_CHECK(ctx);
    return VSLAM_OK;
}

}  // namespace vslam
