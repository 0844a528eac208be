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

// counters layout (ints): [0] staging cursor, [1] staging capacity (entries)
__global__ void __launch_bounds__(1024) seg_scan_k(const int *__restrict__ cnt, int *__restrict__ dst, int n,
                                                   int *__restrict__ d_total, int *__restrict__ d_frame_count) {
    __shared__ int s_warp[32];
    __shared__ int s_run;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    if (tid == 0) s_run = *d_total;
    __syncthreads();
    const int start = s_run;
    for (int base = 0; base < n; base += 1024) {
        const int i = base + tid;
        const int v = i < n ? cnt[i] : 0;
        int incl = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl += t;
        }
        if (lane == 31) s_warp[wid] = incl;
        __syncthreads();
        if (wid == 0) {
            int w = s_warp[lane];
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const int t = __shfl_up_sync(0xffffffffu, w, d);
                if (lane >= d) w += t;
            }
            s_warp[lane] = w;
        }
        __syncthreads();
        const int run = s_run;
        const int excl = run + (wid ? s_warp[wid - 1] : 0) + incl - v;
        if (i < n) dst[i] = excl;
        __syncthreads();
        if (tid == 1023) s_run = run + s_warp[31];
        __syncthreads();
    }
    if (tid == 0) {
        *d_total = s_run;
        if (d_frame_count) *d_frame_count = s_run - start;
    }
}

__global__ void __launch_bounds__(256) seg_gather_k(const int *__restrict__ off, const int *__restrict__ cnt,
                                                    const int *__restrict__ dst, int n, int words,
                                                    const u32 *__restrict__ staging, int staging_cap,
                                                    u32 *__restrict__ out, int capacity) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int c = cnt[i];
    if (!c) return;
    const int o = off[i], d = dst[i];
    for (int e = 0; e < c; ++e) {
        if (o + e >= staging_cap || d + e >= capacity) break;
        const u32 *s = staging + (size_t)(o + e) * words;
        u32 *t = out + (size_t)(d + e) * words;
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
                 int *d_frame_count) {
    const int n = (int)n_segments;
    KernelSpan span(ctx, VSLAM_K_COMPACT);
    seg_scan_k<<<1, 1024, 0, ctx->stream>>>(ctx->seg_cnt.as<int>(), ctx->seg_dst.as<int>(), n, d_total, d_frame_count);
    VS_LAUNCH_CHECK(ctx);
    seg_gather_k<<<(n + 255) / 256, 256, 0, ctx->stream>>>(ctx->seg_off.as<int>(), ctx->seg_cnt.as<int>(),
                                                           ctx->seg_dst.as<int>(), n, entry_bytes / 4,
                                                           ctx->staging.as<u32>(), ctx->staging_entries, (u32 *)dst, capacity);
    VS_LAUNCH_CHECK(ctx);
    return VSLAM_OK;
}

}  // namespace vslam
