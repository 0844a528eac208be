// detect.cu -- scale-space extrema + contrast + edge tests.
//
//   initialKeypointDetection   Diff_of_Gauss.cpp:254-297   stride-`window` grid, half-open 2x2x3 window, ties kept
//   FeaturePointLocalization   Diff_of_Gauss.cpp:223-251   exact IEEE sequence of OpenCV's gemm / invert(3x3)
//   computeEdgeResponse + threshold  :79-109, :326-338     2x2 structure sums on the level's gradients, tr^2/det < 12.1
//
// One segment = 256 consecutive grid samples of one (frame, level, sample row): survivors are compacted in j order with
// ballots, one atomicAdd per segment reserves staging space, compaction.cu orders the segments.  A block walks
// kDetectRows sample rows (one segment each); one launch covers every frame of a batch (grid.z = level x frame).
// Every float/double operation below uses the round-to-nearest intrinsics so that nvcc cannot contract to FMA.
#include "detect.cuh"
#include "detect_math.cuh"

namespace vslam {

// computeEdgeResponse on the gradients of Gaussian level g (Sobel ksize=1 of row E formed on the fly)
__device__ float edge_response(const u8 *__restrict__ g, int rows, int cols, size_t pitch, int y, int x, int pad) {
    float ix2 = 0.f, iy2 = 0.f, ixy = 0.f;
    // window 3 never leaves the map; wider windows do in the reference (padded coordinates on an unpadded Mat: undefined
    // behaviour) -- the defined semantics here: samples outside the map repeat the border sample
    for (int uu = y - pad; uu < y + pad; ++uu)
        for (int vv = x - pad; vv < x + pad; ++vv) {
            const int u = clampi(uu, 0, rows - 1), v = clampi(vv, 0, cols - 1);
            const u8 *row = g + (size_t)u * pitch;
            const float a = (float)((int)row[refl101(v + 1, cols)] - (int)row[refl101(v - 1, cols)]);
            const float b = (float)((int)g[(size_t)refl101(u + 1, rows) * pitch + v] - (int)g[(size_t)refl101(u - 1, rows) * pitch + v]);
            ix2 = __fadd_rn(ix2, __fmul_rn(a, a));
            iy2 = __fadd_rn(iy2, __fmul_rn(b, b));
            ixy = __fadd_rn(ixy, __fmul_rn(a, b));
        }
    const float det = __double2float_rn(__dsub_rn(__dmul_rn((double)ix2, (double)iy2), __dmul_rn((double)ixy, (double)ixy)));
    const float tr = __fadd_rn(ix2, iy2);
    return __fdiv_rn(__fmul_rn(tr, tr), det);
}

constexpr int kDetectRows = 4;              // sample rows per block

__global__ void __launch_bounds__(256) dog_detect_k(DetectArgs a) {
    __shared__ int s_wcnt[2][8];
    __shared__ int s_base[2];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int level = blockIdx.z % 3 + 1, frame = blockIdx.z / 3;
    const int pad = a.pad, win = a.window;
    const int jj = blockIdx.x * 256 + tid;
    const int j = pad + jj * win;
    const int rows = a.rows, cols = a.cols;
    const size_t pitch = a.pitch, fo = (size_t)frame * a.frame_stride;
    const bool valid = j < cols;
    const u8 *d0 = a.dog[level - 1] + fo, *d1 = a.dog[level] + fo, *d2 = a.dog[level + 1] + fo;
    const int si_end = min(a.ni, (int)(blockIdx.y + 1) * kDetectRows);
    for (int si = blockIdx.y * kDetectRows; si < si_end; ++si) {
    const int i = pad + si * win;           // i < rows by construction of the grid
    const int par = si & 1;                 // the shared slots alternate: one barrier less per row

    bool cand = false, pass_c = false, pass_e = false;
    int value = 0, c = 0, dx = 0, dy = 0, ds = 0;
    // padded (u,v) <-> original (clamp(u-pad), clamp(v-pad))   (padOctave = replicate, Diff_of_Gauss.cpp:260)
#define PAD_AT(m, u, v) ((int)(m)[(size_t)clampi((u) - pad, 0, rows - 1) * pitch + clampi((v) - pad, 0, cols - 1)])
    if (valid && pad == 1 && i >= 2 && i < rows && j >= 2 && j < cols) {
        // window 3 away from the border (nearly every sample): the same reads without the clamps.  Centre = original (i - 1, j - 1);
        // window rows i - 2, i - 1, columns j - 2, j - 1
        const size_t o = (size_t)(i - 1) * pitch + (size_t)(j - 1);
        const u8 *p0 = d0 + o, *p1 = d1 + o, *p2 = d2 + o, *q0 = p0 - pitch, *q1 = p1 - pitch, *q2 = p2 - pitch;
        const int a0 = q0[-1], b0 = q0[0], e0 = p0[-1], f0 = p0[0];
        const int a1 = q1[-1], b1 = q1[0], e1 = p1[-1], f1 = p1[0];
        const int a2 = q2[-1], b2 = q2[0], e2 = p2[-1], f2 = p2[0];
        c = f1;
        const int mn = min(min(min(a0, b0), min(e0, f0)), min(min(min(a1, b1), min(e1, f1)), min(min(a2, b2), min(e2, f2))));
        const int mx = max(max(max(a0, b0), max(e0, f0)), max(max(max(a1, b1), max(e1, f1)), max(max(a2, b2), max(e2, f2))));
        cand = (c == mn) || (c == mx);
        if (cand) {
            dx = e1 - (int)p1[1];
            dy = b1 - (int)p1[pitch];
            ds = f0 - f2;
        }
    } else if (valid && pad > 0) {
        c = PAD_AT(d1, i, j);
        int mn = c, mx = c;
        for (int u = i - pad; u < i + pad; ++u)
            for (int v = j - pad; v < j + pad; ++v) {
                const int t0 = PAD_AT(d0, u, v), t1 = PAD_AT(d1, u, v), t2 = PAD_AT(d2, u, v);
                mn = min(mn, min(t0, min(t1, t2)));
                mx = max(mx, max(t0, max(t1, t2)));
            }
        cand = (c == mn) || (c == mx);
        if (cand) {
            dx = PAD_AT(d1, i, j - 1) - PAD_AT(d1, i, j + 1);
            dy = PAD_AT(d1, i - 1, j) - PAD_AT(d1, i + 1, j);
            ds = PAD_AT(d0, i, j) - PAD_AT(d2, i, j);
        }
    }
#undef PAD_AT
    if (cand) {
        const float zhat = localize(dx, dy, ds, c);
        pass_c = zhat > a.contrast_thr;
        if (pass_c) {
            value = (int)__fmul_rn(zhat, 255.0f);                       // Diff_of_Gauss.cpp:246
            pass_e = true;
            if (a.apply_edge) {
                const float resp = edge_response(a.gauss[level] + fo, rows, cols, pitch, i, j, pad);
                pass_e = resp < a.edge_thr;                              // :336 (NaN and +inf fail)
            }
        }
    }
    // ordered compaction inside the block
    const unsigned bal = __ballot_sync(0xffffffffu, pass_e);
    int n_cand = 0, n_con = 0;
    if (a.stage_counts) {                   // (block-uniform)
        n_cand = __syncthreads_count(cand);
        n_con = __syncthreads_count(pass_c);
    }
    if (lane == 0) s_wcnt[par][wid] = __popc(bal);
    __syncthreads();
    int before = 0, total = 0;
#pragma unroll
    for (int w = 0; w < 8; ++w) {
        const int c = s_wcnt[par][w];
        if (w < wid) before += c;
        total += c;
    }
    const int seg = frame * a.segs_per_frame + a.seg_base + ((level - 1) * a.ni + si) * a.nchunks + blockIdx.x;
    if (tid == 0) {
        int off = 0;
        if (total) off = atomicAdd(a.cursor, total);
        s_base[par] = off;
        a.seg_off[seg] = off;
        a.seg_cnt[seg] = total;
        if (a.stage_counts) {
            unsigned long long *sc = a.stage_counts + (size_t)frame * a.stage_stride;
            const int n_valid = min(256, a.nj - blockIdx.x * 256);
            atomicAdd(&sc[0], (unsigned long long)n_valid);
            if (n_cand) atomicAdd(&sc[1], (unsigned long long)n_cand);
            if (n_con) atomicAdd(&sc[2], (unsigned long long)n_con);
            if (total) atomicAdd(&sc[3], (unsigned long long)total);
        }
    }
    if (total) {                            // (block-uniform: every thread summed the same counts)
        __syncthreads();
        if (pass_e) {
            const int pos = s_base[par] + before + __popc(bal & ((1u << lane) - 1u));
            if (pos < a.staging_cap) {
                vslam_keypoint k;
                k.row = i; k.col = j; k.value = value; k.padding = pad; k.octave = a.octave; k.level = level;
                a.staging[pos] = k;
            }
        }
    }
    }
}

int launch_detect(vslam_ctx *ctx, DetectArgs &a, const vslam_dog_params &p) {
    a.window = p.window;
    a.pad = (p.window - 1) / 2;
    a.contrast_thr = p.contrast_thr;
    const float rp1 = p.edge_r + 1.0f;
    a.edge_thr = (rp1 * rp1) / p.edge_r;                                    // Diff_of_Gauss.cpp:332-333
    if (a.n_frames < 1) a.n_frames = 1;
    if (a.n_frames > 21845) VS_FAIL(ctx, VSLAM_E_INVALID, "detector: %d frames per launch", a.n_frames);
    dim3 g(a.nchunks, (a.ni + kDetectRows - 1) / kDetectRows, 3 * a.n_frames);
    KernelSpan span(ctx, VSLAM_K_DETECT);
    dog_detect_k<<<g, 256, 0, ctx->stream>>>(a);
    VS_LAUNCH_CHECK(ctx);
    return VSLAM_OK;
}

void detect_grid(int rows, int cols, int window, int *ni, int *nj, int *nchunks) {
    const int pad = (window - 1) / 2;
    *ni = rows > pad ? (rows - pad + window - 1) / window : 0;
    *nj = cols > pad ? (cols - pad + window - 1) / window : 0;
    *nchunks = (*nj + 255) / 256;
}

}  // namespace vslam
