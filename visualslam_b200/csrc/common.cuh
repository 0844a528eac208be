// common.cuh -- context, error plumbing and small device helpers shared by every kernel file.
// Product code: never includes or links anything from oracle/.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string>
#include <vector>

#include "../../include/vslam.h"

typedef uint8_t u8;
typedef uint16_t u16;
typedef uint32_t u32;

namespace vslam {

void set_global_error(const std::string &s);

// A grow-only device buffer owned by a context / pyramid.
struct DevBuf {
    void *p = nullptr;
    size_t cap = 0;
    int reserve(size_t bytes) {
        if (bytes <= cap) return VSLAM_OK;
        if (p) cudaFree(p);
        p = nullptr; cap = 0;
        size_t want = bytes + bytes / 8 + 256;
        cudaError_t e = cudaMalloc(&p, want);
        if (e != cudaSuccess) { cudaGetLastError(); p = nullptr; return VSLAM_E_NOMEM; }
        cap = want;
        return VSLAM_OK;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
    template <class T> T *as() const { return (T *)p; }
};

}  // namespace vslam

struct vslam_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    std::string err;
    long long launches = 0;
    int sm_count = 148;
    // segment-ordered compaction workspace (detect + harris corners)
    vslam::DevBuf seg_off, seg_cnt, seg_dst, seg_partial, staging, counters, stage_counts;
    int staging_entries = 0;       // capacity of `staging` in entries (per frame)
    // second set of the list workspace: vslam_dog_harris keeps two ordered lists growing at once (swap_list_sets, capi.cu)
    vslam::DevBuf alt_seg_off, alt_seg_cnt, alt_seg_dst, alt_seg_partial, alt_staging, alt_counters, alt_out_dev;
    int alt_staging_entries = 0;
    // deferred completion (vslam_dog_harris with VSLAM_DEFER, vslam_finish): two calls may be in flight; each keeps its device
    // lists, an event behind its last kernel and a pinned snapshot of its per-frame counts until vslam_finish copies the lists out
    struct Deferred {
        bool live = false;
        cudaEvent_t ev = nullptr;
        void *kp_host = nullptr, *co_host = nullptr;
        int *kp_counts = nullptr, *co_counts = nullptr;
        int kp_cap = 0, co_cap = 0, n_frames = 0;
        size_t co_entry = 0;
    } defer[2];
    int defer_next = 0;
    vslam::DevBuf defer_kp[2], defer_co[2];
    // JPEG ingest (jpeg.cu): nvJPEG handle + decoder state (opaque here), device staging of the decoded BGR frames
    void *jpeg_handle = nullptr, *jpeg_state = nullptr;
    vslam::DevBuf jpeg_bgr;
    // throughput-path workspace: one frame's pyramid, reused
    vslam::DevBuf ingest;          // aligned copy of the input frame(s)
    vslam::DevBuf ws_maps;         // octave bases of a chunk of frames (or all planes of one frame)
    vslam::DevBuf out_dev;         // device output list when the caller's list is on the host
    vslam::DevBuf maps_dev;        // device response/r8/mask when the caller's maps are on the host
    // host frames are uploaded in chunks on a second stream; compute on chunk c waits for up_ev[c] only, so the upload
    // of later chunks overlaps the kernels of earlier ones
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t up_start = nullptr;
    std::vector<cudaEvent_t> up_ev;    // chunk events of ingest buffer 0 (`ingest`)
    int up_chunk = 0, up_chunks = 0;   // the batch the CURRENT call consumes: frames per upload chunk, chunks that may still be in flight
    int up_total = 0, up_foff = 0;     //   (0: frames were already on the device); chunks of that upload; first frame the call looks at
    size_t up_ds = 0;                  //   bytes between uploaded frames
    int up_slot = 0;                   //   ingest buffer of that batch
    // Two ingest buffers, used alternately: a caller may start the upload of batch i+1 (vslam_upload_frames) before it calls the
    // detectors on batch i, so the copy runs under batch i's kernels.  An upload into a buffer waits for the compute-stream work
    // enqueued so far (the consumers of the batch that buffer held).
    vslam::DevBuf ingest2;
    std::vector<cudaEvent_t> up_ev2;
    int slot_chunk[2] = {0, 0}, slot_total[2] = {0, 0};
    size_t slot_ds[2] = {0, 0};
    int ingest_next = 0;
    // an upload into a buffer has to wait for the readers of the batch it held, not for everything on the compute stream: every
    // detector call records slot_done[buffer] behind its last kernel (consume_end, capi.cu).  slot_busy: a call has started reading the
    // buffer and has not recorded yet (an entry point without consume_end, or an error path): the upload then waits conservatively.
    cudaEvent_t slot_done[2] = {nullptr, nullptr};
    bool slot_done_set[2] = {false, false}, slot_busy[2] = {false, false};
    int consuming = -1;
    // host result lists leave chunk by chunk on a third stream while later chunks compute (capi.cu: ListDrain)
    cudaStream_t down_stream = nullptr;
    std::vector<cudaEvent_t> down_ev;
    void *pinned = nullptr;        // small pinned host scratch for counters
    size_t pinned_cap = 0;
    // cached tap tables of the fused octave kernel (passed by value as kernel parameters)
    int taps_octaves = 0; double taps_sigma0 = 0; int taps_scale_samples = 0;
    void *h_taps = nullptr;        // FusedTaps[taps_octaves], malloc'd
    bool wide_octave[16] = {false};   // octaves whose kernels the fused kernels do not take: the exact global-memory path (pyramid.cu)
    bool any_wide = false;
    vslam::DevBuf grad_lut;        // magnitude / first-quadrant angle of every (|dx|, |dy|) (pyramid.cu: grad_all_k)
    vslam::DevBuf wide_ws;         // its scratch: the taps of an octave and the 8.8 horizontal sums
    // optional per-kernel CUDA-event timing (vslam_profile): events bracket every launch on ctx->stream
    bool profiling = false;
    std::vector<cudaEvent_t> ev_pool;
    size_t ev_used = 0;
    struct Span { int id; cudaEvent_t a, b; };
    std::vector<Span> spans;
    double prof_ms[VSLAM_KERNEL_COUNT] = {0};
    long long prof_n[VSLAM_KERNEL_COUNT] = {0};
};

namespace vslam {
// RAII bracket around one kernel launch; free when profiling is off.
struct KernelSpan {
    vslam_ctx *c; int id; cudaEvent_t stop = nullptr;
    KernelSpan(vslam_ctx *ctx, int kid) : c(ctx), id(kid) {
        if (!c->profiling) return;
        cudaEvent_t ev[2];
        for (int i = 0; i < 2; ++i) {
            if (c->ev_used == c->ev_pool.size()) { cudaEvent_t e; cudaEventCreate(&e); c->ev_pool.push_back(e); }
            ev[i] = c->ev_pool[c->ev_used++];
        }
        cudaEventRecord(ev[0], c->stream);
        stop = ev[1];
        c->spans.push_back({id, ev[0], ev[1]});
    }
    ~KernelSpan() { if (stop) cudaEventRecord(stop, c->stream); }
};
}  // namespace vslam

#define VS_FAIL(ctx, code, ...)                                  \
    do {                                                         \
        char _b[512];                                            \
        snprintf(_b, sizeof _b, __VA_ARGS__);                    \
        if (ctx) (ctx)->err = _b;                                \
        vslam::set_global_error(_b);                             \
        return (code);                                           \
    } while (0)

#define VS_CUDA(ctx, call)                                                                        \
    do {                                                                                          \
        cudaError_t _e = (call);                                                                  \
        if (_e != cudaSuccess) {                                                                  \
            cudaGetLastError();                                                                   \
            VS_FAIL(ctx, VSLAM_E_CUDA, "%s:%d %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(_e)); \
        }                                                                                         \
    } while (0)

#define VS_TRY(expr)                      \
    do {                                  \
        int _rc = (expr);                 \
        if (_rc != VSLAM_OK) return _rc;  \
    } while (0)

#define VS_LAUNCH_CHECK(ctx)                       \
    do {                                           \
        (ctx)->launches++;                         \
        VS_CUDA(ctx, cudaGetLastError());          \
    } while (0)

namespace vslam {

// cv::borderInterpolate(p, len, BORDER_REFLECT_101) incl. repeated bouncing (kernel wider than the image).
__host__ __device__ __forceinline__ int refl101(int p, int len) {
    if ((unsigned)p < (unsigned)len) return p;
    if (len == 1) return 0;
    do {
        if (p < 0) p = -p;
        else p = 2 * (len - 1) - p;
    } while ((unsigned)p >= (unsigned)len);
    return p;
}
__host__ __device__ __forceinline__ int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }
static inline size_t round_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// ---- fixed-point Gaussian taps (taps.cu) ---------------------------------------------------------------------
constexpr int kMaxKsize = 255;              // radius <= 127
int host_gauss_ksize(double sigma);
int host_gauss_taps(double sigma, int n, int32_t *taps);

// exclusive scan of n ints (compaction.cu); partial: ceil(n / 4096) + 4 ints of workspace
int device_exclusive_scan(vslam_ctx *ctx, const int *cnt, long long n, int *partial, int *dst, int *d_total);

// ---- JPEG ingest (jpeg.cu): releases a context's nvJPEG handle, decoder state and staging
void jpeg_release(vslam_ctx *ctx);

// ---- segment-ordered compaction (compaction.cu) -------------------------------------------------------------
int seg_prepare(vslam_ctx *ctx, size_t n_segments, size_t entry_bytes, size_t entries);
// after the producer kernel: order the staged entries of n_segments segments into dst[dst_base ...];
// d_total[0] accumulates the number written so far (device counter); entry_bytes 12 or 24.
// n_segments = n_frames * segments per frame; d_frame_count (n_frames ints) receives the per-frame counts.
// compact != 0 (corners only): dst receives 8-byte vslam_corner8 entries
int seg_finalise(vslam_ctx *ctx, size_t n_segments, int entry_bytes, void *dst, int capacity, int *d_total,
                 int *d_frame_count, int n_frames, int compact = 0);

}  // namespace vslam
