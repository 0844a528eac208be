// capi.cu -- the C ABI of include/vslam.h: contexts, GaussPyramid build/getters, DoG detection, Harris.
// Host orchestration only; kernels live in pyramid.cu, detect.cu, harris.cu, compaction.cu.
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <functional>
#include <mutex>

#include "descriptor.cuh"
#include "detect.cuh"
#include "fused.cuh"
#include "harris.cuh"
#include "pyramid.cuh"
#include "stages.cuh"

using namespace vslam;

namespace vslam {
static std::mutex g_err_mu;
static std::string g_err;
void set_global_error(const std::string &s) {
    std::lock_guard<std::mutex> l(g_err_mu);
    g_err = s;
}
}  // namespace vslam

struct vslam_pyramid {
    int device = 0;
    int n_frames = 0, rows = 0, cols = 0, num_octaves = 0, scale_samples = 3;
    double sigma0 = 0;
    uint32_t materialise = 0;
    int orows[16], ocols[16];
    size_t opitch[16], ooff[16], oplane[16];   // pitch, byte offset of the octave inside a frame block, plane bytes
    size_t frame_bytes = 0;
    double sigma[16][6];
    DevBuf maps;                                // n_frames x [octave][base, gauss0..5, dog0..4]
    DevBuf grads;                               // materialise != 0: n_frames x [octave][level][gx, gy, mag, orient] dense fp32 planes
    size_t goff[16], grad_frame_floats = 0;     // float offset of an octave inside a frame's gradient block
    float *grad_plane(int f, int o, int l, int k) const { return grads.as<float>() + (size_t)f * grad_frame_floats + goff[o] + ((size_t)l * 4 + k) * (size_t)orows[o] * ocols[o]; }
    u8 *plane(int f, int o, int idx) const { return maps.as<u8>() + (size_t)f * frame_bytes + ooff[o] + oplane[o] * idx; }
};

static const int kCounterBase = 16;
static const int FWO_SAMPLES = kFusedWO / 3;   // sample columns per strip             // counters: [0] staging cursor, [2] running total, [16+f] per frame

extern "C" {

const char *vslam_version(void) { return "visualslam_b200 0.1 (sm_100a)"; }

const char *vslam_last_error(const vslam_ctx *ctx) {
    if (ctx) return ctx->err.c_str();
    static thread_local std::string copy;
    std::lock_guard<std::mutex> l(g_err_mu);
    copy = g_err;
    return copy.c_str();
}

// live contexts per device
static int ctx_live(int device, int delta) {
    static std::mutex mu;
    static int live[64] = {0};
    std::lock_guard<std::mutex> lock(mu);
    if (device < 0 || device >= 64) return 1;
    live[device] += delta;
    return live[device];
}

int vslam_ctx_create_on_stream(int device, void *cuda_stream, vslam_ctx **out) {
    if (!out) VS_FAIL((vslam_ctx *)nullptr, VSLAM_E_INVALID, "vslam_ctx_create: out is null");
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n <= 0) {
        cudaGetLastError();
        VS_FAIL((vslam_ctx *)nullptr, VSLAM_E_CUDA, "no CUDA device available (%s); this library has no CPU fallback",
                e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
    }
    if (device < 0 || device >= n) VS_FAIL((vslam_ctx *)nullptr, VSLAM_E_INVALID, "device %d out of range [0,%d)", device, n);
    vslam_ctx *ctx = new (std::nothrow) vslam_ctx();
    if (!ctx) return VSLAM_E_NOMEM;
    ctx->device = device;
    if (cudaSetDevice(device) != cudaSuccess) { delete ctx; VS_FAIL((vslam_ctx *)nullptr, VSLAM_E_CUDA, "cudaSetDevice(%d) failed", device); }
    if (cuda_stream) {
        ctx->stream = (cudaStream_t)cuda_stream;
    } else {
        if (cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess) {
            delete ctx;
            VS_FAIL((vslam_ctx *)nullptr, VSLAM_E_CUDA, "cudaStreamCreate failed");
        }
        ctx->own_stream = true;
    }
    cudaDeviceGetAttribute(&ctx->sm_count, cudaDevAttrMultiProcessorCount, device);
    ctx->pinned_cap = 1 << 20;
    if (cudaMallocHost(&ctx->pinned, ctx->pinned_cap) != cudaSuccess) { ctx->pinned = nullptr; ctx->pinned_cap = 0; cudaGetLastError(); }
    ctx_live(device, +1);
    *out = ctx;
    return VSLAM_OK;
}

int vslam_ctx_create(int device, vslam_ctx **out) { return vslam_ctx_create_on_stream(device, nullptr, out); }

void vslam_ctx_destroy(vslam_ctx *ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    DevBuf *bufs[] = {&ctx->seg_off, &ctx->seg_cnt, &ctx->seg_dst, &ctx->seg_partial, &ctx->staging, &ctx->counters, &ctx->stage_counts,
                      &ctx->ingest, &ctx->ingest2, &ctx->wide_ws, &ctx->grad_lut, &ctx->ws_maps, &ctx->out_dev, &ctx->maps_dev, &ctx->alt_seg_off, &ctx->alt_seg_cnt, &ctx->alt_seg_dst,
                      &ctx->alt_seg_partial, &ctx->alt_staging, &ctx->alt_counters, &ctx->alt_out_dev, &ctx->defer_kp[0], &ctx->defer_kp[1],
                      &ctx->defer_co[0], &ctx->defer_co[1]};
    for (DevBuf *b : bufs) b->release();
    if (ctx->pinned) cudaFreeHost(ctx->pinned);
    for (cudaEvent_t e : ctx->ev_pool) cudaEventDestroy(e);
    for (cudaEvent_t e : ctx->up_ev) cudaEventDestroy(e);
    for (cudaEvent_t e : ctx->up_ev2) cudaEventDestroy(e);
    for (auto &d : ctx->defer) if (d.ev) cudaEventDestroy(d.ev);
    for (cudaEvent_t e : ctx->slot_done) if (e) cudaEventDestroy(e);
    jpeg_release(ctx);
    if (ctx->up_start) cudaEventDestroy(ctx->up_start);
    if (ctx->copy_stream) cudaStreamDestroy(ctx->copy_stream);
    for (cudaEvent_t e : ctx->down_ev) cudaEventDestroy(e);
    if (ctx->down_stream) cudaStreamDestroy(ctx->down_stream);
    if (ctx->own_stream) cudaStreamDestroy(ctx->stream);
    free(ctx->h_taps);
    // device-wide caches (tap-table images of the octave kernels) go with the last context of the device
    if (ctx_live(ctx->device, -1) == 0) { mma_tab_release(ctx->device); umma_tab_release(ctx->device); }
    delete ctx;
}

int vslam_sync(vslam_ctx *ctx) {
    if (!ctx) return VSLAM_E_INVALID;
    VS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return VSLAM_OK;
}

long long vslam_launch_count(const vslam_ctx *ctx) { return ctx ? ctx->launches : 0; }

static void drain_spans(vslam_ctx *ctx) {
    cudaStreamSynchronize(ctx->stream);
    for (auto &s : ctx->spans) {
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, s.a, s.b) == cudaSuccess) { ctx->prof_ms[s.id] += ms; ctx->prof_n[s.id]++; }
    }
    cudaGetLastError();
    ctx->spans.clear();
    ctx->ev_used = 0;
}

int vslam_profile(vslam_ctx *ctx, int enable) {
    if (!ctx) return VSLAM_E_INVALID;
    drain_spans(ctx);
    for (int i = 0; i < VSLAM_KERNEL_COUNT; ++i) { ctx->prof_ms[i] = 0; ctx->prof_n[i] = 0; }
    ctx->profiling = enable != 0;
    return VSLAM_OK;
}

int vslam_profile_read(vslam_ctx *ctx, int kernel_id, double *total_ms, long long *launches) {
    if (!ctx || kernel_id < 0 || kernel_id >= VSLAM_KERNEL_COUNT) return VSLAM_E_INVALID;
    drain_spans(ctx);
    if (total_ms) *total_ms = ctx->prof_ms[kernel_id];
    if (launches) *launches = ctx->prof_n[kernel_id];
    return VSLAM_OK;
}

double vslam_calculate_sigma(int octave, int level, double sigma0, int scale_samples) {
    const double k = pow(2.0, 1.0 / (double)scale_samples);      // GaussPyramid.hpp:69
    return pow(2, octave) * sigma0 * pow(k, level);               // GaussPyramid.cpp:157,161
}
int vslam_calculate_num_octaves(int rows, int cols) {              // GaussPyramid.cpp:150-152
    return (int)floor(log2((double)(rows < cols ? rows : cols))) - 4;
}
int vslam_gauss_ksize(double sigma) { return host_gauss_ksize(sigma); }
int vslam_gauss_taps(double sigma, int ksize, int32_t *taps) {
    if (!taps) return VSLAM_E_INVALID;
    return host_gauss_taps(sigma, ksize, taps);
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------------------
// internals
// ---------------------------------------------------------------------------------------------------------------
static int ensure_taps(vslam_ctx *ctx, int num_octaves, double sigma0, int scale_samples) {
    if (ctx->h_taps && ctx->taps_octaves == num_octaves && ctx->taps_sigma0 == sigma0 && ctx->taps_scale_samples == scale_samples)
        return VSLAM_OK;
    free(ctx->h_taps);
    ctx->h_taps = malloc(sizeof(FusedTaps) * (size_t)num_octaves);
    if (!ctx->h_taps) return VSLAM_E_NOMEM;
    ctx->taps_octaves = 0;
    FusedTaps *h = (FusedTaps *)ctx->h_taps;
    ctx->any_wide = false;
    for (int o = 0; o < num_octaves; ++o) {
        // kernels wider than 247 taps (or a stack that does not fit shared memory): the exact global-memory path takes the octave
        const int rc = build_fused_taps(sigma0, scale_samples, o, &h[o]);
        ctx->wide_octave[o] = rc != 0 || fused_smem_bytes(h[o]) > 223 * 1024;
        ctx->any_wide = ctx->any_wide || ctx->wide_octave[o];
    }
    ctx->taps_octaves = num_octaves; ctx->taps_sigma0 = sigma0; ctx->taps_scale_samples = scale_samples;
    return VSLAM_OK;
}

struct Geometry {
    int num_octaves;
    int orows[16], ocols[16];
    size_t opitch[16], ooff[16], oplane[16];
    size_t frame_bytes;
};

static int make_geometry(vslam_ctx *ctx, int rows, int cols, const vslam_pyramid_params *pp, Geometry *g, int *num_octaves,
                         double *sigma0) {
    if (rows <= 0 || cols <= 0) VS_FAIL(ctx, VSLAM_E_INVALID, "image size %dx%d", rows, cols);
    if (rows > (1 << 15) || cols > (1 << 15)) VS_FAIL(ctx, VSLAM_E_INVALID, "image larger than 32768 on a side");
    const int ss = pp ? pp->scale_samples : 3;
    if (ss != 3 && ss != 0) VS_FAIL(ctx, VSLAM_E_INVALID, "scale_samples must be 3 (numLevels_ = 6)");
    int no = pp ? pp->num_octaves : 4;
    if (no <= 0) no = vslam_calculate_num_octaves(rows, cols);
    if (no <= 0 || no > 12) VS_FAIL(ctx, VSLAM_E_INVALID, "num_octaves %d", no);
    *sigma0 = pp && pp->sigma0 > 0 ? pp->sigma0 : 1.6;
    *num_octaves = no;
    g->num_octaves = no;
    int r = 2 * rows, c = 2 * cols;
    size_t off = 0;
    for (int o = 0; o < no; ++o) {
        if (r <= 0 || c <= 0) VS_FAIL(ctx, VSLAM_E_INVALID, "octave %d would be empty (%dx%d input)", o, rows, cols);
        g->orows[o] = r; g->ocols[o] = c;
        g->opitch[o] = plane_pitch(c);
        g->oplane[o] = round_up(plane_bytes(r, c), 256);
        g->ooff[o] = off;
        off += g->oplane[o] * 12;
        half_size(r, c, &r, &c);
    }
    g->frame_bytes = off;
    return VSLAM_OK;
}

static void octave_planes(const Geometry &g, u8 *frame_block, int o, OctavePlanes *op) {
    u8 *b = frame_block + g.ooff[o];
    op->rows = g.orows[o]; op->cols = g.ocols[o]; op->pitch = g.opitch[o];
    op->base = b;
    for (int l = 0; l < 6; ++l) op->gauss[l] = b + g.oplane[o] * (1 + l);
    for (int l = 0; l < 5; ++l) op->dog[l] = b + g.oplane[o] * (7 + l);
    if (o + 1 < g.num_octaves) {
        op->next_base = frame_block + g.ooff[o + 1];
        op->next_rows = g.orows[o + 1]; op->next_cols = g.ocols[o + 1]; op->next_pitch = g.opitch[o + 1];
    } else {
        op->next_base = nullptr; op->next_rows = op->next_cols = 0; op->next_pitch = 0;
    }
}

// Frames that live on the host are copied into ctx->ingest (pitch = multiple of 256) in chunks of about 64 MB on the
// context's copy stream; device frames are used in place.  Consumers call wait_ingest() before touching a frame range.
static int ingest_frames(vslam_ctx *ctx, const u8 *frames, int n, int rows, int cols, size_t pitch, size_t frame_stride,
                         int space, const u8 **d_frames, size_t *d_pitch, size_t *d_stride) {
    if (!frames || n <= 0) VS_FAIL(ctx, VSLAM_E_INVALID, "frames null or n_frames <= 0");
    if (pitch < (size_t)cols) VS_FAIL(ctx, VSLAM_E_INVALID, "pitch %zu < cols %d", pitch, cols);
    if (n > 1 && frame_stride < pitch * (size_t)rows) VS_FAIL(ctx, VSLAM_E_INVALID, "frame_stride too small");
    ctx->up_chunks = 0; ctx->up_foff = 0; ctx->consuming = -1;
    if (space == VSLAM_DEVICE) {
        *d_frames = frames; *d_pitch = pitch; *d_stride = frame_stride;
        // frames handed out by vslam_upload_frames: their chunks may still be in flight on the copy stream
        for (int sl = 0; sl < 2; ++sl) {
            const DevBuf &buf = sl ? ctx->ingest2 : ctx->ingest;
            const std::vector<cudaEvent_t> &ev = sl ? ctx->up_ev2 : ctx->up_ev;
            const u8 *b = buf.as<u8>();
            if (!ctx->slot_total[sl] || !b || frames < b || frames >= b + buf.cap || frame_stride != ctx->slot_ds[sl]) continue;
            ctx->up_slot = sl; ctx->up_chunk = ctx->slot_chunk[sl]; ctx->up_total = ctx->slot_total[sl]; ctx->up_ds = ctx->slot_ds[sl];
            ctx->up_foff = (int)((size_t)(frames - b) / ctx->up_ds);
            ctx->slot_busy[sl] = true; ctx->consuming = sl;
            // an upload that has already landed (started under the previous batch's kernels) needs no waits and no short first launches
            if (cudaEventQuery(ev[(size_t)ctx->up_total - 1]) != cudaSuccess) { cudaGetLastError(); ctx->up_chunks = ctx->up_total; }
            break;
        }
        return VSLAM_OK;
    }
    const size_t dp = round_up((size_t)cols, 256), ds = dp * (size_t)rows;
    const int sl = ctx->ingest_next;
    ctx->ingest_next ^= 1;
    DevBuf &ibuf = sl ? ctx->ingest2 : ctx->ingest;
    std::vector<cudaEvent_t> &iev = sl ? ctx->up_ev2 : ctx->up_ev;
    if (ibuf.reserve(ds * (size_t)n)) VS_FAIL(ctx, VSLAM_E_NOMEM, "ingest buffer (%zu bytes)", ds * (size_t)n);
    if (!ctx->copy_stream) {
        VS_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
        VS_CUDA(ctx, cudaEventCreateWithFlags(&ctx->up_start, cudaEventDisableTiming));
    }
    int chunk = (int)(((size_t)16 << 20) / ds);                 // small chunks: the first kernels start after ~0.3 ms of upload
    if (chunk < 1) chunk = 1;
    if (chunk > n) chunk = n;
    const int chunks = (n + chunk - 1) / chunk;
    while ((int)iev.size() < chunks) {
        cudaEvent_t e;
        VS_CUDA(ctx, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        iev.push_back(e);
    }
    // readers of the batch this buffer held may still be running: wait for the event behind the last of them, or -- when a reader has
    // not recorded one -- for everything enqueued on the compute stream so far
    if (!ctx->slot_busy[sl] && ctx->slot_done_set[sl]) {
        VS_CUDA(ctx, cudaStreamWaitEvent(ctx->copy_stream, ctx->slot_done[sl], 0));
    } else {
        VS_CUDA(ctx, cudaEventRecord(ctx->up_start, ctx->stream));
        VS_CUDA(ctx, cudaStreamWaitEvent(ctx->copy_stream, ctx->up_start, 0));
    }
    ctx->slot_busy[sl] = true; ctx->consuming = sl;             // the calling detector reads it (vslam_upload_frames clears this)
    for (int c = 0; c < chunks; ++c) {
        const int f0 = c * chunk, nf = n - f0 < chunk ? n - f0 : chunk;
        if (pitch == (size_t)cols && dp == pitch && frame_stride == ds) {
            VS_CUDA(ctx, cudaMemcpyAsync(ibuf.as<u8>() + ds * f0, frames + frame_stride * f0, ds * (size_t)nf, cudaMemcpyHostToDevice, ctx->copy_stream));
        } else {
            for (int f = f0; f < f0 + nf; ++f)
                VS_CUDA(ctx, cudaMemcpy2DAsync(ibuf.as<u8>() + ds * f, dp, frames + frame_stride * f, pitch, (size_t)cols,
                                               (size_t)rows, cudaMemcpyHostToDevice, ctx->copy_stream));
        }
        VS_CUDA(ctx, cudaEventRecord(iev[c], ctx->copy_stream));
    }
    ctx->slot_chunk[sl] = chunk; ctx->slot_total[sl] = chunks; ctx->slot_ds[sl] = ds;
    ctx->up_slot = sl; ctx->up_chunk = chunk; ctx->up_chunks = chunks; ctx->up_total = chunks; ctx->up_ds = ds;
    *d_frames = ibuf.as<u8>(); *d_pitch = dp; *d_stride = ds;
    return VSLAM_OK;
}

// the compute stream waits for the upload of frames [f0, f0 + nf)
static int wait_ingest(vslam_ctx *ctx, int f0, int nf) {
    if (!ctx->up_chunks) return VSLAM_OK;
    f0 += ctx->up_foff;
    const int c0 = f0 / ctx->up_chunk, c1 = (f0 + nf - 1) / ctx->up_chunk;
    const std::vector<cudaEvent_t> &ev = ctx->up_slot ? ctx->up_ev2 : ctx->up_ev;
    for (int c = c0; c <= c1 && c < ctx->up_chunks; ++c) VS_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ev[c], 0));
    return VSLAM_OK;
}
// behind the last kernel of a detector call: the ingest buffer it read may be overwritten once this event has passed
static int consume_end(vslam_ctx *ctx) {
    const int sl = ctx->consuming;
    if (sl < 0) return VSLAM_OK;
    if (!ctx->slot_done[sl]) VS_CUDA(ctx, cudaEventCreateWithFlags(&ctx->slot_done[sl], cudaEventDisableTiming));
    VS_CUDA(ctx, cudaEventRecord(ctx->slot_done[sl], ctx->stream));
    ctx->slot_done_set[sl] = true; ctx->slot_busy[sl] = false; ctx->consuming = -1;
    return VSLAM_OK;
}
// frames per compute launch when the input is being uploaded: ~128 MB of frames, and a short first launch so that the
// kernels start early (ingest_first)
static long long ingest_cap(const vslam_ctx *ctx, long long fb) {
    if (ctx->up_chunks > 1 && fb > 8LL * ctx->up_chunk) fb = 8LL * ctx->up_chunk;
    return fb;
}
static int ingest_first(const vslam_ctx *ctx, long long fb) {
    return (int)(ctx->up_chunks > 1 && fb > 2LL * ctx->up_chunk ? 2LL * ctx->up_chunk : fb);
}

// an octave on the exact slow path: its taps go up, then launch_octave_wide (pyramid.cu)
static int build_octave_wide(vslam_ctx *ctx, const OctavePlanes &op, size_t frame_stride, int octave, int n_frames) {
    std::vector<int> taps;
    int off[6], ks[6];
    for (int l = 0; l < 6; ++l) {
        const double sigma = vslam_calculate_sigma(octave, l, ctx->taps_sigma0, ctx->taps_scale_samples);
        ks[l] = host_gauss_ksize(sigma);
        off[l] = (int)taps.size();
        taps.resize(taps.size() + (size_t)ks[l]);
        if (host_gauss_taps(sigma, ks[l], taps.data() + off[l])) VS_FAIL(ctx, VSLAM_E_INVALID, "octave %d level %d: no Gaussian kernel for sigma %g", octave, l, sigma);
    }
    const size_t tb = round_up(taps.size() * sizeof(int), 256), hb = (size_t)n_frames * op.rows * op.cols * sizeof(u16);
    if (ctx->wide_ws.reserve(tb + hb)) VS_FAIL(ctx, VSLAM_E_NOMEM, "wide-kernel scratch (%zu bytes)", tb + hb);
    VS_CUDA(ctx, cudaMemcpyAsync(ctx->wide_ws.p, taps.data(), taps.size() * sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
    VS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));            // `taps` is a local (and pageable) buffer
    return launch_octave_wide(ctx, op, frame_stride, ctx->wide_ws.as<int>(), off, ks, (u16 *)(ctx->wide_ws.as<u8>() + tb), n_frames);
}

// rows A-D for n_frames frames into consecutive blocks laid out by Geometry (frame f at block + f * g.frame_bytes):
// one batched upsample launch, then one fused launch per octave with every map materialised.
static int build_frames(vslam_ctx *ctx, const Geometry &g, const u8 *frames, int n_frames, int rows, int cols, size_t pitch,
                        size_t stride, u8 *block) {
    const FusedTaps *h = (const FusedTaps *)ctx->h_taps;
    VS_TRY(wait_ingest(ctx, 0, n_frames));
    for (int f0 = 0; f0 < n_frames; f0 += 32768) {
        const int nf = n_frames - f0 < 32768 ? n_frames - f0 : 32768;
        u8 *blk = block + g.frame_bytes * (size_t)f0;
        VS_TRY(launch_upsample2x(ctx, frames + stride * f0, rows, cols, pitch, stride, blk + g.ooff[0], g.opitch[0], g.frame_bytes, nf));
        for (int o = 0; o < g.num_octaves; ++o) {
            OctavePlanes op;
            octave_planes(g, blk, o, &op);
            if (ctx->wide_octave[o]) {
                VS_TRY(build_octave_wide(ctx, op, g.frame_bytes, o, nf));
                continue;
            }
            FusedArgs a;
            memset(&a, 0, sizeof a);
            a.taps = h[o];
            a.base = op.base; a.base_pitch = op.pitch; a.base_stride = g.frame_bytes;
            a.rows = op.rows; a.cols = op.cols; a.octave = o;
            a.next_base = op.next_base; a.next_pitch = op.next_pitch; a.next_stride = g.frame_bytes;
            a.next_rows = op.next_rows; a.next_cols = op.next_cols;
            for (int l = 0; l < 6; ++l) a.gauss[l] = op.gauss[l];
            for (int l = 0; l < 5; ++l) a.dog[l] = op.dog[l];
            a.map_pitch = op.pitch; a.map_stride = g.frame_bytes;
            // the tensor-pipe kernel when it has a materialising instantiation for this octave (the default pyramid), else dp4a / dp2a
            const int rc = launch_octave_mma(ctx, a, nf, ctx->taps_sigma0, ctx->taps_scale_samples);
            if (rc < 0) return rc;
            if (rc == 1) VS_TRY(launch_octave_fused(ctx, a, nf));
        }
    }
    return VSLAM_OK;
}

// Per-frame staging list: producers reserve space in arbitrary block order, so the staged set must be complete for the
// ordered gather to return the FIRST `capacity` entries.  Sized max(capacity, 1M) entries, never more than the frame can emit.
static size_t staging_entries(int capacity, size_t frame_bound) {
    size_t want = (size_t)capacity > ((size_t)1 << 20) ? (size_t)capacity : ((size_t)1 << 20);
    if (want > frame_bound) want = frame_bound;
    return want + 1;
}

static size_t detect_segments(const Geometry &g, int window, int *seg_base /*[num_octaves]*/) {
    size_t n = 0;
    for (int o = 0; o < g.num_octaves; ++o) {
        int ni, nj, nc;
        detect_grid(g.orows[o], g.ocols[o], window, &ni, &nj, &nc);
        seg_base[o] = (int)n;
        n += (size_t)3 * ni * nc;
    }
    return n;
}

// rows G-J for n_frames frames whose plane blocks start at `block`, g.frame_bytes apart (frames frame .. frame + n_frames - 1 of the
// call: one launch per octave for the batch); appends to d_out in emission order.  n_seg = segments of ONE frame.
static int detect_frame(vslam_ctx *ctx, const Geometry &g, u8 *block, const vslam_dog_params &dp, const int *seg_base,
                        size_t n_seg, vslam_keypoint *d_out, int capacity, int frame, unsigned long long *d_stage,
                        int only_octave = -1, int apply_edge = 1, int n_frames = 1) {
    int *cnt = ctx->counters.as<int>();
    VS_CUDA(ctx, cudaMemsetAsync(cnt, 0, sizeof(int), ctx->stream));
    for (int o = 0; o < g.num_octaves; ++o) {
        if (only_octave >= 0 && o != only_octave) continue;
        OctavePlanes op;
        octave_planes(g, block, o, &op);
        DetectArgs a;
        a.apply_edge = apply_edge;
        for (int l = 0; l < 5; ++l) a.dog[l] = op.dog[l];
        for (int l = 0; l < 6; ++l) a.gauss[l] = op.gauss[l];
        a.rows = op.rows; a.cols = op.cols; a.pitch = op.pitch; a.octave = o;
        detect_grid(op.rows, op.cols, dp.window, &a.ni, &a.nj, &a.nchunks);
        if (a.ni == 0 || a.nchunks == 0) continue;
        a.seg_base = seg_base[o];
        a.seg_off = ctx->seg_off.as<int>(); a.seg_cnt = ctx->seg_cnt.as<int>();
        a.cursor = cnt;
        a.staging = ctx->staging.as<vslam_keypoint>(); a.staging_cap = ctx->staging_entries;
        a.stage_counts = d_stage ? d_stage + ((size_t)frame * g.num_octaves + o) * 4 : nullptr;
        a.n_frames = n_frames; a.frame_stride = g.frame_bytes; a.segs_per_frame = (int)n_seg; a.stage_stride = g.num_octaves * 4;
        VS_TRY(launch_detect(ctx, a, dp));
    }
    return seg_finalise(ctx, n_seg * (size_t)n_frames, (int)sizeof(vslam_keypoint), d_out, capacity, cnt + 2, cnt + kCounterBase + frame, n_frames);
}

static int finish_counts(vslam_ctx *ctx, int n_frames, int capacity, int *counts, long long *total_out) {
    std::vector<int> h((size_t)n_frames);
    VS_CUDA(ctx, cudaMemcpyAsync(h.data(), ctx->counters.as<int>() + kCounterBase, sizeof(int) * (size_t)n_frames,
                                 cudaMemcpyDeviceToHost, ctx->stream));
    VS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    long long total = 0;
    for (int f = 0; f < n_frames; ++f) { if (counts) counts[f] = h[f]; total += h[f]; }
    *total_out = total;
    (void)capacity;
    return VSLAM_OK;
}

static int reserve_counters(vslam_ctx *ctx, int n_frames) {
    const size_t bytes = sizeof(int) * (size_t)(kCounterBase + n_frames);
    if (ctx->counters.reserve(bytes)) VS_FAIL(ctx, VSLAM_E_NOMEM, "counter buffer");
    VS_CUDA(ctx, cudaMemsetAsync(ctx->counters.p, 0, bytes, ctx->stream));
    return VSLAM_OK;
}

extern "C" {

// storage (grow-only: a rebuild allocates nothing) and every map of the pyramid's frames
static int pyramid_fill(vslam_ctx *ctx, vslam_pyramid *p, const Geometry &g, const u8 *d_frames, size_t d_pitch, size_t d_stride) {
    const int n_frames = p->n_frames, no = p->num_octaves;
    if (p->maps.reserve(g.frame_bytes * (size_t)n_frames)) VS_FAIL(ctx, VSLAM_E_NOMEM, "pyramid of %d frames needs %zu bytes", n_frames, g.frame_bytes * (size_t)n_frames);
    VS_TRY(build_frames(ctx, g, d_frames, n_frames, p->rows, p->cols, d_pitch, d_stride, p->maps.as<u8>()));
    // the maps processGradients leaves behind (GaussPyramid.cpp:65-104), when the caller wants them resident
    const unsigned gmask = (p->materialise >> VSLAM_MAP_GRADX) & 15u;
    if (gmask) {
        size_t fl = 0;
        for (int o = 0; o < no; ++o) { p->goff[o] = fl; fl += (size_t)24 * g.orows[o] * g.ocols[o]; }
        p->grad_frame_floats = fl;
        if (p->grads.reserve(fl * 4 * (size_t)n_frames)) VS_FAIL(ctx, VSLAM_E_NOMEM, "gradient maps of %d frames need %zu bytes", n_frames, fl * 4 * (size_t)n_frames);
        for (int o = 0; o < no; ++o)
            VS_TRY(launch_gradients_all(ctx, p->plane(0, o, 1), g.orows[o], g.ocols[o], g.opitch[o], g.oplane[o], g.frame_bytes,
                                        p->grads.as<float>() + p->goff[o], (size_t)g.orows[o] * g.ocols[o], fl, gmask, n_frames));
    }
    return VSLAM_OK;
}

int vslam_pyramid_build(vslam_ctx *ctx, const uint8_t *frames, int n_frames, int rows, int cols, size_t pitch,
                        size_t frame_stride, int space, const vslam_pyramid_params *params, vslam_pyramid **out) {
    if (!ctx || !out) return VSLAM_E_INVALID;
    *out = nullptr;
    VS_CUDA(ctx, cudaSetDevice(ctx->device));
    Geometry g;
    int no; double s0;
    VS_TRY(make_geometry(ctx, rows, cols, params, &g, &no, &s0));
    VS_TRY(ensure_taps(ctx, no, s0, 3));
    const u8 *d_frames; size_t d_pitch, d_stride;
    VS_TRY(ingest_frames(ctx, frames, n_frames, rows, cols, pitch, frame_stride, space, &d_frames, &d_pitch, &d_stride));
    vslam_pyramid *p = new (std::nothrow) vslam_pyramid();
    if (!p) return VSLAM_E_NOMEM;
    p->device = ctx->device; p->n_frames = n_frames; p->rows = rows; p->cols = cols; p->num_octaves = no;
    p->sigma0 = s0; p->materialise = params ? params->materialise : 0;
    for (int o = 0; o < no; ++o) {
        p->orows[o] = g.orows[o]; p->ocols[o] = g.ocols[o]; p->opitch[o] = g.opitch[o]; p->ooff[o] = g.ooff[o];
        p->oplane[o] = g.oplane[o];
        for (int l = 0; l < 6; ++l) p->sigma[o][l] = vslam_calculate_sigma(o, l, s0, 3);
    }
    p->frame_bytes = g.frame_bytes;
    const int rc = pyramid_fill(ctx, p, g, d_frames, d_pitch, d_stride);
    if (rc) { cudaStreamSynchronize(ctx->stream); p->maps.release(); p->grads.release(); delete p; return rc; }
    *out = p;
    return VSLAM_OK;
}

int vslam_pyramid_rebuild(vslam_ctx *ctx, vslam_pyramid *pyr, const uint8_t *frames, size_t pitch, size_t frame_stride, int space) {
    if (!ctx || !pyr) return VSLAM_E_INVALID;
    VS_CUDA(ctx, cudaSetDevice(ctx->device));
    if (pyr->device != ctx->device) VS_FAIL(ctx, VSLAM_E_INVALID, "pyramid lives on device %d, context on %d", pyr->device, ctx->device);
    vslam_pyramid_params pp;
    pp.num_octaves = pyr->num_octaves; pp.scale_samples = 3; pp.sigma0 = pyr->sigma0; pp.materialise = pyr->materialise;
    Geometry g;
    int no; double s0;
    VS_TRY(make_geometry(ctx, pyr->rows, pyr->cols, &pp, &g, &no, &s0));
    VS_TRY(ensure_taps(ctx, no, s0, 3));
    const u8 *d_frames; size_t d_pitch, d_stride;
    VS_TRY(ingest_frames(ctx, frames, pyr->n_frames, pyr->rows, pyr->cols, pitch, frame_stride, space, &d_frames, &d_pitch, &d_stride));
    return pyramid_fill(ctx, pyr, g, d_frames, d_pitch, d_stride);
}

void vslam_pyramid_destroy(vslam_pyramid *pyr) {
    if (!pyr) return;
    cudaSetDevice(pyr->device);
    cudaDeviceSynchronize();
    pyr->maps.release();
    pyr->grads.release();
    delete pyr;
}
int vslam_pyramid_num_octaves(const vslam_pyramid *pyr) { return pyr ? pyr->num_octaves : VSLAM_E_INVALID; }
int vslam_pyramid_num_frames(const vslam_pyramid *pyr) { return pyr ? pyr->n_frames : VSLAM_E_INVALID; }
int vslam_pyramid_octave_size(const vslam_pyramid *pyr, int octave, int *rows, int *cols) {
    if (!pyr || octave < 0 || octave >= pyr->num_octaves) return VSLAM_E_INVALID;
    if (rows) *rows = pyr->orows[octave];
    if (cols) *cols = pyr->ocols[octave];
    return VSLAM_OK;
}
int vslam_pyramid_sigma(const vslam_pyramid *pyr, int octave, int level, double *sigma) {
    if (!pyr || !sigma || octave < 0 || octave >= pyr->num_octaves || level < 0 || level >= 6) return VSLAM_E_INVALID;
    *sigma = pyr->sigma[octave][level];
    return VSLAM_OK;
}

static int map_plane_index(int kind, int level) {
    if (kind == VSLAM_MAP_BASE) return 0;
    if (kind == VSLAM_MAP_GAUSS && level >= 0 && level < 6) return 1 + level;
    if (kind == VSLAM_MAP_DOG && level >= 0 && level < 5) return 7 + level;
    return -1;
}

const uint8_t *vslam_pyramid_device_map(const vslam_pyramid *pyr, int frame, int kind, int octave, int level, size_t *pitch) {
    if (!pyr || frame < 0 || frame >= pyr->n_frames || octave < 0 || octave >= pyr->num_octaves) return nullptr;
    const int idx = map_plane_index(kind, level);
    if (idx < 0) return nullptr;
    if (pitch) *pitch = pyr->opitch[octave];
    return pyr->plane(frame, octave, idx);
}

int vslam_pyramid_get(vslam_ctx *ctx, const vslam_pyramid *pyr, int frame, int kind, int octave, int level, void *dst,
                      size_t dst_pitch, int space) {
    if (!ctx || !pyr || !dst) return VSLAM_E_INVALID;
    if (frame < 0 || frame >= pyr->n_frames || octave < 0 || octave >= pyr->num_octaves)
        VS_FAIL(ctx, VSLAM_E_INVALID, "frame %d / octave %d out of range", frame, octave);
    VS_CUDA(ctx, cudaSetDevice(ctx->device));
    const int r = pyr->orows[octave], c = pyr->ocols[octave];
    const cudaMemcpyKind mk = space == VSLAM_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost;
    if (kind <= VSLAM_MAP_DOG) {
        const int idx = map_plane_index(kind, level);
        if (idx < 0) VS_FAIL(ctx, VSLAM_E_INVALID, "kind %d level %d", kind, level);
        if (dst_pitch < (size_t)c) VS_FAIL(ctx, VSLAM_E_INVALID, "dst_pitch %zu < %d", dst_pitch, c);
        VS_CUDA(ctx, cudaMemcpy2DAsync(dst, dst_pitch, pyr->plane(frame, octave, idx), pyr->opitch[octave], (size_t)c, (size_t)r, mk, ctx->stream));
    } else if (kind <= VSLAM_MAP_ORIENT) {
        if (level < 0 || level >= 6) VS_FAIL(ctx, VSLAM_E_INVALID, "level %d", level);
        if (dst_pitch < (size_t)c * 4) VS_FAIL(ctx, VSLAM_E_INVALID, "dst_pitch %zu < %d floats", dst_pitch, c);
        if (pyr->materialise >> kind & 1u) {      // resident since vslam_pyramid_build
            VS_CUDA(ctx, cudaMemcpy2DAsync(dst, dst_pitch, pyr->grad_plane(frame, octave, level, kind - VSLAM_MAP_GRADX), (size_t)c * 4, (size_t)c * 4, (size_t)r, mk, ctx->stream));
            if (space != VSLAM_DEVICE) VS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
            return VSLAM_OK;
        }
        // otherwise gradient maps are formed from the resident Gaussian level (row E); one dense fp32 scratch plane
        float *tmp = (float *)dst;
        size_t tp = dst_pitch / 4;
        if (space != VSLAM_DEVICE || (dst_pitch & 3)) {
            if (ctx->maps_dev.reserve((size_t)r * c * 4)) VS_FAIL(ctx, VSLAM_E_NOMEM, "gradient scratch");
            tmp = ctx->maps_dev.as<float>(); tp = (size_t)c;
        }
        const u8 *gl = pyr->plane(frame, octave, 1 + level);
        VS_TRY(launch_gradients(ctx, gl, r, c, pyr->opitch[octave], kind == VSLAM_MAP_GRADX ? tmp : nullptr,
                                kind == VSLAM_MAP_GRADY ? tmp : nullptr, kind == VSLAM_MAP_MAG ? tmp : nullptr,
                                kind == VSLAM_MAP_ORIENT ? tmp : nullptr, tp));
        if ((void *)tmp != dst)
            VS_CUDA(ctx, cudaMemcpy2DAsync(dst, dst_pitch, tmp, tp * 4, (size_t)c * 4, (size_t)r, mk, ctx->stream));
    } else {
        VS_FAIL(ctx, VSLAM_E_INVALID, "unknown map kind %d", kind);
    }
    if (space != VSLAM_DEVICE) VS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return VSLAM_OK;
}

static int check_dog_params(vslam_ctx *ctx, const vslam_dog_params *in, vslam_dog_params *dp) {
    dp->window = in ? in->window : 3;
    dp->contrast_thr = in ? in->contrast_thr : 0.03f;
    dp->edge_r = in ? in->edge_r : 10.0f;
    if (dp->window < 1 || dp->window > 15) VS_FAIL(ctx, VSLAM_E_INVALID, "window %d", dp->window);
    if (!(dp->edge_r > 0)) VS_FAIL(ctx, VSLAM_E_INVALID, "edge_r %g", dp->edge_r);
    return VSLAM_OK;
}

// A device list that grows chunk by chunk (seg_finalise appends at the running total) streams to the caller's host
// buffer while later chunks compute: after every chunk the running total is snapshot into pinned memory behind an event;
// once the NEXT chunk has been queued, the host waits for that event and queues the chunk's slice on a third stream.
struct ListDrain {
    u8 *host = nullptr;
    const u8 *dev = nullptr;
    size_t entry = 0;
    long long cap = 0, copied = 0;
    int n = 0, slot0 = 0;          // chunks so far; first pinned slot / event of this drain (two drains may be open at once)
    bool on = false;
};
static const int kDrainMax = 1024;

static int drain_begin(vslam_ctx *ctx, ListDrain &d, void *out, int out_space, const void *d_list, size_t entry, int capacity, long long chunks,
                       int slot0 = 0) {
    d = ListDrain();
    if (out_space == VSLAM_DEVICE || !out || chunks < 2 || chunks > kDrainMax || !ctx->pinned || ctx->pinned_cap < 2 * kDrainMax * sizeof(int)) return VSLAM_OK;
    if (!ctx->down_stream) VS_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->down_stream, cudaStreamNonBlocking));
    d.slot0 = slot0;
    while ((long long)ctx->down_ev.size() < slot0 + chunks) {
        cudaEvent_t e;
        VS_CUDA(ctx, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        ctx->down_ev.push_back(e);
    }
    d.host = (u8 *)out; d.dev = (const u8 *)d_list; d.entry = entry; d.cap = capacity; d.on = true;
    return VSLAM_OK;
}
static int drain_settle(vslam_ctx *ctx, ListDrain &d, int j) {
    VS_CUDA(ctx, cudaEventSynchronize(ctx->down_ev[d.slot0 + j]));
    long long t = ((const int *)ctx->pinned)[d.slot0 + j];
    if (t > d.cap) t = d.cap;
    if (t > d.copied) {
        VS_CUDA(ctx, cudaMemcpyAsync(d.host + d.copied * d.entry, d.dev + d.copied * d.entry, (size_t)(t - d.copied) * d.entry,
                                     cudaMemcpyDeviceToHost, ctx->down_stream));
        d.copied = t;
    }
    return VSLAM_OK;
}
// after a chunk's seg_finalise
static int drain_chunk(vslam_ctx *ctx, ListDrain &d, const int *d_total) {
    if (!d.on) return VSLAM_OK;
    VS_CUDA(ctx, cudaMemcpyAsync((int *)ctx->pinned + d.slot0 + d.n, d_total, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    VS_CUDA(ctx, cudaEventRecord(ctx->down_ev[d.slot0 + d.n], ctx->stream));
    if (d.n > 0) VS_TRY(drain_settle(ctx, d, d.n - 1));
    ++d.n;
    return VSLAM_OK;
}
static int drain_end(vslam_ctx *ctx, ListDrain &d) {
    if (!d.on) return VSLAM_OK;
    if (d.n > 0) VS_TRY(drain_settle(ctx, d, d.n - 1));
    VS_CUDA(ctx, cudaStreamSynchronize(ctx->down_stream));
    return VSLAM_OK;
}

static int emit_list(vslam_ctx *ctx, void *out, int out_space, const void *d_list, size_t entry, long long total, int capacity,
                     bool drained = false) {
    if (out_space != VSLAM_DEVICE && out && !drained) {
        const size_t n = (size_t)(total < capacity ? total : capacity);
        if (n) VS_CUDA(ctx, cudaMemcpyAsync(out, d_list, n * entry, cudaMemcpyDeviceToHost, ctx->stream));
        VS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    }
    if (total > capacity) VS_FAIL(ctx, VSLAM_E_CAPACITY, "output list holds %d entries, %lld produced", capacity, total);
    return VSLAM_OK;
}

int vslam_dog_detect(vslam_ctx *ctx, const vslam_pyramid *pyr, const vslam_dog_params *params, vslam_keypoint *out,
                     int capacity, int space, int *counts, long long *stage_counts) {
    if (!ctx || !pyr || capacity < 0 || (!out && capacity > 0)) return VSLAM_E_INVALID;
    VS_CUDA(ctx, cudaSetDevice(ctx->device));
    vslam_dog_params dp;
    VS_TRY(check_dog_params(ctx, params, &dp));
    Geometry g;
    g.num_octaves = pyr->num_octaves; g.frame_bytes = pyr->frame_bytes;
    for (int o = 0; o < g.num_octaves; ++o) {
        g.orows[o] = pyr->orows[o]; g.ocols[o] = pyr->ocols[o]; g.opitch[o] = pyr->opitch[o]; g.ooff[o] = pyr->ooff[o];
        g.oplane[o] = pyr->oplane[o];
    }
    int seg_base[16];
    const size_t n_seg = detect_segments(g, dp.window, seg_base);
    // frames per launch: grid.z = 3 x frames <= 65535, segment arrays <= 2^28 entries
    int fb = pyr->n_frames < 16384 ? pyr->n_frames : 16384;
    if (n_seg && (size_t)fb > ((size_t)1 << 28) / n_seg) fb = (int)(((size_t)1 << 28) / n_seg);
    if (fb < 1) fb = 1;
    VS_TRY(seg_prepare(ctx, (n_seg ? n_seg : 1) * (size_t)fb, sizeof(vslam_keypoint), staging_entries(capacity, n_seg * 256 * (size_t)fb)));
    VS_TRY(reserve_counters(ctx, pyr->n_frames));
    vslam_keypoint *d_out = out;
    if (space != VSLAM_DEVICE) {
        if (ctx->out_dev.reserve(sizeof(vslam_keypoint) * (size_t)(capacity + 1))) VS_FAIL(ctx, VSLAM_E_NOMEM, "device keypoint list");
        d_out = ctx->out_dev.as<vslam_keypoint>();
    }
    unsigned long long *d_stage = nullptr;
    const size_t stage_n = (size_t)pyr->n_frames * g.num_octaves * 4;
    if (stage_counts) {
        if (ctx->stage_counts.reserve(stage_n * 8)) VS_FAIL(ctx, VSLAM_E_NOMEM, "stage counters");
        d_stage = ctx->stage_counts.as<unsigned long long>();
        VS_CUDA(ctx, cudaMemsetAsync(d_stage, 0, stage_n * 8, ctx->stream));
    }
    for (int f = 0; f < pyr->n_frames; f += fb)
        VS_TRY(detect_frame(ctx, g, pyr->maps.as<u8>() + pyr->frame_bytes * f, dp, seg_base, n_seg, d_out, capacity, f, d_stage, -1, 1,
                            pyr->n_frames - f < fb ? pyr->n_frames - f : fb));
    long long total = 0;
    if (stage_counts) VS_CUDA(ctx, cudaMemcpyAsync(stage_counts, d_stage, stage_n * 8, cudaMemcpyDeviceToHost, ctx->stream));
    VS_TRY(finish_counts(ctx, pyr->n_frames, capacity, counts, &total));
    return emit_list(ctx, out, space, d_out, sizeof(vslam_keypoint), total, capacity);
}

// Throughput path: batched fused pipeline.  Per chunk of frames: one upsample launch, one fused launch per octave
// (blur + DoG + detector; only the octave bases live in HBM), then the segment scan / ordered gather.
static int dog_keypoints_fused(vslam_ctx *ctx, const Geometry &g, const u8 *d_frames, int n_frames, int rows, int cols,
                               size_t d_pitch, size_t d_stride, const vslam_dog_params &dp, vslam_keypoint *d_out, int capacity,
                               ListDrain *drain, void *out, int out_space, const std::function<int(int, int)> *per_chunk = nullptr) {
    const FusedTaps *h = (const FusedTaps *)ctx->h_taps;
    // bases-only workspace layout of one frame; every row carries the reflect-101 margins its octave's kernel stages from
    size_t boff[16], bpitch[16], bbytes = 0;
    int bpad[16], bpadl[16], brp[16];
    for (int o = 0; o < g.num_octaves; ++o) {
        bpad[o] = octave_base_pad(ctx, ctx->taps_sigma0, ctx->taps_scale_samples, o);
        bpadl[o] = octave_base_pad_left(ctx, ctx->taps_sigma0, ctx->taps_scale_samples, o);   // what of the left margin is written (the layout stays symmetric)
        brp[o] = octave_base_row_pad(ctx, ctx->taps_sigma0, ctx->taps_scale_samples, o);     // tcgen05 kernel: margins above / below too
        bpitch[o] = bpad[o] ? round_up((size_t)g.ocols[o] + 2 * (size_t)bpad[o] + 4, 16) : g.opitch[o];   // + 4: the right margin is written in words
        boff[o] = bbytes + (size_t)brp[o] * bpitch[o] + bpad[o];
        bbytes += round_up(bpitch[o] * (size_t)(g.orows[o] + 2 * brp[o]) + 256, 256);     // + slack: staging may read past the last row's margin
    }
    // segments of one frame: (octave, level 1..3, sample row, strip)
    int seg_base[16], ni[16], nj[16];
    long long spf = 0;
    for (int o = 0; o < g.num_octaves; ++o) {
        int nc;
        detect_grid(g.orows[o], g.ocols[o], 3, &ni[o], &nj[o], &nc);
        seg_base[o] = (int)spf;
        spf += 3LL * ni[o] * octave_strips(ctx, ctx->taps_sigma0, ctx->taps_scale_samples, o, g.ocols[o]);
    }
    if (spf == 0) spf = 1;
    // chunk size: workspace <= 6 GB, segment arrays <= 2^30 entries
    long long fb = (long long)(6e9 / (double)(bbytes + spf * 12));
    if (fb > (1LL << 30) / spf) fb = (1LL << 30) / spf;
    if (fb > 32768) fb = 32768;
    // frames still being uploaded: launches double in size (16, 32, 64, ... frames at 1080p) -- every launch's frames have
    // landed by the time the previous launch ends (the upload is ~2.5x faster than these kernels), the first kernels
    // start after two upload chunks, and most frames run in large launches.  The lists are small: no need for many chunks.
    const bool doubling = ctx->up_chunks > 1;
    if (fb > n_frames) fb = n_frames;
    // host lists and a batch that is already resident: four launches instead of one, so that three quarters of the lists leave under
    // later launches (136 MB per 1080p x 256 step would otherwise follow the last kernel: 2.6 ms at PCIe speed alone, 3-4x that with 8
    // GPUs on one host).  More launches cost kernel time (wave quantisation of long-running CTAs): measured 1 / 2 / 4 / 6 / 12 launches =
    // 31.7 / 31.2 / 31.5 / 32.5 / 36.4 ms per step at one GPU (VSLAM_HOST_CHUNKS overrides)
    static const int host_chunks = [] { const char *e = getenv("VSLAM_HOST_CHUNKS"); return e && atoi(e) > 0 ? atoi(e) : 4; }();
    if (!doubling && out && out_space != VSLAM_DEVICE && n_frames >= 48 && fb > (n_frames + host_chunks - 1) / host_chunks)
        fb = (n_frames + host_chunks - 1) / host_chunks;
    if (fb < 1) fb = 1;
    if (ctx->ws_maps.reserve(bbytes * (size_t)fb)) VS_FAIL(ctx, VSLAM_E_NOMEM, "octave-base workspace (%zu bytes)", bbytes * (size_t)fb);
    const size_t bound = (size_t)spf * (FWO_SAMPLES) * (size_t)fb;
    VS_TRY(seg_prepare(ctx, (size_t)spf * fb, sizeof(vslam_keypoint), staging_entries(capacity, bound)));
    int *cnt = ctx->counters.as<int>();
    const float rp1 = dp.edge_r + 1.0f;
    // FeaturePointLocalization with a singular A*A^T (any zero derivative): dog_zhat = float(value)/255f, kept iff
    // > contrast_thr, new value (int)(dog_zhat*255f) -- tabulated with the same IEEE float ops (Diff_of_Gauss.cpp:243-246)
    int cmin = 256;
    u8 val_lut[256];
    for (int c = 255; c >= 0; --c) {
        volatile float v = (float)c / 255.0f;
        volatile float w = v * 255.0f;
        val_lut[c] = (u8)(int)w;
        if (v > dp.contrast_thr) cmin = c;
    }
    VS_TRY(drain_begin(ctx, *drain, out, out_space, d_out, sizeof(vslam_keypoint), capacity, (n_frames + fb - 1) / fb + 32));
    const int first = doubling ? ingest_first(ctx, fb) : (int)fb;
    for (int f0 = 0, step = first; f0 < n_frames; f0 += step, step = (int)(doubling && 2LL * step < fb ? 2LL * step : fb)) {
        // host lists leave chunk by chunk: a short LAST chunk keeps the part of the lists that can only leave after the last kernel small
        if (drain->on && doubling && step >= n_frames - f0 && n_frames - f0 > 2 * first) step = n_frames - f0 - first;
        const int nf = n_frames - f0 < step ? n_frames - f0 : step;
        const size_t n_seg = (size_t)spf * nf;
        if (f0) VS_CUDA(ctx, cudaMemsetAsync(ctx->seg_cnt.p, 0, n_seg * 4, ctx->stream));
        VS_CUDA(ctx, cudaMemsetAsync(cnt, 0, sizeof(int), ctx->stream));
        u8 *ws = ctx->ws_maps.as<u8>();
        VS_TRY(wait_ingest(ctx, f0, nf));
        if (per_chunk) VS_TRY((*per_chunk)(f0, nf));             // vslam_dog_harris: the Harris launch of the same frames goes first
        VS_TRY(launch_upsample2x(ctx, d_frames + d_stride * f0, rows, cols, d_pitch, d_stride, ws + boff[0], bpitch[0], bbytes, nf));
        for (int o = 0; o < g.num_octaves; ++o) {
            FusedArgs a;
            memset(&a, 0, sizeof a);
            a.taps = h[o];
            a.base = ws + boff[o]; a.base_pitch = bpitch[o]; a.base_stride = bbytes; a.base_pad = bpad[o]; a.base_row_pad = brp[o];
            VS_TRY(launch_reflect_pad(ctx, ws + boff[o], g.orows[o], g.ocols[o], bpitch[o], bbytes, bpad[o], brp[o], nf, bpadl[o],
                                      octave_base_pad_right(ctx, ctx->taps_sigma0, ctx->taps_scale_samples, o, g.ocols[o])));
            a.rows = g.orows[o]; a.cols = g.ocols[o]; a.octave = o;
            if (o + 1 < g.num_octaves) {
                a.next_base = ws + boff[o + 1]; a.next_pitch = bpitch[o + 1]; a.next_stride = bbytes;
                a.next_rows = g.orows[o + 1]; a.next_cols = g.ocols[o + 1];
            }
            a.detect = 1; a.cmin = cmin; memcpy(a.val_lut, val_lut, 256); a.contrast_thr = dp.contrast_thr; a.edge_thr = (rp1 * rp1) / dp.edge_r;   // Diff_of_Gauss.cpp:332-333
            a.ni = ni[o]; a.nj = nj[o]; a.seg_base = seg_base[o]; a.segs_per_frame = (int)spf;
            a.seg_off = ctx->seg_off.as<int>(); a.seg_cnt = ctx->seg_cnt.as<int>();
            a.cursor = cnt; a.staging = ctx->staging.as<vslam_keypoint>(); a.staging_cap = ctx->staging_entries;
            int rc = launch_octave_umma(ctx, a, nf, ctx->taps_sigma0, ctx->taps_scale_samples);         // tcgen05 + TMA (octaves 0, 1)
            if (rc == 1) rc = launch_octave_mma(ctx, a, nf, ctx->taps_sigma0, ctx->taps_scale_samples);  // mma.sync blur
            if (rc < 0) return rc;
            if (rc == 1) VS_TRY(launch_octave_fused(ctx, a, nf));                                         // dp4a/dp2a blur
        }
        VS_TRY(seg_finalise(ctx, n_seg, (int)sizeof(vslam_keypoint), d_out, capacity, cnt + 2, cnt + kCounterBase + f0, nf));
        VS_TRY(drain_chunk(ctx, *drain, cnt + 2));
    }
    return VSLAM_OK;
}

int vslam_dog_keypoints(vslam_ctx *ctx, const uint8_t *frames, int n_frames, int rows, int cols, size_t pitch,
                        size_t frame_stride, int space, const vslam_pyramid_params *pparams,
                        const vslam_dog_params *dparams, vslam_keypoint *out, int capacity, int out_space, int *counts) {
    if (!ctx || capacity < 0 || (!out && capacity > 0)) return VSLAM_E_INVALID;
    VS_CUDA(ctx, cudaSetDevice(ctx->device));
    vslam_dog_params dp;
    VS_TRY(check_dog_params(ctx, dparams, &dp));
    Geometry g;
    int no; double s0;
    VS_TRY(make_geometry(ctx, rows, cols, pparams, &g, &no, &s0));
    VS_TRY(ensure_taps(ctx, no, s0, 3));
    const u8 *d_frames; size_t d_pitch, d_stride;
    VS_TRY(ingest_frames(ctx, frames, n_frames, rows, cols, pitch, frame_stride, space, &d_frames, &d_pitch, &d_stride));
    VS_TRY(reserve_counters(ctx, n_frames));
    vslam_keypoint *d_out = out;
    if (out_space != VSLAM_DEVICE) {
        if (ctx->out_dev.reserve(sizeof(vslam_keypoint) * (size_t)(capacity + 1))) VS_FAIL(ctx, VSLAM_E_NOMEM, "device keypoint list");
        d_out = ctx->out_dev.as<vslam_keypoint>();
    }
    ListDrain drain;
    if (dp.window == 3 && !ctx->any_wide) {
        VS_TRY(dog_keypoints_fused(ctx, g, d_frames, n_frames, rows, cols, d_pitch, d_stride, dp, d_out, capacity, &drain, out, out_space));
    } else {
        // (also: pyramids with an octave on the exact slow path, i.e. kernels wider than 247 taps)
        // other grid strides: materialise one frame at a time and run the stand-alone detector (detect.cu)
        if (ctx->ws_maps.reserve(g.frame_bytes)) VS_FAIL(ctx, VSLAM_E_NOMEM, "frame workspace (%zu bytes)", g.frame_bytes);
        int seg_base[16];
        const size_t n_seg = detect_segments(g, dp.window, seg_base);
        VS_TRY(seg_prepare(ctx, n_seg ? n_seg : 1, sizeof(vslam_keypoint), staging_entries(capacity, n_seg * 256)));
        VS_TRY(wait_ingest(ctx, 0, n_frames));
        ctx->up_chunks = 0;
        for (int f = 0; f < n_frames; ++f) {
            VS_TRY(build_frames(ctx, g, d_frames + d_stride * f, 1, rows, cols, d_pitch, d_stride, ctx->ws_maps.as<u8>()));
            VS_TRY(detect_frame(ctx, g, ctx->ws_maps.as<u8>(), dp, seg_base, n_seg, d_out, capacity, f, nullptr));
        }
    }
    VS_TRY(consume_end(ctx));
    long long total = 0;
    VS_TRY(finish_counts(ctx, n_frames, capacity, counts, &total));
    VS_TRY(drain_end(ctx, drain));
    return emit_list(ctx, out, out_space, d_out, sizeof(vslam_keypoint), total, capacity, drain.on);
}

// vslam_harris in three pieces, so that vslam_dog_harris can run one Harris launch per chunk of the DoG loop
struct HarrisJob {
    const u8 *d_frames; size_t d_pitch, d_stride;
    int n_frames, rows, cols, capacity, out_space, compact, w2;
    float k;
    size_t centry, spf, fb, px;
    void *corners, *d_out;
    float *response, *d_resp; u8 *r8, *mask3, *d_r8, *d_m3;
    ListDrain drain;
};

// arguments, workspaces of the CURRENT list set (ctx->seg_*, staging, counters, out_dev) and the list drain.  max_chunk: the
// largest number of frames one harris_chunk call will be given (0: this function picks the launch size, job.fb)
static int harris_begin(vslam_ctx *ctx, HarrisJob &j, const u8 *d_frames, size_t d_pitch, size_t d_stride, int n_frames, int rows, int cols,
                        const vslam_harris_params *params, float *response, uint8_t *r8, uint8_t *mask3, void *corners, int capacity,
                        int out_space, int max_chunk, int drain_slot0) {
    if (rows <= 0 || cols <= 0 || rows > (1 << 15) || cols > (1 << 15)) VS_FAIL(ctx, VSLAM_E_INVALID, "image size %dx%d", rows, cols);
    j = HarrisJob();
    j.compact = (out_space & VSLAM_COMPACT) ? 1 : 0;             // `corners` is a vslam_corner8 buffer
    out_space &= ~VSLAM_COMPACT;
    if (out_space != VSLAM_HOST && out_space != VSLAM_DEVICE) VS_FAIL(ctx, VSLAM_E_INVALID, "out_space %d", out_space);
    j.out_space = out_space;
    j.centry = j.compact ? sizeof(vslam_corner8) : sizeof(vslam_corner);
    j.k = params ? params->k : 0.04f;
    j.w2 = params ? params->nms2_window : 5;
    if (j.w2 != 3 && j.w2 != 5) VS_FAIL(ctx, VSLAM_E_INVALID, "nms2_window must be 3 or 5 (got %d)", j.w2);
    j.d_frames = d_frames; j.d_pitch = d_pitch; j.d_stride = d_stride; j.n_frames = n_frames; j.rows = rows; j.cols = cols;
    j.capacity = capacity; j.corners = corners; j.response = response; j.r8 = r8; j.mask3 = mask3;
    j.spf = (size_t)rows * harris_strips(cols);                  // segments per frame: (row, strip)
    size_t fb = ((size_t)1 << 30) / j.spf;                       // frames per launch
    if (fb > 32768) fb = 32768;
    if (max_chunk > 0) {
        if (fb < (size_t)max_chunk) VS_FAIL(ctx, VSLAM_E_INVALID, "image too large for %d frames per Harris launch", max_chunk);
        fb = (size_t)max_chunk;
    } else {
        fb = (size_t)ingest_cap(ctx, (long long)fb);
    }
    if (fb > (size_t)n_frames) fb = (size_t)n_frames;
    if (fb < 1) VS_FAIL(ctx, VSLAM_E_INVALID, "image too large for the segment table");
    j.fb = fb;
    VS_TRY(seg_prepare(ctx, j.spf * fb, sizeof(vslam_corner), staging_entries(capacity, (size_t)rows * cols * fb)));
    VS_TRY(reserve_counters(ctx, n_frames));
    j.d_out = corners;
    if (out_space != VSLAM_DEVICE) {
        if (ctx->out_dev.reserve(j.centry * (size_t)(capacity + 1))) VS_FAIL(ctx, VSLAM_E_NOMEM, "device corner list");
        j.d_out = ctx->out_dev.p;
    }
    j.px = (size_t)rows * cols;
    j.d_resp = response; j.d_r8 = r8; j.d_m3 = mask3;
    if (out_space != VSLAM_DEVICE && (response || r8 || mask3)) {
        if (ctx->maps_dev.reserve(j.px * 6 * (size_t)n_frames)) VS_FAIL(ctx, VSLAM_E_NOMEM, "device maps (%zu bytes)", j.px * 6 * (size_t)n_frames);
        j.d_resp = response ? ctx->maps_dev.as<float>() : nullptr;
        j.d_r8 = r8 ? ctx->maps_dev.as<u8>() + j.px * 4 * n_frames : nullptr;
        j.d_m3 = mask3 ? ctx->maps_dev.as<u8>() + j.px * 5 * n_frames : nullptr;
    }
    const long long chunks = max_chunk > 0 ? kDrainMax / 2 : ((long long)n_frames + (long long)fb - 1) / (long long)fb + 1;
    VS_TRY(drain_begin(ctx, j.drain, corners, out_space, j.d_out, j.centry, capacity, chunks, drain_slot0));
    return VSLAM_OK;
}

// frames [f0, f0 + nf) (nf <= job.fb): one launch, the ordered gather behind it, the chunk's slice of the list on its way out
static int harris_chunk(vslam_ctx *ctx, HarrisJob &j, int f0, int nf) {
    int *cnt = ctx->counters.as<int>();
    if (f0) VS_CUDA(ctx, cudaMemsetAsync(ctx->seg_cnt.p, 0, j.spf * nf * 4, ctx->stream));
    VS_CUDA(ctx, cudaMemsetAsync(cnt, 0, sizeof(int), ctx->stream));
    HarrisArgs a;
    memset(&a, 0, sizeof a);
    a.frames = j.d_frames + j.d_stride * f0; a.pitch = j.d_pitch; a.frame_stride = j.d_stride;
    a.rows = j.rows; a.cols = j.cols; a.k = j.k; a.nms2_pad = (j.w2 - 1) / 2;
    a.response = j.d_resp ? j.d_resp + j.px * f0 : nullptr;
    a.r8 = j.d_r8 ? j.d_r8 + j.px * f0 : nullptr;
    a.mask3 = j.d_m3 ? j.d_m3 + j.px * f0 : nullptr;
    a.seg_off = ctx->seg_off.as<int>(); a.seg_cnt = ctx->seg_cnt.as<int>();
    a.cursor = cnt; a.staging = ctx->staging.as<vslam_corner>(); a.staging_cap = ctx->staging_entries;
    a.want_corners = 1;
    VS_TRY(launch_harris(ctx, a, nf));
    VS_TRY(seg_finalise(ctx, j.spf * nf, (int)sizeof(vslam_corner), j.d_out, j.capacity, cnt + 2, cnt + kCounterBase + f0, nf, j.compact));
    VS_TRY(drain_chunk(ctx, j.drain, cnt + 2));
    return VSLAM_OK;
}

static int harris_end(vslam_ctx *ctx, HarrisJob &j, int *counts) {
    if (j.out_space != VSLAM_DEVICE) {
        if (j.response) VS_CUDA(ctx, cudaMemcpyAsync(j.response, j.d_resp, j.px * 4 * j.n_frames, cudaMemcpyDeviceToHost, ctx->stream));
        if (j.r8) VS_CUDA(ctx, cudaMemcpyAsync(j.r8, j.d_r8, j.px * j.n_frames, cudaMemcpyDeviceToHost, ctx->stream));
        if (j.mask3) VS_CUDA(ctx, cudaMemcpyAsync(j.mask3, j.d_m3, j.px * j.n_frames, cudaMemcpyDeviceToHost, ctx->stream));
    }
    long long total = 0;
    VS_TRY(finish_counts(ctx, j.n_frames, j.capacity, counts, &total));
    VS_TRY(drain_end(ctx, j.drain));
    return emit_list(ctx, j.corners, j.out_space, j.d_out, j.centry, total, j.capacity, j.drain.on);
}

int vslam_harris(vslam_ctx *ctx, const uint8_t *frames, int n_frames, int rows, int cols, size_t pitch,
                 size_t frame_stride, int space, const vslam_harris_params *params, float *response, uint8_t *r8,
                 uint8_t *mask3, vslam_corner *corners, int capacity, int out_space, int *counts) {
    if (!ctx || capacity < 0 || (!corners && capacity > 0)) return VSLAM_E_INVALID;
    VS_CUDA(ctx, cudaSetDevice(ctx->device));
    if (rows <= 0 || cols <= 0 || rows > (1 << 15) || cols > (1 << 15)) VS_FAIL(ctx, VSLAM_E_INVALID, "image size %dx%d", rows, cols);
    const u8 *d_frames; size_t d_pitch, d_stride;
    VS_TRY(ingest_frames(ctx, frames, n_frames, rows, cols, pitch, frame_stride, space, &d_frames, &d_pitch, &d_stride));
    HarrisJob j;
    VS_TRY(harris_begin(ctx, j, d_frames, d_pitch, d_stride, n_frames, rows, cols, params, response, r8, mask3, corners, capacity, out_space, 0, 0));
    const int first = ingest_first(ctx, (long long)j.fb);
    for (int f0 = 0, step = first; f0 < n_frames; f0 += step, step = (int)j.fb) {
        const int nf = n_frames - f0 < step ? n_frames - f0 : step;
        VS_TRY(wait_ingest(ctx, f0, nf));
        VS_TRY(harris_chunk(ctx, j, f0, nf));
    }
    VS_TRY(consume_end(ctx));
    return harris_end(ctx, j, counts);
}

// the two list workspaces of a context trade places: code written against ctx->seg_* / staging / counters / out_dev then works on
// the other list.  RAII: the sets are swapped back on every exit path.
struct ListSetSwap {
    vslam_ctx *c;
    explicit ListSetSwap(vslam_ctx *ctx) : c(ctx) { flip(); }
    ~ListSetSwap() { flip(); }
    void flip() {
        std::swap(c->seg_off, c->alt_seg_off); std::swap(c->seg_cnt, c->alt_seg_cnt); std::swap(c->seg_dst, c->alt_seg_dst);
        std::swap(c->seg_partial, c->alt_seg_partial); std::swap(c->staging, c->alt_staging); std::swap(c->counters, c->alt_counters);
        std::swap(c->out_dev, c->alt_out_dev); std::swap(c->staging_entries, c->alt_staging_entries);
    }
};

// pinned snapshot of a deferred call: [gen][kp counts n | co counts n] behind the drain slots
static int *defer_pinned(vslam_ctx *ctx, int gen) { return (int *)((u8 *)ctx->pinned + (16 << 10) + (size_t)gen * (320 << 10)); }

// VSLAM_DEFER: everything enqueued, nothing waited for.  The lists stay in the call's own device buffers; vslam_finish copies them.
static int dog_harris_deferred(vslam_ctx *ctx, const Geometry &g, const u8 *d_frames, size_t d_pitch, size_t d_stride, int n_frames, int rows,
                               int cols, const vslam_dog_params &dp, const vslam_harris_params *hparams, vslam_keypoint *keypoints, int kp_capacity,
                               int *kp_counts, vslam_corner *corners, int co_capacity, int *co_counts, int out_space) {
    const int gen = ctx->defer_next;
    vslam_ctx::Deferred &D = ctx->defer[gen];
    if (D.live) VS_FAIL(ctx, VSLAM_E_INVALID, "two deferred calls are pending: vslam_finish first");
    const int compact = (out_space & VSLAM_COMPACT) ? 1 : 0;
    const size_t centry = compact ? sizeof(vslam_corner8) : sizeof(vslam_corner);
    if (ctx->defer_kp[gen].reserve(sizeof(vslam_keypoint) * (size_t)(kp_capacity + 1)) || ctx->defer_co[gen].reserve(centry * (size_t)(co_capacity + 1)))
        VS_FAIL(ctx, VSLAM_E_NOMEM, "device lists of a deferred call");
    if (!D.ev) VS_CUDA(ctx, cudaEventCreateWithFlags(&D.ev, cudaEventDisableTiming));
    if (!ctx->down_stream) VS_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->down_stream, cudaStreamNonBlocking));
    // A deferred caller runs ahead of the GPU: the previous call's kernels are usually still queued, so this batch's upload (started
    // under them) will have landed before these launches reach the GPU.  One wait for the whole upload, then full-size launches --
    // not the doubling launch sizes that chase an upload in progress.
    VS_TRY(wait_ingest(ctx, 0, n_frames));
    ctx->up_chunks = 0;
    HarrisJob hj;
    {
        ListSetSwap sw(ctx);
        VS_TRY(harris_begin(ctx, hj, d_frames, d_pitch, d_stride, n_frames, rows, cols, hparams, nullptr, nullptr, nullptr, ctx->defer_co[gen].p,
                            co_capacity, VSLAM_DEVICE | (compact ? VSLAM_COMPACT : 0), n_frames, kDrainMax));
    }
    const std::function<int(int, int)> hook = [&](int f0, int nf) -> int {
        ListSetSwap sw(ctx);
        return harris_chunk(ctx, hj, f0, nf);
    };
    ListDrain kdrain;
    VS_TRY(dog_keypoints_fused(ctx, g, d_frames, n_frames, rows, cols, d_pitch, d_stride, dp, ctx->defer_kp[gen].as<vslam_keypoint>(), kp_capacity,
                               &kdrain, nullptr, VSLAM_DEVICE, &hook));
    VS_TRY(consume_end(ctx));
    int *snap = defer_pinned(ctx, gen);
    VS_CUDA(ctx, cudaMemcpyAsync(snap, ctx->counters.as<int>() + kCounterBase, sizeof(int) * (size_t)n_frames, cudaMemcpyDeviceToHost, ctx->stream));
    VS_CUDA(ctx, cudaMemcpyAsync(snap + n_frames, ctx->alt_counters.as<int>() + kCounterBase, sizeof(int) * (size_t)n_frames, cudaMemcpyDeviceToHost,
                                 ctx->stream));
    VS_CUDA(ctx, cudaEventRecord(D.ev, ctx->stream));
    D.live = true; D.kp_host = keypoints; D.co_host = corners; D.kp_counts = kp_counts; D.co_counts = co_counts;
    D.kp_cap = kp_capacity; D.co_cap = co_capacity; D.n_frames = n_frames; D.co_entry = centry;
    ctx->defer_next ^= 1;
    return VSLAM_OK;
}

int vslam_finish(vslam_ctx *ctx) {
    if (!ctx) return VSLAM_E_INVALID;
    int gen = ctx->defer_next;                                  // with two calls pending, the slot the next call would take is the older one
    if (!ctx->defer[gen].live) gen ^= 1;
    vslam_ctx::Deferred &D = ctx->defer[gen];
    if (!D.live) return VSLAM_OK;
    VS_CUDA(ctx, cudaSetDevice(ctx->device));
    D.live = false;
    VS_CUDA(ctx, cudaEventSynchronize(D.ev));
    const int *snap = defer_pinned(ctx, gen);
    long long nk = 0, nc = 0;
    for (int f = 0; f < D.n_frames; ++f) { D.kp_counts[f] = snap[f]; nk += snap[f]; D.co_counts[f] = snap[D.n_frames + f]; nc += snap[D.n_frames + f]; }
    const size_t ck = (size_t)(nk < D.kp_cap ? nk : D.kp_cap), cc = (size_t)(nc < D.co_cap ? nc : D.co_cap);
    if (ck) VS_CUDA(ctx, cudaMemcpyAsync(D.kp_host, ctx->defer_kp[gen].p, ck * sizeof(vslam_keypoint), cudaMemcpyDeviceToHost, ctx->down_stream));
    if (cc) VS_CUDA(ctx, cudaMemcpyAsync(D.co_host, ctx->defer_co[gen].p, cc * D.co_entry, cudaMemcpyDeviceToHost, ctx->down_stream));
    VS_CUDA(ctx, cudaStreamSynchronize(ctx->down_stream));
    if (nk > D.kp_cap || nc > D.co_cap)
        VS_FAIL(ctx, VSLAM_E_CAPACITY, "deferred call: lists hold %d / %d entries, %lld / %lld produced", D.kp_cap, D.co_cap, nk, nc);
    return VSLAM_OK;
}

int vslam_dog_harris(vslam_ctx *ctx, const uint8_t *frames, int n_frames, int rows, int cols, size_t pitch, size_t frame_stride, int space,
                     const vslam_pyramid_params *pparams, const vslam_dog_params *dparams, const vslam_harris_params *hparams,
                     vslam_keypoint *keypoints, int kp_capacity, int *kp_counts, vslam_corner *corners, int co_capacity, int *co_counts,
                     int out_space) {
    if (!ctx || kp_capacity < 0 || co_capacity < 0 || (!keypoints && kp_capacity > 0) || (!corners && co_capacity > 0)) return VSLAM_E_INVALID;
    VS_CUDA(ctx, cudaSetDevice(ctx->device));
    vslam_dog_params dp;
    VS_TRY(check_dog_params(ctx, dparams, &dp));
    const bool defer = (out_space & VSLAM_DEFER) != 0;
    out_space &= ~VSLAM_DEFER;
    const int kp_space = out_space & ~VSLAM_COMPACT;
    if (defer && kp_space != VSLAM_HOST) VS_FAIL(ctx, VSLAM_E_INVALID, "VSLAM_DEFER goes with host lists (device lists are never waited for)");
    if (defer && (!kp_counts || !co_counts || !ctx->pinned)) VS_FAIL(ctx, VSLAM_E_INVALID, "VSLAM_DEFER needs both count arrays");
    const size_t h_spf = (size_t)(rows > 0 ? rows : 1) * harris_strips(cols > 0 ? cols : 1);
    Geometry g;
    int no; double s0;
    VS_TRY(make_geometry(ctx, rows, cols, pparams, &g, &no, &s0));
    VS_TRY(ensure_taps(ctx, no, s0, 3));
    if (dp.window != 3 || ctx->any_wide || (size_t)n_frames > ((size_t)1 << 30) / h_spf || n_frames > 32768) {
        if (defer) VS_FAIL(ctx, VSLAM_E_INVALID, "VSLAM_DEFER: window 3, kernels of at most 247 taps and at most 32768 frames per call");
        // shapes the interleaved loop does not take: the two calls one after the other (same results)
        const int rk = vslam_dog_keypoints(ctx, frames, n_frames, rows, cols, pitch, frame_stride, space, pparams, dparams, keypoints, kp_capacity,
                                           kp_space, kp_counts);
        if (rk != VSLAM_OK && rk != VSLAM_E_CAPACITY) return rk;
        const int rc = vslam_harris(ctx, frames, n_frames, rows, cols, pitch, frame_stride, space, hparams, nullptr, nullptr, nullptr, corners,
                                    co_capacity, out_space, co_counts);
        return rk != VSLAM_OK ? rk : rc;
    }
    const u8 *d_frames; size_t d_pitch, d_stride;
    VS_TRY(ingest_frames(ctx, frames, n_frames, rows, cols, pitch, frame_stride, space, &d_frames, &d_pitch, &d_stride));
    VS_TRY(reserve_counters(ctx, n_frames));
    if (defer) return dog_harris_deferred(ctx, g, d_frames, d_pitch, d_stride, n_frames, rows, cols, dp, hparams, keypoints, kp_capacity, kp_counts,
                                          corners, co_capacity, co_counts, out_space);
    vslam_keypoint *d_kp = keypoints;
    if (kp_space != VSLAM_DEVICE) {
        if (ctx->out_dev.reserve(sizeof(vslam_keypoint) * (size_t)(kp_capacity + 1))) VS_FAIL(ctx, VSLAM_E_NOMEM, "device keypoint list");
        d_kp = ctx->out_dev.as<vslam_keypoint>();
    }
    HarrisJob hj;
    {
        ListSetSwap sw(ctx);
        VS_TRY(harris_begin(ctx, hj, d_frames, d_pitch, d_stride, n_frames, rows, cols, hparams, nullptr, nullptr, nullptr, corners, co_capacity,
                            out_space, n_frames, kDrainMax));
    }
    const std::function<int(int, int)> hook = [&](int f0, int nf) -> int {
        ListSetSwap sw(ctx);
        return harris_chunk(ctx, hj, f0, nf);
    };
    ListDrain kdrain;
    VS_TRY(dog_keypoints_fused(ctx, g, d_frames, n_frames, rows, cols, d_pitch, d_stride, dp, d_kp, kp_capacity, &kdrain, keypoints, kp_space, &hook));
    VS_TRY(consume_end(ctx));
    long long total = 0;
    VS_TRY(finish_counts(ctx, n_frames, kp_capacity, kp_counts, &total));
    VS_TRY(drain_end(ctx, kdrain));
    const int rk = emit_list(ctx, keypoints, kp_space, d_kp, sizeof(vslam_keypoint), total, kp_capacity, kdrain.on);
    if (rk != VSLAM_OK && rk != VSLAM_E_CAPACITY) return rk;
    int rc;
    {
        ListSetSwap sw(ctx);
        rc = harris_end(ctx, hj, co_counts);
    }
    return rk != VSLAM_OK ? rk : rc;
}

int vslam_upload_frames(vslam_ctx *ctx, const uint8_t *frames, int n_frames, int rows, int cols, size_t pitch, size_t frame_stride,
                        const uint8_t **d_frames, size_t *d_pitch, size_t *d_frame_stride) {
    if (!ctx || !d_frames || !d_pitch || !d_frame_stride) return VSLAM_E_INVALID;
    VS_CUDA(ctx, cudaSetDevice(ctx->device));
    if (rows <= 0 || cols <= 0) VS_FAIL(ctx, VSLAM_E_INVALID, "image size %dx%d", rows, cols);
    VS_TRY(ingest_frames(ctx, frames, n_frames, rows, cols, pitch, frame_stride, VSLAM_HOST, d_frames, d_pitch, d_frame_stride));
    ctx->up_chunks = 0;                                          // the consumers re-arm the waits from the returned pointer
    if (ctx->consuming >= 0) { ctx->slot_busy[ctx->consuming] = false; ctx->consuming = -1; }   // nobody reads the new batch yet
    return VSLAM_OK;
}

// ---- per-function entry points (host buffers; the reference's own function granularity, stages.cu) --------------------
// Scratch: ctx->maps_dev carved into 256-byte aligned regions.
static int stage_scratch(vslam_ctx *ctx, const size_t *bytes, int n, u8 **ptr) {
    size_t tot = 0;
    for (int i = 0; i < n; ++i) tot += round_up(bytes[i], 256);
    if (ctx->maps_dev.reserve(tot)) VS_FAIL(ctx, VSLAM_E_NOMEM, "stage scratch (%zu bytes)", tot);
    u8 *p = ctx->maps_dev.as<u8>();
    for (int i = 0; i < n; ++i) { ptr[i] = p; p += round_up(bytes[i], 256); }
    return VSLAM_OK;
}
#define H2D(dst, src, n) VS_CUDA(ctx, cudaMemcpyAsync(dst, src, n, cudaMemcpyHostToDevice, ctx->stream))
#define D2H(dst, src, n) VS_CUDA(ctx, cudaMemcpyAsync(dst, src, n, cudaMemcpyDeviceToHost, ctx->stream))
static int stage_check(vslam_ctx *ctx, const void *a, const void *b, int rows, int cols) {
    if (!ctx) return VSLAM_E_INVALID;
    if (!a || !b) VS_FAIL(ctx, VSLAM_E_INVALID, "null buffer");
    if (rows <= 0 || cols <= 0 || rows > (1 << 15) || cols > (1 << 15)) VS_FAIL(ctx, VSLAM_E_INVALID, "image size %dx%d", rows, cols);
    VS_CUDA(ctx, cudaSetDevice(ctx->device));
    return VSLAM_OK;
}

int vslam_bgr_to_gray(vslam_ctx *ctx, const uint8_t *bgr, int n_frames, int rows, int cols, size_t pitch, size_t frame_stride,
                      uint8_t *gray, size_t gray_pitch, size_t gray_stride) {
    if (!ctx || !bgr || !gray || n_frames <= 0 || rows <= 0 || cols <= 0) return VSLAM_E_INVALID;
    VS_CUDA(ctx, cudaSetDevice(ctx->device));
    if (n_frames > 65535 || rows > 65535) VS_FAIL(ctx, VSLAM_E_INVALID, "bgr_to_gray: %d frames of %d rows per call", n_frames, rows);
    if (pitch < 3 * (size_t)cols || gray_pitch < (size_t)cols) VS_FAIL(ctx, VSLAM_E_INVALID, "pitch smaller than a row");
    if (n_frames > 1 && (frame_stride < pitch * (size_t)rows || gray_stride < gray_pitch * (size_t)rows)) VS_FAIL(ctx, VSLAM_E_INVALID, "frame stride too small");
    const size_t sp = round_up(3 * (size_t)cols, 4), ss = sp * rows, dp = round_up((size_t)cols, 4), ds = dp * rows;
    const size_t b[2] = {ss * n_frames, ds * n_frames};
    u8 *d[2];
    VS_TRY(stage_scratch(ctx, b, 2, d));
    for (int f = 0; f < n_frames; ++f)
        VS_CUDA(ctx, cudaMemcpy2DAsync(d[0] + ss * f, sp, bgr + frame_stride * f, pitch, 3 * (size_t)cols, rows, cudaMemcpyHostToDevice, ctx->stream));
    VS_TRY(launch_bgr2gray(ctx, d[0], rows, cols, sp, ss, d[1], dp, ds, n_frames));
    for (int f = 0; f < n_frames; ++f)
        VS_CUDA(ctx, cudaMemcpy2DAsync(gray + gray_stride * f, gray_pitch, d[1] + ds * f, dp, cols, rows, cudaMemcpyDeviceToHost, ctx->stream));
    return vslam_sync(ctx);
}

static int rotate_common(vslam_ctx *ctx, const uint8_t *src, int rows, int cols, size_t pitch, int cx, int cy, float c, float s, int x0,
                         int y0, int ow, int oh, uint8_t *dst, size_t dst_pitch, int mode) {
    VS_TRY(stage_check(ctx, src, dst, rows, cols));
    if (pitch < (size_t)cols || dst_pitch < (size_t)ow || ow <= 0 || oh <= 0 || oh > 65535) VS_FAIL(ctx, VSLAM_E_INVALID, "rotate: bad pitch or window");
    const size_t b[2] = {(size_t)rows * cols, (size_t)ow * oh};
    u8 *d[2];
    VS_TRY(stage_scratch(ctx, b, 2, d));
    VS_CUDA(ctx, cudaMemcpy2DAsync(d[0], cols, src, pitch, cols, rows, cudaMemcpyHostToDevice, ctx->stream));
    VS_TRY(launch_rotate_gather(ctx, d[0], rows, cols, cols, cx, cy, c, s, x0, y0, ow, oh, d[1], ow, mode));
    VS_CUDA(ctx, cudaMemcpy2DAsync(dst, dst_pitch, d[1], ow, ow, oh, cudaMemcpyDeviceToHost, ctx->stream));
    return vslam_sync(ctx);
}

int vslam_rotate_mat_ccw(vslam_ctx *ctx, const uint8_t *src, int rows, int cols, size_t pitch, int center_x, int center_y, float cos_a,
                         float sin_a, uint8_t *dst, size_t dst_pitch) {
    return rotate_common(ctx, src, rows, cols, pitch, center_x, center_y, cos_a, sin_a, 0, 0, cols, rows, dst, dst_pitch, 0);
}

int vslam_rotated_window(vslam_ctx *ctx, const uint8_t *src, int rows, int cols, size_t pitch, int center_x, int center_y, int window,
                         float cos_a, float sin_a, uint8_t *dst, size_t dst_pitch) {
    if (window <= 0) return VSLAM_E_INVALID;
    const int pad = window / 2;
    return rotate_common(ctx, src, rows, cols, pitch, center_x, center_y, cos_a, sin_a, center_x - pad, center_y - pad, window, window, dst,
                         dst_pitch, 1);
}

int vslam_get_rotation_matrix_2d(float center_x, float center_y, double angle_deg, double scale, double *m) {
    if (!m) return VSLAM_E_INVALID;
    double angle = angle_deg;
    angle *= 3.1415926535897932384626433832795 / 180;                        // CV_PI / 180
    const double alpha = cos(angle) * scale, beta = sin(angle) * scale;
    m[0] = alpha; m[1] = beta; m[2] = (1 - alpha) * center_x - beta * center_y;
    m[3] = -beta; m[4] = alpha; m[5] = beta * center_x + (1 - alpha) * center_y;
    return VSLAM_OK;
}

int vslam_warp_affine(vslam_ctx *ctx, const uint8_t *src, int rows, int cols, size_t pitch, const double *m, uint8_t *dst, int dst_rows,
                      int dst_cols, size_t dst_pitch) {
    VS_TRY(stage_check(ctx, src, dst, rows, cols));
    if (!m || dst_rows <= 0 || dst_cols <= 0 || dst_rows > 65535 || dst_cols > (1 << 15)) VS_FAIL(ctx, VSLAM_E_INVALID, "warp_affine: destination %dx%d", dst_rows, dst_cols);
    if (pitch < (size_t)cols || dst_pitch < (size_t)dst_cols) VS_FAIL(ctx, VSLAM_E_INVALID, "warp_affine: pitch smaller than a row");
    // imgwarp.cpp: without WARP_INVERSE_MAP the matrix is inverted first (doubles, each operation on its own)
    double M[6];
    for (int i = 0; i < 6; ++i) M[i] = m[i];
    double D = M[0] * M[4] - M[1] * M[3];
    D = D != 0 ? 1. / D : 0;
    const double A11 = M[4] * D, A22 = M[0] * D;
    M[0] = A11; M[1] *= -D; M[3] *= -D; M[4] = A22;
    const double b1 = -M[0] * M[2] - M[1] * M[5], b2 = -M[3] * M[2] - M[4] * M[5];
    M[2] = b1; M[5] = b2;
    const size_t b[2] = {(size_t)rows * cols, (size_t)dst_rows * dst_cols};
    u8 *d[2];
    VS_TRY(stage_scratch(ctx, b, 2, d));
    VS_CUDA(ctx, cudaMemcpy2DAsync(d[0], cols, src, pitch, cols, rows, cudaMemcpyHostToDevice, ctx->stream));
    VS_TRY(launch_warp_affine(ctx, d[0], rows, cols, cols, M, d[1], dst_rows, dst_cols, dst_cols));
    VS_CUDA(ctx, cudaMemcpy2DAsync(dst, dst_pitch, d[1], dst_cols, dst_cols, dst_rows, cudaMemcpyDeviceToHost, ctx->stream));
    return vslam_sync(ctx);
}

int vslam_gaussian_blur3(vslam_ctx *ctx, const uint8_t *src, int rows, int cols, size_t pitch, uint8_t *dst, size_t dst_pitch) {
    VS_TRY(stage_check(ctx, src, dst, rows, cols));
    if (pitch < (size_t)cols || dst_pitch < (size_t)cols) VS_FAIL(ctx, VSLAM_E_INVALID, "pitch smaller than cols");
    const size_t n = (size_t)rows * cols, b[2] = {n, n};
    u8 *d[2];
    VS_TRY(stage_scratch(ctx, b, 2, d));
    VS_CUDA(ctx, cudaMemcpy2DAsync(d[0], cols, src, pitch, cols, rows, cudaMemcpyHostToDevice, ctx->stream));
    VS_TRY(launch_blur3(ctx, d[0], rows, cols, cols, d[1], cols));
    VS_CUDA(ctx, cudaMemcpy2DAsync(dst, dst_pitch, d[1], cols, cols, rows, cudaMemcpyDeviceToHost, ctx->stream));
    return vslam_sync(ctx);
}

int vslam_sobel1(vslam_ctx *ctx, const uint8_t *src, int rows, int cols, size_t pitch, float *gx, float *gy) {
    VS_TRY(stage_check(ctx, src, gx ? (void *)gx : (void *)gy, rows, cols));
    if (pitch < (size_t)cols) VS_FAIL(ctx, VSLAM_E_INVALID, "pitch smaller than cols");
    const size_t n = (size_t)rows * cols, b[3] = {n, n * 4, n * 4};
    u8 *d[3];
    VS_TRY(stage_scratch(ctx, b, 3, d));
    VS_CUDA(ctx, cudaMemcpy2DAsync(d[0], cols, src, pitch, cols, rows, cudaMemcpyHostToDevice, ctx->stream));
    VS_TRY(launch_sobel1(ctx, d[0], rows, cols, cols, gx ? (float *)d[1] : nullptr, gy ? (float *)d[2] : nullptr));
    if (gx) D2H(gx, d[1], n * 4);
    if (gy) D2H(gy, d[2], n * 4);
    return vslam_sync(ctx);
}

int vslam_harris_response(vslam_ctx *ctx, const float *ix, const float *iy, int rows, int cols, int window, float k, float *response) {
    VS_TRY(stage_check(ctx, ix, iy, rows, cols));
    if (!response || window < 1 || !(window & 1) || window > 31) VS_FAIL(ctx, VSLAM_E_INVALID, "window %d", window);
    const size_t n = (size_t)rows * cols * 4, b[3] = {n, n, n};
    u8 *d[3];
    VS_TRY(stage_scratch(ctx, b, 3, d));
    H2D(d[0], ix, n); H2D(d[1], iy, n);
    VS_TRY(launch_harris_grad(ctx, (float *)d[0], (float *)d[1], rows, cols, (window - 1) / 2, k, (float *)d[2]));
    D2H(response, d[2], n);
    return vslam_sync(ctx);
}

int vslam_convert_scale_abs(vslam_ctx *ctx, const float *src, int rows, int cols, uint8_t *dst) {
    VS_TRY(stage_check(ctx, src, dst, rows, cols));
    const size_t n = (size_t)rows * cols, b[2] = {n * 4, n};
    u8 *d[2];
    VS_TRY(stage_scratch(ctx, b, 2, d));
    H2D(d[0], src, n * 4);
    VS_TRY(launch_cvtabs(ctx, (float *)d[0], n, d[1]));
    D2H(dst, d[1], n);
    return vslam_sync(ctx);
}

int vslam_nms3(vslam_ctx *ctx, const uint8_t *r8, int rows, int cols, uint8_t *mask) {
    VS_TRY(stage_check(ctx, r8, mask, rows, cols));
    const size_t n = (size_t)rows * cols, b[2] = {n, n};
    u8 *d[2];
    VS_TRY(stage_scratch(ctx, b, 2, d));
    H2D(d[0], r8, n);
    VS_TRY(launch_nms3(ctx, d[0], rows, cols, d[1]));
    D2H(mask, d[1], n);
    return vslam_sync(ctx);
}

int vslam_nms2(vslam_ctx *ctx, const float *response, int rows, int cols, int window, float *out_map) {
    VS_TRY(stage_check(ctx, response, out_map, rows, cols));
    if (window < 1 || !(window & 1) || window > 31) VS_FAIL(ctx, VSLAM_E_INVALID, "window %d", window);
    const size_t n = (size_t)rows * cols * 4, b[2] = {n, n};
    u8 *d[2];
    VS_TRY(stage_scratch(ctx, b, 2, d));
    H2D(d[0], response, n);
    VS_TRY(launch_nms2(ctx, (float *)d[0], rows, cols, (window - 1) / 2, (float *)d[1]));
    D2H(out_map, d[1], n);
    return vslam_sync(ctx);
}

int vslam_dog_detect_octave(vslam_ctx *ctx, const vslam_pyramid *pyr, int frame, int octave, const vslam_dog_params *params,
                            int apply_edge, vslam_keypoint *out, int capacity, int *count) {
    if (!ctx || !pyr || capacity < 0 || (!out && capacity > 0) || !count) return VSLAM_E_INVALID;
    if (frame < 0 || frame >= pyr->n_frames || octave < 0 || octave >= pyr->num_octaves) VS_FAIL(ctx, VSLAM_E_INVALID, "frame %d / octave %d out of range", frame, octave);
    VS_CUDA(ctx, cudaSetDevice(ctx->device));
    vslam_dog_params dp;
    VS_TRY(check_dog_params(ctx, params, &dp));
    Geometry g;
    g.num_octaves = pyr->num_octaves; g.frame_bytes = pyr->frame_bytes;
    for (int o = 0; o < g.num_octaves; ++o) {
        g.orows[o] = pyr->orows[o]; g.ocols[o] = pyr->ocols[o]; g.opitch[o] = pyr->opitch[o]; g.ooff[o] = pyr->ooff[o];
        g.oplane[o] = pyr->oplane[o];
    }
    int seg_base[16];
    const size_t n_seg = detect_segments(g, dp.window, seg_base);
    VS_TRY(seg_prepare(ctx, n_seg ? n_seg : 1, sizeof(vslam_keypoint), staging_entries(capacity, n_seg * 256)));
    VS_TRY(reserve_counters(ctx, 1));
    if (ctx->out_dev.reserve(sizeof(vslam_keypoint) * (size_t)(capacity + 1))) VS_FAIL(ctx, VSLAM_E_NOMEM, "device keypoint list");
    vslam_keypoint *d_out = ctx->out_dev.as<vslam_keypoint>();
    VS_TRY(detect_frame(ctx, g, pyr->maps.as<u8>() + pyr->frame_bytes * frame, dp, seg_base, n_seg, d_out, capacity, 0, nullptr, octave, apply_edge));
    long long total = 0;
    VS_TRY(finish_counts(ctx, 1, capacity, count, &total));
    return emit_list(ctx, out, VSLAM_HOST, d_out, sizeof(vslam_keypoint), total, capacity);
}

int vslam_edge_responses(vslam_ctx *ctx, const vslam_pyramid *pyr, int frame, int octave, const vslam_keypoint *keypoints, int n,
                         float *responses) {
    if (!ctx || !pyr || n < 0 || (n > 0 && (!keypoints || !responses))) return VSLAM_E_INVALID;
    if (frame < 0 || frame >= pyr->n_frames || octave < 0 || octave >= pyr->num_octaves) VS_FAIL(ctx, VSLAM_E_INVALID, "frame %d / octave %d out of range", frame, octave);
    if (n == 0) return VSLAM_OK;
    VS_CUDA(ctx, cudaSetDevice(ctx->device));
    const size_t b[3] = {sizeof(vslam_keypoint) * (size_t)n, 4 * (size_t)n, 6 * sizeof(void *)};
    u8 *d[3];
    VS_TRY(stage_scratch(ctx, b, 3, d));
    const u8 *gl[6];
    for (int l = 0; l < 6; ++l) gl[l] = pyr->plane(frame, octave, 1 + l);
    H2D(d[0], keypoints, b[0]);
    H2D(d[2], gl, b[2]);
    VS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));           // gl lives on this stack frame
    VS_TRY(launch_edge_list(ctx, (const vslam_keypoint *)d[0], n, (const u8 *const *)d[2], pyr->orows[octave], pyr->ocols[octave],
                            pyr->opitch[octave], (float *)d[1]));
    D2H(responses, d[1], b[1]);
    return vslam_sync(ctx);
}

int vslam_gauss_ksize_f32(double sigma) { return sigma > 0 ? ((int)nearbyint(sigma * 4 * 2 + 1) | 1) : -1; }

int vslam_gauss_taps_f32(double sigma, int ksize, float *taps) {
    if (!taps || ksize <= 0 || !(ksize & 1) || !(sigma > 0)) return VSLAM_E_INVALID;
    std::vector<double> t((size_t)ksize);
    const double scale2x = -0.5 / (sigma * sigma);
    double sum = 0;
    for (int i = 0; i < ksize; ++i) {
        const double x = i - (ksize - 1) * 0.5;
        t[i] = exp(scale2x * x * x);
        sum += t[i];
    }
    sum = 1. / sum;
    for (int i = 0; i < ksize; ++i) taps[i] = (float)(t[i] * sum);
    return VSLAM_OK;
}

// Kernel-side description of the pyramid + the tap tables of sigma = 1.5 * getSigmaAt(o, l) (Diff_of_Gauss.cpp:346, :610) for the
// (octave, level) pairs the list uses; octave >= 0 reads every keypoint in that octave.
static int desc_pyr(vslam_ctx *ctx, const vslam_pyramid *pyr, int frame, int octave, const vslam_keypoint *kp, long long n, std::vector<float> &h_taps,
                    DescPyr &P, int &max_r) {
    if (pyr->num_octaves > kDescMaxOctaves) VS_FAIL(ctx, VSLAM_E_INVALID, "descriptor stage supports up to %d octaves", kDescMaxOctaves);
    bool used[kDescMaxOctaves][6] = {};
    for (long long i = 0; i < n; ++i) {
        const int o = octave >= 0 ? octave : kp[i].octave;
        if (o >= 0 && o < pyr->num_octaves && kp[i].level >= 0 && kp[i].level < 6) used[o][kp[i].level] = true;
    }
    memset(&P, 0, sizeof P);
    P.maps = pyr->maps.as<u8>(); P.frame_bytes = pyr->frame_bytes; P.num_octaves = pyr->num_octaves; P.octave = octave; P.frame0 = frame;
    max_r = 0;
    h_taps.clear();
    for (int o = 0; o < pyr->num_octaves; ++o) {
        P.ooff[o] = pyr->ooff[o]; P.oplane[o] = pyr->oplane[o]; P.opitch[o] = pyr->opitch[o]; P.orows[o] = pyr->orows[o]; P.ocols[o] = pyr->ocols[o];
        for (int l = 0; l < 6; ++l) {
            if (!used[o][l]) continue;
            const double sigma = 1.5 * pyr->sigma[o][l];
            const int ks = vslam_gauss_ksize_f32(sigma);
            if (ks <= 0) VS_FAIL(ctx, VSLAM_E_INVALID, "sigma of octave %d level %d is not positive", o, l);
            P.tap_off[o][l] = (int)h_taps.size();
            P.ksize[o][l] = ks;
            h_taps.resize(h_taps.size() + ks);
            vslam_gauss_taps_f32(sigma, ks, h_taps.data() + P.tap_off[o][l]);
            max_r = ks / 2 > max_r ? ks / 2 : max_r;
        }
    }
    if (h_taps.empty()) h_taps.push_back(1.f);
    return VSLAM_OK;
}

// orientation stage of n keypoints (frame_of: host list of their frames or nullptr); masks[i] bit b = bin b is above 0.8 of the peak
static int run_orientations(vslam_ctx *ctx, const vslam_pyramid *pyr, int frame, int octave, const vslam_keypoint *kp, const int *frame_of, int n,
                            std::vector<unsigned long long> &masks, float *histograms) {
    std::vector<float> h_taps;
    DescPyr P;
    int max_r = 0;
    VS_TRY(desc_pyr(ctx, pyr, frame, octave, kp, n, h_taps, P, max_r));
    // Keypoints are ordered widest kernel first (up to 309 taps at octave 3 of the default pyramid, 25 at octave 0) and
    // launched in width classes, each with the shared memory its widest member needs: the narrow classes, which hold most of
    // the keypoints, then fit many more CTAs per SM than a launch sized for the widest kernel would allow
    std::vector<std::pair<int, int>> by_r((size_t)n);                      // (-r, index): stable order inside one width
    std::vector<int> order((size_t)n);
    {   // counting sort on the radius (a dozen distinct values): the list order is kept inside one width
        std::vector<int> rad((size_t)n);
        int rmax = 0;
        for (int i = 0; i < n; ++i) {
            const int o = octave >= 0 ? octave : kp[i].octave;
            const int ks = (o >= 0 && o < pyr->num_octaves && kp[i].level >= 0 && kp[i].level < 6) ? P.ksize[o][kp[i].level] : 1;
            rad[(size_t)i] = ks / 2;
            rmax = ks / 2 > rmax ? ks / 2 : rmax;
        }
        std::vector<int> start((size_t)rmax + 2, 0);
        for (int i = 0; i < n; ++i) ++start[(size_t)(rmax - rad[(size_t)i]) + 1];          // bucket = rmax - r: widest first
        for (int b2 = 0; b2 <= rmax; ++b2) start[(size_t)b2 + 1] += start[(size_t)b2];
        for (int i = 0; i < n; ++i) {
            const int pos = start[(size_t)(rmax - rad[(size_t)i])]++;
            by_r[(size_t)pos] = {-rad[(size_t)i], i};
            order[(size_t)pos] = i;
        }
    }
    const size_t b[6] = {sizeof(vslam_keypoint) * (size_t)n, 8 * (size_t)n, histograms ? 144 * (size_t)n : 0, 4 * h_taps.size(), frame_of ? 4 * (size_t)n : 0,
                         4 * (size_t)n};
    u8 *d[6];
    VS_TRY(stage_scratch(ctx, b, 6, d));
    H2D(d[0], kp, b[0]);
    H2D(d[3], h_taps.data(), b[3]);
    if (frame_of) H2D(d[4], frame_of, b[4]);
    H2D(d[5], order.data(), b[5]);
    P.taps = (const float *)d[3];
    const int *d_frame = frame_of ? (const int *)d[4] : nullptr;
    static const int kClassR[] = {80, 40, 20, 0};                          // class c holds the radii above kClassR[c]
    for (int c = 0, i0 = 0; c < 4 && i0 < n; ++c) {
        int i1 = i0;
        while (i1 < n && -by_r[(size_t)i1].first > kClassR[c]) ++i1;
        if (c == 3) i1 = n;
        if (i1 > i0)
            VS_TRY(launch_orient_hist(ctx, (const vslam_keypoint *)d[0], d_frame, (const int *)d[5] + i0, i1 - i0, P, -by_r[(size_t)i0].first,
                                      (unsigned long long *)d[1], histograms ? (float *)d[2] : nullptr));
        i0 = i1;
    }
    (void)max_r;
    masks.resize((size_t)n);
    D2H(masks.data(), d[1], b[1]);
    if (histograms) D2H(histograms, d[2], b[2]);
    return vslam_sync(ctx);                                                // h_taps outlives the copies
}

static int run_sift(vslam_ctx *ctx, const vslam_pyramid *pyr, int frame, int octave, const vslam_keypoint *kp, const int *frame_of, int n,
                    float *descriptors, uint8_t *valid) {
    std::vector<float> h_taps;
    DescPyr P;
    int max_r = 0;
    VS_TRY(desc_pyr(ctx, pyr, frame, octave, kp, n, h_taps, P, max_r));
    std::vector<float2> cs((size_t)n);
    {   // Rotation::cos_sin_of_angle (rotation.cpp:9-17); the orientation stage only emits bin * 10 degrees: 36 distinct angles
        auto cos_sin = [](int value) {
            const float angle = (float)((double)(float)value * (3.1415926535897932384626433832795 / (double)180.0f));
            return make_float2((float)cos((double)angle), (float)sin((double)angle));
        };
        float2 lut[36];
        for (int b2 = 0; b2 < 36; ++b2) lut[b2] = cos_sin(10 * b2);
        for (int i = 0; i < n; ++i) {
            const int v = kp[i].value;
            cs[(size_t)i] = (v >= 0 && v < 360 && v % 10 == 0) ? lut[v / 10] : cos_sin(v);
        }
    }
    const size_t b[6] = {sizeof(vslam_keypoint) * (size_t)n, 8 * (size_t)n, 512 * (size_t)n, (size_t)n, 4 * h_taps.size(), frame_of ? 4 * (size_t)n : 0};
    u8 *d[6];
    VS_TRY(stage_scratch(ctx, b, 6, d));
    H2D(d[0], kp, b[0]);
    H2D(d[1], cs.data(), b[1]);
    H2D(d[4], h_taps.data(), b[4]);
    if (frame_of) H2D(d[5], frame_of, b[5]);
    P.taps = (const float *)d[4];
    VS_TRY(launch_sift(ctx, (const vslam_keypoint *)d[0], frame_of ? (const int *)d[5] : nullptr, (const float2 *)d[1], n, P, max_r, (float *)d[2], d[3]));
    D2H(descriptors, d[2], b[2]);
    D2H(valid, d[3], b[3]);
    return vslam_sync(ctx);                                                // the host vectors above outlive the copies
}

// :362-368: bins in ascending order per keypoint; returns the number of points (only the first `capacity` are stored)
static long long expand_orientations(const vslam_keypoint *kp, const unsigned long long *masks, long long n, vslam_keypoint *out, long long capacity) {
    long long m = 0;
    for (long long i = 0; i < n; ++i)
        for (unsigned long long bits = masks[i] & ((1ull << 36) - 1); bits; bits &= bits - 1) {      // set bits in ascending order
            const int bin = __builtin_ctzll(bits);
            if (m < capacity) out[m] = vslam_keypoint{kp[i].row, kp[i].col, bin * 10, 0, kp[i].octave, kp[i].level};
            ++m;
        }
    return m;
}

int vslam_assign_orientations(vslam_ctx *ctx, const vslam_pyramid *pyr, int frame, int octave, const vslam_keypoint *keypoints, int n,
                              vslam_keypoint *out, int capacity, int *count, float *histograms) {
    if (!ctx || !pyr || !count || n < 0 || capacity < 0 || (n > 0 && !keypoints) || (capacity > 0 && !out)) return VSLAM_E_INVALID;
    if (frame < 0 || frame >= pyr->n_frames || octave < 0 || octave >= pyr->num_octaves) VS_FAIL(ctx, VSLAM_E_INVALID, "frame %d / octave %d out of range", frame, octave);
    *count = 0;
    if (n == 0) return VSLAM_OK;
    VS_CUDA(ctx, cudaSetDevice(ctx->device));
    std::vector<unsigned long long> masks;
    VS_TRY(run_orientations(ctx, pyr, frame, octave, keypoints, nullptr, n, masks, histograms));
    const long long m = expand_orientations(keypoints, masks.data(), n, out, capacity);
    if (m > 0x7fffffff) VS_FAIL(ctx, VSLAM_E_INVALID, "oriented keypoint count overflows int");
    *count = (int)m;
    return VSLAM_OK;
}

int vslam_sift_descriptors(vslam_ctx *ctx, const vslam_pyramid *pyr, int frame, int octave, const vslam_keypoint *oriented, int n,
                           float *descriptors, uint8_t *valid) {
    if (!ctx || !pyr || n < 0 || (n > 0 && (!oriented || !descriptors || !valid))) return VSLAM_E_INVALID;
    if (frame < 0 || frame >= pyr->n_frames || octave < 0 || octave >= pyr->num_octaves) VS_FAIL(ctx, VSLAM_E_INVALID, "frame %d / octave %d out of range", frame, octave);
    if (n == 0) return VSLAM_OK;
    VS_CUDA(ctx, cudaSetDevice(ctx->device));
    return run_sift(ctx, pyr, frame, octave, oriented, nullptr, n, descriptors, valid);
}

// Every frame and octave of a pyramid: orientation histograms in width classes, the expansion of their masks ON THE DEVICE (count,
// scan, scatter: the lists do not visit the host between the two kernels), SIFT in four launches whose descriptors leave on the
// drain stream while the next launch computes.  One small synchronisation in the middle (the number of oriented points sizes the grid).
int vslam_describe_keypoints(vslam_ctx *ctx, const vslam_pyramid *pyr, const vslam_keypoint *keypoints, const int *counts, vslam_keypoint *oriented,
                             int capacity, int *oriented_counts, float *descriptors, uint8_t *valid) {
    if (!ctx || !pyr || !counts || !oriented_counts || capacity < 0 || (capacity > 0 && (!oriented || !descriptors || !valid))) return VSLAM_E_INVALID;
    long long n64 = 0;
    for (int f = 0; f < pyr->n_frames; ++f) {
        if (counts[f] < 0) VS_FAIL(ctx, VSLAM_E_INVALID, "negative count for frame %d", f);
        n64 += counts[f];
        oriented_counts[f] = 0;
    }
    if (n64 > 0x7fffffff / 36) VS_FAIL(ctx, VSLAM_E_INVALID, "too many keypoints for one call (%lld)", n64);
    if (n64 == 0) return VSLAM_OK;
    if (!keypoints) return VSLAM_E_INVALID;
    VS_CUDA(ctx, cudaSetDevice(ctx->device));
    const int n = (int)n64;
    std::vector<int> frame_of((size_t)n);
    {
        long long i = 0;
        for (int f = 0; f < pyr->n_frames; ++f)
            for (int c = 0; c < counts[f]; ++c) frame_of[(size_t)i++] = f;
    }
    std::vector<float> h_taps;
    DescPyr P;
    int max_r = 0;
    VS_TRY(desc_pyr(ctx, pyr, 0, -1, keypoints, n, h_taps, P, max_r));
    // widest kernel first, list order inside one width (counting sort on the radius): the launch classes of run_orientations
    std::vector<int> order((size_t)n), rad((size_t)n);
    for (int i = 0; i < n; ++i) {
        const int o = keypoints[i].octave, l = keypoints[i].level;
        rad[(size_t)i] = (o >= 0 && o < pyr->num_octaves && l >= 0 && l < 6 ? P.ksize[o][l] : 1) / 2;
    }
    std::vector<int> start((size_t)max_r + 2, 0);
    for (int i = 0; i < n; ++i) ++start[(size_t)(max_r - rad[(size_t)i]) + 1];
    for (int b2 = 0; b2 <= max_r; ++b2) start[(size_t)b2 + 1] += start[(size_t)b2];
    const std::vector<int> bucket0(start.begin(), start.end());            // first position of radius max_r - b
    for (int i = 0; i < n; ++i) order[(size_t)start[(size_t)(max_r - rad[(size_t)i])]++] = i;
    OrientLut lut;
    for (int b2 = 0; b2 < 36; ++b2) {                                       // Rotation::cos_sin_of_angle (rotation.cpp:9-17) of bin * 10 degrees
        const float angle = (float)((double)(float)(10 * b2) * (3.1415926535897932384626433832795 / (double)180.0f));
        lut.cs[b2] = make_float2((float)cos((double)angle), (float)sin((double)angle));
    }
    const size_t cap = (size_t)capacity, nb = (size_t)n / 4096 + 8;
    const size_t b[14] = {sizeof(vslam_keypoint) * (size_t)n, 8 * (size_t)n, 4 * h_taps.size(), 4 * (size_t)n, 4 * (size_t)n, 4 * (size_t)n + 64,
                          4 * (size_t)n + 64, 4 * nb, 64, sizeof(vslam_keypoint) * cap, 8 * cap, 4 * cap, 512 * cap, cap};
    u8 *d[14];
    VS_TRY(stage_scratch(ctx, b, 14, d));
    const vslam_keypoint *d_kp = (const vslam_keypoint *)d[0];
    unsigned long long *d_masks = (unsigned long long *)d[1];
    const int *d_frame = (const int *)d[3], *d_order = (const int *)d[4];
    int *d_cnt = (int *)d[5], *d_pos = (int *)d[6], *d_partial = (int *)d[7], *d_total = (int *)d[8];
    vslam_keypoint *d_ori = (vslam_keypoint *)d[9];
    float2 *d_cs = (float2 *)d[10];
    int *d_oframe = (int *)d[11];
    float *d_desc = (float *)d[12];
    u8 *d_valid = d[13];
    H2D(d[0], keypoints, b[0]);
    H2D(d[2], h_taps.data(), b[2]);
    H2D(d[3], frame_of.data(), b[3]);
    H2D(d[4], order.data(), b[4]);
    P.taps = (const float *)d[2];
    static const int kClassR[] = {80, 40, 20, 0};                           // class c holds the radii above kClassR[c]
    for (int c = 0, i0 = 0; c < 4 && i0 < n; ++c) {
        // keypoints with radius > kClassR[c]: the buckets of radii max_r .. kClassR[c] + 1
        int i1 = c == 3 || kClassR[c] >= max_r ? (c == 3 ? n : i0) : bucket0[(size_t)(max_r - kClassR[c])];
        if (i1 < i0) i1 = i0;
        if (i1 > i0) VS_TRY(launch_orient_hist(ctx, d_kp, d_frame, d_order + i0, i1 - i0, P, rad[(size_t)order[(size_t)i0]], d_masks, nullptr));
        i0 = i1;
    }
    VS_TRY(launch_orient_expand(ctx, d_kp, d_frame, d_masks, n, capacity, lut, d_cnt, d_pos, d_partial, d_total, d_ori, d_cs, d_oframe, 0));
    int *h_total = (int *)ctx->pinned + 3000;                               // pinned: behind the drain slots (2 x 1024 ints), before the deferred snapshots (16 KB)
    if (!ctx->pinned) VS_FAIL(ctx, VSLAM_E_NOMEM, "pinned staging");
    VS_CUDA(ctx, cudaMemcpyAsync(h_total, d_total, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    VS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));                       // the host vectors above are free from here on
    const long long m = *h_total;
    const int stored = (int)(m < capacity ? m : capacity);
    std::vector<int> h_pos((size_t)n);
    if (stored > 0) {
        VS_TRY(launch_orient_expand(ctx, d_kp, d_frame, d_masks, n, capacity, lut, d_cnt, d_pos, d_partial, d_total, d_ori, d_cs, d_oframe, 1));
        if (!ctx->down_stream) VS_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->down_stream, cudaStreamNonBlocking));
        while (ctx->down_ev.size() < 8) {
            cudaEvent_t e;
            VS_CUDA(ctx, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
            ctx->down_ev.push_back(e);
        }
        VS_CUDA(ctx, cudaEventRecord(ctx->down_ev[0], ctx->stream));
        VS_CUDA(ctx, cudaStreamWaitEvent(ctx->down_stream, ctx->down_ev[0], 0));
        VS_CUDA(ctx, cudaMemcpyAsync(oriented, d_ori, sizeof(vslam_keypoint) * (size_t)stored, cudaMemcpyDeviceToHost, ctx->down_stream));
        const int parts = stored >= 4096 ? 4 : 1;
        for (int q = 0; q < parts; ++q) {
            const int i0 = (int)((long long)stored * q / parts), i1 = (int)((long long)stored * (q + 1) / parts);
            if (i1 <= i0) continue;
            VS_TRY(launch_sift(ctx, d_ori + i0, d_oframe + i0, d_cs + i0, i1 - i0, P, max_r, d_desc + (size_t)i0 * 128, d_valid + i0));
            VS_CUDA(ctx, cudaEventRecord(ctx->down_ev[1 + q], ctx->stream));
            VS_CUDA(ctx, cudaStreamWaitEvent(ctx->down_stream, ctx->down_ev[1 + q], 0));
            VS_CUDA(ctx, cudaMemcpyAsync(descriptors + (size_t)i0 * 128, d_desc + (size_t)i0 * 128, 512 * (size_t)(i1 - i0), cudaMemcpyDeviceToHost, ctx->down_stream));
            VS_CUDA(ctx, cudaMemcpyAsync(valid + i0, d_valid + i0, (size_t)(i1 - i0), cudaMemcpyDeviceToHost, ctx->down_stream));
        }
    }
    // per-frame counts from the scanned positions (while the SIFT launches run)
    VS_CUDA(ctx, cudaMemcpyAsync(h_pos.data(), d_pos, sizeof(int) * (size_t)n, cudaMemcpyDeviceToHost, ctx->stream));
    VS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    {
        long long i0 = 0;
        for (int f = 0; f < pyr->n_frames; ++f) {
            const long long i1 = i0 + counts[f];
            const long long lo = i0 < n ? h_pos[(size_t)i0] : m, hi = i1 < n ? h_pos[(size_t)i1] : m;
            oriented_counts[f] = (int)(hi - lo);
            i0 = i1;
        }
    }
    if (ctx->down_stream) VS_CUDA(ctx, cudaStreamSynchronize(ctx->down_stream));
    return m > capacity ? VSLAM_E_CAPACITY : VSLAM_OK;
}

int vslam_write_descriptor_file(const char *path, const float *descriptors, int n) {
    if (!path || n < 0 || (n > 0 && !descriptors)) return VSLAM_E_INVALID;
    FILE *f = fopen(path, "wb");
    if (!f) VS_FAIL((vslam_ctx *)nullptr, VSLAM_E_INVALID, "cannot write %s", path);
    const int32_t head[3] = {n, 128, 24};                                  // :842-848: frontSize = sizeof(std::vector<float>)
    bool ok = fwrite(head, 4, 3, f) == 3;
    if (n > 0) ok = ok && fwrite(descriptors, 512, (size_t)n, f) == (size_t)n;
    ok = (fclose(f) == 0) && ok;
    if (!ok) VS_FAIL((vslam_ctx *)nullptr, VSLAM_E_INVALID, "short write to %s", path);
    return VSLAM_OK;
}

int vslam_localize(vslam_ctx *ctx, const int32_t *candidates, int n, float *dog_zhat) {
    if (!ctx || n < 0 || (n > 0 && (!candidates || !dog_zhat))) return VSLAM_E_INVALID;
    if (n == 0) return VSLAM_OK;
    VS_CUDA(ctx, cudaSetDevice(ctx->device));
    const size_t b[2] = {16 * (size_t)n, 4 * (size_t)n};
    u8 *d[2];
    VS_TRY(stage_scratch(ctx, b, 2, d));
    H2D(d[0], candidates, b[0]);
    VS_TRY(launch_localize_list(ctx, (const int *)d[0], n, (float *)d[1]));
    D2H(dog_zhat, d[1], b[1]);
    return vslam_sync(ctx);
}

int vslam_edge_from_windows(vslam_ctx *ctx, const float *gx, const float *gy, int n, int window_len, float *responses, float *sums) {
    if (!ctx || n < 0 || window_len <= 0 || (n > 0 && (!gx || !gy || (!responses && !sums)))) return VSLAM_E_INVALID;
    if (n == 0) return VSLAM_OK;
    VS_CUDA(ctx, cudaSetDevice(ctx->device));
    const size_t w = 4 * (size_t)n * window_len, b[4] = {w, w, 4 * (size_t)n, 12 * (size_t)n};
    u8 *d[4];
    VS_TRY(stage_scratch(ctx, b, 4, d));
    H2D(d[0], gx, w); H2D(d[1], gy, w);
    VS_TRY(launch_edge_windows(ctx, (const float *)d[0], (const float *)d[1], n, window_len, responses ? (float *)d[2] : nullptr,
                               sums ? (float *)d[3] : nullptr));
    if (responses) D2H(responses, d[2], b[2]);
    if (sums) D2H(sums, d[3], b[3]);
    return vslam_sync(ctx);
}
#undef H2D
#undef D2H

}  // extern "C"
