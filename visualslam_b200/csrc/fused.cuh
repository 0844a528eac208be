// fused.cuh -- arguments of the fused octave kernel (fused.cu): rows B, C, D, G, H, I, J of SURVEY.md 8(a) in one pass.
#pragma once
#include "common.cuh"

namespace vslam {

constexpr int kFusedHWords = 64;    // ceil(245 / 4) = 62, padded to a multiple of 2 (+ slack)
constexpr int kFusedVPairs = 128;   // r + 1 <= 123, padded to a multiple of 2 (+ slack)

// One Gaussian level, packed for the two passes of octave_fused_k.
struct FusedLevel {
    int n, r;
    int nw;                         // H pass: tap words (4 byte taps each), padded to a multiple of 2
    int jp;                         // V pass: tap pairs, padded to a multiple of 2
    int hoff;                       // word offset of this level's h buffer inside the CTA's shared memory
    int np;                         // pairs of 8.8 rows the buffer holds: r + R/2
    u32 hq[kFusedHWords];           // taps t[4w .. 4w+3], zero padded
    u32 eo[kFusedVPairs];           // (t[2j], t[2j+1]) | (t[2j-1], t[2j]) << 16
};

struct FusedTaps {
    FusedLevel lv[6];
    int rmin, rmax, rpad;           // rpad >= rmax, rpad == 2 (mod 4): left halo of the staged base rows
    int stage_rows, stage_pitch;    // staged base tile: rows, bytes between rows (multiple of 16)
    int spq_magic;                  // ceil(2^20 / (stage_pitch / 16)): flattened index -> row without a division
    int hbuf_words;                 // total words of the six h buffers (+ spare pair rows); multiple of 4
};

// Launch description.  All planes of one octave share `pitch`; frame f's plane is at ptr + f * frame_stride.
struct OctaveIO {
    const u8 *base;                 // octave input
    size_t base_pitch, base_stride;
    int base_pad;                   // columns left and right of every base row that hold its reflect-101 continuation (0: none)
    int base_row_pad;               // rows above and below the plane that hold its reflect-101 continuation (0: none)
    int rows, cols, octave;
    int n_strips, n_segs, seg_rows; // grid = (strips, vertical segments, frames); seg_rows is a multiple of R
    // next octave's base: level 3 decimated by two (GaussPyramid.cpp:123-126), or nullptr
    u8 *next_base;
    size_t next_pitch, next_stride;
    int next_rows, next_cols;
    // optional materialised planes (GaussPyramid::octaveBlur / octaveDiff), or nullptr
    u8 *gauss[6];
    u8 *dog[5];
    size_t map_pitch, map_stride;
    // keypoint detection (window 3 only), or detect == 0
    int detect;
    float contrast_thr, edge_thr;
    int cmin;                       // zero-derivative contrast test: value/255 > contrast_thr  <=>  value >= cmin
    u8 val_lut[256];                // (int)((value/255f) * 255f), Diff_of_Gauss.cpp:246, for the same case
    int ni, nj;                     // sample rows / columns of this octave
    int seg_base;                   // first segment of this octave inside a frame's segment block
    int segs_per_frame;
    int *seg_off, *seg_cnt;
    int *cursor;                    // staging cursor (one per launch batch)
    vslam_keypoint *staging;
    int staging_cap;
};
struct FusedArgs : OctaveIO {
    FusedTaps taps;
};

// geometry shared by host and device
constexpr int kFusedW = 64;         // columns a CTA computes
constexpr int kFusedWO = 60;        // columns a CTA owns (2 halo columns each side; 60 = 20 sample columns of stride 3)
constexpr int kFusedR = 32;         // rows per marching step

int build_fused_taps(double sigma0, int scale_samples, int octave, FusedTaps *out);
size_t fused_smem_bytes(const FusedTaps &t);
int fused_strips(int cols);
// Launch one octave for n_frames frames.  The kernel is chosen by the shared-memory footprint of the octave.
int launch_octave_fused(vslam_ctx *ctx, FusedArgs &a, int n_frames);
// Tensor-pipe variant (fused_mma.cu): same results; used for keypoints-only launches whose tap tables fit.
// Returns 1 when the launch was not taken (caller falls back to launch_octave_fused), 0 on success, < 0 on error.
// strips per frame of the kernel that will run the octave on the detect path (fused_mma.cu)
int octave_strips(vslam_ctx *ctx, double sigma0, int scale_samples, int octave, int cols);
// reflect-101 columns the kernel of that octave wants next to every base row (a multiple of 16; 0 for the dp4a/dp2a kernel)
int octave_base_pad(vslam_ctx *ctx, double sigma0, int scale_samples, int octave);
// rows above / below the plane the kernel of that octave wants (0 for all but the tcgen05 kernel)
int octave_base_row_pad(vslam_ctx *ctx, double sigma0, int scale_samples, int octave);
// fills them: base[y][-pad .. -1] and base[y][cols .. cols + pad - 1] of every frame, then rows -row_pad .. -1 and
// rows .. rows + row_pad - 1 (whole padded rows, so the corners are reflections in both axes)
int launch_reflect_pad(vslam_ctx *ctx, u8 *base, int rows, int cols, size_t pitch, size_t stride, int pad, int row_pad, int n_frames, int pad_left = 0,
                       int pad_right = 0);
int octave_base_pad_left(vslam_ctx *ctx, double sigma0, int scale_samples, int octave);
int octave_base_pad_right(vslam_ctx *ctx, double sigma0, int scale_samples, int octave, int cols);
int launch_octave_mma(vslam_ctx *ctx, const OctaveIO &io, int n_frames, double sigma0, int scale_samples);
// tap-table images cached per device: freed when the last context of the device goes (vslam_ctx_destroy)
void mma_tab_release(int device);
void umma_tab_release(int device);
// tcgen05 / TMA variant (fused_umma.cu): same contract as launch_octave_mma; takes the octaves whose kernels have at most 63 taps
int launch_octave_umma(vslam_ctx *ctx, const OctaveIO &io, int n_frames, double sigma0, int scale_samples);
bool octave_uses_umma(vslam_ctx *ctx, double sigma0, int scale_samples, int octave);
int umma_base_pad(vslam_ctx *ctx, double sigma0, int scale_samples, int octave);

}  // namespace vslam
