// pyramid.cuh -- device-side description of one octave and the launchers in pyramid.cu
#pragma once
#include "common.cuh"

namespace vslam {

// u8 planes of one octave of one frame, all with the same pitch.
struct OctavePlanes {
    int rows, cols;
    size_t pitch;            // bytes between rows of every u8 plane (multiple of 128)
    u8 *base;                // octave input  (GaussPyramid::octaveImage)
    u8 *gauss[6];            // octaveBlur
    u8 *dog[5];              // octaveDiff
    u8 *next_base;           // level-3 Gaussian decimated by 2 (nullptr on the last octave)
    int next_rows, next_cols;
    size_t next_pitch;
};

inline size_t plane_pitch(int cols) { return round_up((size_t)cols, 128); }
inline size_t plane_bytes(int rows, int cols) { return plane_pitch(cols) * (size_t)rows; }
inline void half_size(int r, int c, int *orows, int *ocols) {
    // cv::resize dsize = saturate_cast<int>(size*0.5) = cvRound: round half to even (GaussPyramid.cpp:126)
    auto rne = [](int v) { int q = v / 2; return (v & 1) ? ((q & 1) ? q + 1 : q) : q; };
    *orows = rne(r); *ocols = rne(c);
}

// row A
// n_frames frames per launch: frame f at src + f * src_stride -> dst + f * dst_stride
int launch_upsample2x(vslam_ctx *ctx, const u8 *src, int rows, int cols, size_t src_pitch, size_t src_stride, u8 *dst,
                      size_t dst_pitch, size_t dst_stride, int n_frames);
// row E on demand: any output may be null.
int launch_gradients(vslam_ctx *ctx, const u8 *g, int rows, int cols, size_t pitch, float *gx, float *gy, float *mag,
                     float *orient, size_t out_pitch_elems);

// row E for every level of one octave of a batch into resident planes: per frame and level four dense fp32 planes
// [gx, gy, magnitude, orientation] of `oplane` floats; mask bit k selects plane k
int launch_gradients_all(vslam_ctx *ctx, const u8 *g, int rows, int cols, size_t pitch, size_t gplane, size_t gstride, float *out,
                         size_t oplane, size_t ostride, unsigned mask, int n_frames);

// exact global-memory path for octaves whose kernels are wider than the fused kernels take (pyramid.cu): every map of the octave for
// n_frames frames; d_taps: device ints, level l at d_taps + tap_off[l], ksize[l] taps; d_hsum: n_frames * rows * cols u16 of scratch
int launch_octave_wide(vslam_ctx *ctx, const OctavePlanes &op, size_t frame_stride, const int *d_taps, const int *tap_off, const int *ksize,
                       u16 *d_hsum, int n_frames);

}  // namespace vslam
