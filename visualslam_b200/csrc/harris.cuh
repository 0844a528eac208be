// harris.cuh -- arguments of the fused Harris kernel (harris.cu)
#pragma once
#include "common.cuh"

namespace vslam {

constexpr int kHarrisStrip = 112;   // columns a warp owns (28 lanes x 4 pixels; 2 halo lanes each side)
constexpr int kHarrisBand = 120;    // rows a warp marches over (9 rows of pipeline warm-up per band)

struct HarrisArgs {
    const u8 *frames;             // frame f at frames + f * frame_stride
    size_t pitch, frame_stride;
    int rows, cols;
    float k;
    int nms2_pad;                 // (nms2_window - 1) / 2, 1 or 2
    float *response;              // dense n_frames x rows x cols, or nullptr
    u8 *r8, *mask3;               // dense n_frames x rows x cols, or nullptr
    int n_strips, n_bands;
    // ordered emission: segment = (frame, row, strip)
    int *seg_off, *seg_cnt;
    int *cursor;
    vslam_corner *staging;
    int staging_cap;
    int want_corners;
};

int harris_strips(int cols);
int launch_harris(vslam_ctx *ctx, HarrisArgs &a, int n_frames);

}  // namespace vslam
