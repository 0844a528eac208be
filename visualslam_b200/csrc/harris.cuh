// harris.cuh -- arguments of the fused Harris kernel (harris.cu)
#pragma once
#include "common.cuh"

namespace vslam {

struct HarrisArgs {
    const u8 *frame;
    int rows, cols;
    size_t pitch;
    float k;
    int nms2_pad;                 // (nms2_window - 1) / 2, 1 or 2
    float *response;              // dense rows x cols, or nullptr
    u8 *r8, *mask3;               // dense rows x cols, or nullptr
    int seg_base;                 // segment = seg_base + row * ntx + tile column
    int *seg_off, *seg_cnt;
    int *cursor;
    vslam_corner *staging;
    int staging_cap;
};

void harris_grid(int rows, int cols, int *ntx, int *nty);
int launch_harris(vslam_ctx *ctx, const HarrisArgs &a);

}  // namespace vslam
