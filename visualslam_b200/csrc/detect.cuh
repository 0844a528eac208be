// detect.cuh -- arguments of the DoG detector kernel (detect.cu)
#pragma once
#include "common.cuh"

namespace vslam {

struct DetectArgs {
    const u8 *dog[5];
    const u8 *gauss[6];
    int rows, cols;
    size_t pitch;
    int octave;
    int window, pad;
    float contrast_thr, edge_thr;
    int apply_edge;                   // 0: stop after FeaturePointLocalization (initialKeypointDetection's own output)
    int ni, nj, nchunks;              // sample rows, sample columns, 256-sample chunks per row
    int seg_base;                     // first segment index of this octave
    int *seg_off, *seg_cnt;
    int *cursor;                      // staging cursor (atomic)
    vslam_keypoint *staging;
    int staging_cap;
    unsigned long long *stage_counts; // 4 counters of this (frame, octave) or nullptr
    // batched launches: grid.z = 3 levels x n_frames; frame f's planes, segments and counters follow frame 0's at these strides
    int n_frames;
    size_t frame_stride;              // bytes between the frames' plane blocks
    int segs_per_frame;
    int stage_stride;                 // counters (u64) between the frames' stage_counts
};

void detect_grid(int rows, int cols, int window, int *ni, int *nj, int *nchunks);
int launch_detect(vslam_ctx *ctx, DetectArgs &a, const vslam_dog_params &p);

}  // namespace vslam
