// descriptor.cuh -- launchers of the orientation / SIFT descriptor kernels (descriptor.cu); device pointers throughout.
#pragma once
#include "common.cuh"

namespace vslam {
// the six levels of one octave of one frame: u8 Gaussians, and the fp32 blur taps of sigma = 1.5 * getSigmaAt(octave, level)
struct DescLevels {
    const u8 *gauss[6];
    const float *taps;     // the six tap tables back to back
    int tap_off[6];        // first tap of level l
    int ksize[6];          // cvRound(sigma * 4 * 2 + 1) | 1
};
size_t orient_smem_bytes(int r);
size_t sift_smem_bytes(int r);
int launch_orient_hist(vslam_ctx *ctx, const vslam_keypoint *kp, int n, const DescLevels &lv, int max_r, int rows, int cols, size_t pitch,
                       unsigned long long *masks, float *histos);
int launch_sift(vslam_ctx *ctx, const vslam_keypoint *kp, const float2 *cs, int n, const DescLevels &lv, int max_r, int rows, int cols,
                size_t pitch, float *desc, u8 *valid);
}  // namespace vslam
