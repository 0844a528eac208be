// descriptor.cuh -- launchers of the orientation / SIFT descriptor kernels (descriptor.cu); device pointers throughout.
#pragma once
#include "common.cuh"

namespace vslam {
// Where the kernels find a (frame, octave, level) Gaussian of a device pyramid and the fp32 blur taps of
// sigma = 1.5 * getSigmaAt(octave, level); passed by value.
static const int kDescMaxOctaves = 8;
struct DescPyr {
    const u8 *maps;                       // vslam_pyramid::maps
    size_t frame_bytes;
    size_t ooff[kDescMaxOctaves], oplane[kDescMaxOctaves], opitch[kDescMaxOctaves];
    int orows[kDescMaxOctaves], ocols[kDescMaxOctaves];
    int num_octaves;
    int octave;                           // >= 0: every keypoint is read in this octave (per-octave entry points); -1: its own
    int frame0;                           // frame of every keypoint when no per-keypoint frame list is given
    const float *taps;                    // the tap tables back to back
    int tap_off[kDescMaxOctaves][6];      // first tap of (octave, level)
    int ksize[kDescMaxOctaves][6];        // cvRound(sigma * 4 * 2 + 1) | 1; 0 = level not prepared
};
size_t orient_smem_bytes(int r);
size_t sift_smem_bytes(int r);
// frame_of: device list of the frame of every keypoint, or nullptr (all in P.frame0)
// order: device list of the n keypoint indices this launch covers (nullptr: keypoints 0..n-1); max_r sizes the shared memory
int launch_orient_hist(vslam_ctx *ctx, const vslam_keypoint *kp, const int *frame_of, const int *order, int n, const DescPyr &P, int max_r,
                       unsigned long long *masks, float *histos);
int launch_sift(vslam_ctx *ctx, const vslam_keypoint *kp, const int *frame_of, const float2 *cs, int n, const DescPyr &P, int max_r, float *desc,
                u8 *valid);
// filterKeypoints' expansion of the orientation masks on the device (:362-368).  phase 0: per-keypoint counts + exclusive scan
// (pos, *d_total); phase 1: oriented points, their cos / sin (lut[bin], computed on the host like Rotation::cos_sin_of_angle) and
// frames, at most `capacity` of them.  partial: ceil(n / 4096) + 4 ints.
struct OrientLut { float2 cs[36]; };
int launch_orient_expand(vslam_ctx *ctx, const vslam_keypoint *kp, const int *frame_of, const unsigned long long *masks, int n, int capacity,
                         const OrientLut &lut, int *cnt, int *pos, int *partial, int *d_total, vslam_keypoint *out, float2 *cs, int *oframe,
                         int phase);
}  // namespace vslam
