// stages.cuh -- launchers of the per-function kernels (stages.cu); all pointers are device pointers, maps are dense.
#pragma once
#include "common.cuh"

namespace vslam {
int launch_bgr2gray(vslam_ctx *ctx, const u8 *src, int rows, int cols, size_t pitch, size_t stride, u8 *dst, size_t dpitch,
                    size_t dstride, int n_frames);
int launch_rotate_gather(vslam_ctx *ctx, const u8 *src, int rows, int cols, size_t pitch, int cx, int cy, float c, float s, int x0,
                         int y0, int ow, int oh, u8 *dst, size_t dpitch, int mode);
int launch_warp_affine(vslam_ctx *ctx, const u8 *src, int rows, int cols, size_t pitch, const double *minv, u8 *dst, int drows, int dcols,
                       size_t dpitch);
int launch_blur3(vslam_ctx *ctx, const u8 *src, int rows, int cols, size_t pitch, u8 *dst, size_t dpitch);
int launch_sobel1(vslam_ctx *ctx, const u8 *src, int rows, int cols, size_t pitch, float *gx, float *gy);
int launch_harris_grad(vslam_ctx *ctx, const float *ix, const float *iy, int rows, int cols, int pad, float k, float *out);
int launch_cvtabs(vslam_ctx *ctx, const float *src, size_t n, u8 *dst);
int launch_nms3(vslam_ctx *ctx, const u8 *r8, int rows, int cols, u8 *mask);
int launch_nms2(vslam_ctx *ctx, const float *r, int rows, int cols, int pad, float *out);
int launch_edge_list(vslam_ctx *ctx, const vslam_keypoint *kp, int n, const u8 *const *gauss, int rows, int cols, size_t pitch, float *resp);
int launch_localize_list(vslam_ctx *ctx, const int *q, int n, float *zhat);
int launch_edge_windows(vslam_ctx *ctx, const float *gx, const float *gy, int n, int len, float *resp, float *sums);
}  // namespace vslam
