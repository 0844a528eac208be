// grad_math.cuh -- cv::phase(degrees) / cv::magnitude arithmetic shared by grad_k (pyramid.cu) and the descriptor kernels
// (descriptor.cu): OpenCV's fastAtan polynomial evaluated without FMA, magnitude with a correctly rounded square root.
#pragma once
#include "common.cuh"

namespace vslam {
__device__ __forceinline__ float fast_atan2_deg(float y, float x) {
    const float scale = (float)(180.0 / 3.14159265358979323846);
    const float p1 = 0.9997878412794807f * scale, p3 = -0.3258083974640975f * scale,
                p5 = 0.1555786518463281f * scale, p7 = -0.04432655554792128f * scale;
    const float ax = fabsf(x), ay = fabsf(y);
    const float eps = (float)2.2204460492503131e-16;
    float a, c, c2;
    if (ax >= ay) {
        c = __fdiv_rn(ay, __fadd_rn(ax, eps));
        c2 = __fmul_rn(c, c);
        a = __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c);
    } else {
        c = __fdiv_rn(ax, __fadd_rn(ay, eps));
        c2 = __fmul_rn(c, c);
        a = __fsub_rn(90.f, __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c));
    }
    if (x < 0) a = __fsub_rn(180.f, a);
    if (y < 0) a = __fsub_rn(360.f, a);
    return a;
}

// one sample of GaussPyramid::processGradients (GaussPyramid.cpp:86-101) formed from a level's u8 Gaussian: Sobel ksize 1
// = central differences with reflect-101 borders
__device__ __forceinline__ void grad_at(const u8 *__restrict__ g, int rows, int cols, size_t pitch, int y, int x, float &dx, float &dy) {
    const u8 *r = g + (size_t)y * pitch;
    dx = (float)((int)r[refl101(x + 1, cols)] - (int)r[refl101(x - 1, cols)]);
    dy = (float)((int)g[(size_t)refl101(y + 1, rows) * pitch + x] - (int)g[(size_t)refl101(y - 1, rows) * pitch + x]);
}
__device__ __forceinline__ float mag_at(const u8 *__restrict__ g, int rows, int cols, size_t pitch, int y, int x) {
    float dx, dy;
    grad_at(g, rows, cols, pitch, y, x, dx, dy);
    return __fsqrt_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)));
}
}  // namespace vslam
