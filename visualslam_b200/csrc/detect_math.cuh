// detect_math.cuh -- the exact IEEE sequences of the contrast and edge tests, shared by detect.cu and fused.cu.
// Every float/double operation uses the round-to-nearest intrinsics so that nvcc cannot contract to FMA.
#pragma once
#include "common.cuh"

namespace vslam {

// FeaturePointLocalization's value: dog_zhat = value/255 + 0.5 * A^T * (-(A A^T)^-1) * A   (row H)
// a0, a1, a2 = float(dx)/255f, float(dy)/255f, float(ds)/255f (all non-zero), v = float(value)/255f
static __device__ __noinline__ float localize_f(float a0, float a1, float a2, float v) {
    const double A0 = (double)a0, A1 = (double)a1, A2 = (double)a2;
    // B = A * A^T through gemm's generic path: double product, stored as float
    const double b00 = (double)__double2float_rn(__dmul_rn(A0, A0)), b01 = (double)__double2float_rn(__dmul_rn(A0, A1)),
                 b02 = (double)__double2float_rn(__dmul_rn(A0, A2)), b11 = (double)__double2float_rn(__dmul_rn(A1, A1)),
                 b12 = (double)__double2float_rn(__dmul_rn(A1, A2)), b22 = (double)__double2float_rn(__dmul_rn(A2, A2));
    const double b10 = b01, b20 = b02, b21 = b12;
#define MUL(a, b) __dmul_rn(a, b)
#define SUB(a, b) __dsub_rn(a, b)
    // det3, left to right (OpenCV lapack.cpp)
    double d = __dadd_rn(SUB(MUL(b00, SUB(MUL(b11, b22), MUL(b12, b21))), MUL(b01, SUB(MUL(b10, b22), MUL(b12, b20)))),
                         MUL(b02, SUB(MUL(b10, b21), MUL(b11, b20))));
    if (d == 0.0) return v;
    d = __ddiv_rn(1.0, d);
    // M = -(B^-1), entries rounded to float (negation is exact)
    const float m0 = -__double2float_rn(MUL(SUB(MUL(b11, b22), MUL(b12, b21)), d));
    const float m1 = -__double2float_rn(MUL(SUB(MUL(b02, b21), MUL(b01, b22)), d));
    const float m2 = -__double2float_rn(MUL(SUB(MUL(b01, b12), MUL(b02, b11)), d));
    const float m3 = -__double2float_rn(MUL(SUB(MUL(b12, b20), MUL(b10, b22)), d));
    const float m4 = -__double2float_rn(MUL(SUB(MUL(b00, b22), MUL(b02, b20)), d));
    const float m5 = -__double2float_rn(MUL(SUB(MUL(b02, b10), MUL(b00, b12)), d));
    const float m6 = -__double2float_rn(MUL(SUB(MUL(b10, b21), MUL(b11, b20)), d));
    const float m7 = -__double2float_rn(MUL(SUB(MUL(b01, b20), MUL(b00, b21)), d));
    const float m8 = -__double2float_rn(MUL(SUB(MUL(b00, b11), MUL(b01, b10)), d));
#undef MUL
#undef SUB
    // z = M * A through gemm's small-matrix path: float accumulate
    const float z0 = __fadd_rn(__fadd_rn(__fmul_rn(m0, a0), __fmul_rn(m1, a1)), __fmul_rn(m2, a2));
    const float z1 = __fadd_rn(__fadd_rn(__fmul_rn(m3, a0), __fmul_rn(m4, a1)), __fmul_rn(m5, a2));
    const float z2 = __fadd_rn(__fadd_rn(__fmul_rn(m6, a0), __fmul_rn(m7, a1)), __fmul_rn(m8, a2));
    // 0.5 * A^T * z through gemm's generic path: double accumulate, alpha = 0.5
    const double acc = __dadd_rn(__dadd_rn(__dmul_rn(A0, (double)z0), __dmul_rn(A1, (double)z1)), __dmul_rn(A2, (double)z2));
    const float q = __double2float_rn(__dmul_rn(0.5, acc));
    return __fadd_rn(v, q);
}

static __device__ __forceinline__ float localize(int dx, int dy, int ds, int value) {
    const float v = __fdiv_rn((float)value, 255.0f);
    // any zero derivative -> a zero row and column in B -> det3 == 0 exactly -> cv::invert yields zeros
    if (dx == 0 || dy == 0 || ds == 0) return v;
    return localize_f(__fdiv_rn((float)dx, 255.0f), __fdiv_rn((float)dy, 255.0f), __fdiv_rn((float)ds, 255.0f), v);
}


// computeEdgeResponse (Diff_of_Gauss.cpp:79-109) from the three structure sums (exact integers in fp32)
__device__ __forceinline__ float edge_from_sums(float ix2, float iy2, float ixy) {
    const float det = __double2float_rn(__dsub_rn(__dmul_rn((double)ix2, (double)iy2), __dmul_rn((double)ixy, (double)ixy)));
    const float tr = __fadd_rn(ix2, iy2);
    return __fdiv_rn(__fmul_rn(tr, tr), det);
}

}  // namespace vslam
