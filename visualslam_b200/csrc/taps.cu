// taps.cu -- host side: OpenCV's 8U Gaussian kernel (what GaussianBlur(Size(0,0), sigma) builds at
// GaussPyramid.cpp:177) as 8.8 fixed-point taps (packed for the kernel by build_fused_taps in fused.cu).
#include <math.h>
#include <string.h>

#include "common.cuh"

namespace vslam {

// cvRound(sigma*3*2+1)|1 for CV_8U sources (imgproc smooth: createGaussianKernels); round half to even.
int host_gauss_ksize(double sigma) { return (int)nearbyint(sigma * 3 * 2 + 1) | 1; }

// getGaussianKernelBitExact + getGaussianKernelFixedPoint_ED (fractionBits = 8) restated with host doubles.
// The taps always sum to 256; the centre tap absorbs the rounding residue.
int host_gauss_taps(double sigma, int n, int32_t *taps) {
    if (n <= 0 || !(n & 1)) return VSLAM_E_INVALID;
    if (sigma <= 0) {
        static const int s1[] = {256}, s3[] = {64, 128, 64}, s5[] = {16, 64, 96, 64, 16},
                         s7[] = {8, 28, 56, 72, 56, 28, 8};
        const int *s = n == 1 ? s1 : n == 3 ? s3 : n == 5 ? s5 : n == 7 ? s7 : nullptr;
        if (s) {
            for (int i = 0; i < n; ++i) taps[i] = s[i];
            return VSLAM_OK;
        }
        sigma = n * 0.15 + 0.35;
    }
    const int half = (n - 1) / 2;
    const double scale2x = -0.125 / (sigma * sigma);
    std::vector<double> e((size_t)half + 1);
    double total = 0.0;
    for (int i = 0; i < half; ++i) {
        const int x = 1 - n + 2 * i;
        e[i] = exp((double)(x * x) * scale2x);
        total += e[i];
    }
    total = total * 2.0 + 1.0;
    const double norm = 1.0 / total;
    double carry = 0.0;
    long long used = 0;
    for (int i = 0; i < half; ++i) {
        const double want = e[i] * norm * 256.0 + carry;
        const long long q = (long long)nearbyint(want);
        carry = want - (double)q;
        taps[i] = taps[n - 1 - i] = (int32_t)q;
        used += q;
    }
    taps[half] = (int32_t)(256 - 2 * used);
    return VSLAM_OK;
}

}  // namespace vslam
