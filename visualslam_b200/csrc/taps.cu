// taps.cu -- host side: OpenCV's 8U Gaussian kernel (what GaussianBlur(Size(0,0), sigma) builds at
// GaussPyramid.cpp:177) as 8.8 fixed-point taps, and their packing for the dp4a / dp2a blur kernels.
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

// Pack one level's taps for the kernels in pyramid.cu.
//  H pass (dp4a): the block stages pixels from X0 - hp; a thread's first tap sits (hp - r) pixels into its
//  staged window = wofs whole words + delta bytes, so the byte taps are stored with `delta` leading zeros.
//  V pass (dp2a): staged rows are paired (2p, 2p+1); even outputs use E[j] = (t[2j], t[2j+1]), odd outputs
//  O[j] = (t[2j-1], t[2j]).
static int pack_level(const int32_t *t, int n, int hp, LevelTaps *lv) {
    memset(lv, 0, sizeof *lv);
    const int r = n / 2;
    lv->n = n; lv->r = r;
    lv->delta = (hp - r) & 3;
    lv->wofs = (hp - r) >> 2;
    const int nb = n + lv->delta;
    int kq = (nb + 3) / 4;
    kq = (kq + 3) / 4 * 4;
    if (kq > kHWords) return VSLAM_E_INVALID;
    lv->kq = kq;
    for (int i = 0; i < n; ++i) {
        if (t[i] < 0 || t[i] > 255) return VSLAM_E_INVALID;
        const int b = i + lv->delta;
        lv->hq[b >> 2] |= (u32)t[i] << (8 * (b & 3));
    }
    int jp = r + 1;
    jp = (jp + 7) / 8 * 8;
    if (jp > kVPairs) return VSLAM_E_INVALID;
    lv->jp = jp;
    auto tap = [&](int i) -> u32 { return (i >= 0 && i < n) ? (u32)t[i] : 0u; };
    for (int j = 0; j <= r; ++j)
        lv->eo[j] = tap(2 * j) | tap(2 * j + 1) << 8 | tap(2 * j - 1) << 16 | tap(2 * j) << 24;
    return VSLAM_OK;
}

int build_octave_taps(double sigma0, int scale_samples, int octave, OctaveTaps *out) {
    int32_t t[kMaxKsize];
    int ns[6], rmax = 0;
    double sig[6];
    const double k = pow(2.0, 1.0 / (double)scale_samples);          // GaussPyramid.hpp:69
    for (int l = 0; l < 6; ++l) {
        sig[l] = pow(2, octave) * sigma0 * pow(k, l);                // GaussPyramid.cpp:157
        ns[l] = host_gauss_ksize(sig[l]);
        if (ns[l] > kMaxKsize) return VSLAM_E_INVALID;
        if (ns[l] / 2 > rmax) rmax = ns[l] / 2;
    }
    out->rmax = rmax;
    out->hp = (rmax + 3) / 4 * 4;
    for (int l = 0; l < 6; ++l) {
        int rc = host_gauss_taps(sig[l], ns[l], t);
        if (rc) return rc;
        rc = pack_level(t, ns[l], out->hp, &out->lv[l]);
        if (rc) return rc;
    }
    return VSLAM_OK;
}

}  // namespace vslam
