// TEST INFRASTRUCTURE ONLY -- the numeric half of the OpenCV stand-in (opencv2/shim.hpp).  Every imgproc routine maps
// onto a primitive of the plain-C oracle (oracle/vslam_oracle.c, linked in), which tests/test_oracle_golden.py pins to
// cv2 4.13; the small core routines (gemm, invert, determinant, trace) restate OpenCV's code paths for the shapes the
// reference uses (core/src/matmul.simd.hpp gemmImpl small-matrix branch, GEMMSingleMul; lapack.cpp det2/det3/invert).
// Built with -ffp-contract=off: every float operation rounds on its own, as in OpenCV's scalar code.
#include <opencv2/shim.hpp>

extern "C" {
void vo_upsample2x(const uint8_t *src, int h, int w, uint8_t *dst);
int vo_gaussian_blur_8u(const uint8_t *src, int h, int w, double sigma, uint8_t *dst);
int vo_blur_8u(const uint8_t *src, int h, int w, const int *taps, int n, uint8_t *dst);
void vo_downsample_nearest(const uint8_t *src, int h, int w, uint8_t *dst);
void vo_half_size(int h, int w, int *oh, int *ow);
void vo_sobel1(const uint8_t *g, int h, int w, float *gx, float *gy);
void vo_mag_orient(const float *gx, const float *gy, size_t n, float *mag, float *orient);
int vo_gauss_ksize_32f(double sigma);
int vo_gauss_taps_32f(double sigma, int n, float *taps);
int vo_refl101(int p, int len);
int vo_warp_affine_u8(const uint8_t *src, int h, int w, size_t pitch, const double m[6], uint8_t *dst, int dh, int dw);
}

namespace cv {

long &shim_oob() { static long n = 0; return n; }

static void need(bool ok, const char *what) { if (!ok) throw std::runtime_error(std::string("opencv shim: unsupported ") + what); }

// dense copy of a (possibly strided) 8U / 32F Mat
template <class T> static std::vector<T> dense(const Mat &m) {
    std::vector<T> v((size_t)m.rows * m.cols);
    for (int i = 0; i < m.rows; ++i) std::memcpy(v.data() + (size_t)i * m.cols, m.data + (size_t)i * m.step, (size_t)m.cols * sizeof(T));
    return v;
}
template <class T> static void undense(const std::vector<T> &v, Mat &m) {
    for (int i = 0; i < m.rows; ++i) std::memcpy(m.data + (size_t)i * m.step, v.data() + (size_t)i * m.cols, (size_t)m.cols * sizeof(T));
}

// ---- core ---------------------------------------------------------------------------------------------------------
// cv::gemm(A, B, alpha, noArray(), 0, D).  CV_32F: inner length 2..4 equal to a side of D takes the small-matrix branch
// (float products and sums, then (float)(t * alpha)); everything else is GEMMSingleMul with a double accumulator.
Mat gemm_(const Mat &a, const Mat &b, double alpha) {
    need(a.type() == b.type() && (a.type() == CV_32F || a.type() == CV_64F) && a.cols == b.rows, "gemm operands");
    const int len = a.cols, dr = a.rows, dc = b.cols;
    Mat d = Mat::zeros(dr, dc, a.type());
    if (a.type() == CV_64F) {
        for (int i = 0; i < dr; ++i)
            for (int j = 0; j < dc; ++j) {
                double s = 0;
                for (int k = 0; k < len; ++k) s += a.at<double>(i, k) * b.at<double>(k, j);
                d.at<double>(i, j) = s * alpha;
            }
        return d;
    }
    if (len >= 2 && len <= 4 && (len == dc || len == dr)) {
        for (int i = 0; i < dr; ++i)
            for (int j = 0; j < dc; ++j) {
                volatile float t = a.at<float>(i, 0) * b.at<float>(0, j);
                for (int k = 1; k < len; ++k) {
                    volatile float p = a.at<float>(i, k) * b.at<float>(k, j);
                    t = t + p;
                }
                d.at<float>(i, j) = (float)((double)t * alpha);
            }
        return d;
    }
    for (int i = 0; i < dr; ++i)
        for (int j = 0; j < dc; ++j) {
            double s = 0;
            for (int k = 0; k < len; ++k) s += (double)a.at<float>(i, k) * (double)b.at<float>(k, j);
            d.at<float>(i, j) = (float)(s * alpha);
        }
    return d;
}

// m + Scalar on CV_32F: the scalar is converted to float, the sum is a float add
Mat add_scalar_(const Mat &m, double s) {
    need(m.type() == CV_32F, "Mat + scalar depth");
    Mat d = m.clone();
    const float fs = (float)s;
    for (int i = 0; i < d.rows; ++i)
        for (int j = 0; j < d.cols; ++j) { volatile float v = d.at<float>(i, j) + fs; d.at<float>(i, j) = v; }
    return d;
}

Mat negate_(const Mat &m) {
    Mat d = m.clone();
    for (int i = 0; i < d.rows; ++i)
        for (int j = 0; j < d.cols; ++j) {
            if (m.type() == CV_32F) d.at<float>(i, j) = -d.at<float>(i, j);
            else if (m.type() == CV_64F) d.at<double>(i, j) = -d.at<double>(i, j);
            else need(false, "negation depth");
        }
    return d;
}

// cv::subtract on CV_8U saturates (GaussPyramid.cpp:197)
Mat subtract_(const Mat &a, const Mat &b) {
    need(a.type() == CV_8U && b.type() == CV_8U && a.rows == b.rows && a.cols == b.cols, "Mat - Mat operands");
    Mat d(a.rows, a.cols, CV_8U);
    for (int i = 0; i < a.rows; ++i)
        for (int j = 0; j < a.cols; ++j) {
            const int v = (int)a.at<uchar>(i, j) - (int)b.at<uchar>(i, j);
            d.at<uchar>(i, j) = (uchar)(v < 0 ? 0 : v);
        }
    return d;
}

Mat compare_gt_(const Mat &a, const Mat &b) {
    need(a.type() == CV_8U && b.type() == CV_8U && a.rows == b.rows && a.cols == b.cols, "Mat > Mat operands");
    Mat d(a.rows, a.cols, CV_8U);
    for (int i = 0; i < a.rows; ++i)
        for (int j = 0; j < a.cols; ++j) d.at<uchar>(i, j) = a.at<uchar>(i, j) > b.at<uchar>(i, j) ? 255 : 0;
    return d;
}

// lapack.cpp: determinant of a 2x2 / 3x3 CV_32F matrix is formed in double (det2 / det3)
double determinant(const Mat &m) {
    need(m.type() == CV_32F && m.rows == m.cols && (m.rows == 2 || m.rows == 3), "determinant shape");
    if (m.rows == 2) return (double)m.at<float>(0, 0) * m.at<float>(1, 1) - (double)m.at<float>(0, 1) * m.at<float>(1, 0);
    const double b00 = m.at<float>(0, 0), b01 = m.at<float>(0, 1), b02 = m.at<float>(0, 2), b10 = m.at<float>(1, 0), b11 = m.at<float>(1, 1),
                 b12 = m.at<float>(1, 2), b20 = m.at<float>(2, 0), b21 = m.at<float>(2, 1), b22 = m.at<float>(2, 2);
    return b00 * (b11 * b22 - b12 * b21) - b01 * (b10 * b22 - b12 * b20) + b02 * (b10 * b21 - b11 * b20);
}

Scalar trace(const Mat &m) {
    double s = 0;
    for (int i = 0; i < std::min(m.rows, m.cols); ++i) s += m.type() == CV_32F ? (double)m.at<float>(i, i) : m.at<double>(i, i);
    return Scalar(s);
}

// cv::invert(DECOMP_LU) for 3x3 CV_32F: det3 in double; a zero determinant gives the zero matrix
Mat Mat::inv() const {
    need(type() == CV_32F && rows == 3 && cols == 3, "inv shape");
    const Mat &m = *this;
    Mat d = Mat::zeros(3, 3, CV_32F);
    double det = determinant(m);
    if (det != 0.0) {
        const double b00 = m.at<float>(0, 0), b01 = m.at<float>(0, 1), b02 = m.at<float>(0, 2), b10 = m.at<float>(1, 0), b11 = m.at<float>(1, 1),
                     b12 = m.at<float>(1, 2), b20 = m.at<float>(2, 0), b21 = m.at<float>(2, 1), b22 = m.at<float>(2, 2);
        det = 1.0 / det;
        d.at<float>(0, 0) = (float)((b11 * b22 - b12 * b21) * det);
        d.at<float>(0, 1) = (float)((b02 * b21 - b01 * b22) * det);
        d.at<float>(0, 2) = (float)((b01 * b12 - b02 * b11) * det);
        d.at<float>(1, 0) = (float)((b12 * b20 - b10 * b22) * det);
        d.at<float>(1, 1) = (float)((b00 * b22 - b02 * b20) * det);
        d.at<float>(1, 2) = (float)((b02 * b10 - b00 * b12) * det);
        d.at<float>(2, 0) = (float)((b10 * b21 - b11 * b20) * det);
        d.at<float>(2, 1) = (float)((b01 * b20 - b00 * b21) * det);
        d.at<float>(2, 2) = (float)((b00 * b11 - b01 * b10) * det);
    }
    return d;
}

void transpose(const Mat &src, Mat &dst) {
    Mat d(src.cols, src.rows, src.type());
    const size_t es = src.elemSize();
    for (int i = 0; i < src.rows; ++i)
        for (int j = 0; j < src.cols; ++j) std::memcpy(d.data + (size_t)j * d.step + (size_t)i * es, src.data + (size_t)i * src.step + (size_t)j * es, es);
    dst = std::move(d);
}

void copyMakeBorder(const Mat &src, Mat &dst, int top, int bottom, int left, int right, int borderType) {
    need(borderType == BORDER_REPLICATE, "copyMakeBorder border type");
    Mat d(src.rows + top + bottom, src.cols + left + right, src.type());
    const size_t es = src.elemSize();
    for (int i = 0; i < d.rows; ++i) {
        const int si = std::min(std::max(i - top, 0), src.rows - 1);
        for (int j = 0; j < d.cols; ++j) {
            const int sj = std::min(std::max(j - left, 0), src.cols - 1);
            std::memcpy(d.data + (size_t)i * d.step + (size_t)j * es, src.data + (size_t)si * src.step + (size_t)sj * es, es);
        }
    }
    dst = std::move(d);
}

// ---- imgproc ------------------------------------------------------------------------------------------------------
void GaussianBlur(const Mat &src, Mat &dst, Size ksize, double sigmaX, double, int borderType) {
    need(borderType == BORDER_DEFAULT && !src.empty(), "GaussianBlur border / empty source");
    if (src.type() == CV_8U) {
        need(!src.is_view(), "8U GaussianBlur on a ROI");
        const std::vector<uchar> s = dense<uchar>(src);
        std::vector<uchar> o(s.size());
        if (ksize.width == 0 && ksize.height == 0 && sigmaX > 0) {
            need(vo_gaussian_blur_8u(s.data(), src.rows, src.cols, sigmaX, o.data()) == 0, "8U GaussianBlur sigma");
        } else {
            // Size(3,3), sigma 0 (Harris_corners.cpp:158): getGaussianKernel's small table {0.25, 0.5, 0.25} -> 64, 128, 64
            need(ksize.width == 3 && ksize.height == 3 && sigmaX == 0, "8U GaussianBlur kernel");
            static const int taps[3] = {64, 128, 64};
            need(vo_blur_8u(s.data(), src.rows, src.cols, taps, 3, o.data()) == 0, "8U 3x3 blur");
        }
        Mat d(src.rows, src.cols, CV_8U);
        undense(o, d);
        dst = std::move(d);
        return;
    }
    // CV_32F, Size(0,0): n = cvRound(sigma * 4 * 2 + 1) | 1; separable engine: row filter taps left to right, symmetric
    // column filter (centre, then pairs), fused multiply-adds.  A ROI reads its parent around it; reflect-101 applies
    // at the PARENT's edges (FilterEngine with a non-isolated ROI).
    need(src.type() == CV_32F && ksize.width == 0 && ksize.height == 0 && sigmaX > 0, "32F GaussianBlur arguments");
    const int n = vo_gauss_ksize_32f(sigmaX), r = n / 2, R = src.rows, C = src.cols;
    std::vector<float> k((size_t)n);
    need(vo_gauss_taps_32f(sigmaX, n, k.data()) == 0, "32F taps");
    const uchar *origin = src.data - (size_t)src.ofy * src.step - (size_t)src.ofx * sizeof(float);
    auto px = [&](int py, int pxx) -> float {      // parent coordinates, reflect-101 at the parent's edges
        return *(const float *)(origin + (size_t)vo_refl101(py, src.wrows) * src.step + (size_t)vo_refl101(pxx, src.wcols) * sizeof(float));
    };
    std::vector<float> h((size_t)(R + 2 * r) * C);
    for (int q = 0; q < R + 2 * r; ++q)
        for (int j = 0; j < C; ++j) {
            const int py = src.ofy + q - r;
            float s = k[0] * px(py, src.ofx + j - r);
            for (int u = 1; u < n; ++u) s = std::fmaf(k[u], px(py, src.ofx + j + u - r), s);
            h[(size_t)q * C + j] = s;
        }
    Mat d(R, C, CV_32F);
    for (int i = 0; i < R; ++i)
        for (int j = 0; j < C; ++j) {
            float s = k[r] * h[(size_t)(i + r) * C + j];
            for (int v = 1; v <= r; ++v) s = std::fmaf(k[r + v], h[(size_t)(i + r + v) * C + j] + h[(size_t)(i + r - v) * C + j], s);
            d.at<float>(i, j) = s;
        }
    dst = std::move(d);
}

void Sobel(const Mat &src, Mat &dst, int ddepth, int dx, int dy, int ksize, double scale, double delta, int borderType) {
    need(src.type() == CV_8U && ddepth == CV_32F && ksize == 1 && scale == 1 && delta == 0 && borderType == BORDER_DEFAULT && dx + dy == 1 && !src.is_view(), "Sobel arguments");
    const std::vector<uchar> s = dense<uchar>(src);
    std::vector<float> gx(s.size()), gy(s.size());
    vo_sobel1(s.data(), src.rows, src.cols, gx.data(), gy.data());
    dst.create(src.rows, src.cols, CV_32F);
    undense(dx ? gx : gy, dst);
}

void magnitude(const Mat &x, const Mat &y, Mat &mag) {
    need(x.type() == CV_32F && y.type() == CV_32F && x.rows == y.rows && x.cols == y.cols, "magnitude operands");
    const std::vector<float> a = dense<float>(x), b = dense<float>(y);
    std::vector<float> m(a.size());
    vo_mag_orient(a.data(), b.data(), a.size(), m.data(), nullptr);
    mag.create(x.rows, x.cols, CV_32F);
    undense(m, mag);
}

void phase(const Mat &x, const Mat &y, Mat &angle, bool angleInDegrees) {
    need(angleInDegrees && x.type() == CV_32F && y.type() == CV_32F && x.rows == y.rows && x.cols == y.cols, "phase operands");
    const std::vector<float> a = dense<float>(x), b = dense<float>(y);
    std::vector<float> o(a.size());
    vo_mag_orient(a.data(), b.data(), a.size(), nullptr, o.data());
    angle.create(x.rows, x.cols, CV_32F);
    undense(o, angle);
}

void resize(const Mat &src, Mat &dst, Size dsize, double fx, double fy, int interpolation) {
    need(src.type() == CV_8U && dsize.width == 0 && dsize.height == 0 && !src.is_view(), "resize arguments");
    const std::vector<uchar> s = dense<uchar>(src);
    if (fx == 2 && fy == 2 && interpolation == INTER_LINEAR) {
        std::vector<uchar> o((size_t)4 * s.size());
        vo_upsample2x(s.data(), src.rows, src.cols, o.data());
        Mat d(2 * src.rows, 2 * src.cols, CV_8U);
        undense(o, d);
        dst = std::move(d);
    } else {
        need(fx == 0.5 && fy == 0.5 && interpolation == INTER_NEAREST, "resize factors");
        int oh, ow;
        vo_half_size(src.rows, src.cols, &oh, &ow);
        std::vector<uchar> o((size_t)oh * ow);
        vo_downsample_nearest(s.data(), src.rows, src.cols, o.data());
        Mat d(oh, ow, CV_8U);
        undense(o, d);
        dst = std::move(d);
    }
}

// morphological dilation, default border (out-of-image neighbours do not take part), anchor at the kernel centre
void dilate(const Mat &src, Mat &dst, const Mat &kernel) {
    need(src.type() == CV_8U && kernel.type() == CV_8U, "dilate depth");
    Mat d = Mat::zeros(src.rows, src.cols, CV_8U);
    const int ay = kernel.rows / 2, ax = kernel.cols / 2;
    for (int i = 0; i < src.rows; ++i)
        for (int j = 0; j < src.cols; ++j) {
            int m = 0;
            for (int u = 0; u < kernel.rows; ++u)
                for (int v = 0; v < kernel.cols; ++v) {
                    const int y = i + u - ay, x = j + v - ax;
                    if (!kernel.at<uchar>(u, v) || y < 0 || x < 0 || y >= src.rows || x >= src.cols) continue;
                    m = std::max(m, (int)src.at<uchar>(y, x));
                }
            d.at<uchar>(i, j) = (uchar)m;
        }
    dst = std::move(d);
}

// saturate_cast<uchar>(cvRound(|v|)); cvRound is cvtss2si: INT_MIN for |v| >= 2^31, which saturates to 0
void convertScaleAbs(const Mat &src, Mat &dst, double alpha, double beta) {
    need(alpha == 1 && beta == 0 && (src.type() == CV_32F || src.type() == CV_8U), "convertScaleAbs arguments");
    Mat d(src.rows, src.cols, CV_8U);
    for (int i = 0; i < src.rows; ++i)
        for (int j = 0; j < src.cols; ++j) {
            if (src.type() == CV_8U) { d.at<uchar>(i, j) = src.at<uchar>(i, j); continue; }
            const double a = std::fabs((double)src.at<float>(i, j));
            const int v = a >= 2147483648.0 ? 0 : (a >= 255.5 ? 255 : (int)std::nearbyint(a));
            d.at<uchar>(i, j) = (uchar)(v > 255 ? 255 : v);
        }
    dst = std::move(d);
}

// imgproc/src/imgwarp.cpp getRotationMatrix2D
Mat getRotationMatrix2D(Point2f center, double angle, double scale) {
    angle *= CV_PI / 180;
    const double alpha = std::cos(angle) * scale, beta = std::sin(angle) * scale;
    Mat M(2, 3, CV_64F);
    M.at<double>(0, 0) = alpha; M.at<double>(0, 1) = beta; M.at<double>(0, 2) = (1 - alpha) * center.x - beta * center.y;
    M.at<double>(1, 0) = -beta; M.at<double>(1, 1) = alpha; M.at<double>(1, 2) = beta * center.x + (1 - alpha) * center.y;
    return M;
}

// cv::warpAffine with the defaults the reference uses (INTER_LINEAR, BORDER_CONSTANT 0) on 8UC1: the oracle's restatement,
// pinned to cv2 by tests/test_oracle_golden.py.  A ROI view is warped as an image of its own (OpenCV does not look outside it).
void warpAffine(const Mat &src, Mat &dst, const Mat &M, Size dsize) {
    need(src.type() == CV_8U && M.type() == CV_64F && M.rows == 2 && M.cols == 3 && src.rows > 0 && src.cols > 0 && dsize.width > 0 && dsize.height > 0,
         "warpAffine operands");
    double m[6];
    for (int i = 0; i < 2; ++i)
        for (int j = 0; j < 3; ++j) m[3 * i + j] = M.at<double>(i, j);
    Mat d(dsize.height, dsize.width, CV_8U);
    std::vector<uchar> out((size_t)dsize.height * dsize.width);
    need(vo_warp_affine_u8(src.data, src.rows, src.cols, src.step, m, out.data(), dsize.height, dsize.width) == 0, "warpAffine size");
    undense(out, d);
    dst = std::move(d);
}

}  // namespace cv
