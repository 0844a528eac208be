// TEST INFRASTRUCTURE ONLY -- a minimal stand-in for the OpenCV C++ API, just large enough to compile the reference's
// own translation units UNMODIFIED (KeyPointDetection/Harris_corners.cpp, Diff_of_Gauss.cpp,
// include/src/GaussPyramid/GaussPyramid.cpp, include/src/Rotation/rotation.cpp) in an image that has no OpenCV C++
// headers or libraries.  See oracle/Makefile (target _ref) and oracle/ref_driver.cpp.
//
// What is real and what is a stub:
//   * cv::Mat (shared buffer, ROI views that remember their parent, move leaves the source empty), Point/Size/Rect/
//     Scalar/Vec, at<T>() -- real.
//   * the imgproc / core numerics the hot path calls (resize x2 / x0.5, 8U and 32F GaussianBlur, Sobel ksize 1,
//     magnitude, phase, copyMakeBorder, dilate, convertScaleAbs, determinant, trace, invert, gemm, transpose, compare)
//     -- implemented in shim.cpp on top of the plain-C oracle's primitives (oracle/vslam_oracle.c), each of which is
//     pinned to cv2 4.13 by tests/test_oracle_golden.py.  The LOOPS that consume them are the reference's own code.
//   * display, codecs, drawing (imshow, waitKey, imread, line, circle, ...) -- empty inline no-ops: the driver never
//     calls the reference's main().
// Typed access that disagrees with the Mat's element size (the reference's NMS2 writes float into a CV_8U Mat,
// Harris_corners.cpp:86,118 -- undefined behaviour in OpenCV) goes to a float "shadow" plane of the same shape, i.e. the
// intended semantics; out-of-buffer access (HarrisCorner's swapped indices on non-square images, rotateImageSection) is
// counted in cv::shim_oob() and redirected to a dummy element instead of corrupting memory.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <iostream>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#define CV_8U 0
#define CV_32F 5
#define CV_64F 6
#define CV_8UC1 0
#define CV_8UC3 16
#define CV_32FC1 5
#define CV_PI 3.1415926535897932384626433832795

namespace cv {

// OpenCV's core/base.hpp pulls these into namespace cv; the reference relies on it (min(img.rows, img.cols), pow, ...)
using std::abs;
using std::exp;
using std::log;
using std::max;
using std::min;
using std::pow;
using std::sqrt;
using std::swap;

typedef unsigned char uchar;

enum { BORDER_CONSTANT = 0, BORDER_REPLICATE = 1, BORDER_REFLECT = 2, BORDER_WRAP = 3, BORDER_REFLECT_101 = 4, BORDER_DEFAULT = 4 };
enum { INTER_NEAREST = 0, INTER_LINEAR = 1 };
enum { IMREAD_GRAYSCALE = 0, IMREAD_COLOR = 1 };
enum { COLOR_BGR2GRAY = 6 };
enum { NORM_MINMAX = 32 };
enum { LINE_8 = 8 };

template <class T> struct Point_ {
    T x, y;
    Point_() : x(0), y(0) {}
    Point_(T x_, T y_) : x(x_), y(y_) {}
    template <class U> Point_(const Point_<U> &p) : x((T)p.x), y((T)p.y) {}
    Point_ &operator-=(const Point_ &o) { x -= o.x; y -= o.y; return *this; }
    Point_ &operator+=(const Point_ &o) { x += o.x; y += o.y; return *this; }
};
template <class T> Point_<T> operator-(const Point_<T> &a, const Point_<T> &b) { return Point_<T>(a.x - b.x, a.y - b.y); }
template <class T> Point_<T> operator+(const Point_<T> &a, const Point_<T> &b) { return Point_<T>(a.x + b.x, a.y + b.y); }
typedef Point_<int> Point2i;
typedef Point_<float> Point2f;
typedef Point_<double> Point2d;
typedef Point2i Point;

struct Size {
    int width, height;
    Size() : width(0), height(0) {}
    Size(int w, int h) : width(w), height(h) {}
};
inline std::ostream &operator<<(std::ostream &o, const Size &s) { return o << "[" << s.width << " x " << s.height << "]"; }

struct Rect {
    int x, y, width, height;
    Rect() : x(0), y(0), width(0), height(0) {}
    Rect(int x_, int y_, int w, int h) : x(x_), y(y_), width(w), height(h) {}
    Rect(const Point &a, const Point &b) : x(std::min(a.x, b.x)), y(std::min(a.y, b.y)), width(std::max(a.x, b.x) - std::min(a.x, b.x)), height(std::max(a.y, b.y) - std::min(a.y, b.y)) {}
};

struct Scalar {
    double val[4];
    Scalar(double a = 0, double b = 0, double c = 0, double d = 0) : val{a, b, c, d} {}
    double operator[](int i) const { return val[i]; }
};

template <class T, int N> struct Vec {
    T val[N];
    Vec() { for (int i = 0; i < N; ++i) val[i] = T(); }
    Vec(T a, T b, T c) { static_assert(N == 3, "3-vector"); val[0] = a; val[1] = b; val[2] = c; }
    T &operator[](int i) { return val[i]; }
    const T &operator[](int i) const { return val[i]; }
};
typedef Vec<float, 3> Vec3f;
typedef Vec<double, 3> Vec3d;

template <class T> struct DataType;
template <> struct DataType<uchar> { enum { type = CV_8U }; };
template <> struct DataType<float> { enum { type = CV_32F }; };
template <> struct DataType<double> { enum { type = CV_64F }; };

long &shim_oob();   // out-of-buffer typed accesses seen so far (shim.cpp)

class Mat {
  public:
    int rows = 0, cols = 0;
    size_t step = 0;
    uchar *data = nullptr;
    // the view's parent: GaussianBlur on a ROI reads the parent around it (OpenCV: Mat::locateROI)
    int wrows = 0, wcols = 0, ofy = 0, ofx = 0;

    Mat() {}
    Mat(int r, int c, int type) { create(r, c, type); }
    Mat(Size s, int type) { create(s.height, s.width, type); }
    Mat(const Mat &m, const Rect &roi) { *this = m; view(roi); }
    template <class T, int N> explicit Mat(const Vec<T, N> &v) {
        create(N, 1, DataType<T>::type);
        for (int i = 0; i < N; ++i) at<T>(i, 0) = v[i];
    }
    Mat(const Mat &) = default;
    Mat &operator=(const Mat &) = default;
    Mat(Mat &&o) noexcept { *this = std::move(o); }
    Mat &operator=(Mat &&o) noexcept {
        if (this != &o) {
            rows = o.rows; cols = o.cols; step = o.step; data = o.data; wrows = o.wrows; wcols = o.wcols; ofy = o.ofy; ofx = o.ofx;
            type_ = o.type_; buf_ = std::move(o.buf_); shadow_ = std::move(o.shadow_);
            o.rows = o.cols = 0; o.step = 0; o.data = nullptr; o.wrows = o.wcols = o.ofy = o.ofx = 0;
        }
        return *this;
    }

    // cv::Mat::create: keeps the buffer when shape and type already match
    void create(int r, int c, int type) {
        if (data && r == rows && c == cols && type == type_ && !is_view()) return;
        type_ = type; rows = wrows = r; cols = wcols = c; ofy = ofx = 0;
        step = (size_t)c * elemSize();
        buf_ = std::make_shared<std::vector<uchar>>((size_t)r * step + 64, (uchar)0);
        data = buf_->data();
        shadow_.reset();
    }
    void create(Size s, int type) { create(s.height, s.width, type); }
    static Mat zeros(int r, int c, int type) { Mat m; m.create(r, c, type); std::memset(m.data, 0, (size_t)r * m.step); return m; }
    static Mat zeros(Size s, int type) { return zeros(s.height, s.width, type); }
    static Mat ones(int r, int c, int type) {
        Mat m = zeros(r, c, type);
        for (int i = 0; i < r; ++i)
            for (int j = 0; j < c; ++j) {
                if (type == CV_8U) m.at<uchar>(i, j) = 1;
                else if (type == CV_32F) m.at<float>(i, j) = 1.f;
                else m.at<double>(i, j) = 1.0;
            }
        return m;
    }
    Mat clone() const {
        Mat m;
        if (!data) return m;
        m.create(rows, cols, type_);
        for (int i = 0; i < rows; ++i) std::memcpy(m.data + (size_t)i * m.step, data + (size_t)i * step, (size_t)cols * elemSize());
        if (shadow_) m.shadow_ = std::make_shared<std::vector<float>>(*shadow_);
        return m;
    }
    int type() const { return type_; }
    int depth() const { return type_ & 7; }
    int channels() const { return (type_ >> 3) + 1; }
    size_t elemSize() const { const int d = type_ & 7; return (size_t)(d == CV_8U ? 1 : d == CV_32F ? 4 : 8) * ((type_ >> 3) + 1); }
    bool empty() const { return !data || rows == 0 || cols == 0; }
    Size size() const { return Size(cols, rows); }
    bool is_view() const { return rows != wrows || cols != wcols; }
    Mat inv() const;                       // shim.cpp: cv::invert(DECOMP_LU) for 2x2 / 3x3
    const float *shadow() const { return shadow_ ? shadow_->data() : nullptr; }

    template <class T> T &at(int i, int j) { return *(T *)elem<T>(i, j); }
    template <class T> const T &at(int i, int j) const { return *(const T *)const_cast<Mat *>(this)->elem<T>(i, j); }
    template <class T> T &at(int i) { return rows == 1 ? at<T>(0, i) : at<T>(i, 0); }
    template <class T> const T &at(int i) const { return rows == 1 ? at<T>(0, i) : at<T>(i, 0); }

  private:
    int type_ = 0;
    std::shared_ptr<std::vector<uchar>> buf_;
    std::shared_ptr<std::vector<float>> shadow_;

    void view(const Rect &r) {
        if (r.x < 0 || r.y < 0 || r.width < 0 || r.height < 0 || r.x + r.width > cols || r.y + r.height > rows)
            throw std::out_of_range("cv::Mat ROI outside the matrix");
        data += (size_t)r.y * step + (size_t)r.x * elemSize();
        ofy += r.y; ofx += r.x; rows = r.height; cols = r.width;
    }
    template <class T> void *elem(int i, int j) {
        static thread_local double dummy[2];
        if (sizeof(T) != elemSize()) {                      // type-confused access: the float shadow plane
            if (!shadow_) shadow_ = std::make_shared<std::vector<float>>((size_t)rows * cols + 1, 0.f);
            if (sizeof(T) != 4 || i < 0 || j < 0 || (size_t)i * cols + j >= (size_t)rows * cols) { ++shim_oob(); dummy[0] = 0; return dummy; }
            return shadow_->data() + (size_t)i * cols + j;
        }
        // release-mode cv::Mat::at does no bounds check: data + i * step + j * sizeof(T); stay inside the allocation
        const long long off = (long long)i * (long long)step + (long long)j * (long long)sizeof(T);
        const uchar *lo = buf_ ? buf_->data() : data;
        const long long base = data - lo, cap = buf_ ? (long long)buf_->size() - 64 : (long long)rows * (long long)step;
        if (base + off < 0 || base + off + (long long)sizeof(T) > cap) { ++shim_oob(); dummy[0] = 0; return dummy; }
        return data + off;
    }
};

// ---- matrix expressions the reference writes (Diff_of_Gauss.cpp:238-241, rotation.cpp:136) -------------------------
struct ScaledMat { const Mat &m; double alpha; };
Mat gemm_(const Mat &a, const Mat &b, double alpha);                 // shim.cpp
Mat add_scalar_(const Mat &m, double s);
Mat negate_(const Mat &m);
Mat subtract_(const Mat &a, const Mat &b);
Mat compare_gt_(const Mat &a, const Mat &b);
inline Mat operator*(const Mat &a, const Mat &b) { return gemm_(a, b, 1.0); }
inline ScaledMat operator*(double s, const Mat &m) { return ScaledMat{m, s}; }
inline Mat operator*(const ScaledMat &a, const Mat &b) { return gemm_(a.m, b, a.alpha); }
inline Mat operator+(double s, const Mat &m) { return add_scalar_(m, s); }
inline Mat operator-(const Mat &m) { return negate_(m); }
inline Mat operator-(const Mat &a, const Mat &b) { return subtract_(a, b); }
inline Mat operator>(const Mat &a, const Mat &b) { return compare_gt_(a, b); }

// ---- core / imgproc numerics (shim.cpp) --------------------------------------------------------------------------
double determinant(const Mat &m);
Scalar trace(const Mat &m);
void transpose(const Mat &src, Mat &dst);
void copyMakeBorder(const Mat &src, Mat &dst, int top, int bottom, int left, int right, int borderType);
void GaussianBlur(const Mat &src, Mat &dst, Size ksize, double sigmaX, double sigmaY = 0, int borderType = BORDER_DEFAULT);
void Sobel(const Mat &src, Mat &dst, int ddepth, int dx, int dy, int ksize = 3, double scale = 1, double delta = 0, int borderType = BORDER_DEFAULT);
void magnitude(const Mat &x, const Mat &y, Mat &mag);
void phase(const Mat &x, const Mat &y, Mat &angle, bool angleInDegrees = false);
void resize(const Mat &src, Mat &dst, Size dsize, double fx = 0, double fy = 0, int interpolation = INTER_LINEAR);
void dilate(const Mat &src, Mat &dst, const Mat &kernel);
void convertScaleAbs(const Mat &src, Mat &dst, double alpha = 1, double beta = 0);
Mat getRotationMatrix2D(Point2f center, double angle, double scale);
void warpAffine(const Mat &src, Mat &dst, const Mat &M, Size dsize);

// ---- display, codecs, drawing: not on any path the driver runs -----------------------------------------------------
namespace samples { inline std::string findFile(const std::string &p) { return p; } }
inline Mat imread(const std::string &, int = IMREAD_COLOR) { return Mat(); }
inline bool imwrite(const std::string &, const Mat &) { return false; }
inline void imshow(const std::string &, const Mat &) {}
inline void moveWindow(const std::string &, int, int) {}
inline int waitKey(int = 0) { return -1; }
inline void cvtColor(const Mat &, Mat &, int) { throw std::runtime_error("shim: cvtColor is not on the driver's path"); }
inline void normalize(const Mat &, Mat &, double = 1, double = 0, int = 0, int = -1) {}
inline void minMaxLoc(const Mat &, double *, double *, Point *, Point *) {}
inline void vconcat(const Mat &, const Mat &, Mat &) {}
inline void line(Mat &, Point, Point, const Scalar &, int = 1, int = LINE_8) {}
inline void circle(Mat &, Point, int, const Scalar &, int = 1) {}

}  // namespace cv
