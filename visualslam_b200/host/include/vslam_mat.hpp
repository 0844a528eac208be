// vslam_mat.hpp -- the image type of the C++17 host layer.
//
// The reference's surface is written against cv::Mat (GaussPyramid.hpp:2-6); OpenCV C++ does not exist on the target
// box, so the drop-in keeps the reference's names and call shapes on this minimal ref-counted matrix: rows, cols, type
// (CV_8U / CV_32F; CV_64F for the 2 x 3 matrices of SLAM::Rotation), at<T>(r, c), clone(), zeros(), plus PGM (8U) / PFM (32F) file IO instead of imread/imshow.
// No arithmetic lives here: every operator of the hot path goes through the C ABI (include/vslam.h).
#pragma once
#include <cstdint>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

namespace vslam {

constexpr int CV_8U = 0, CV_32F = 5, CV_64F = 6, CV_8UC3 = 16;   // OpenCV's type codes, so ported call sites read the same
constexpr int IMREAD_GRAYSCALE = 0, IMREAD_COLOR = 1, COLOR_BGR2GRAY = 6;

struct Size {
    int width = 0, height = 0;
};
struct Point {
    int x = 0, y = 0;
    Point() = default;
    Point(int x_, int y_) : x(x_), y(y_) {}
};

class Mat {
  public:
    int rows = 0, cols = 0;
    Mat() = default;
    Mat(int r, int c, int type) { create(r, c, type); }
    static Mat zeros(int r, int c, int type) { return Mat(r, c, type); }   // create() zero-fills
    void create(int r, int c, int type) {
        if (r < 0 || c < 0 || (type != CV_8U && type != CV_32F && type != CV_64F && type != CV_8UC3)) throw std::invalid_argument("vslam::Mat: bad shape or type");
        rows = r; cols = c; type_ = type;
        step = (size_t)c * elemSize();
        buf_ = std::shared_ptr<uint8_t>(new uint8_t[step * (size_t)r + 1](), std::default_delete<uint8_t[]>());
        data = buf_.get();
    }
    int type() const { return type_; }
    size_t elemSize() const { return type_ == CV_32F ? 4 : type_ == CV_64F ? 8 : type_ == CV_8UC3 ? 3 : 1; }
    int channels() const { return type_ == CV_8UC3 ? 3 : 1; }
    bool empty() const { return !data || rows == 0 || cols == 0; }
    Size size() const { return Size{cols, rows}; }
    size_t total() const { return (size_t)rows * cols; }
    Mat clone() const {
        Mat m(rows, cols, type_);
        if (!empty()) std::memcpy(m.data, data, step * (size_t)rows);
        return m;
    }
    template <class T> T &at(int r, int c) { return *reinterpret_cast<T *>(data + (size_t)r * step + (size_t)c * sizeof(T)); }
    template <class T> const T &at(int r, int c) const { return *reinterpret_cast<const T *>(data + (size_t)r * step + (size_t)c * sizeof(T)); }
    template <class T> T *ptr(int r = 0) { return reinterpret_cast<T *>(data + (size_t)r * step); }
    template <class T> const T *ptr(int r = 0) const { return reinterpret_cast<const T *>(data + (size_t)r * step); }

    uint8_t *data = nullptr;
    size_t step = 0;                              // bytes between rows (dense)

  private:
    int type_ = CV_8U;
    std::shared_ptr<uint8_t> buf_;
};

// headless IO: binary PGM (P5, 8U) and PFM (Pf, little endian, 32F; rows stored top to bottom with a negative scale)
Mat imread_pgm(const std::string &path);                 // throws std::runtime_error
// imread(path, IMREAD_COLOR / IMREAD_GRAYSCALE) for binary PNM files: P6 -> CV_8UC3 in BGR order (what cv::imread hands to
// cvtColor at Diff_of_Gauss.cpp:730-739), P5 -> CV_8U (IMREAD_COLOR replicates it into three channels, as cv::imread does)
Mat imread(const std::string &path, int flags = IMREAD_COLOR);
// cvtColor(src, dst, COLOR_BGR2GRAY) on the GPU (vslam_bgr_to_gray); the only conversion the reference uses
void cvtColor(const Mat &src, Mat &dst, int code);
// cv::getRotationMatrix2D / cv::warpAffine as SLAM::Rotation::doubleCrop calls them (rotation.cpp:152,185,198): a 2 x 3 CV_64F
// matrix (host arithmetic behind vslam_get_rotation_matrix_2d) and the bilinear warp of an 8U image on the GPU (vslam_warp_affine)
Mat getRotationMatrix2D(float center_x, float center_y, double angle, double scale);
void warpAffine(const Mat &src, Mat &dst, const Mat &M, Size dsize);
void imwrite_pgm(const std::string &path, const Mat &m);
void imwrite_pfm(const std::string &path, const Mat &m);

}  // namespace vslam
