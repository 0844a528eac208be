// rotation.h -- SLAM::Rotation of the reference (KeyPointDetection/include/src/Rotation/rotation.h:11-26), same static
// methods, on vslam::Mat.  The point helpers are host geometry exactly as written in rotation.cpp:5-62 (float products of
// int offsets, truncation); the two image gathers run on the GPU through the C ABI (vslam_rotate_mat_ccw,
// vslam_rotated_window).  drawRotated / doubleCrop (rotation.cpp:132-205): the point arithmetic is host code as written
// (M * [x y 1]^T in double, truncation), cv::warpAffine runs on the GPU (vslam_warp_affine); drawRotated's circles are not
// drawn (no HighGUI on this box).
#pragma once
#include <vector>

#include "../GaussPyramid.hpp"

struct Point2i {
    int x = 0, y = 0;
    Point2i() = default;
    Point2i(int x_, int y_) : x(x_), y(y_) {}
};
struct Point2f {
    float x = 0, y = 0;
    Point2f() = default;
    Point2f(float x_, float y_) : x(x_), y(y_) {}
};

namespace SLAM {
class Transform {};

class Rotation : public Transform {
  public:
    static float convertToRadians(float theta);
    static const Point2f cos_sin_of_angle(float theta, bool degrees = true);
    static Point2i rotate_pt_CW(const Point2i &pt, const Point2i &center, const Point2f &angles);
    static Point2i rotate_pt_CCW(const Point2i &pt, const Point2i &center, const Point2f &angles);
    static Point2i rotate_pt_CCW(const Point2i &pt, const Point2i &center, float theta, bool degrees = true);
    static Point2i rotate_pt_CW(const Point2i &pt, const Point2i &center, float theta, bool degrees = true);
    static Mat rotate_mat_CCW(Mat &I, const Point2i &center, const Point2f &angles);
    static Mat getRotatedWindow(Mat &I, const Point2i &center, int windowSize, float theta, bool degrees = true);
    static std::vector<Point2i> getRotatedWindowPoints(Mat &I, const Point2i &center, int windowSize, float theta, bool degrees = true);
    static vslam::Point drawRotated(vslam::Point &pt, Mat &src, Mat &roi_rotation_mat, bool draw = true);
    static Mat doubleCrop(Mat &src, const Point2i &center, int windowSize, double angle);
};
}  // namespace SLAM
