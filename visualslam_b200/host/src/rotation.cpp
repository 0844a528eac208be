// rotation.cpp -- host side of SLAM::Rotation (reference: include/src/Rotation/rotation.cpp).
#include "Rotation/rotation.h"

#include <cmath>
#include <stdexcept>

namespace SLAM {

static const double kPi = 3.1415926535897932384626433832795;      // CV_PI

float Rotation::convertToRadians(float theta) { return theta * (kPi / 180.0f); }       // rotation.cpp:5-7

const Point2f Rotation::cos_sin_of_angle(float theta, bool degrees) {                  // :9-17
    float angle = degrees ? convertToRadians(theta) : theta;
    return Point2f((float)std::cos((double)angle), (float)std::sin((double)angle));
}

Point2i Rotation::rotate_pt_CW(const Point2i &pt, const Point2i &center, const Point2f &angles) {     // :19-28
    const int rx = pt.x - center.x, ry = pt.y - center.y;
    volatile float a = rx * angles.x, b = ry * angles.y, c = rx * angles.y, d = ry * angles.x;          // products rounded to float
    return Point2i((int)(a - b) + center.x, (int)(c + d) + center.y);
}

Point2i Rotation::rotate_pt_CCW(const Point2i &pt, const Point2i &center, const Point2f &angles) {    // :30-39
    const int rx = pt.x - center.x, ry = pt.y - center.y;
    volatile float a = rx * angles.x, b = ry * angles.y, c = -rx * angles.y, d = ry * angles.x;
    return Point2i((int)(a + b) + center.x, (int)(c + d) + center.y);
}

Point2i Rotation::rotate_pt_CCW(const Point2i &pt, const Point2i &center, float theta, bool degrees) {  // :42-52
    return rotate_pt_CCW(pt, center, cos_sin_of_angle(theta, degrees));
}
Point2i Rotation::rotate_pt_CW(const Point2i &pt, const Point2i &center, float theta, bool degrees) {   // :54-64
    return rotate_pt_CW(pt, center, cos_sin_of_angle(theta, degrees));
}

Mat Rotation::rotate_mat_CCW(Mat &I, const Point2i &center, const Point2f &angles) {                     // :66-84
    if (I.type() != vslam::CV_8U) throw std::invalid_argument("rotate_mat_CCW: CV_8U only (as the reference)");
    Mat rotated = Mat::zeros(I.rows, I.cols, vslam::CV_8U);
    if (!I.empty())
        vslam::throw_on(vslam_rotate_mat_ccw(vslam::host_context(), I.data, I.rows, I.cols, I.step, center.x, center.y, angles.x, angles.y,
                                             rotated.data, rotated.step), "vslam_rotate_mat_ccw");
    return rotated;
}

Mat Rotation::getRotatedWindow(Mat &I, const Point2i &center, int windowSize, float theta, bool degrees) {   // :87-108
    if (I.type() != vslam::CV_8U || windowSize <= 0) throw std::invalid_argument("getRotatedWindow: CV_8U image, positive window");
    Mat ROI = Mat::zeros(windowSize, windowSize, vslam::CV_8U);
    const Point2f angles = cos_sin_of_angle(theta, degrees);
    vslam::throw_on(vslam_rotated_window(vslam::host_context(), I.data, I.rows, I.cols, I.step, center.x, center.y, windowSize, angles.x,
                                         angles.y, ROI.data, ROI.step), "vslam_rotated_window");
    return ROI;
}

std::vector<Point2i> Rotation::getRotatedWindowPoints(Mat &, const Point2i &center, int windowSize, float theta, bool degrees) {   // :112-130
    std::vector<Point2i> rotatedPoints;
    const int padding = windowSize / 2;
    const Point2f angles = cos_sin_of_angle(theta, degrees);
    rotatedPoints.reserve((size_t)windowSize * windowSize);
    for (int i = center.y - padding; i <= center.y + padding; ++i)          // '<=': (windowSize + 1)^2 points, as the reference
        for (int j = center.x - padding; j <= center.x + padding; ++j) rotatedPoints.emplace_back(rotate_pt_CW(Point2i(j, i), center, angles));
    return rotatedPoints;
}

Mat Rotation::doubleCrop(Mat &, const Point2i &, int, double) {
    throw std::runtime_error("Rotation::doubleCrop needs cv::getRotationMatrix2D / cv::warpAffine: not part of this build");
}

}  // namespace SLAM
