// rotation.cpp -- host side of SLAM::Rotation (reference: include/src/Rotation/rotation.cpp).
#include "Rotation/rotation.h"

#include <algorithm>
#include <cmath>
#include <cstring>
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

// drawRotated (:133-145): roi_rotation_mat (2 x 3, CV_64F) * [x y 1]^T -- cv::gemm sums the three products left to right in
// double -- and (int) truncation.  The circles of draw = true are not drawn (no HighGUI here).
vslam::Point Rotation::drawRotated(vslam::Point &pt, Mat &, Mat &m, bool) {
    if (m.type() != vslam::CV_64F || m.rows != 2 || m.cols != 3) throw std::invalid_argument("drawRotated: 2 x 3 CV_64F matrix expected");
    volatile double s0 = 0, s1 = 0;
    s0 = s0 + m.at<double>(0, 0) * (double)pt.x; s0 = s0 + m.at<double>(0, 1) * (double)pt.y; s0 = s0 + m.at<double>(0, 2) * 1.0;
    s1 = s1 + m.at<double>(1, 0) * (double)pt.x; s1 = s1 + m.at<double>(1, 1) * (double)pt.y; s1 = s1 + m.at<double>(1, 2) * 1.0;
    return vslam::Point((int)s0, (int)s1);
}

// a Rect of a Mat as a Mat of its own (cv::Mat(src, Rect) is a view; warpAffine never looks outside it, so a copy is equivalent);
// throws where OpenCV's constructor asserts
static Mat crop_of(const Mat &src, int x, int y, int w, int h) {
    if (x < 0 || y < 0 || w < 0 || h < 0 || x + w > src.cols || y + h > src.rows) throw std::out_of_range("Rotation: Rect outside the matrix");
    Mat r(h, w, vslam::CV_8U);
    for (int i = 0; i < h; ++i) std::memcpy(r.data + (size_t)i * r.step, src.data + (size_t)(y + i) * src.step + x, (size_t)w);
    return r;
}

Mat Rotation::doubleCrop(Mat &src, const Point2i &center, int windowSize, double angle) {               // :147-205
    if (src.type() != vslam::CV_8U) throw std::invalid_argument("doubleCrop: CV_8U only");
    const int padding = windowSize / 2;
    const double scale = 1.0;
    Mat rotation_mat = vslam::getRotationMatrix2D((float)center.x, (float)center.y, angle, scale);       // :152
    vslam::Point pos1(center.x - padding, center.y - padding), pos2(center.x + padding, center.y - padding);   // :156-166
    vslam::Point pos3(center.x - padding, center.y + padding), pos4(center.x + padding, center.y + padding);
    const vslam::Point r[4] = {drawRotated(pos1, src, rotation_mat, false), drawRotated(pos2, src, rotation_mat, false),
                               drawRotated(pos3, src, rotation_mat, false), drawRotated(pos4, src, rotation_mat, false)};
    int min_x = r[0].x, max_x = r[0].x, min_y = r[0].y, max_y = r[0].y;                                  // :169-175
    for (int i = 1; i < 4; ++i) {
        min_x = std::min(min_x, r[i].x); max_x = std::max(max_x, r[i].x);
        min_y = std::min(min_y, r[i].y); max_y = std::max(max_y, r[i].y);
    }
    Mat ROI = crop_of(src, min_x, min_y, max_x - min_x, max_y - min_y);                                  // :181
    Mat roi_rotation_mat = vslam::getRotationMatrix2D((float)(ROI.cols / 2), (float)(ROI.rows / 2), angle, scale);   // :184-185
    vslam::Point roi_pos1(ROI.cols / 2, 0), roi_pos4(ROI.cols / 2, ROI.rows);                            // :187, :190
    const vslam::Point p1 = drawRotated(roi_pos1, ROI, roi_rotation_mat, false), p4 = drawRotated(roi_pos4, ROI, roi_rotation_mat, false);
    Mat rotate_dst;
    vslam::warpAffine(ROI, rotate_dst, roi_rotation_mat, ROI.size());                                    // :198
    const int x0 = std::min(p1.x, p4.x), y0 = std::min(p1.y, p4.y);                                      // :201 Rect(p1, p4)
    return crop_of(rotate_dst, x0, y0, std::max(p1.x, p4.x) - x0, std::max(p1.y, p4.y) - y0);
}

}  // namespace SLAM
