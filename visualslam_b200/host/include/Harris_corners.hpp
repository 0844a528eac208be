// Harris_corners.hpp -- the Harris free functions of the reference (KeyPointDetection/Harris_corners.cpp), same names
// and argument order; plus the OpenCV calls of its main() (GaussianBlur 3x3, Sobel ksize 1, convertScaleAbs) under
// their OpenCV names, and the fused one-call path.  All arithmetic runs on the GPU through the C ABI.
//   StructureMatrix :10-29, HarrisCorner :31-68, NonMaximumSuppression :70-81, NMS2 :83-129
// Reference quirks that are NOT reproduced: HarrisCorner's swapped row/column loops (defined only for
// square images) and NMS2 writing floats into a CV_8U Mat; the functions below return H x W CV_32F maps.
#pragma once
#include <vector>

#include "GaussPyramid.hpp"

struct HarrisCornerPoint {
    int row, col;
    float value;
};

void StructureMatrix(Mat &M, Mat &Ix, Mat &Iy, int padding, int i, int j);
Mat HarrisCorner(Mat &Ix, Mat &Iy);
Mat NonMaximumSuppression(Mat &response, int windowSize);
Mat NMS2(Mat &response, int windowSize);

// the OpenCV calls of Harris_corners.cpp:157-164, 176 on 8U / 32F images
void GaussianBlur3x3(const Mat &src, Mat &dst);                       // GaussianBlur(src, dst, Size(3,3), 0, 0, BORDER_DEFAULT)
void Sobel1(const Mat &src, Mat &grad_x, Mat &grad_y);                // Sobel(src, ., CV_32F, 1,0 / 0,1, ksize = 1)
void convertScaleAbs(const Mat &src, Mat &dst);

// main()'s pipeline :157-179 in one fused kernel (vslam_harris): response, 8-bit response, 3x3 NMS mask, NMS2 corners
std::vector<HarrisCornerPoint> harrisFromFrame(const Mat &img, Mat *response = nullptr, Mat *abs_response = nullptr,
                                               Mat *nms_mask = nullptr, int nms2_window = 5);
