// Harris_corners.cpp -- host side of the Harris free functions (reference: KeyPointDetection/Harris_corners.cpp).
#include "Harris_corners.hpp"

#include <stdexcept>

using namespace vslam;

static void need(const Mat &m, int type, const char *what) {
    if (m.empty() || m.type() != type) throw std::invalid_argument(std::string(what) + ": unexpected image type");
}

// Harris_corners.cpp:10-29: the closed window around (i, j) of the (already padded) gradient maps goes to the GPU, which
// accumulates Ix2, Iy2, IxIy in the reference's order.  HarrisCorner below does not call this (its kernel forms the sums
// of every pixel itself); it is here because the reference exposes it.
void StructureMatrix(Mat &M, Mat &Ix, Mat &Iy, int padding, int i, int j) {
    need(Ix, CV_32F, "StructureMatrix"); need(Iy, CV_32F, "StructureMatrix"); need(M, CV_32F, "StructureMatrix");
    std::vector<float> gx, gy;
    for (int u = i - padding; u <= i + padding; u++)
        for (int v = j - padding; v <= j + padding; v++) {
            gx.push_back(Ix.at<float>(u, v));
            gy.push_back(Iy.at<float>(u, v));
        }
    float sums[3] = {0, 0, 0};
    throw_on(vslam_edge_from_windows(host_context(), gx.data(), gy.data(), 1, (int)gx.size(), nullptr, sums), "vslam_edge_from_windows");
    M.at<float>(0, 0) = sums[0]; M.at<float>(0, 1) = sums[2];
    M.at<float>(1, 0) = sums[2]; M.at<float>(1, 1) = sums[1];
}

// Harris_corners.cpp:31-68 (windowSize 3, k 0.04)
Mat HarrisCorner(Mat &Ix, Mat &Iy) {
    need(Ix, CV_32F, "HarrisCorner"); need(Iy, CV_32F, "HarrisCorner");
    if (Ix.rows != Iy.rows || Ix.cols != Iy.cols) throw std::invalid_argument("HarrisCorner: gradient maps differ in size");
    Mat image = Mat::zeros(Ix.rows, Ix.cols, CV_32F);
    throw_on(vslam_harris_response(host_context(), Ix.ptr<float>(), Iy.ptr<float>(), Ix.rows, Ix.cols, 3, 0.04f, image.ptr<float>()),
             "vslam_harris_response");
    return image;
}

// Harris_corners.cpp:70-81 (the reference ignores windowSize and dilates with a 3 x 3 kernel)
Mat NonMaximumSuppression(Mat &response, int windowSize) {
    (void)windowSize;
    need(response, CV_8U, "NonMaximumSuppression");
    Mat mask(response.rows, response.cols, CV_8U);
    throw_on(vslam_nms3(host_context(), response.data, response.rows, response.cols, mask.data), "vslam_nms3");
    return mask;
}

// Harris_corners.cpp:83-129
Mat NMS2(Mat &response, int windowSize) {
    need(response, CV_32F, "NMS2");
    Mat out = Mat::zeros(response.rows, response.cols, CV_32F);
    throw_on(vslam_nms2(host_context(), response.ptr<float>(), response.rows, response.cols, windowSize, out.ptr<float>()), "vslam_nms2");
    return out;
}

void GaussianBlur3x3(const Mat &src, Mat &dst) {
    need(src, CV_8U, "GaussianBlur3x3");
    Mat out(src.rows, src.cols, CV_8U);
    throw_on(vslam_gaussian_blur3(host_context(), src.data, src.rows, src.cols, src.step, out.data, out.step), "vslam_gaussian_blur3");
    dst = out;
}

void Sobel1(const Mat &src, Mat &grad_x, Mat &grad_y) {
    need(src, CV_8U, "Sobel1");
    Mat gx(src.rows, src.cols, CV_32F), gy(src.rows, src.cols, CV_32F);
    throw_on(vslam_sobel1(host_context(), src.data, src.rows, src.cols, src.step, gx.ptr<float>(), gy.ptr<float>()), "vslam_sobel1");
    grad_x = gx; grad_y = gy;
}

void convertScaleAbs(const Mat &src, Mat &dst) {
    need(src, CV_32F, "convertScaleAbs");
    Mat out(src.rows, src.cols, CV_8U);
    throw_on(vslam_convert_scale_abs(host_context(), src.ptr<float>(), src.rows, src.cols, out.data), "vslam_convert_scale_abs");
    dst = out;
}

std::vector<HarrisCornerPoint> harrisFromFrame(const Mat &img, Mat *response, Mat *abs_response, Mat *nms_mask, int nms2_window) {
    need(img, CV_8U, "harrisFromFrame");
    vslam_harris_params hp;
    hp.k = 0.04f; hp.nms2_window = nms2_window;
    if (response) response->create(img.rows, img.cols, CV_32F);
    if (abs_response) abs_response->create(img.rows, img.cols, CV_8U);
    if (nms_mask) nms_mask->create(img.rows, img.cols, CV_8U);
    std::vector<vslam_corner> buf(img.total() + 1);
    int count = 0;
    throw_on(vslam_harris(host_context(), img.data, 1, img.rows, img.cols, img.step, img.step * (size_t)img.rows, VSLAM_HOST, &hp,
                          response ? response->ptr<float>() : nullptr, abs_response ? abs_response->data : nullptr,
                          nms_mask ? nms_mask->data : nullptr, buf.data(), (int)img.total(), VSLAM_HOST, &count),
             "vslam_harris");
    std::vector<HarrisCornerPoint> out((size_t)count);
    for (int i = 0; i < count; ++i) out[i] = HarrisCornerPoint{buf[i].row, buf[i].col, buf[i].value};
    return out;
}
