// TEST INFRASTRUCTURE ONLY -- C entry points into the reference's OWN code, compiled unmodified from /root/reference
// against the OpenCV stand-in of oracle/ref_shim/ (see oracle/Makefile, target _ref -> oracle/_ref/libvslam_ref_*.so).
// Nothing in this file restates an algorithm: it builds cv::Mat arguments, calls the reference's functions in the order
// its main() does, and copies the results out.  Used by tools/make_ref_golden.py (fixtures under tests/golden/) and,
// when oracle/_ref exists, by tests/test_oracle_ref.py to pin the C oracle's loops to the reference itself.
//
// One .so per reference executable (both define free functions at global scope):
//   -DREF_HARRIS: linked with KeyPointDetection/Harris_corners.cpp          (-Dmain=harris_ref_main)
//   -DREF_DOG:    linked with KeyPointDetection/Diff_of_Gauss.cpp (-Dmain=dog_ref_main), GaussPyramid.cpp, rotation.cpp
#include <opencv2/shim.hpp>

using namespace cv;

#define REF_API extern "C" __attribute__((visibility("default")))

static Mat wrap_u8(const uint8_t *img, int h, int w) {
    Mat m(h, w, CV_8U);
    for (int i = 0; i < h; ++i) std::memcpy(m.data + (size_t)i * m.step, img + (size_t)i * w, (size_t)w);
    return m;
}

REF_API long ref_oob_count() { return cv::shim_oob(); }

#ifdef REF_HARRIS
// Harris_corners.cpp:10,31,70,83
void StructureMatrix(Mat &M, Mat &Ix, Mat &Iy, int padding, int i, int j);
Mat HarrisCorner(Mat &Ix, Mat &Iy);
Mat NonMaximumSuppression(Mat &response, int windowSize);
Mat NMS2(Mat &response, int windowSize);

// main() of Harris_corners.cpp:146-181 without the codec and display calls.  SQUARE images only: HarrisCorner swaps rows and
// columns (:40,47-48) and reads outside its matrices otherwise.  Outputs are n x n, row-major:
// response fp32, r8 = convertScaleAbs(response), mask3 = NonMaximumSuppression(r8, 3), nms2 = the values NMS2 writes
// (float plane; the reference writes them into a CV_8U Mat -- see shim.hpp).  Returns the number of out-of-buffer
// accesses the run made (0 on a square image).
REF_API long ref_harris(const uint8_t *img, int n, float *response, uint8_t *r8, uint8_t *mask3, float *nms2) {
    const long oob0 = cv::shim_oob();
    Mat image = wrap_u8(img, n, n);
    Mat blurred;
    GaussianBlur(image, blurred, Size(3, 3), 0, 0, BORDER_DEFAULT);                 // :158
    Mat grad_x, grad_y;
    Sobel(blurred, grad_x, CV_32F, 1, 0, 1, 1, 0, BORDER_DEFAULT);                   // :163
    Sobel(blurred, grad_y, CV_32F, 0, 1, 1, 1, 0, BORDER_DEFAULT);                   // :164
    Mat HResponse = HarrisCorner(grad_x, grad_y);                                    // :172
    Mat abs_HResponse;
    convertScaleAbs(HResponse, abs_HResponse);                                       // :176
    Mat nms = NonMaximumSuppression(abs_HResponse, 3);                               // :178
    Mat nms2m = NMS2(HResponse, 5);                                                  // :179
    for (int i = 0; i < n; ++i)
        for (int j = 0; j < n; ++j) {
            response[(size_t)i * n + j] = HResponse.at<float>(i, j);
            r8[(size_t)i * n + j] = abs_HResponse.at<uint8_t>(i, j);
            mask3[(size_t)i * n + j] = nms.at<uint8_t>(i, j);
            nms2[(size_t)i * n + j] = nms2m.shadow() ? nms2m.shadow()[(size_t)i * n + j] : 0.f;
        }
    return cv::shim_oob() - oob0;
}

// StructureMatrix on the replicate-padded gradient maps of an image, at padded position (i, j): the four entries
REF_API void ref_structure_matrix(const uint8_t *img, int n, int i, int j, float *m4) {
    Mat image = wrap_u8(img, n, n), blurred, gx, gy, pgx, pgy;
    GaussianBlur(image, blurred, Size(3, 3), 0, 0, BORDER_DEFAULT);
    Sobel(blurred, gx, CV_32F, 1, 0, 1, 1, 0, BORDER_DEFAULT);
    Sobel(blurred, gy, CV_32F, 0, 1, 1, 1, 0, BORDER_DEFAULT);
    copyMakeBorder(gx, pgx, 1, 1, 1, 1, BORDER_REPLICATE);
    copyMakeBorder(gy, pgy, 1, 1, 1, 1, BORDER_REPLICATE);
    Mat M = Mat::zeros(2, 2, CV_32F);
    StructureMatrix(M, pgx, pgy, 1, i, j);
    m4[0] = M.at<float>(0, 0); m4[1] = M.at<float>(0, 1); m4[2] = M.at<float>(1, 0); m4[3] = M.at<float>(1, 1);
}
#endif  // REF_HARRIS

#ifdef REF_DOG
#include <GaussPyramid/GaussPyramid.hpp>
#include <Rotation/rotation.h>

// Diff_of_Gauss.cpp:27-35 -- the type is local to that translation unit; the same declaration is needed to call its functions
namespace SLAM {
struct point {
    point(int row_, int col_, int value_, int padding_, int octave_, int level_) : row(row_), col(col_), value(value_), padding(padding_), octave(octave_), level(level_) {}
    int row;
    int col;
    int value;
    int padding;
    int octave;
    int level;
};
}  // namespace SLAM

// Diff_of_Gauss.cpp:79,223,254,301,561
float computeEdgeResponse(const SLAM::point &keypoint, const Mat &grad_x, const Mat &grad_y);
bool FeaturePointLocalization(std::vector<Mat> &dogs_padded, std::vector<SLAM::point> &keypoints, int level, SLAM::point &point);
void initialKeypointDetection(std::vector<SLAM::point> &keypoints, GaussPyramid &pyramid, int octave, int windowSize);
void filterKeypoints(GaussPyramid &pyramid, int octave, std::vector<SLAM::point> &keypoints, std::vector<SLAM::point> &reducedKeypoints);
void SIFT(std::vector<SLAM::point> &reducedKeypoints, std::vector<std::vector<float>> &featureDescriptors_vec, GaussPyramid &pyramid, int octave);

static void put(int32_t *out, long cap, long &n, const SLAM::point &p) {
    if (n < cap) { int32_t *q = out + 6 * n; q[0] = p.row; q[1] = p.col; q[2] = p.value; q[3] = p.padding; q[4] = p.octave; q[5] = p.level; }
    ++n;
}

// The hot path of Diff_of_Gauss.cpp main(): GaussPyramid{img, numOctaves, sigma} (:746), then per octave
// initialKeypointDetection (:785) and the edge test of filterKeypoints (:326-336: computeEdgeResponse < (r+1)^2/r).
// contrast[] <- keypoints after initialKeypointDetection (6 ints each), edge[] <- those that pass the edge test;
// counts[2 * octave + {0,1}] <- their numbers per octave.  Returns 0, or -1 on an exception from the stand-in.
REF_API int ref_dog_keypoints(const uint8_t *img, int h, int w, int num_octaves, double sigma, int window, int32_t *contrast, long contrast_cap,
                              int32_t *edge, long edge_cap, long *counts) {
    try {
        Mat image = wrap_u8(img, h, w);
        GaussPyramid pyramid{image, num_octaves, sigma};
        long nc = 0, ne = 0;
        for (const auto &kv : pyramid.pyramidGauss()) {
            const int octave = kv.first;
            std::vector<SLAM::point> keypoints;
            initialKeypointDetection(keypoints, pyramid, octave, window);
            long c0 = nc, e0 = ne;
            for (const auto &keypoint : keypoints) {
                put(contrast, contrast_cap, nc, keypoint);
                // Diff_of_Gauss.cpp:329-336, the reference's own expressions
                float response = computeEdgeResponse(keypoint, pyramid.octaveGradX(octave).at(keypoint.level), pyramid.octaveGradY(octave).at(keypoint.level));
                float r = 10.0f;
                float threshold = ((r + 1.0f) * (r + 1.0f)) / r;
                if (response < threshold) put(edge, edge_cap, ne, keypoint);
            }
            counts[2 * octave] = nc - c0;
            counts[2 * octave + 1] = ne - e0;
        }
        return 0;
    } catch (const std::exception &e) {
        std::cerr << "ref_dog_keypoints: " << e.what() << std::endl;
        return -1;
    }
}

// One map of the pyramid the reference's constructor builds.  kind: 0 octave image, 1 gauss, 2 dog (u8); 3 gradX, 4 gradY,
// 5 magnitude, 6 orientation (fp32).  dims <- rows, cols.  out may be null (sizes only).
REF_API int ref_pyramid_map(const uint8_t *img, int h, int w, int num_octaves, double sigma, int kind, int octave, int level, void *out, int *dims) {
    try {
        Mat image = wrap_u8(img, h, w);
        GaussPyramid pyramid{image, num_octaves, sigma};
        const Mat &m = kind == 0 ? pyramid.octaveImage(octave) : kind == 1 ? pyramid.octaveBlur(octave).at(level) : kind == 2 ? pyramid.octaveDiff(octave).at(level)
                       : kind == 3 ? pyramid.octaveGradX(octave).at(level) : kind == 4 ? pyramid.octaveGradY(octave).at(level)
                       : kind == 5 ? pyramid.octaveGradMag(octave).at(level) : pyramid.octaveGradOrient(octave).at(level);
        dims[0] = m.rows; dims[1] = m.cols;
        if (out)
            for (int i = 0; i < m.rows; ++i) std::memcpy((uint8_t *)out + (size_t)i * m.cols * m.elemSize(), m.data + (size_t)i * m.step, (size_t)m.cols * m.elemSize());
        return 0;
    } catch (const std::exception &e) {
        std::cerr << "ref_pyramid_map: " << e.what() << std::endl;
        return -1;
    }
}

REF_API double ref_sigma_at(const uint8_t *img, int h, int w, int num_octaves, double sigma, int octave, int level) {
    Mat image = wrap_u8(img, h, w);
    GaussPyramid pyramid{image, num_octaves, sigma};
    return pyramid.getSigmaAt(octave, level);
}

// FeaturePointLocalization on explicit derivatives: three 3x3 padded maps are built so that the reference's own reads
// (:226-228) produce dx, dy, ds, and the function runs as written.  Returns dog_zhat's decision and the new value.
REF_API int ref_localize(int dx, int dy, int ds, int value, int *new_value) {
    std::vector<Mat> dogs(3);
    for (auto &m : dogs) m = Mat::zeros(3, 3, CV_8U);
    auto split = [](int d, uint8_t &a, uint8_t &b) { a = (uint8_t)(d > 0 ? d : 0); b = (uint8_t)(d > 0 ? 0 : -d); };   // a - b == d
    uint8_t a, b;
    split(dx, a, b); dogs[1].at<uint8_t>(1, 0) = a; dogs[1].at<uint8_t>(1, 2) = b;
    split(dy, a, b); dogs[1].at<uint8_t>(0, 1) = a; dogs[1].at<uint8_t>(2, 1) = b;
    split(ds, a, b); dogs[0].at<uint8_t>(1, 1) = a; dogs[2].at<uint8_t>(1, 1) = b;
    std::vector<SLAM::point> kps;
    SLAM::point p(1, 1, value, 1, 0, 1);
    const bool kept = FeaturePointLocalization(dogs, kps, 1, p);
    *new_value = kept ? kps.at(0).value : value;
    return kept ? 1 : 0;
}

// The stage after the hot path, as main() runs it per octave (:785-791): filterKeypoints (edge test + orientation
// assignment) and SIFT.  oriented[] <- reducedKeypoints (row, col, angle, 0, octave, level); desc[] <- 128 floats per
// oriented keypoint; oob[] <- out-of-buffer reads SIFT made for that keypoint (rotateImageSection's swapped indices,
// :549-554: the reference's result is undefined there).  SIFT runs one keypoint at a time so that the count is per keypoint.
REF_API long ref_describe(const uint8_t *img, int h, int w, int num_octaves, double sigma, int32_t *oriented, float *desc, int32_t *oob, long cap) {
    try {
        Mat image = wrap_u8(img, h, w);
        GaussPyramid pyramid{image, num_octaves, sigma};
        long n = 0;
        for (const auto &kv : pyramid.pyramidGauss()) {
            const int octave = kv.first;
            std::vector<SLAM::point> keypoints, reduced;
            initialKeypointDetection(keypoints, pyramid, octave, 3);
            filterKeypoints(pyramid, octave, keypoints, reduced);
            for (const auto &p : reduced) {
                std::vector<SLAM::point> one{p};
                std::vector<std::vector<float>> d;
                const long o0 = cv::shim_oob();
                SIFT(one, d, pyramid, octave);
                if (n < cap) {
                    int32_t *q = oriented + 6 * n;
                    q[0] = p.row; q[1] = p.col; q[2] = p.value; q[3] = p.padding; q[4] = p.octave; q[5] = p.level;
                    for (int k = 0; k < 128; ++k) desc[128 * n + k] = d.at(0).at(k);
                    oob[n] = (int32_t)(cv::shim_oob() - o0);
                }
                ++n;
            }
        }
        return n;
    } catch (const std::exception &e) {
        std::cerr << "ref_describe: " << e.what() << std::endl;
        return -1;
    }
}

// SLAM::Rotation (rotation.cpp:66-108): the two image gathers on a u8 image
REF_API void ref_rotate_mat_ccw(const uint8_t *img, int h, int w, int cx, int cy, float theta_deg, uint8_t *out) {
    Mat I = wrap_u8(img, h, w);
    const Point2f angles = SLAM::Rotation::cos_sin_of_angle(theta_deg, true);
    Mat r = SLAM::Rotation::rotate_mat_CCW(I, Point2i(cx, cy), angles);
    for (int i = 0; i < h; ++i) std::memcpy(out + (size_t)i * w, r.data + (size_t)i * r.step, (size_t)w);
}
REF_API long ref_rotated_window(const uint8_t *img, int h, int w, int cx, int cy, int window, float theta_deg, uint8_t *out) {
    const long o0 = cv::shim_oob();
    Mat I = wrap_u8(img, h, w);
    Mat r = SLAM::Rotation::getRotatedWindow(I, Point2i(cx, cy), window, theta_deg, true);
    for (int i = 0; i < window; ++i) std::memcpy(out + (size_t)i * window, r.data + (size_t)i * r.step, (size_t)window);
    return cv::shim_oob() - o0;
}
REF_API void ref_rotate_pt(int x, int y, int cx, int cy, float theta_deg, int ccw, int *xr, int *yr) {
    const Point2i p = ccw ? SLAM::Rotation::rotate_pt_CCW(Point2i(x, y), Point2i(cx, cy), theta_deg, true)
                          : SLAM::Rotation::rotate_pt_CW(Point2i(x, y), Point2i(cx, cy), theta_deg, true);
    *xr = p.x; *yr = p.y;
}
// SLAM::Rotation::doubleCrop (rotation.cpp:147-205) as written; returns 0, or -3 where the stand-in's Mat(src, Rect) / warpAffine
// reject what OpenCV would reject (a Rect outside its Mat), -4 when out is too small
REF_API int ref_double_crop(const uint8_t *img, int h, int w, int cx, int cy, int window, double angle, uint8_t *out, int out_cap, int *oh, int *ow) {
    try {
        Mat I = wrap_u8(img, h, w);
        Mat r = SLAM::Rotation::doubleCrop(I, Point2i(cx, cy), window, angle);
        *oh = r.rows; *ow = r.cols;
        if ((long long)r.rows * r.cols > out_cap) return -4;
        for (int i = 0; i < r.rows; ++i) std::memcpy(out + (size_t)i * r.cols, r.data + (size_t)i * r.step, (size_t)r.cols);
        return 0;
    } catch (const std::exception &e) {
        return -3;
    }
}
#endif  // REF_DOG
