// Diff_of_Gauss.cpp -- host side of the DoG free functions (reference: KeyPointDetection/Diff_of_Gauss.cpp).
#include "Diff_of_Gauss.hpp"

#include <cstring>
#include <stdexcept>

using namespace vslam;

static_assert(sizeof(SLAM::point) == sizeof(vslam_keypoint), "SLAM::point and vslam_keypoint share one layout");

// Diff_of_Gauss.cpp:79-109: the 2 x 2 half-open window of the two gradient maps goes to the GPU, which accumulates the
// structure sums and forms tr^2 / det with the reference's rounding sequence.
float computeEdgeResponse(const SLAM::point &keypoint, const Mat &grad_x, const Mat &grad_y) {
    const int x = keypoint.col, y = keypoint.row, padding = 1;
    float gx[4], gy[4];
    int n = 0;
    for (int u = y - padding; u < y + padding; u++)
        for (int v = x - padding; v < x + padding; v++) {
            gx[n] = grad_x.at<float>(u, v);
            gy[n] = grad_y.at<float>(u, v);
            ++n;
        }
    float response = 0.f;
    throw_on(vslam_edge_from_windows(host_context(), gx, gy, 1, 4, &response, nullptr), "vslam_edge_from_windows");
    return response;
}

// Diff_of_Gauss.cpp:223-251
bool FeaturePointLocalization(std::vector<Mat> &dogs_padded, std::vector<SLAM::point> &keypoints, int level, SLAM::point &point) {
    const int i = point.row, j = point.col;
    int32_t q[4];
    q[0] = (int)dogs_padded[level].at<uint8_t>(i, j - 1) - (int)dogs_padded[level].at<uint8_t>(i, j + 1);
    q[1] = (int)dogs_padded[level].at<uint8_t>(i - 1, j) - (int)dogs_padded[level].at<uint8_t>(i + 1, j);
    q[2] = (int)dogs_padded[level - 1].at<uint8_t>(i, j) - (int)dogs_padded[level + 1].at<uint8_t>(i, j);
    q[3] = point.value;
    float dog_zhat = 0.f;
    throw_on(vslam_localize(host_context(), q, 1, &dog_zhat), "vslam_localize");
    if (dog_zhat > 0.03f) {
        point.value = (int)(dog_zhat * 255.0f);
        keypoints.emplace_back(point);
        return true;
    }
    return false;
}

// Diff_of_Gauss.cpp:254-297: extrema on the stride-windowSize grid + FeaturePointLocalization of every candidate, one octave
void initialKeypointDetection(std::vector<SLAM::point> &keypoints, GaussPyramid &pyramid, int octave, int windowSize) {
    vslam_ctx *ctx = host_context();
    vslam_dog_params dp;
    dp.window = windowSize; dp.contrast_thr = 0.03f; dp.edge_r = 10.0f;
    int rows = 0, cols = 0;
    throw_on(vslam_pyramid_octave_size(pyramid.devicePyramid(), octave, &rows, &cols), "vslam_pyramid_octave_size");
    const int pad = (windowSize - 1) / 2;
    const size_t ni = rows > pad ? (size_t)(rows - pad + windowSize - 1) / windowSize : 0;
    const size_t nj = cols > pad ? (size_t)(cols - pad + windowSize - 1) / windowSize : 0;
    const size_t cap = 3 * ni * nj;                        // every grid sample of the three inner levels
    std::vector<vslam_keypoint> buf(cap ? cap : 1);
    int count = 0;
    throw_on(vslam_dog_detect_octave(ctx, pyramid.devicePyramid(), 0, octave, &dp, 0, buf.data(), (int)cap, &count), "vslam_dog_detect_octave");
    keypoints.reserve(keypoints.size() + (size_t)count);
    for (int i = 0; i < count; ++i)
        keypoints.emplace_back(buf[i].row, buf[i].col, buf[i].value, buf[i].padding, buf[i].octave, buf[i].level);
}

// Diff_of_Gauss.cpp:326-338: computeEdgeResponse on the level's gradients, keep response < (r + 1)^2 / r, r = 10 (end of the hot path)
void filterKeypointsEdge(GaussPyramid &pyramid, int octave, std::vector<SLAM::point> &keypoints, std::vector<SLAM::point> &edgeKeypoints) {
    if (keypoints.empty()) return;
    std::vector<float> resp(keypoints.size());
    throw_on(vslam_edge_responses(host_context(), pyramid.devicePyramid(), 0, octave, reinterpret_cast<const vslam_keypoint *>(keypoints.data()),
                                  (int)keypoints.size(), resp.data()),
             "vslam_edge_responses");
    const float r = 10.0f;
    const float threshold = ((r + 1.0f) * (r + 1.0f)) / r;
    for (size_t i = 0; i < keypoints.size(); ++i)
        if (resp[i] < threshold) edgeKeypoints.emplace_back(keypoints[i]);
}

// Diff_of_Gauss.cpp:340-369: 16x16 window, Gaussian-weighted magnitudes, 36-bin histogram, one point per bin above 0.8 of the peak
void assignOrientations(GaussPyramid &pyramid, int octave, const std::vector<SLAM::point> &edgeKeypoints, std::vector<SLAM::point> &reducedKeypoints) {
    if (edgeKeypoints.empty()) return;
    std::vector<vslam_keypoint> buf(edgeKeypoints.size() * 36);
    int count = 0;
    throw_on(vslam_assign_orientations(host_context(), pyramid.devicePyramid(), 0, octave, reinterpret_cast<const vslam_keypoint *>(edgeKeypoints.data()),
                                       (int)edgeKeypoints.size(), buf.data(), (int)buf.size(), &count, nullptr),
             "vslam_assign_orientations");
    reducedKeypoints.reserve(reducedKeypoints.size() + (size_t)count);
    for (int i = 0; i < count; ++i)
        reducedKeypoints.emplace_back(buf[i].row, buf[i].col, buf[i].value, buf[i].padding, buf[i].octave, buf[i].level);
}

// Diff_of_Gauss.cpp:301-369, the whole function: edge threshold, then orientation assignment
void filterKeypoints(GaussPyramid &pyramid, int octave, std::vector<SLAM::point> &keypoints, std::vector<SLAM::point> &reducedKeypoints) {
    std::vector<SLAM::point> edgeKeypoints;
    filterKeypointsEdge(pyramid, octave, keypoints, edgeKeypoints);
    assignOrientations(pyramid, octave, edgeKeypoints, reducedKeypoints);
}

// Diff_of_Gauss.cpp:561-693: one 128-vector per oriented keypoint, appended to featureDescriptors_vec
void SIFT(std::vector<SLAM::point> &reducedKeypoints, std::vector<std::vector<float>> &featureDescriptors_vec, GaussPyramid &pyramid, int octave,
          std::vector<uint8_t> *valid) {
    if (reducedKeypoints.empty()) return;
    const size_t n = reducedKeypoints.size();
    std::vector<float> desc(n * 128);
    std::vector<uint8_t> ok(n);
    throw_on(vslam_sift_descriptors(host_context(), pyramid.devicePyramid(), 0, octave, reinterpret_cast<const vslam_keypoint *>(reducedKeypoints.data()),
                                    (int)n, desc.data(), ok.data()),
             "vslam_sift_descriptors");
    featureDescriptors_vec.reserve(featureDescriptors_vec.size() + n);
    for (size_t i = 0; i < n; ++i) featureDescriptors_vec.emplace_back(desc.begin() + i * 128, desc.begin() + (i + 1) * 128);
    if (valid) valid->insert(valid->end(), ok.begin(), ok.end());
}

// Diff_of_Gauss.cpp:837-863
bool writeFeatureDescriptors(const std::string &file_name, const std::vector<std::vector<float>> &featureDescriptors_vec) {
    std::vector<float> flat;
    flat.reserve(featureDescriptors_vec.size() * 128);
    for (const auto &d : featureDescriptors_vec) {
        if (d.size() != 128) throw std::invalid_argument("writeFeatureDescriptors: 128 values per descriptor");
        flat.insert(flat.end(), d.begin(), d.end());
    }
    return vslam_write_descriptor_file(file_name.c_str(), flat.data(), (int)featureDescriptors_vec.size()) == VSLAM_OK;
}

std::vector<SLAM::point> detectKeypoints(Mat &img, int numOctaves, double sigma, int windowSize) {
    if (img.empty() || img.type() != CV_8U) throw std::invalid_argument("detectKeypoints: 8-bit single-channel image expected");
    vslam_pyramid_params pp;
    pp.num_octaves = numOctaves; pp.scale_samples = 3; pp.sigma0 = sigma; pp.materialise = 0;
    vslam_dog_params dp;
    dp.window = windowSize; dp.contrast_thr = 0.03f; dp.edge_r = 10.0f;
    size_t cap = (size_t)img.rows * img.cols * 2 + 1024;      // 3 levels x (4/9) (1 + 1/4 + ...) samples per input pixel < 1.8
    std::vector<vslam_keypoint> buf(cap);
    int count = 0;
    throw_on(vslam_dog_keypoints(host_context(), img.data, 1, img.rows, img.cols, img.step, img.step * (size_t)img.rows, VSLAM_HOST, &pp, &dp,
                                 buf.data(), (int)cap, VSLAM_HOST, &count),
             "vslam_dog_keypoints");
    std::vector<SLAM::point> out;
    out.reserve((size_t)count);
    for (int i = 0; i < count; ++i) out.emplace_back(buf[i].row, buf[i].col, buf[i].value, buf[i].padding, buf[i].octave, buf[i].level);
    return out;
}
