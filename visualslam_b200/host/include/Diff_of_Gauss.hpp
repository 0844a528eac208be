// Diff_of_Gauss.hpp -- the DoG free functions of the reference (KeyPointDetection/Diff_of_Gauss.cpp keeps them in the
// .cpp; promoted to a header so that other code can call them), same names, argument order and results:
//   SLAM::point                         Diff_of_Gauss.cpp:25-35
//   computeEdgeResponse                 :79-109
//   FeaturePointLocalization            :223-251
//   initialKeypointDetection            :254-297
//   filterKeypoints                     :301-369   = filterKeypointsEdge (:326-338, the end of the hot path) + assignOrientations (:340-369)
//   SIFT                                :561-693   (valid: optional per-descriptor flags, see include/vslam.h)
//   writeFeatureDescriptors             :837-863   featureDescriptors.dat
// All arithmetic runs on the GPU through the C ABI (include/vslam.h); these functions only move data.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

#include "GaussPyramid.hpp"

namespace SLAM {
struct point {
    point(int row_, int col_, int value_, int padding_, int octave_, int level_)
        : row(row_), col(col_), value(value_), padding(padding_), octave(octave_), level(level_) {}
    int row;
    int col;
    int value;
    int padding;
    int octave;
    int level;
};
}  // namespace SLAM

float computeEdgeResponse(const SLAM::point &keypoint, const Mat &grad_x, const Mat &grad_y);
bool FeaturePointLocalization(std::vector<Mat> &dogs_padded, std::vector<SLAM::point> &keypoints, int level, SLAM::point &point);
void initialKeypointDetection(std::vector<SLAM::point> &keypoints, GaussPyramid &pyramid, int octave, int windowSize);
void filterKeypointsEdge(GaussPyramid &pyramid, int octave, std::vector<SLAM::point> &keypoints, std::vector<SLAM::point> &edgeKeypoints);
void assignOrientations(GaussPyramid &pyramid, int octave, const std::vector<SLAM::point> &edgeKeypoints, std::vector<SLAM::point> &reducedKeypoints);
void filterKeypoints(GaussPyramid &pyramid, int octave, std::vector<SLAM::point> &keypoints, std::vector<SLAM::point> &reducedKeypoints);
void SIFT(std::vector<SLAM::point> &reducedKeypoints, std::vector<std::vector<float>> &featureDescriptors_vec, GaussPyramid &pyramid, int octave,
          std::vector<uint8_t> *valid = nullptr);
bool writeFeatureDescriptors(const std::string &file_name, const std::vector<std::vector<float>> &featureDescriptors_vec);

// The whole hot path of ./DoG for one image in one call (vslam_dog_keypoints: fused kernels, nothing but keypoints
// leaves the GPU): every octave's keypoints after the edge threshold, in the reference's order.
std::vector<SLAM::point> detectKeypoints(Mat &img, int numOctaves, double sigma, int windowSize = 3);
