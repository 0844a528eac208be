// GaussPyramid.hpp -- C++17 mirror of the reference's GaussPyramid (KeyPointDetection/include/src/GaussPyramid/
// GaussPyramid.hpp:14-71): same class name, constructors, getters and return types, with cv::Mat replaced by vslam::Mat
// and all arithmetic behind the C ABI (include/vslam.h): the constructor builds the pyramid on the GPU
// (vslam_pyramid_build = createPyramid, GaussPyramid.cpp:106-131), the getters copy maps to host images on first use.
// There is no CPU implementation: construction throws std::runtime_error when no CUDA device is available.
#pragma once
#include <cmath>
#include <iostream>
#include <map>
#include <string>
#include <vector>

#include "vslam.h"
#include "vslam_mat.hpp"

using vslam::Mat;
using vslam::Point;

namespace vslam {
// the process-wide context the host layer runs on (device $VSLAM_DEVICE, default 0); throws when there is no GPU
vslam_ctx *host_context();
void throw_on(int status, const char *what);
}  // namespace vslam

class GaussPyramid {
  public:
    GaussPyramid(Mat &img, int numOctaves, double sigma) : numOctaves_{numOctaves}, sigma_{sigma} { this->createPyramid(img); }
    GaussPyramid(Mat &img, double sigma) : sigma_{sigma} {
        this->calculateNumOctaves(img);
        this->createPyramid(img);
    }
    ~GaussPyramid();
    GaussPyramid(const GaussPyramid &) = delete;
    GaussPyramid &operator=(const GaussPyramid &) = delete;
    // member variable getters
    int getNumOctaves() { return this->numOctaves_; }
    int getNumLevels() { return this->numLevels_; }
    int getNumScaleSamples() { return this->scaleSamples_; }
    // octave getters (std::out_of_range on a bad octave, as map::at / vector::at in the reference)
    double getSigmaAt(int octave, int level);
    double calculateSigma(int octave, int level);
    const Mat &octaveImage(int octave) { return this->img_pyramid.at(octave); }
    const std::vector<double> &octaveSigma(int octave) { return this->sigmas_pyramid.at(octave); }
    const std::vector<Mat> &octaveBlur(int octave) { return fetch(gauss_pyramid, VSLAM_MAP_GAUSS, octave); }
    const std::vector<Mat> &octaveDiff(int octave) { return fetch(diff_pyramid, VSLAM_MAP_DOG, octave); }
    const std::vector<Mat> &octaveGradX(int octave) { return fetch(grad_x_pyramid, VSLAM_MAP_GRADX, octave); }
    const std::vector<Mat> &octaveGradY(int octave) { return fetch(grad_y_pyramid, VSLAM_MAP_GRADY, octave); }
    const std::vector<Mat> &octaveGradMag(int octave) { return fetch(grad_mag_pyramid, VSLAM_MAP_MAG, octave); }
    const std::vector<Mat> &octaveGradOrient(int octave) { return fetch(grad_orient_pyramid, VSLAM_MAP_ORIENT, octave); }
    // pyramid getters
    const std::vector<Mat> &imagePyramid() { return this->img_pyramid; }
    const std::map<int, std::vector<Mat>> &pyramidGauss() { return fetch_all(gauss_pyramid, VSLAM_MAP_GAUSS); }
    const std::map<int, std::vector<Mat>> &pyramidDiff() { return fetch_all(diff_pyramid, VSLAM_MAP_DOG); }
    const std::map<int, std::vector<Mat>> &pyramidGradX() { return fetch_all(grad_x_pyramid, VSLAM_MAP_GRADX); }
    const std::map<int, std::vector<Mat>> &pyramidGradY() { return fetch_all(grad_y_pyramid, VSLAM_MAP_GRADY); }
    const std::map<int, std::vector<Mat>> &pyramidGradMag() { return fetch_all(grad_mag_pyramid, VSLAM_MAP_MAG); }
    const std::map<int, std::vector<Mat>> &pyramidGradOrient() { return fetch_all(grad_orient_pyramid, VSLAM_MAP_ORIENT); }
    static std::vector<Mat> padOctave(int padding, const std::vector<Mat> &images);
    // display functions: headless -- the images are written as octave_<name>_<level>.pgm / .pfm into $VSLAM_DUMP_DIR
    // (nothing happens when the variable is unset); HighGUI is unavailable on this box
    static void showOctave(const std::vector<Mat> images, const std::string window_name, const Point pos = Point(0, 0));
    static void showPyramid(const std::map<int, std::vector<Mat>> pyramid);
    // extension: the device-resident pyramid, for the detector entry points of Diff_of_Gauss.hpp
    const vslam_pyramid *devicePyramid() const { return pyr_; }

  private:
    void calculateNumOctaves(Mat &img);
    double calculateSigma(int level);
    void createPyramid(Mat &img);
    const std::vector<Mat> &fetch(std::map<int, std::vector<Mat>> &store, int kind, int octave);
    const std::map<int, std::vector<Mat>> &fetch_all(std::map<int, std::vector<Mat>> &store, int kind);
    std::vector<Mat> img_pyramid;
    std::map<int, std::vector<double>> sigmas_pyramid;
    std::map<int, std::vector<Mat>> gauss_pyramid;
    std::map<int, std::vector<Mat>> diff_pyramid;
    std::map<int, std::vector<Mat>> grad_x_pyramid;
    std::map<int, std::vector<Mat>> grad_y_pyramid;
    std::map<int, std::vector<Mat>> grad_mag_pyramid;
    std::map<int, std::vector<Mat>> grad_orient_pyramid;
    int numOctaves_ = 0;
    int scaleSamples_ = 3;
    int numLevels_ = scaleSamples_ + 3;
    int currentOctave_ = 0;
    double sigma_ = 0.0f;
    double k_ = pow(2.0f, 1.0f / double(scaleSamples_));
    vslam_pyramid *pyr_ = nullptr;
};
