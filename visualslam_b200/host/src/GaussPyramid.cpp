// GaussPyramid.cpp -- host side of the GaussPyramid mirror (reference: include/src/GaussPyramid/GaussPyramid.cpp).
// createPyramid :106-131 -> vslam_pyramid_build; processGradients :65-104 -> on-demand gradient maps of vslam_pyramid_get;
// calculateNumOctaves :150-152, calculateSigma :156-162, getSigmaAt :187-189 -> the C ABI's sigma table.
#include "GaussPyramid.hpp"

#include <cstdlib>
#include <mutex>
#include <stdexcept>

namespace vslam {

void throw_on(int status, const char *what) {
    if (status == VSLAM_OK) return;
    throw std::runtime_error(std::string(what) + ": " + vslam_last_error(nullptr));
}

vslam_ctx *host_context() {
    static std::mutex mu;
    static vslam_ctx *ctx = nullptr;
    std::lock_guard<std::mutex> l(mu);
    if (!ctx) {
        const char *dev = std::getenv("VSLAM_DEVICE");
        throw_on(vslam_ctx_create(dev ? std::atoi(dev) : 0, &ctx), "vslam_ctx_create");
    }
    return ctx;
}

// cvtColor(img_color, img, COLOR_BGR2GRAY): Diff_of_Gauss.cpp:739, tests/GaussPyramid_Test.cpp:81
void cvtColor(const Mat &src, Mat &dst, int code) {
    if (code != COLOR_BGR2GRAY || src.type() != CV_8UC3) throw std::invalid_argument("cvtColor: COLOR_BGR2GRAY of a CV_8UC3 image only");
    Mat out(src.rows, src.cols, CV_8U);
    if (!src.empty())
        throw_on(vslam_bgr_to_gray(host_context(), src.data, 1, src.rows, src.cols, src.step, src.step * (size_t)src.rows, out.data, out.step,
                                   out.step * (size_t)out.rows), "vslam_bgr_to_gray");
    dst = out;
}

// cv::getRotationMatrix2D(Point2f center, angle, scale): rotation.cpp:152,185
Mat getRotationMatrix2D(float center_x, float center_y, double angle, double scale) {
    Mat m(2, 3, CV_64F);
    throw_on(vslam_get_rotation_matrix_2d(center_x, center_y, angle, scale, m.ptr<double>()), "vslam_get_rotation_matrix_2d");
    return m;
}

// cv::warpAffine(src, dst, M, dsize) with the defaults (INTER_LINEAR, BORDER_CONSTANT 0): rotation.cpp:198
void warpAffine(const Mat &src, Mat &dst, const Mat &M, Size dsize) {
    if (src.type() != CV_8U || src.empty() || M.type() != CV_64F || M.rows != 2 || M.cols != 3 || dsize.width <= 0 || dsize.height <= 0)
        throw std::invalid_argument("warpAffine: non-empty CV_8U image, 2 x 3 CV_64F matrix, positive size");
    Mat out(dsize.height, dsize.width, CV_8U);
    throw_on(vslam_warp_affine(host_context(), src.data, src.rows, src.cols, src.step, M.ptr<double>(), out.data, out.rows, out.cols, out.step),
             "vslam_warp_affine");
    dst = out;
}

}  // namespace vslam

using namespace vslam;

GaussPyramid::~GaussPyramid() { vslam_pyramid_destroy(pyr_); }

void GaussPyramid::calculateNumOctaves(Mat &img) { this->numOctaves_ = vslam_calculate_num_octaves(img.rows, img.cols); }

double GaussPyramid::calculateSigma(int level) { return vslam_calculate_sigma(this->currentOctave_, level, this->sigma_, this->scaleSamples_); }
double GaussPyramid::calculateSigma(int octave, int level) { return vslam_calculate_sigma(octave, level, this->sigma_, this->scaleSamples_); }
double GaussPyramid::getSigmaAt(int octave, int level) { return this->sigmas_pyramid.at(octave).at(level); }

void GaussPyramid::createPyramid(Mat &img) {
    if (img.empty() || img.type() != CV_8U) throw std::invalid_argument("GaussPyramid: 8-bit single-channel image expected");
    vslam_ctx *ctx = host_context();
    vslam_pyramid_params pp;
    pp.num_octaves = this->numOctaves_;
    pp.scale_samples = this->scaleSamples_;
    pp.sigma0 = this->sigma_;
    pp.materialise = 0;                                  // gradient maps are formed on demand by the getters
    throw_on(vslam_pyramid_build(ctx, img.data, 1, img.rows, img.cols, img.step, img.step * (size_t)img.rows, VSLAM_HOST, &pp, &pyr_),
             "vslam_pyramid_build");
    this->numOctaves_ = vslam_pyramid_num_octaves(pyr_);
    for (int o = 0; o < this->numOctaves_; ++o) {
        this->currentOctave_ = o;
        int r = 0, c = 0;
        throw_on(vslam_pyramid_octave_size(pyr_, o, &r, &c), "vslam_pyramid_octave_size");
        Mat base(r, c, CV_8U);
        throw_on(vslam_pyramid_get(ctx, pyr_, 0, VSLAM_MAP_BASE, o, 0, base.data, base.step, VSLAM_HOST), "vslam_pyramid_get");
        this->img_pyramid.push_back(base);
        std::vector<double> sig;
        for (int l = 0; l < this->numLevels_; ++l) {
            double s = 0;
            throw_on(vslam_pyramid_sigma(pyr_, o, l, &s), "vslam_pyramid_sigma");
            sig.push_back(s);
        }
        this->sigmas_pyramid[o] = sig;
    }
}

const std::vector<Mat> &GaussPyramid::fetch(std::map<int, std::vector<Mat>> &store, int kind, int octave) {
    if (octave < 0 || octave >= this->numOctaves_) throw std::out_of_range("GaussPyramid: octave out of range");
    auto it = store.find(octave);
    if (it != store.end()) return it->second;
    vslam_ctx *ctx = host_context();
    int r = 0, c = 0;
    throw_on(vslam_pyramid_octave_size(pyr_, octave, &r, &c), "vslam_pyramid_octave_size");
    const int n = kind == VSLAM_MAP_DOG ? this->numLevels_ - 1 : this->numLevels_;
    std::vector<Mat> v;
    for (int l = 0; l < n; ++l) {
        Mat m(r, c, (kind == VSLAM_MAP_GAUSS || kind == VSLAM_MAP_DOG) ? CV_8U : CV_32F);
        throw_on(vslam_pyramid_get(ctx, pyr_, 0, kind, octave, l, m.data, m.step, VSLAM_HOST), "vslam_pyramid_get");
        v.push_back(m);
    }
    return store.emplace(octave, std::move(v)).first->second;
}

const std::map<int, std::vector<Mat>> &GaussPyramid::fetch_all(std::map<int, std::vector<Mat>> &store, int kind) {
    for (int o = 0; o < this->numOctaves_; ++o) fetch(store, kind, o);
    return store;
}

// GaussPyramid.cpp:133-141: copyMakeBorder(BORDER_REPLICATE) + clone of every image.  A host-side copy: the GPU
// detector clamps indices instead (include/vslam.h), this exists for callers that want the padded images themselves.
std::vector<Mat> GaussPyramid::padOctave(int padding, const std::vector<Mat> &images) {
    std::vector<Mat> out;
    for (const Mat &m : images) {
        Mat p(m.rows + 2 * padding, m.cols + 2 * padding, m.type());
        const size_t es = m.elemSize();
        for (int r = 0; r < p.rows; ++r) {
            const int sr = std::min(std::max(r - padding, 0), m.rows - 1);
            for (int c = 0; c < p.cols; ++c) {
                const int sc = std::min(std::max(c - padding, 0), m.cols - 1);
                std::memcpy(p.data + (size_t)r * p.step + (size_t)c * es, m.data + (size_t)sr * m.step + (size_t)sc * es, es);
            }
        }
        out.push_back(p);
    }
    return out;
}

void GaussPyramid::showOctave(const std::vector<Mat> images, const std::string window_name, const Point) {
    const char *dir = std::getenv("VSLAM_DUMP_DIR");
    if (!dir) return;
    std::string name = window_name;
    for (char &ch : name)
        if (ch == ' ' || ch == '/') ch = '_';
    for (size_t i = 0; i < images.size(); ++i) {
        const std::string base = std::string(dir) + "/" + name + "_" + std::to_string(i);
        if (images[i].type() == CV_8U) imwrite_pgm(base + ".pgm", images[i]);
        else imwrite_pfm(base + ".pfm", images[i]);
    }
}

void GaussPyramid::showPyramid(const std::map<int, std::vector<Mat>> pyramid) {
    for (const auto &kv : pyramid) showOctave(kv.second, "Octave " + std::to_string(kv.first));
}
