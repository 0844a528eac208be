// vslam_mat.cpp -- PGM / PFM file IO of vslam::Mat (the headless stand-in for imread / imwrite / imshow).
#include "vslam_mat.hpp"

#include <cstdio>

namespace vslam {

static int pnm_int(FILE *f) {
    int c = fgetc(f);
    for (;;) {
        while (c == ' ' || c == '\n' || c == '\r' || c == '\t') c = fgetc(f);
        if (c != '#') break;
        while (c != '\n' && c != EOF) c = fgetc(f);
    }
    int v = 0, n = 0;
    while (c >= '0' && c <= '9') { v = v * 10 + (c - '0'); c = fgetc(f); ++n; }
    if (!n) return -1;
    return v;                                       // the single whitespace after the token has been consumed
}

Mat imread_pgm(const std::string &path) {
    FILE *f = fopen(path.c_str(), "rb");
    if (!f) throw std::runtime_error("cannot open " + path);
    char magic[3] = {0, 0, 0};
    if (fread(magic, 1, 2, f) != 2 || magic[0] != 'P' || magic[1] != '5') { fclose(f); throw std::runtime_error(path + ": not a binary PGM (P5)"); }
    const int w = pnm_int(f), h = pnm_int(f), mx = pnm_int(f);
    if (w <= 0 || h <= 0 || mx <= 0 || mx > 255) { fclose(f); throw std::runtime_error(path + ": unsupported PGM header"); }
    Mat m(h, w, CV_8U);
    const size_t got = fread(m.data, 1, (size_t)w * h, f);
    fclose(f);
    if (got != (size_t)w * h) throw std::runtime_error(path + ": truncated PGM");
    return m;
}

Mat imread(const std::string &path, int flags) {
    FILE *f = fopen(path.c_str(), "rb");
    if (!f) throw std::runtime_error("cannot open " + path);
    char magic[3] = {0, 0, 0};
    if (fread(magic, 1, 2, f) != 2 || magic[0] != 'P' || (magic[1] != '5' && magic[1] != '6')) { fclose(f); throw std::runtime_error(path + ": not a binary PNM (P5 / P6)"); }
    const int ch = magic[1] == '6' ? 3 : 1;
    const int w = pnm_int(f), h = pnm_int(f), mx = pnm_int(f);
    if (w <= 0 || h <= 0 || mx <= 0 || mx > 255) { fclose(f); throw std::runtime_error(path + ": unsupported PNM header"); }
    std::vector<uint8_t> raw((size_t)w * h * ch);
    const size_t got = fread(raw.data(), 1, raw.size(), f);
    fclose(f);
    if (got != raw.size()) throw std::runtime_error(path + ": truncated PNM");
    if (flags == IMREAD_GRAYSCALE) {
        if (ch != 1) throw std::runtime_error(path + ": IMREAD_GRAYSCALE of a colour PNM is the decoder's own conversion; read it IMREAD_COLOR and cvtColor");
        Mat m(h, w, CV_8U);
        std::memcpy(m.data, raw.data(), raw.size());
        return m;
    }
    Mat m(h, w, CV_8UC3);
    for (size_t i = 0; i < (size_t)w * h; ++i) {
        const uint8_t r = raw[ch * i], g = raw[ch * i + (ch == 3 ? 1 : 0)], b = raw[ch * i + (ch == 3 ? 2 : 0)];
        m.data[3 * i] = b; m.data[3 * i + 1] = g; m.data[3 * i + 2] = r;
    }
    return m;
}

void imwrite_pgm(const std::string &path, const Mat &m) {
    if (m.type() != CV_8U) throw std::invalid_argument("imwrite_pgm: CV_8U only");
    FILE *f = fopen(path.c_str(), "wb");
    if (!f) throw std::runtime_error("cannot write " + path);
    fprintf(f, "P5\n%d %d\n255\n", m.cols, m.rows);
    for (int r = 0; r < m.rows; ++r) fwrite(m.ptr<uint8_t>(r), 1, (size_t)m.cols, f);
    fclose(f);
}

void imwrite_pfm(const std::string &path, const Mat &m) {
    if (m.type() != CV_32F) throw std::invalid_argument("imwrite_pfm: CV_32F only");
    FILE *f = fopen(path.c_str(), "wb");
    if (!f) throw std::runtime_error("cannot write " + path);
    fprintf(f, "Pf\n%d %d\n-1.0\n", m.cols, m.rows);
    for (int r = 0; r < m.rows; ++r) fwrite(m.ptr<float>(r), 4, (size_t)m.cols, f);
    fclose(f);
}

}  // namespace vslam
