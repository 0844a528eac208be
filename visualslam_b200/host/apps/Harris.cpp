// Harris -- the reference's ./Harris executable (KeyPointDetection/Harris_corners.cpp:146-191): 3x3 pre-blur, Sobel,
// HarrisCorner, convertScaleAbs + NonMaximumSuppression, NMS2, function by function as main() does, or with --fused in
// one kernel.  Prints the reference's stdout lines of NMS2 (:88-89 rows/cols, :127 largest value); the display tail
// (circles, imshow, imwrite on 's') is replaced by optional map dumps: HighGUI is unavailable on this box.
//   usage: Harris [image.pgm] [--fused] [--dump prefix]           (default image: ../images/chessboard.pgm)
#include <cstdio>
#include <cstring>
#include <iostream>

#include "Harris_corners.hpp"

int main(int argc, char **argv) {
    std::string path = "../images/chessboard.pgm", dump;
    bool fused = false;
    for (int i = 1; i < argc; ++i) {
        if (!strcmp(argv[i], "--fused")) fused = true;
        else if (!strcmp(argv[i], "--dump") && i + 1 < argc) dump = argv[++i];
        else if (!strcmp(argv[i], "--help")) { printf("usage: Harris [image.pgm] [--fused] [--dump prefix]\n"); return 0; }
        else path = argv[i];
    }
    try {
        Mat img = vslam::imread_pgm(path);
        Mat HResponse, abs_HResponse, nms, nms2;
        size_t corners = 0;
        float largest = 0.f;
        if (fused) {
            std::vector<HarrisCornerPoint> list = harrisFromFrame(img, &HResponse, &abs_HResponse, &nms, 5);
            corners = list.size();
            for (const auto &c : list) largest = c.value > largest ? c.value : largest;
        } else {
            Mat blurred, grad_x, grad_y;
            GaussianBlur3x3(img, blurred);                     // Harris_corners.cpp:158
            Sobel1(blurred, grad_x, grad_y);                   // :161-164
            HResponse = HarrisCorner(grad_x, grad_y);          // :173
            convertScaleAbs(HResponse, abs_HResponse);         // :176
            nms = NonMaximumSuppression(abs_HResponse, 3);     // :178
            nms2 = NMS2(HResponse, 5);                         // :179
            for (int r = 0; r < nms2.rows; ++r)
                for (int c = 0; c < nms2.cols; ++c) {
                    const float v = nms2.at<float>(r, c);
                    if (v > 0.f) { ++corners; largest = v > largest ? v : largest; }
                }
        }
        std::cout << "Response rows: " << HResponse.rows << std::endl;         // Harris_corners.cpp:88
        std::cout << "Response cols: " << HResponse.cols << std::endl;         // :89
        std::cout << "largest value: " << largest << std::endl;                // :127
        std::cout << "Number of corners (NMS2): " << corners << std::endl;
        size_t isolated = 0;
        for (int r = 0; r < nms.rows; ++r)
            for (int c = 0; c < nms.cols; ++c) isolated += nms.at<uint8_t>(r, c) != 0;
        std::cout << "Number of 3x3 non-maximum-suppressed pixels: " << isolated << std::endl;
        if (!dump.empty()) {
            vslam::imwrite_pfm(dump + "_response.pfm", HResponse);
            vslam::imwrite_pgm(dump + "_abs.pgm", abs_HResponse);
            vslam::imwrite_pgm(dump + "_nms.pgm", nms);
        }
    } catch (const std::exception &e) {
        fprintf(stderr, "Harris: %s\n", e.what());
        return 1;
    }
    return 0;
}
