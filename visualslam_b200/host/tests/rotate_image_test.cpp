// RotateImgTest -- the reference's tests/rotate_image_test.cpp:55-84 on the GPU-backed SLAM::Rotation, headless: the
// timing line of the rotated 16x16 window (the only numbers the reference's test prints), plus the rotated image and
// window written as PGM files instead of imshow / trackbars, and the double-crop comparison (:85-97) with its own timing line.
//   usage: RotateImgTest [image.pgm] [--angle deg] [--dump prefix]          (default image: ../images/blox.pgm)
#include <chrono>
#include <cstring>
#include <iostream>

#include "Rotation/rotation.h"

using namespace std;

int main(int argc, char **argv) {
    std::string path = "../images/blox.pgm", dump;
    float angle = 45.0f;
    for (int i = 1; i < argc; ++i) {
        if (!strcmp(argv[i], "--angle") && i + 1 < argc) angle = (float)atof(argv[++i]);
        else if (!strcmp(argv[i], "--dump") && i + 1 < argc) dump = argv[++i];
        else if (!strcmp(argv[i], "--help")) { printf("usage: RotateImgTest [image.pgm] [--angle deg] [--dump prefix]\n"); return 0; }
        else path = argv[i];
    }
    try {
        Mat img = vslam::imread(path, vslam::IMREAD_GRAYSCALE);            // rotate_image_test.cpp:25
        const Point2i center(img.cols / 2, img.rows / 2);                 // :35
        const int windowSize = 16, times = 100;                           // :38, :58
        Mat ROI = SLAM::Rotation::getRotatedWindow(img, center, 100, angle);        // :73
        auto t0 = chrono::steady_clock::now();
        for (int i = 0; i < times; i++) SLAM::Rotation::getRotatedWindow(img, center, windowSize, angle);   // :75-78
        double t = chrono::duration<double, milli>(chrono::steady_clock::now() - t0).count() / times;
        cout << "Time rotating 16x16 ROI (averaged for " << times << " runs): " << t << " milliseconds." << endl;      // :81
        Mat ROI2 = SLAM::Rotation::doubleCrop(img, center, 100, (double)angle);                 // :86
        t0 = chrono::steady_clock::now();
        for (int i = 0; i < times; i++) SLAM::Rotation::doubleCrop(img, center, windowSize, (double)angle);   // :88-91
        t = chrono::duration<double, milli>(chrono::steady_clock::now() - t0).count() / times;
        cout << "Time rotating 16x16 ROI with the double crop method (averaged for " << times << " runs): " << t << " milliseconds." << endl;   // :94
        Mat rotated = SLAM::Rotation::rotate_mat_CCW(img, center, SLAM::Rotation::cos_sin_of_angle(angle));   // Rotation_Demo :44-48
        const Point2i rot = SLAM::Rotation::rotate_pt_CCW(Point2i(img.cols / 2, 0), center, angle);           // Rotation_Point_Demo :51-56
        cout << "Rotated point: (" << rot.x << ", " << rot.y << ")" << endl;
        if (!dump.empty()) {
            vslam::imwrite_pgm(dump + "_rotated.pgm", rotated);
            vslam::imwrite_pgm(dump + "_roi.pgm", ROI);
            if (!ROI2.empty()) vslam::imwrite_pgm(dump + "_roi2.pgm", ROI2);
        }
    } catch (const std::exception &e) {
        fprintf(stderr, "RotateImgTest: %s\n", e.what());
        return 1;
    }
    return 0;
}
