// Pyramid_Test -- the reference's tests/GaussPyramid_Test.cpp:63-115 on the GPU-backed GaussPyramid: prints the same
// lines (octave / level counts, octave sizes, the three sigmas, the sigma count) and then checks them (exit code 1 on
// a mismatch, which the reference leaves to the reader's eye).
//   usage: Pyramid_Test [image.pgm]                                (default image: ../images/building.pgm)
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <iostream>

#include "GaussPyramid.hpp"

using std::cout;
using std::endl;

int main(int argc, char **argv) {
    const std::string path = argc > 1 ? argv[1] : "../images/building.pgm";
    try {
        cout << "vslam version : " << vslam_version() << endl;
        Mat img_color = vslam::imread(path, vslam::IMREAD_COLOR);      // tests/GaussPyramid_Test.cpp:78-81
        Mat img;
        vslam::cvtColor(img_color, img, vslam::COLOR_BGR2GRAY);
        int numOctaves = 4;
        double sigma = 1.6;
        GaussPyramid pyramid{img, numOctaves, sigma};
        int bad = 0;
        // 0. the pyramid has as many levels as octaves
        cout << "Number of expected octaves: " << pyramid.getNumOctaves() << endl;
        cout << "Number of expected levels per octave: " << pyramid.getNumLevels() << endl;
        bad += pyramid.getNumOctaves() != 4 || pyramid.getNumLevels() != 6;
        int w = 2 * img.cols, h = 2 * img.rows;
        for (const auto &kv : pyramid.pyramidGauss()) {
            cout << "Pyramid level: " << kv.first << endl;
            cout << "number of levels: " << kv.second.size() << endl;
            cout << "image size: width.height (" << kv.second.at(0).cols << ", " << kv.second.at(0).rows << ")" << endl;
            bad += kv.second.size() != 6 || kv.second.at(0).cols != w || kv.second.at(0).rows != h;
            w = (int)std::nearbyint(w * 0.5); h = (int)std::nearbyint(h * 0.5);      // cvRound, GaussPyramid.cpp:126
        }
        cout << endl;
        // 1. the sigma values
        double mySigma = pyramid.getSigmaAt(0, 0);
        cout << "First sigma value of pyramid : " << mySigma << endl;
        double mySecondSigma = pyramid.getSigmaAt(0, pyramid.getNumLevels() - 1);
        cout << "Last sigma value of 1st octave : " << mySecondSigma << endl;
        double mySigma_doubled = pyramid.getSigmaAt(0, pyramid.getNumLevels() - 3);
        cout << "double sigma value : " << mySigma_doubled << endl;
        bad += std::fabs(mySigma - 1.6) > 1e-12 || std::fabs(mySecondSigma - 5.0797) > 1e-4 || std::fabs(mySigma_doubled - 3.2) > 1e-12;
        // 2. the number of sigma values
        cout << "Number of sigma values per octave: " << pyramid.octaveSigma(numOctaves - 1).size() << endl;
        bad += pyramid.octaveSigma(numOctaves - 1).size() != 6;
        // 3. the gradient maps exist and have the octave's size
        Mat gradX = pyramid.octaveGradX(1).at(0);
        Mat gradY = pyramid.octaveGradY(1).at(0);
        Mat mag = pyramid.octaveGradMag(1).at(0);
        Mat orient = pyramid.octaveGradOrient(1).at(0);
        cout << "gradient maps of octave 1: " << gradX.cols << " x " << gradX.rows << endl;
        bad += gradX.cols != img.cols || gradY.rows != img.rows || mag.cols != img.cols || orient.rows != img.rows;
        bool threw = false;
        try { pyramid.octaveBlur(7); } catch (const std::out_of_range &) { threw = true; }
        bad += !threw;
        // 4. the two-argument constructor (GaussPyramid.hpp:21): calculateNumOctaves = floor(log2(min side)) - 4 (GaussPyramid.cpp:150-152);
        //    octaves beyond the fourth have kernels wider than 247 taps and run on the exact global-memory path
        {
            GaussPyramid autoPyr{img, sigma};
            const int expect = (int)std::floor(std::log2((double)std::min(img.rows, img.cols))) - 4;
            cout << "Number of octaves of GaussPyramid(img, sigma): " << autoPyr.getNumOctaves() << endl;
            bad += autoPyr.getNumOctaves() != expect || (int)autoPyr.pyramidGauss().size() != expect;
            if (expect > 4) bad += autoPyr.octaveBlur(expect - 1).size() != 6 || autoPyr.octaveBlur(expect - 1).at(5).empty();
        }
        cout << (bad ? "FAILED" : "OK") << endl;
        return bad ? 1 : 0;
    } catch (const std::exception &e) {
        fprintf(stderr, "Pyramid_Test: %s\n", e.what());
        return 2;
    }
}
