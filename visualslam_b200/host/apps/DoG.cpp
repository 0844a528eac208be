// DoG -- the reference's ./DoG executable (KeyPointDetection/Diff_of_Gauss.cpp:725-870) up to the end of the hot path:
// pyramid (4 octaves, sigma 1.6), per octave initialKeypointDetection + filterKeypoints' edge threshold, with the
// reference's stdout lines (:784, :786).  Orientation / SIFT descriptors / drawing follow the hot path in the reference
// and are out of scope; HighGUI is unavailable on this box, so keypoints go to a text file instead.
//   usage: DoG [image.pgm|image.ppm] [--fused] [--out keypoints.txt]      (default image: ../images/home.pgm, like the reference's
//   ../images/home.jpg; JPEG decoding is not part of the path -- convert once with any tool)
#include <cstdio>
#include <cstring>
#include <iostream>

#include "Diff_of_Gauss.hpp"

using std::cout;
using std::endl;

int main(int argc, char **argv) {
    std::string path = "../images/home.pgm", out_path;
    bool fused = false;
    for (int i = 1; i < argc; ++i) {
        if (!strcmp(argv[i], "--fused")) fused = true;
        else if (!strcmp(argv[i], "--out") && i + 1 < argc) out_path = argv[++i];
        else if (!strcmp(argv[i], "--help")) { printf("usage: DoG [image.pgm|image.ppm] [--fused] [--out keypoints.txt]\n"); return 0; }
        else path = argv[i];
    }
    try {
        // Diff_of_Gauss.cpp:730-739: imread(IMREAD_COLOR) + cvtColor(COLOR_BGR2GRAY); a grey PGM is taken as it is
        Mat img_color = vslam::imread(path, vslam::IMREAD_COLOR);
        Mat img;
        vslam::cvtColor(img_color, img, vslam::COLOR_BGR2GRAY);
        int numOctaves = 4;          // Diff_of_Gauss.cpp:742
        double pyr_sigma = 1.6;      // :743
        int windowSize = 3;          // :772
        std::vector<SLAM::point> reducedKeypoints_AllScales;
        if (fused) {
            // one call, fused kernels: same keypoints, nothing but the list leaves the GPU
            reducedKeypoints_AllScales = detectKeypoints(img, numOctaves, pyr_sigma, windowSize);
        } else {
            GaussPyramid pyramid{img, numOctaves, pyr_sigma};
            for (const auto &kv : pyramid.pyramidGauss()) {
                std::vector<SLAM::point> keypoints;
                std::vector<SLAM::point> reducedKeypoints;
                int octave = kv.first;
                cout << "Pyramid octave: " << kv.first << endl;
                initialKeypointDetection(keypoints, pyramid, octave, windowSize);
                cout << "image size: width.height (" << kv.second.at(0).cols << ", " << kv.second.at(0).rows << ")" << endl;
                filterKeypoints(pyramid, octave, keypoints, reducedKeypoints);
                cout << "keypoints after contrast test: " << keypoints.size() << ", after edge test: " << reducedKeypoints.size() << endl;
                for (auto &point : reducedKeypoints) reducedKeypoints_AllScales.emplace_back(point);
            }
        }
        cout << "Number of keypoints after edge response: " << reducedKeypoints_AllScales.size() << endl;
        if (!out_path.empty()) {
            FILE *f = fopen(out_path.c_str(), "w");
            if (!f) { fprintf(stderr, "cannot write %s\n", out_path.c_str()); return 1; }
            for (const auto &k : reducedKeypoints_AllScales) fprintf(f, "%d %d %d %d %d %d\n", k.row, k.col, k.value, k.padding, k.octave, k.level);
            fclose(f);
        }
    } catch (const std::exception &e) {
        fprintf(stderr, "DoG: %s\n", e.what());
        return 1;
    }
    return 0;
}
