// DoG -- the reference's ./DoG executable (KeyPointDetection/Diff_of_Gauss.cpp:725-870) without its windows:
// pyramid (4 octaves, sigma 1.6), per octave initialKeypointDetection + filterKeypoints (edge threshold, orientation
// assignment) + SIFT, the reference's stdout lines (:784, :786, :815, :834) and featureDescriptors.dat (:837-863).
// Drawing / HighGUI are unavailable on this box; the keypoint lists go to text files instead.
//   usage: DoG [image.pgm|image.ppm] [--fused] [--out keypoints.txt] [--oriented oriented.txt] [--desc featureDescriptors.dat]
//          [--no-descriptors]                                             (default image: ../images/home.pgm, like the reference's
//   ../images/home.jpg; JPEG decoding is not part of the path -- convert once with any tool)
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>

#include "Diff_of_Gauss.hpp"

using std::cout;
using std::endl;

int main(int argc, char **argv) {
    std::string path = "../images/home.pgm", out_path, oriented_path, desc_path = "featureDescriptors.dat";
    bool fused = false, descriptors = true;
    for (int i = 1; i < argc; ++i) {
        if (!strcmp(argv[i], "--fused")) fused = true;
        else if (!strcmp(argv[i], "--out") && i + 1 < argc) out_path = argv[++i];
        else if (!strcmp(argv[i], "--oriented") && i + 1 < argc) oriented_path = argv[++i];
        else if (!strcmp(argv[i], "--desc") && i + 1 < argc) desc_path = argv[++i];
        else if (!strcmp(argv[i], "--no-descriptors")) descriptors = false;
        else if (!strcmp(argv[i], "--help")) {
            printf("usage: DoG [image.pgm|image.ppm] [--fused] [--out keypoints.txt] [--oriented oriented.txt] [--desc featureDescriptors.dat] [--no-descriptors]\n");
            return 0;
        }
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
        std::vector<SLAM::point> edgeKeypoints_AllScales;        // the hot path's list (after the edge threshold)
        std::vector<SLAM::point> reducedKeypoints_AllScales;     // :777, after orientation assignment
        std::vector<std::vector<float>> featureDescriptors_vec;  // :779
        if (fused) {
            // one call, fused kernels: same keypoints, nothing but the list leaves the GPU
            edgeKeypoints_AllScales = detectKeypoints(img, numOctaves, pyr_sigma, windowSize);
            if (descriptors) {
                GaussPyramid pyramid{img, numOctaves, pyr_sigma};
                for (int octave = 0; octave < numOctaves; ++octave) {
                    std::vector<SLAM::point> edgeKeypoints, reducedKeypoints;
                    for (const auto &k : edgeKeypoints_AllScales)
                        if (k.octave == octave) edgeKeypoints.emplace_back(k);
                    assignOrientations(pyramid, octave, edgeKeypoints, reducedKeypoints);
                    SIFT(reducedKeypoints, featureDescriptors_vec, pyramid, octave);
                    for (auto &point : reducedKeypoints) reducedKeypoints_AllScales.emplace_back(point);
                }
            }
        } else {
            GaussPyramid pyramid{img, numOctaves, pyr_sigma};
            for (const auto &kv : pyramid.pyramidGauss()) {
                std::vector<SLAM::point> keypoints;
                std::vector<SLAM::point> edgeKeypoints;
                std::vector<SLAM::point> reducedKeypoints;
                int octave = kv.first;
                cout << "Pyramid octave: " << kv.first << endl;
                initialKeypointDetection(keypoints, pyramid, octave, windowSize);
                cout << "image size: width.height (" << kv.second.at(0).cols << ", " << kv.second.at(0).rows << ")" << endl;
                filterKeypointsEdge(pyramid, octave, keypoints, edgeKeypoints);          // filterKeypoints (:787) in its two halves
                cout << "keypoints after contrast test: " << keypoints.size() << ", after edge test: " << edgeKeypoints.size() << endl;
                for (auto &point : edgeKeypoints) edgeKeypoints_AllScales.emplace_back(point);
                if (descriptors) {
                    assignOrientations(pyramid, octave, edgeKeypoints, reducedKeypoints);
                    SIFT(reducedKeypoints, featureDescriptors_vec, pyramid, octave);      // :791
                    for (auto &point : reducedKeypoints) reducedKeypoints_AllScales.emplace_back(point);
                }
            }
        }
        cout << "Number of keypoints after edge response: " << edgeKeypoints_AllScales.size() << endl;
        auto dump = [](const std::string &p, const std::vector<SLAM::point> &list) {
            FILE *f = fopen(p.c_str(), "w");
            if (!f) { fprintf(stderr, "cannot write %s\n", p.c_str()); return false; }
            for (const auto &k : list) fprintf(f, "%d %d %d %d %d %d\n", k.row, k.col, k.value, k.padding, k.octave, k.level);
            fclose(f);
            return true;
        };
        if (!out_path.empty() && !dump(out_path, edgeKeypoints_AllScales)) return 1;
        if (descriptors) {
            cout << "Number of features found in image: " << featureDescriptors_vec.size() << endl;                         // :815
            std::cout << "Number of keypoints after edge orientation: " << reducedKeypoints_AllScales.size() << std::endl;   // :834
            if (!oriented_path.empty() && !dump(oriented_path, reducedKeypoints_AllScales)) return 1;
            if (!writeFeatureDescriptors(desc_path, featureDescriptors_vec)) return EXIT_FAILURE;                           // :837-863
        }
    } catch (const std::exception &e) {
        fprintf(stderr, "DoG: %s\n", e.what());
        return 1;
    }
    return 0;
}
