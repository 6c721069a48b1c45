// A caller written the way Homework7/hw7/main.cpp:81-103 uses the reference's ISSKeypoint
// (myISSDetector, minus PCL's loader and viewer): reads xyz float32 triples from a raw file,
// runs the detector with the reference's parameters and prints the keypoint count and the
// first keypoints.  Built against include/pcnn_iss_compat.hpp by tests (GPU box only).
#include <cstdio>
#include <iostream>
#include <vector>

#include "pcnn_iss_compat.hpp"

int main(int argc, char** argv) {
    if (argc < 3) return 2;
    const size_t num_points = (size_t)std::stoul(argv[2]);
    std::vector<float> raw(num_points * 3);
    FILE* f = std::fopen(argv[1], "rb");
    if (!f || std::fread(raw.data(), sizeof(float), raw.size(), f) != raw.size()) return 3;
    std::fclose(f);
    MyPCDType point_cloud(num_points, std::vector<float>(3));
    for (size_t i = 0; i < num_points; i++)
        for (int a = 0; a < 3; a++) point_cloud[i][a] = raw[3 * i + a];

    float test = 0.02;
    ISSKeypoint iss_detector;
    iss_detector.useWeightedCovMat(true);
    iss_detector.setInputPointCloud(point_cloud);
    iss_detector.setLocalRadius(6 * test);
    iss_detector.setNonMaxRadius(4 * test);
    iss_detector.setThreshold(0.9, 0.9);
    iss_detector.setMinNeighbors(5);
    MyPCDType keypoints;
    iss_detector.compute(keypoints);
    std::cout << "key points size : " << keypoints.size() << std::endl;
    for (size_t k = 0; k < keypoints.size(); k++)
        std::printf("%.9g %.9g %.9g\n", keypoints[k][0], keypoints[k][1], keypoints[k][2]);
    return 0;
}
