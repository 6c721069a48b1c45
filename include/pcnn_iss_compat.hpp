// pcnn_iss_compat.hpp -- the reference's ISSKeypoint class (Homework7/hw7/include/iss_detector.hpp
// :15-37) over the C ABI of include/pcnn.h: same type names, method names and argument meaning,
// so `#include "iss_detector.hpp"` can be replaced by this header and the calling code of
// Homework7/hw7/main.cpp:81-103 compiles and runs unchanged (link with -lpcnn).
//
//   useWeightedCovMat / setLocalRadius / setNonMaxRadius / setThreshold / setMinNeighbors /
//   setInputPointCloud      iss_detector.cpp:7-36   (stored, as there)
//   compute(keypoints)      iss_detector.cpp:38-111 (index build + pcnn_iss_detect_host; keypoints
//                           are the cloud's points in ascending index, as :104-105 emits them)
// Errors of the C ABI become std::runtime_error with pcnn_last_error() as the message.
#ifndef PCNN_ISS_COMPAT_HPP
#define PCNN_ISS_COMPAT_HPP

#include <cstdint>
#include <stdexcept>
#include <string>
#include <vector>

#include "pcnn.h"

typedef std::vector<std::vector<float>> MyPCDType;

class ISSKeypoint {
public:
    void useWeightedCovMat(bool use) { use_w_cov_mat = use; }
    void setLocalRadius(float r) { local_radius = r; }
    void setNonMaxRadius(float r) { non_max_radius = r; }
    void setThreshold(float g21, float g32) { gama21 = g21; gama32 = g32; }
    void setMinNeighbors(int n) { min_neighbors = n; }
    void setInputPointCloud(MyPCDType& input_point_cloud) { point_cloud = input_point_cloud; }
    void setDevice(int cuda_device) { device = cuda_device; }  // not in the reference: which GPU

    void compute(MyPCDType& keypoints) {
        keypoints.clear();
        const int64_t n = (int64_t)point_cloud.size();
        std::vector<float> xyz((size_t)n * 3);
        for (int64_t i = 0; i < n; i++) {
            if (point_cloud[(size_t)i].size() < 3) throw std::runtime_error("ISSKeypoint: a point has fewer than 3 coordinates");
            for (int a = 0; a < 3; a++) xyz[(size_t)(3 * i + a)] = point_cloud[(size_t)i][(size_t)a];
        }
        pcnn_index* ix = nullptr;
        // KDTreeConstruction(point_cloud, 12), iss_detector.cpp:46
        check(pcnn_index_build_host(&ix, PCNN_KDTREE, xyz.data(), n, 3, 12, 1e-4, device), "pcnn_index_build_host");
        pcnn_iss_params prm;
        prm.local_radius = local_radius;
        prm.non_max_radius = non_max_radius;
        prm.gamma21 = gama21;
        prm.gamma32 = gama32;
        prm.min_neighbors = min_neighbors;
        prm.weighted = use_w_cov_mat ? 1 : 0;
        std::vector<int32_t> idx((size_t)(n > 0 ? n : 1));
        int64_t count = 0;
        const int rc = pcnn_iss_detect_host(ix, &prm, idx.data(), n, &count, nullptr);
        pcnn_index_destroy(ix);  // KDTreeDestruction(), :110
        check(rc, "pcnn_iss_detect_host");
        keypoints.reserve((size_t)count);
        for (int64_t k = 0; k < count; k++) keypoints.emplace_back(point_cloud[(size_t)idx[(size_t)k]]);
    }

private:
    static void check(int rc, const char* what) {
        if (rc != PCNN_OK) throw std::runtime_error(std::string(what) + ": " + pcnn_last_error());
    }
    bool use_w_cov_mat = false;
    float local_radius = 0.f;
    float non_max_radius = 0.f;
    float gama21 = 0.f, gama32 = 0.f;
    int min_neighbors = 0;
    int device = 0;
    MyPCDType point_cloud;
};

#endif  // PCNN_ISS_COMPAT_HPP
