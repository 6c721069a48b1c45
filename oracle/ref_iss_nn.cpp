// oracle/ref_iss_nn.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT.
//
// Thin C-ABI driver around the *unmodified* float KD-tree of Homework7, compiled where it lies
// under /root/reference (Homework7/hw7/include/kdtree.hpp, resultSet.hpp, src/kdtree.cpp,
// src/resultSet.cpp; never copied into this repo) into oracle/_ref/libpcnn_ref_hw7.so by
// oracle/Makefile (target _ref).  It performs the first step of ISSKeypoint::compute
// (Homework7/hw7/src/iss_detector.cpp:46-56): KDTreeConstruction(point_cloud, 12), then a
// RadiusNNResultSet + KDTreeRadiusNNSearch for every point, and returns the neighbour lists.
// The rest of the detector needs Eigen (iss_detector.hpp:11), which is not available here:
// the neighbour SETS of the ISS oracle and of the CUDA path are pinned against this library,
// the eigenvalue step is not.
#include <cstdint>
#include <vector>

#include "kdtree.hpp"

extern "C" {

// offsets [n + 1]; idx may be NULL (count only) or hold offsets[n] entries
int ref7_radius_all(const float* pts, int64_t n, float radius, int64_t* offsets, int32_t* idx) {
    std::vector<std::vector<float>> db((size_t)n, std::vector<float>(3));
    for (int64_t i = 0; i < n; i++)
        for (int a = 0; a < 3; a++) db[(size_t)i][a] = pts[3 * i + a];
    Node* root = KDTreeConstruction(db, 12);  // iss_detector.cpp:46
    int64_t run = 0;
    for (int64_t i = 0; i < n; i++) {
        RadiusNNResultSet rs(radius);  // :50
        KDTreeRadiusNNSearch(root, db, rs, db[(size_t)i]);  // :51
        offsets[i] = run;
        if (idx)
            for (const auto& di : rs.distIndexList) idx[run++] = di.index;  // :53-55
        else
            run += rs.size();
    }
    offsets[n] = run;
    KDTreeDestruction();  // :110
    Node::address_set.clear();  // the reference never clears it (kdtree.cpp:400-404): a second call would free twice
    return 0;
}

}  // extern "C"
