// oracle/ref_icp.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT.  PARITY UNPINNED BY THE REFERENCE
// (see icp_loop.hpp).  Same restated ICP loop as port_icp.cpp, but the 1-NN is the
// reference's own vendored nanoflann 1.3.2, included in place from
// /root/reference/Homework9/hw9/include/nanoflann.hpp (never copied), instantiated the way
// registration.cpp:751-754 / :903-906 does: metric_L2 (L2_Adaptor<float>), run-time
// dimension 3, leaf_max_size = 2, k = 1, SearchParams(10) (:776-780, :929-933).
// Eigen is not installed, so KDTreeEigenMatrixAdaptor itself cannot be instantiated; the
// ten-line adaptor below exposes the same three callbacks over a packed float array.
// Built by `make -C oracle _ref` into oracle/_ref/libpcnn_ref_icp.so.
#include "icp_loop.hpp"
#include "nanoflann.hpp"

namespace {

struct Cloud {
    const float* xyz;
    int64_t n;
    inline size_t kdtree_get_point_count() const { return (size_t)n; }
    inline float kdtree_get_pt(const long idx, size_t dim) const { return xyz[3 * idx + (long)dim]; }
    template <class BBOX> bool kdtree_get_bbox(BBOX&) const { return false; }
};

typedef nanoflann::KDTreeSingleIndexAdaptor<nanoflann::L2_Adaptor<float, Cloud>, Cloud, -1, long> Index;

struct NanoNN {
    const Index* index;
    void operator()(const float* q, int64_t& idx, float& d2) const {
        long ri = -1;
        float rd = 0;
        nanoflann::KNNResultSet<float, long> rs(1);
        rs.init(&ri, &rd);
        index->findNeighbors(rs, q, nanoflann::SearchParams(10));
        idx = ri;
        d2 = rd;
    }
};

}  // namespace

extern "C" {

int pcref_icp(const float* src, int64_t n_src, const float* tar, int64_t n_tar,
              const float* tar_normals, const float* init_T, int max_iter, float max_corres_dist,
              float loss_epsilon, int mode, int accum, int nthreads, float* out_T,
              int64_t* out_info, float* out_loss, int32_t* out_corr) {
    Cloud cloud{tar, n_tar};
    Index index(3, cloud, nanoflann::KDTreeSingleIndexAdaptorParams(2 /* max leaf */));
    index.buildIndex();
    NanoNN nn{&index};
    icp_oracle::Params prm{max_iter, max_corres_dist, loss_epsilon, mode, accum, nthreads};
    icp_oracle::Result r = icp_oracle::run(nn, src, n_src, tar, n_tar, tar_normals, init_T, prm, out_corr);
    std::memcpy(out_T, r.T, sizeof(r.T));
    if (out_info) { out_info[0] = r.iters; out_info[1] = r.converged; out_info[2] = r.status; out_info[3] = r.last_pairs; }
    if (out_loss) *out_loss = r.last_loss;
    return 0;
}

int pcref_nn1(const float* tar, int64_t n_tar, const float* q, int64_t nq, int32_t* out_idx,
              float* out_d2, int nthreads) {
    Cloud cloud{tar, n_tar};
    Index index(3, cloud, nanoflann::KDTreeSingleIndexAdaptorParams(2));
    index.buildIndex();
    NanoNN nn{&index};
    icp_oracle::parallel_for(nq, nthreads, [&](int64_t lo, int64_t hi) {
        for (int64_t i = lo; i < hi; i++) {
            int64_t id; float d;
            nn(q + 3 * i, id, d);
            out_idx[i] = (int32_t)id;
            out_d2[i] = d;
        }
    });
    return 0;
}

}  // extern "C"
