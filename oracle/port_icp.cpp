// oracle/port_icp.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT.  PARITY UNPINNED BY THE REFERENCE
// (see icp_loop.hpp): the ICP loop restated from Homework9/hw9/src/registration.cpp:710-1011
// with an own exact float32 1-NN standing in for nanoflann 1.3.2
// (Homework9/hw9/include/nanoflann.hpp:1222-1242, :1347-1410, :142-205, :375-414).
//
// The 1-NN here returns, for query q, the target index minimising
//     d2 = ((dx*dx) + dy*dy) + dz*dz,  dx = q.x - p.x, all float32, no FMA
// (L2_Adaptor::evalMetric, nanoflann.hpp:383-408) with ties resolved to the LOWEST index.
// nanoflann resolves an exact tie to the first point it visits (strict <, :1354-1366, :184);
// the two agree whenever no exact float tie occurs, which tests count.
#include "icp_loop.hpp"

namespace {

struct F3 { float x, y, z; };

inline float d2f(const float* q, const F3& p) {
    float dx = q[0] - p.x, dy = q[1] - p.y, dz = q[2] - p.z;
    float s = dx * dx;
    s = s + dy * dy;
    s = s + dz * dz;
    return s;
}

// Median-split KD-tree over float32 points, leaves of <= 8 points.
struct Kd1 {
    struct N { int axis; float split; int left, right, begin, end; };
    std::vector<F3> pts;
    std::vector<int> order;
    std::vector<N> nodes;

    static float coord(const F3& p, int a) { return a == 0 ? p.x : (a == 1 ? p.y : p.z); }

    int build(int begin, int end) {
        int me = (int)nodes.size();
        nodes.push_back(N{-1, 0.f, -1, -1, begin, end});
        if (end - begin > 8) {
            F3 mn = pts[order[begin]], mx = mn;
            for (int i = begin; i < end; i++) {
                const F3& p = pts[order[i]];
                mn.x = std::min(mn.x, p.x); mx.x = std::max(mx.x, p.x);
                mn.y = std::min(mn.y, p.y); mx.y = std::max(mx.y, p.y);
                mn.z = std::min(mn.z, p.z); mx.z = std::max(mx.z, p.z);
            }
            float ex = mx.x - mn.x, ey = mx.y - mn.y, ez = mx.z - mn.z;
            int axis = (ex >= ey && ex >= ez) ? 0 : (ey >= ez ? 1 : 2);
            int mid = begin + (end - begin) / 2;
            std::nth_element(order.begin() + begin, order.begin() + mid, order.begin() + end,
                             [&](int a, int b) { return coord(pts[a], axis) < coord(pts[b], axis); });
            float split = coord(pts[order[mid]], axis);
            int l = build(begin, mid);
            int r = build(mid, end);
            nodes[me].axis = axis; nodes[me].split = split;
            nodes[me].left = l; nodes[me].right = r;
        }
        return me;
    }

    void search(int n, const float* q, int64_t& best, float& best_d2) const {
        const N& nd = nodes[n];
        if (nd.axis < 0) {
            for (int i = nd.begin; i < nd.end; i++) {
                int id = order[i];
                float d = d2f(q, pts[id]);
                if (d < best_d2 || (d == best_d2 && id < best)) { best_d2 = d; best = id; }
            }
            return;
        }
        // left holds coord <= split, right holds coord >= split
        float diff = q[nd.axis] - nd.split;
        int near = diff < 0 ? nd.left : nd.right, far = diff < 0 ? nd.right : nd.left;
        search(near, q, best, best_d2);
        // fl(diff*diff) is a lower bound of every far-side d2 under monotone rounding;
        // visit on equality so that a lower-index tie is not missed.
        if (diff * diff <= best_d2) search(far, q, best, best_d2);
    }

    void operator()(const float* q, int64_t& idx, float& d2) const {
        idx = -1;
        d2 = INFINITY;
        if (!nodes.empty()) search(0, q, idx, d2);
    }
};

}  // namespace

extern "C" {

// src/tar: packed xyz float32 (3 floats per point); tar_normals: packed xyz or NULL (p2p).
// out: T[16] row-major, info[0]=iters applied, info[1]=converged, info[2]=status,
// info[3]=pairs kept in the last iteration; out_loss = last loss; out_corr optional [n_src].
int pcport_icp(const float* src, int64_t n_src, const float* tar, int64_t n_tar,
               const float* tar_normals, const float* init_T, int max_iter,
               float max_corres_dist, float loss_epsilon, int mode, int accum, int nthreads,
               float* out_T, int64_t* out_info, float* out_loss, int32_t* out_corr) {
    Kd1 kd;
    kd.pts.resize((size_t)n_tar);
    kd.order.resize((size_t)n_tar);
    for (int64_t i = 0; i < n_tar; i++) {
        kd.pts[(size_t)i] = F3{tar[3 * i], tar[3 * i + 1], tar[3 * i + 2]};
        kd.order[(size_t)i] = (int)i;
    }
    if (n_tar > 0) kd.build(0, (int)n_tar);
    icp_oracle::Params prm{max_iter, max_corres_dist, loss_epsilon, mode, accum, nthreads};
    icp_oracle::Result r = icp_oracle::run(kd, src, n_src, tar, n_tar, tar_normals, init_T, prm, out_corr);
    std::memcpy(out_T, r.T, sizeof(r.T));
    if (out_info) { out_info[0] = r.iters; out_info[1] = r.converged; out_info[2] = r.status; out_info[3] = r.last_pairs; }
    if (out_loss) *out_loss = r.last_loss;
    return 0;
}

// The oracle's own stand-ins for Eigen (JacobiSVD<Matrix3f>, 6x6 inverse), exposed so a test can
// compare them with LAPACK directly, and the per-iteration record of what the loop gave them:
// rows [cap][64] doubles (layout: icp_loop.hpp Trace); set before pcport_icp on the same thread,
// pcport_icp_trace(NULL, 0) returns the number of rows written and switches the record off.
void pcport_svd3(const double* A, double* U, double* s, double* V) { icp_oracle::svd3(A, U, s, V); }
int pcport_solve6(const double* N, const double* g, double* x) { return icp_oracle::solve6(N, g, x) ? 1 : 0; }
int pcport_icp_trace(double* rows, int cap) {
    static thread_local icp_oracle::Trace tr;
    const int written = tr.n;
    tr.rows = rows; tr.cap = rows ? cap : 0; tr.n = 0;
    icp_oracle::trace_slot() = rows ? &tr : nullptr;
    return written;
}

// Exact float32 1-NN of every query (same key as above); used to check the CUDA 1-NN alone.
int pcport_nn1(const float* tar, int64_t n_tar, const float* q, int64_t nq, int32_t* out_idx,
               float* out_d2, int nthreads) {
    Kd1 kd;
    kd.pts.resize((size_t)n_tar);
    kd.order.resize((size_t)n_tar);
    for (int64_t i = 0; i < n_tar; i++) {
        kd.pts[(size_t)i] = F3{tar[3 * i], tar[3 * i + 1], tar[3 * i + 2]};
        kd.order[(size_t)i] = (int)i;
    }
    if (n_tar > 0) kd.build(0, (int)n_tar);
    icp_oracle::parallel_for(nq, nthreads, [&](int64_t lo, int64_t hi) {
        for (int64_t i = lo; i < hi; i++) {
            int64_t id; float d;
            kd(q + 3 * i, id, d);
            out_idx[i] = (int32_t)id;
            out_d2[i] = d;
        }
    });
    return 0;
}

}  // extern "C"
