// oracle/port_iss.cpp -- CPU restatement of the reference's ISS keypoint detector.
// TEST INFRASTRUCTURE ONLY: tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may
// call it, as the checker; nothing in the product path links or loads it.
//
// Follows Homework7/hw7/src/iss_detector.cpp:
//   compute()        :38-111   neighbour lists at local_radius, candidates, non-max filter
//   getEigenvalues() :114-152  scatter matrix about the point itself, weights 1/|N(j)|
// and the float KD-tree it searches with (Homework7/hw7/src/kdtree.cpp:347-355,
// include/resultSet.hpp:127-136): key accumulated x, y, z, every step formed in double and
// rounded to float (`ElemType diff; diff += pow(db - query, 2)`), dist = sqrtf(key), hit iff
// dist <= radius.
// The neighbour lists are found here by a uniform-grid search (the SET is defined by the rule
// above, not by the tree that finds it; oracle/_ref/libpcnn_ref_hw7.so, the reference's own
// tree, pins the sets in tests/test_oracle_cpu.py).
// PARITY UNPINNED for the eigenvalues: the reference uses Eigen::SelfAdjointEigenSolver<Matrix3f>
// on a float matrix accumulated in traversal order, and Eigen is not available here; this
// restatement accumulates in double in index order and uses cyclic Jacobi rotations in double.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <thread>
#include <unordered_map>
#include <vector>

namespace {

inline bool hw7_hit(const float* p, const float* q, float radius) {
    float diff = 0;  // ElemType diff = 0 (kdtree.cpp:348)
    for (int i = 0; i < 3; i++) diff += std::pow(p[i] - q[i], 2);  // double pow, float accumulate (:351)
    diff = std::sqrt(diff);                                       // float sqrt (:353)
    return diff <= radius;                                        // resultSet.hpp:130
}

struct Grid {
    const float* pts;
    int64_t n;
    float cell;
    float lo[3];
    std::unordered_map<uint64_t, std::vector<int>> cells;
    static uint64_t key(int64_t x, int64_t y, int64_t z) {
        return ((uint64_t)(x & 0x1fffff) << 42) | ((uint64_t)(y & 0x1fffff) << 21) | (uint64_t)(z & 0x1fffff);
    }
    void coord(const float* p, int64_t* c) const {
        for (int a = 0; a < 3; a++) c[a] = (int64_t)std::floor((p[a] - lo[a]) / cell);
    }
    Grid(const float* pts_, int64_t n_, float r) : pts(pts_), n(n_) {
        for (int a = 0; a < 3; a++) lo[a] = 0.f;
        float hi[3] = {0, 0, 0};
        for (int64_t i = 0; i < n; i++)
            for (int a = 0; a < 3; a++) {
                if (i == 0 || pts[3 * i + a] < lo[a]) lo[a] = pts[3 * i + a];
                if (i == 0 || pts[3 * i + a] > hi[a]) hi[a] = pts[3 * i + a];
            }
        float ext = std::max(hi[0] - lo[0], std::max(hi[1] - lo[1], hi[2] - lo[2]));
        cell = std::max(r * 1.001f, ext / 1000000.f);
        if (!(cell > 0.f)) cell = 1.f;
        for (int64_t i = 0; i < n; i++) {
            int64_t c[3];
            coord(pts + 3 * i, c);
            cells[key(c[0], c[1], c[2])].push_back((int)i);
        }
    }
    // neighbours of q in ascending index order
    void query(const float* q, float r, std::vector<int>& out) const {
        out.clear();
        int64_t c[3];
        coord(q, c);
        for (int64_t x = c[0] - 1; x <= c[0] + 1; x++)
            for (int64_t y = c[1] - 1; y <= c[1] + 1; y++)
                for (int64_t z = c[2] - 1; z <= c[2] + 1; z++) {
                    auto it = cells.find(key(x, y, z));
                    if (it == cells.end()) continue;
                    for (int j : it->second)
                        if (hw7_hit(pts + 3 * (int64_t)j, q, r)) out.push_back(j);
                }
        std::sort(out.begin(), out.end());
    }
};

// cyclic Jacobi, ascending eigenvalues (compiled with -ffp-contract=off: every product rounded)
void sym3_eigenvalues(const double* m, double* ev) {
    double a[3][3] = {{m[0], m[1], m[2]}, {m[1], m[3], m[4]}, {m[2], m[4], m[5]}};
    for (int sweep = 0; sweep < 12; sweep++) {
        double off = std::fabs(a[0][1]) + std::fabs(a[0][2]) + std::fabs(a[1][2]);
        if (off == 0.0) break;
        for (int p = 0; p < 2; p++)
            for (int q = p + 1; q < 3; q++) {
                double apq = a[p][q];
                if (apq == 0.0) continue;
                double theta = (a[q][q] - a[p][p]) / (2.0 * apq);
                double t = (theta >= 0.0 ? 1.0 : -1.0) / (std::fabs(theta) + std::sqrt(theta * theta + 1.0));
                double c = 1.0 / std::sqrt(t * t + 1.0);
                double s = t * c;
                int r = 3 - p - q;
                double arp = a[r][p], arq = a[r][q];
                a[p][p] -= t * apq;
                a[q][q] += t * apq;
                a[p][q] = a[q][p] = 0.0;
                a[r][p] = a[p][r] = c * arp - s * arq;
                a[r][q] = a[q][r] = s * arp + c * arq;
            }
    }
    double e[3] = {a[0][0], a[1][1], a[2][2]};
    std::sort(e, e + 3);
    ev[0] = e[0]; ev[1] = e[1]; ev[2] = e[2];
}

template <typename F>
void parallel_for(int64_t n, int nthreads, F f) {
    if (nthreads < 1) nthreads = 1;
    std::vector<std::thread> th;
    for (int t = 0; t < nthreads; t++)
        th.emplace_back([=]() {
            for (int64_t i = t; i < n; i += nthreads) f(i);
        });
    for (auto& x : th) x.join();
}

}  // namespace

extern "C" {

// neighbour counts at `radius` for every point (the sizes of rnn_idx, iss_detector.cpp:48-56)
int pco_iss_counts(const float* pts, int64_t n, float radius, int32_t* out_counts, int nthreads) {
    Grid g(pts, n, radius);
    parallel_for(n, nthreads, [&](int64_t i) {
        std::vector<int> nb;
        g.query(pts + 3 * i, radius, nb);
        out_counts[i] = (int32_t)nb.size();
    });
    return 0;
}

// ISSKeypoint::compute.  out_lambda3 [n] (the reference's lambda3_vec_forall), out_flag [n].
int pco_iss(const float* pts, int64_t n, float local_radius, float non_max_radius, float g21, float g32,
            int min_neighbors, int weighted, float* out_lambda3, uint8_t* out_flag, int nthreads) {
    Grid g(pts, n, local_radius);
    std::vector<std::vector<int>> rnn((size_t)n);
    parallel_for(n, nthreads, [&](int64_t i) { g.query(pts + 3 * i, local_radius, rnn[(size_t)i]); });  // :48-56
    parallel_for(n, nthreads, [&](int64_t i) {
        out_lambda3[i] = -1.f;  // :69
        const std::vector<int>& nb = rnn[(size_t)i];
        if (nb.size() < 3) return;  // :72
        const float* c = pts + 3 * i;
        double m[6] = {0, 0, 0, 0, 0, 0}, wsum = 0.0;
        for (int j : nb) {
            const float wf = weighted ? 1.f / (float)rnn[(size_t)j].size() : 1.f;  // :128
            const double w = (double)wf;
            const float* p = pts + 3 * (int64_t)j;
            const double dx = (double)(p[0] - c[0]), dy = (double)(p[1] - c[1]), dz = (double)(p[2] - c[2]);  // :131-133
            wsum += w;
            m[0] += w * dx * dx; m[1] += w * dx * dy; m[2] += w * dx * dz;
            m[3] += w * dy * dy; m[4] += w * dy * dz; m[5] += w * dz * dz;
        }
        if (weighted)
            for (double& v : m) v /= wsum;  // :135
        double ev[3];
        sym3_eigenvalues(m, ev);
        const float l1 = (float)ev[2], l2 = (float)ev[1], l3 = (float)ev[0];  // :75-77
        if (l2 / l1 < g21 && l3 / l2 < g32 && l3 > 0) out_lambda3[i] = l3;  // :79-80
    });
    Grid g2(pts, n, non_max_radius);
    parallel_for(n, nthreads, [&](int64_t i) {
        out_flag[i] = 0;
        if (out_lambda3[i] == -1.f) return;  // :91
        std::vector<int> nb;
        g2.query(pts + 3 * i, non_max_radius, nb);  // :93-94
        if ((int)nb.size() < min_neighbors) return;  // :95
        for (int j : nb)
            if (out_lambda3[i] < out_lambda3[j]) return;  // :98-103
        out_flag[i] = 1;
    });
    return 0;
}

}  // extern "C"
