// oracle/port_ransac.cpp -- CPU restatement of the reference's descriptor matching and RANSAC.
// TEST INFRASTRUCTURE ONLY: tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may
// call it, as the checker; nothing in the product path links or loads it.
//
// Follows Homework9/hw9/src/registration.cpp:
//   findRANSACCorrespondencesInter :437-527, findRANSACCorrespondencesUnion :530-611, RANSAC :288-434
// The 1-NN searches are nanoflann's there (exact; L2_Simple_Adaptor: `result += diff * diff` per
// dimension in float32): restated as a linear scan with that arithmetic, lowest index on ties.
// The reference seeds std::mt19937 from std::random_device (:298-299): its hypotheses are not
// reproducible, so the generator is the ABI's (splitmix64 keyed by (seed, iteration)), and the
// 3x3 SVD is this oracle's own (icp_loop.hpp, Jacobi in double) where the reference uses
// Eigen::JacobiSVD<Matrix3f>.  PARITY UNPINNED by the reference (no Eigen, no PCL descriptors, no
// golden output); compiled with -ffp-contract=off so float expressions round as written.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <numeric>
#include <thread>
#include <vector>

#include "icp_loop.hpp"

namespace {

void nn1_scan(const float* a, int64_t na, const float* b, int64_t nb, int dim, int32_t* idx, float* d2, int nthreads) {
    auto work = [&](int64_t i0, int64_t i1) {
        for (int64_t i = i0; i < i1; i++) {
            float best = INFINITY;
            int32_t bj = -1;
            for (int64_t j = 0; j < nb; j++) {
                float r = 0.f;
                for (int d = 0; d < dim; d++) {
                    const float diff = a[i * dim + d] - b[j * dim + d];
                    r += diff * diff;
                }
                if (r < best) { best = r; bj = (int32_t)j; }
            }
            idx[i] = bj;
            d2[i] = best;
        }
    };
    if (nthreads < 1) nthreads = 1;
    std::vector<std::thread> th;
    const int64_t step = (na + nthreads - 1) / nthreads;
    for (int t = 0; t < nthreads; t++) th.emplace_back(work, std::min(na, t * step), std::min(na, (t + 1) * step));
    for (auto& x : th) x.join();
}

inline uint64_t splitmix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}

struct Hyp { float R[9], t[3]; bool ok; };

void hypothesis(const float* src, const float* tar, const int32_t* pairs, int64_t np, uint64_t seed, uint64_t it,
                float coplanar_min, Hyp& h) {
    uint64_t ctr = 0;
    const uint64_t key = splitmix64(seed ^ splitmix64(it));
    auto draw = [&]() -> int64_t { return (int64_t)(splitmix64(key + (ctr++)) % (uint64_t)np); };
    int64_t id[4];
    h.ok = false;
    for (int attempt = 0; attempt < 4096 && !h.ok; attempt++) {
        id[0] = draw();                                   // :324
        for (int i = 1; i < 4; i++) {
            id[i] = draw();                               // :327
            for (int j = 0; j < i; j++)
                while (id[i] == id[j]) id[i] = draw();    // :330-331
        }
        const float* s0 = src + 3 * (int64_t)pairs[2 * id[0]];
        float p[3][3];
        for (int k = 1; k < 4; k++) {
            const float* sk = src + 3 * (int64_t)pairs[2 * id[k]];
            for (int a = 0; a < 3; a++) p[k - 1][a] = sk[a] - s0[a];   // :341-351
        }
        const float nx = p[0][1] * p[1][2] - p[0][2] * p[1][1];      // :353-355
        const float ny = p[0][2] * p[1][0] - p[0][0] * p[1][2];
        const float nz = p[0][0] * p[1][1] - p[0][1] * p[1][0];
        const float len = std::sqrt(nx * nx + ny * ny + nz * nz);
        const float dist = (nx * p[2][0] + ny * p[2][1] + nz * p[2][2]) / len;
        if (dist > coplanar_min) h.ok = true;                          // :358
    }
    if (!h.ok) return;
    float S[3][4], T[3][4], sc[3], tc[3];
    for (int i = 0; i < 4; i++)
        for (int a = 0; a < 3; a++) {
            S[a][i] = src[3 * (int64_t)pairs[2 * id[i]] + a];
            T[a][i] = tar[3 * (int64_t)pairs[2 * id[i] + 1] + a];
        }
    for (int a = 0; a < 3; a++) {
        sc[a] = (((S[a][0] + S[a][1]) + S[a][2]) + S[a][3]) / 4.f;     // :376-377
        tc[a] = (((T[a][0] + T[a][1]) + T[a][2]) + T[a][3]) / 4.f;
        for (int i = 0; i < 4; i++) { S[a][i] -= sc[a]; T[a][i] -= tc[a]; }
    }
    double H[9], U[9], sv[3], V[9];
    for (int r = 0; r < 3; r++)
        for (int c = 0; c < 3; c++) {
            float acc = 0.f;
            for (int i = 0; i < 4; i++) acc += T[r][i] * S[c][i];      // target_mat * source_mat^T, :382
            H[r * 3 + c] = (double)acc;
        }
    icp_oracle::svd3(H, U, sv, V);
    double Rd[9];
    for (int r = 0; r < 3; r++)
        for (int c = 0; c < 3; c++) Rd[r * 3 + c] = U[r * 3 + 0] * V[c * 3 + 0] + U[r * 3 + 1] * V[c * 3 + 1] + U[r * 3 + 2] * V[c * 3 + 2];
    float Rf[9];
    for (int i = 0; i < 9; i++) Rf[i] = (float)Rd[i];
    for (int a = 0; a < 3; a++) h.t[a] = tc[a] - ((Rf[a * 3] * sc[0] + Rf[a * 3 + 1] * sc[1]) + Rf[a * 3 + 2] * sc[2]);  // :386
    const double det = Rd[0] * (Rd[4] * Rd[8] - Rd[5] * Rd[7]) - Rd[1] * (Rd[3] * Rd[8] - Rd[5] * Rd[6]) + Rd[2] * (Rd[3] * Rd[7] - Rd[4] * Rd[6]);
    if (det < 0)  // :390-396
        for (int r = 0; r < 3; r++)
            for (int c = 0; c < 3; c++) Rf[r * 3 + c] = (float)(V[r * 3 + 0] * U[c * 3 + 0] + V[r * 3 + 1] * U[c * 3 + 1] + det * V[r * 3 + 2] * U[c * 3 + 2]);
    for (int i = 0; i < 9; i++) h.R[i] = Rf[i];
}

inline bool inlier(const float* R, const float* t, const float* s, const float* q, float thr) {
    float acc = 0.f;
    for (int a = 0; a < 3; a++) {
        const float m = ((R[a * 3] * s[0] + R[a * 3 + 1] * s[1]) + R[a * 3 + 2] * s[2]) + t[a];
        const float d = q[a] - m;
        acc += d * d;
    }
    return std::sqrt(acc) <= thr;   // :411, :424
}

}  // namespace

extern "C" {

int pco_desc_nn1(const float* a, int64_t na, const float* b, int64_t nb, int dim, int32_t* idx, float* d2, int nthreads) {
    nn1_scan(a, na, b, nb, dim, idx, d2, nthreads);
    return 0;
}

// returns the number of pairs kept; out_pairs must hold 2 * (mode ? ns + nt : ns) ints
int64_t pco_match(const float* sd, int64_t ns, const float* td, int64_t nt, int dim, int mode, float rate,
                  int32_t* out_pairs, int nthreads) {
    if (ns == 0 || nt == 0) return 0;
    std::vector<int32_t> s2t((size_t)ns), t2s((size_t)nt);
    std::vector<float> ds((size_t)ns), dt((size_t)nt);
    nn1_scan(td, nt, sd, ns, dim, t2s.data(), dt.data(), nthreads);
    nn1_scan(sd, ns, td, nt, dim, s2t.data(), ds.data(), nthreads);
    struct P { int32_t s, t; float d; };
    std::vector<P> c;
    if (mode == 0) {
        for (int64_t i = 0; i < ns; i++)
            if (t2s[(size_t)s2t[(size_t)i]] == i) c.push_back({(int32_t)i, s2t[(size_t)i], ds[(size_t)i]});  // :495-506
    } else {
        for (int64_t j = 0; j < nt; j++) c.push_back({t2s[(size_t)j], (int32_t)j, dt[(size_t)j]});   // :566
        for (int64_t i = 0; i < ns; i++) c.push_back({(int32_t)i, s2t[(size_t)i], ds[(size_t)i]});   // :584
    }
    std::stable_sort(c.begin(), c.end(), [](const P& a, const P& b) { return a.d < b.d; });            // :512-514 / :597-601
    const int64_t keep = (int64_t)std::floor((1.f - rate) * (float)c.size());                         // :516 / :603
    for (int64_t i = 0; i < keep; i++) { out_pairs[2 * i] = c[(size_t)i].s; out_pairs[2 * i + 1] = c[(size_t)i].t; }
    return keep;
}

// out: R[9] row-major, t[3]; returns consensus; *best_iter = -1 when nothing had an inlier
int64_t pco_ransac(const float* src, const float* tar, const int32_t* pairs, int64_t np, int max_iter, float thr,
                   float coplanar_min, uint64_t seed, float* R, float* t, int64_t* best_iter, int nthreads) {
    if (nthreads < 1) nthreads = 1;
    std::vector<int64_t> cons((size_t)max_iter, 0);
    auto work = [&](int t0) {
        for (int it = t0; it < max_iter; it += nthreads) {
            Hyp h;
            hypothesis(src, tar, pairs, np, seed, (uint64_t)it, coplanar_min, h);
            if (!h.ok) continue;
            int64_t c = 0;
            for (int64_t k = 0; k < np; k++)
                c += inlier(h.R, h.t, src + 3 * (int64_t)pairs[2 * k], tar + 3 * (int64_t)pairs[2 * k + 1], thr) ? 1 : 0;
            cons[(size_t)it] = c;
        }
    };
    std::vector<std::thread> th;
    for (int i = 0; i < nthreads; i++) th.emplace_back(work, i);
    for (auto& x : th) x.join();
    int64_t best = 0;
    *best_iter = -1;
    for (int it = 0; it < max_iter; it++)
        if (cons[(size_t)it] > best) { best = cons[(size_t)it]; *best_iter = it; }   // strict >, :423
    for (int i = 0; i < 9; i++) R[i] = (i % 4 == 0) ? 1.f : 0.f;
    for (int i = 0; i < 3; i++) t[i] = 0.f;
    if (*best_iter >= 0) {
        Hyp h;
        hypothesis(src, tar, pairs, np, seed, (uint64_t)*best_iter, coplanar_min, h);
        for (int i = 0; i < 9; i++) R[i] = h.R[i];
        for (int i = 0; i < 3; i++) t[i] = h.t[i];
    }
    return best;
}

}  // extern "C"
