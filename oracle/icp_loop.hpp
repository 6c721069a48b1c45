// oracle/icp_loop.hpp -- TEST INFRASTRUCTURE, NOT PRODUCT.
//
// CPU restatement of the reference's ICP loops, shared by port_icp.cpp (own exact 1-NN)
// and ref_icp.cpp (the reference's vendored nanoflann 1.3.2 for the 1-NN).  PARITY UNPINNED
// BY THE REFERENCE: registration.cpp needs Eigen + PCL, which are neither under
// /root/reference nor installed, so it cannot be compiled here, and the reference ships no
// ICP fixture with inputs (Homework9/hw9/my_reg_result_*.txt lack the clouds and depend on
// std::random_device).  The restatement is pinned instead by known-transform recovery on
// synthetic pairs (tests/test_oracle_cpu.py) and by agreeing with itself when the 1-NN is
// swapped for the reference's nanoflann.
//
// Follows Homework9/hw9/src/registration.cpp:
//   transform_inplace      :165-178  transformCloudInplace, float32, p <- R*p + t
//   icp_point2point        :862-1011 ICPpoint2point
//   icp_point2plane        :710-859  ICPpoint2plane
// Quirks kept as written: the gate compares nanoflann's SQUARED distance with
// max_corres_dist (:783, :936); the p2p loss is ASSIGNED d2*d2 of the last kept pair (:939);
// convergence needs unchanged_count > 15 and the counter never resets (:828-838, :948-958);
// the det<0 repair is V*B*U^T with B(2,2)=det (:990-996); the linearised point-to-plane
// R_delta is not re-orthonormalised (:843-844); the source cloud is rewritten in float32
// every iteration (:1003).  Normal-space sampling (:728-729, :880-881) is outside the loop
// and not restated: callers pass the clouds the loop should run on (equivalent to the
// reference's commented-out bypass at :731-732 / :883-884).
//
// accum = 0: sums carried in float32 in pair order ("as written" -- Eigen's own reduction
//            order is not recoverable without Eigen, so this is a stand-in for it);
// accum = 1: the same per-pair float32 terms summed in double (what the CUDA path does).
// Dense linear algebra Eigen would do (3x3 SVD, 6x6 inverse) is done here in double.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <thread>
#include <vector>
#include <algorithm>

namespace icp_oracle {

struct Params {
    int max_iter;
    float max_corres_dist;  // compared against d^2
    float loss_epsilon;
    int mode;               // 0 point-to-point, 1 point-to-plane
    int accum;              // 0 float32 as written, 1 double
    int nthreads;
};

struct Result {
    float T[16];        // row-major 4x4 T_total
    int iters;          // iterations whose update was applied
    int converged;      // 1 if the unchanged_count rule fired
    float last_loss;
    int64_t last_pairs; // correspondences kept in the last executed iteration
    int status;         // 0 ok, 1 = an iteration had no correspondences / singular system
};

inline void transform_inplace(std::vector<float>& c, const float R[9], const float t[3]) {
    for (size_t i = 0; i + 2 < c.size(); i += 3) {
        float x = c[i], y = c[i + 1], z = c[i + 2];
        c[i + 0] = ((R[0] * x + R[1] * y) + R[2] * z) + t[0];
        c[i + 1] = ((R[3] * x + R[4] * y) + R[5] * z) + t[1];
        c[i + 2] = ((R[6] * x + R[7] * y) + R[8] * z) + t[2];
    }
}

// T_total <- T_delta * T_total, float32 4x4 (registration.cpp:849, :1002)
inline void compose(float T[16], const float R[9], const float t[3]) {
    float D[16] = {R[0], R[1], R[2], t[0], R[3], R[4], R[5], t[1],
                   R[6], R[7], R[8], t[2], 0, 0, 0, 1};
    float O[16];
    for (int r = 0; r < 4; r++)
        for (int c = 0; c < 4; c++)
            O[r * 4 + c] = ((D[r * 4 + 0] * T[0 * 4 + c] + D[r * 4 + 1] * T[1 * 4 + c]) +
                            D[r * 4 + 2] * T[2 * 4 + c]) + D[r * 4 + 3] * T[3 * 4 + c];
    std::memcpy(T, O, sizeof(O));
}

// One-sided (Hestenes) Jacobi SVD of a 3x3 matrix in double: A = U diag(s) V^T, singular
// values sorted descending as Eigen's JacobiSVD reports them.
inline void svd3(const double A[9], double U[9], double s[3], double V[9]) {
    double W[9];  // working copy, columns get orthogonalised: W = A * V
    std::memcpy(W, A, sizeof(W));
    for (int i = 0; i < 9; i++) V[i] = (i % 4 == 0) ? 1.0 : 0.0;
    for (int sweep = 0; sweep < 60; sweep++) {
        double off = 0;
        for (int p = 0; p < 2; p++)
            for (int q = p + 1; q < 3; q++) {
                double a = 0, b = 0, c = 0;
                for (int r = 0; r < 3; r++) {
                    a += W[r * 3 + p] * W[r * 3 + p];
                    b += W[r * 3 + q] * W[r * 3 + q];
                    c += W[r * 3 + p] * W[r * 3 + q];
                }
                off = std::max(off, std::fabs(c) / std::sqrt(std::max(a * b, 1e-300)));
                if (std::fabs(c) <= 1e-300) continue;
                double zeta = (b - a) / (2.0 * c);
                double tt = (zeta >= 0 ? 1.0 : -1.0) / (std::fabs(zeta) + std::sqrt(1.0 + zeta * zeta));
                double cs = 1.0 / std::sqrt(1.0 + tt * tt), sn = cs * tt;
                for (int r = 0; r < 3; r++) {
                    double wp = W[r * 3 + p], wq = W[r * 3 + q];
                    W[r * 3 + p] = cs * wp - sn * wq;
                    W[r * 3 + q] = sn * wp + cs * wq;
                    double vp = V[r * 3 + p], vq = V[r * 3 + q];
                    V[r * 3 + p] = cs * vp - sn * vq;
                    V[r * 3 + q] = sn * vp + cs * vq;
                }
            }
        if (off < 1e-15) break;
    }
    int order[3] = {0, 1, 2};
    double nrm[3];
    for (int c = 0; c < 3; c++)
        nrm[c] = std::sqrt(W[c] * W[c] + W[3 + c] * W[3 + c] + W[6 + c] * W[6 + c]);
    std::sort(order, order + 3, [&](int a, int b) { return nrm[a] > nrm[b]; });
    double Vs[9];
    for (int c = 0; c < 3; c++) {
        int o = order[c];
        s[c] = nrm[o];
        for (int r = 0; r < 3; r++) {
            U[r * 3 + c] = nrm[o] > 0 ? W[r * 3 + o] / nrm[o] : 0.0;
            Vs[r * 3 + c] = V[r * 3 + o];
        }
    }
    std::memcpy(V, Vs, sizeof(Vs));
    // complete a rank-deficient U (never needed for proper rigid data)
    if (s[2] <= 0 || s[2] < 1e-300) {
        U[2] = U[3] * U[7] - U[6] * U[4];
        U[5] = U[6] * U[1] - U[0] * U[7];
        U[8] = U[0] * U[4] - U[3] * U[1];
    }
}

// Solve N x = g for symmetric positive definite 6x6 N by Cholesky; false if not SPD.
inline bool solve6(const double N[36], const double g[6], double x[6]) {
    double L[36] = {0};
    for (int i = 0; i < 6; i++)
        for (int j = 0; j <= i; j++) {
            double s = N[i * 6 + j];
            for (int k = 0; k < j; k++) s -= L[i * 6 + k] * L[j * 6 + k];
            if (i == j) {
                if (!(s > 0)) return false;
                L[i * 6 + i] = std::sqrt(s);
            } else
                L[i * 6 + j] = s / L[j * 6 + j];
        }
    double y[6];
    for (int i = 0; i < 6; i++) {
        double s = g[i];
        for (int k = 0; k < i; k++) s -= L[i * 6 + k] * y[k];
        y[i] = s / L[i * 6 + i];
    }
    for (int i = 5; i >= 0; i--) {
        double s = y[i];
        for (int k = i + 1; k < 6; k++) s -= L[k * 6 + i] * x[k];
        x[i] = s / L[i * 6 + i];
    }
    return true;
}

// Optional record of the small dense systems the loop hands to svd3 / solve6, one row of 64
// doubles per executed iteration, so a test can give the SAME matrices to LAPACK
// (numpy.linalg.svd / solve) and check this stand-in for Eigen independently of the end pose:
//   p2p     [0..8] H  [9..17] U  [18..20] s  [21..29] V  [30..38] R_delta (double, after the det repair)
//   p2plane [0..35] N = A^T A  [36..41] g = A^T b  [42..47] x  [48] ok
struct Trace {
    double* rows = nullptr;
    int cap = 0, n = 0;
};
inline Trace*& trace_slot() {
    static thread_local Trace* t = nullptr;
    return t;
}

template <class F>
inline void parallel_for(int64_t n, int nthreads, F f) {
    if (nthreads <= 1 || n < 2) { f(0, n); return; }
    std::vector<std::thread> th;
    int64_t chunk = (n + nthreads - 1) / nthreads;
    for (int t = 0; t < nthreads; t++) {
        int64_t lo = t * chunk, hi = std::min<int64_t>(n, lo + chunk);
        if (lo >= hi) break;
        th.emplace_back([=] { f(lo, hi); });
    }
    for (auto& t : th) t.join();
}

// The iteration loop.  NN is any exact 1-NN functor over the target cloud:
//   void operator()(const float q[3], int64_t& idx, float& d2) const
// with d2 = ((dx*dx)+dy*dy)+dz*dz carried in float32, dx = q - p (nanoflann.hpp:383-408,
// the 0-3 trailing-component loop is the one a 3-D point takes).
// out_corr (optional, n_src entries): 1-NN index per source point in the last executed
// iteration, -1 where the gate rejected it.
template <class NN>
Result run(const NN& nn, const float* src, int64_t n_src, const float* tar, int64_t n_tar,
           const float* tar_normals, const float init_T[16], const Params& prm,
           int32_t* out_corr) {
    (void)n_tar;
    Result res;
    std::memset(&res, 0, sizeof(res));
    float R0[9] = {init_T[0], init_T[1], init_T[2], init_T[4], init_T[5], init_T[6],
                   init_T[8], init_T[9], init_T[10]};
    float t0[3] = {init_T[3], init_T[7], init_T[11]};
    std::vector<float> cur(src, src + 3 * n_src);
    transform_inplace(cur, R0, t0);  // :872-875
    float T[16] = {R0[0], R0[1], R0[2], t0[0], R0[3], R0[4], R0[5], t0[1],
                   R0[6], R0[7], R0[8], t0[2], 0, 0, 0, 1};  // :910-913
    std::vector<int64_t> nn_idx((size_t)n_src);
    std::vector<float> nn_d2((size_t)n_src);
    float last_loss = 0;
    size_t unchanged = 0;

    for (int iter = 0; iter < prm.max_iter; iter++) {
        parallel_for(n_src, prm.nthreads, [&](int64_t lo, int64_t hi) {
            for (int64_t i = lo; i < hi; i++) nn(&cur[3 * i], nn_idx[i], nn_d2[i]);
        });
        int64_t M = 0;
        float loss = 0;
        float R[9], t[3];
        bool ok = true;

        if (prm.mode == 0) {
            // ---- point-to-point, :917-1004
            double sp[3] = {0, 0, 0}, sq[3] = {0, 0, 0};
            float spf[3] = {0, 0, 0}, sqf[3] = {0, 0, 0};
            for (int64_t i = 0; i < n_src; i++)
                if (nn_d2[i] < prm.max_corres_dist) {
                    M++;
                    loss = nn_d2[i] * nn_d2[i];  // assigned, not accumulated (:939)
                    const float* p = &cur[3 * i];
                    const float* q = &tar[3 * nn_idx[i]];
                    for (int a = 0; a < 3; a++) {
                        sp[a] += p[a]; sq[a] += q[a];
                        spf[a] += p[a]; sqf[a] += q[a];
                    }
                }
            res.last_pairs = M;
            if (std::fabs(last_loss - loss) < prm.loss_epsilon) unchanged++;
            if (unchanged > 15) { res.converged = 1; res.last_loss = loss; break; }
            last_loss = loss;
            if (M == 0) { res.status = 1; break; }
            double H[9] = {0};
            double pc[3], qc[3];
            if (prm.accum == 0) {
                float pcf[3], qcf[3];
                for (int a = 0; a < 3; a++) { pcf[a] = spf[a] / (float)M; qcf[a] = sqf[a] / (float)M; }
                float Hf[9] = {0};
                for (int64_t i = 0; i < n_src; i++)
                    if (nn_d2[i] < prm.max_corres_dist) {
                        const float* p = &cur[3 * i];
                        const float* q = &tar[3 * nn_idx[i]];
                        for (int r = 0; r < 3; r++)
                            for (int c = 0; c < 3; c++)
                                Hf[r * 3 + c] += (q[r] - qcf[r]) * (p[c] - pcf[c]);  // tar_c * src_c^T
                    }
                for (int a = 0; a < 9; a++) H[a] = Hf[a];
                for (int a = 0; a < 3; a++) { pc[a] = pcf[a]; qc[a] = qcf[a]; }
            } else {
                for (int a = 0; a < 3; a++) { pc[a] = sp[a] / (double)M; qc[a] = sq[a] / (double)M; }
                for (int64_t i = 0; i < n_src; i++)
                    if (nn_d2[i] < prm.max_corres_dist) {
                        const float* p = &cur[3 * i];
                        const float* q = &tar[3 * nn_idx[i]];
                        for (int r = 0; r < 3; r++)
                            for (int c = 0; c < 3; c++)
                                H[r * 3 + c] += ((double)q[r] - qc[r]) * ((double)p[c] - pc[c]);
                    }
            }
            double U[9], s[3], V[9], Rd[9];
            svd3(H, U, s, V);
            for (int r = 0; r < 3; r++)
                for (int c = 0; c < 3; c++)
                    Rd[r * 3 + c] = U[r * 3 + 0] * V[c * 3 + 0] + U[r * 3 + 1] * V[c * 3 + 1] +
                                    U[r * 3 + 2] * V[c * 3 + 2];  // U * V^T
            double det = Rd[0] * (Rd[4] * Rd[8] - Rd[5] * Rd[7]) - Rd[1] * (Rd[3] * Rd[8] - Rd[5] * Rd[6]) +
                         Rd[2] * (Rd[3] * Rd[7] - Rd[4] * Rd[6]);
            if (det < 0) {  // :990-996 as written: V * B * U^T, B = diag(1,1,det)
                double B[3] = {1, 1, det};
                for (int r = 0; r < 3; r++)
                    for (int c = 0; c < 3; c++)
                        Rd[r * 3 + c] = V[r * 3 + 0] * B[0] * U[c * 3 + 0] + V[r * 3 + 1] * B[1] * U[c * 3 + 1] +
                                        V[r * 3 + 2] * B[2] * U[c * 3 + 2];
            }
            if (Trace* tr = trace_slot())
                if (tr->n < tr->cap) {
                    double* w = tr->rows + 64 * (size_t)tr->n++;
                    std::memset(w, 0, 64 * sizeof(double));
                    std::memcpy(w, H, sizeof(H)); std::memcpy(w + 9, U, sizeof(U)); std::memcpy(w + 18, s, sizeof(s));
                    std::memcpy(w + 21, V, sizeof(V)); std::memcpy(w + 30, Rd, sizeof(Rd));
                }
            for (int a = 0; a < 9; a++) R[a] = (float)Rd[a];
            for (int r = 0; r < 3; r++) {
                if (prm.accum == 0) {
                    float pcf[3] = {(float)pc[0], (float)pc[1], (float)pc[2]};
                    t[r] = (float)qc[r] - ((R[r * 3] * pcf[0] + R[r * 3 + 1] * pcf[1]) + R[r * 3 + 2] * pcf[2]);
                } else
                    t[r] = (float)(qc[r] - (Rd[r * 3] * pc[0] + Rd[r * 3 + 1] * pc[1] + Rd[r * 3 + 2] * pc[2]));
            }
        } else {
            // ---- point-to-plane, :769-852
            double N[36] = {0}, g[6] = {0}, bb = 0;
            float Nf[36] = {0}, gf[6] = {0}, bbf = 0;
            for (int64_t i = 0; i < n_src; i++)
                if (nn_d2[i] < prm.max_corres_dist) {
                    M++;
                    const float* p = &cur[3 * i];
                    const float* q = &tar[3 * nn_idx[i]];
                    const float* n = &tar_normals[3 * nn_idx[i]];
                    float a[6] = {n[2] * p[1] - n[1] * p[2], n[0] * p[2] - n[2] * p[0],
                                  n[1] * p[0] - n[0] * p[1], n[0], n[1], n[2]};  // :808-812
                    float b = n[0] * q[0] + n[1] * q[1] + n[2] * q[2] - n[0] * p[0] - n[1] * p[1] -
                              n[2] * p[2];  // :814
                    for (int r = 0; r < 6; r++) {
                        for (int c = 0; c < 6; c++) {
                            N[r * 6 + c] += (double)a[r] * (double)a[c];
                            Nf[r * 6 + c] += a[r] * a[c];
                        }
                        g[r] += (double)a[r] * (double)b;
                        gf[r] += a[r] * b;
                    }
                    bb += (double)b * (double)b;
                    bbf += b * b;
                }
            res.last_pairs = M;
            if (prm.accum == 0) {
                for (int a = 0; a < 36; a++) N[a] = Nf[a];
                for (int a = 0; a < 6; a++) g[a] = gf[a];
                bb = bbf;
            }
            double x[6] = {0};
            ok = M > 0 && solve6(N, g, x);  // x = (A^T A)^-1 A^T b, :818
            if (Trace* tr = trace_slot())
                if (tr->n < tr->cap) {
                    double* w = tr->rows + 64 * (size_t)tr->n++;
                    std::memset(w, 0, 64 * sizeof(double));
                    std::memcpy(w, N, sizeof(N)); std::memcpy(w + 36, g, sizeof(g)); std::memcpy(w + 42, x, sizeof(x));
                    w[48] = ok ? 1.0 : 0.0;
                }
            if (ok) {
                double xNx = 0, xg = 0;
                for (int r = 0; r < 6; r++) {
                    for (int c = 0; c < 6; c++) xNx += x[r] * N[r * 6 + c] * x[c];
                    xg += x[r] * g[r];
                }
                loss = (float)(xNx - 2.0 * xg + bb);  // = ||A x - b||^2, :820
            }
            if (std::fabs(last_loss - loss) < prm.loss_epsilon) unchanged++;
            if (unchanged > 15) { res.converged = 1; res.last_loss = loss; break; }
            last_loss = loss;
            if (!ok) { res.status = 1; break; }
            float xf[6];
            for (int a = 0; a < 6; a++) xf[a] = (float)x[a];
            float Rl[9] = {1, -xf[2], xf[1], xf[2], 1, -xf[0], -xf[1], xf[0], 1};  // :843
            std::memcpy(R, Rl, sizeof(Rl));
            t[0] = xf[3]; t[1] = xf[4]; t[2] = xf[5];
        }
        if (out_corr)
            for (int64_t i = 0; i < n_src; i++)
                out_corr[i] = nn_d2[i] < prm.max_corres_dist ? (int32_t)nn_idx[i] : -1;
        compose(T, R, t);
        transform_inplace(cur, R, t);
        res.iters = iter + 1;
        res.last_loss = loss;
    }
    std::memcpy(res.T, T, sizeof(T));
    return res;
}

}  // namespace icp_oracle
