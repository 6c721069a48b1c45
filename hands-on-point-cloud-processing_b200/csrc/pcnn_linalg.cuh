// pcnn_linalg.cuh -- small dense linear algebra shared by the ICP and RANSAC kernels (device only).
#pragma once

namespace pcnn {

// One-sided Jacobi SVD, 3x3, double: A = U diag(s) V^T, s descending.
__device__ inline void svd3_dev(const double* A, double* U, double* s, double* V) {
    double W[9];
    for (int i = 0; i < 9; i++) { W[i] = A[i]; V[i] = (i % 4 == 0) ? 1.0 : 0.0; }
    for (int sweep = 0; sweep < 60; sweep++) {
        double off = 0;
#pragma unroll
        for (int p = 0; p < 2; p++)
#pragma unroll
            for (int q = p + 1; q < 3; q++) {
                double a = 0, b = 0, c = 0;
#pragma unroll
                for (int r = 0; r < 3; r++) {
                    a += W[r * 3 + p] * W[r * 3 + p];
                    b += W[r * 3 + q] * W[r * 3 + q];
                    c += W[r * 3 + p] * W[r * 3 + q];
                }
                off = fmax(off, fabs(c) / sqrt(fmax(a * b, 1e-300)));
                if (fabs(c) <= 1e-300) continue;
                const double zeta = (b - a) / (2.0 * c);
                const double tt = (zeta >= 0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
                const double cs = 1.0 / sqrt(1.0 + tt * tt), sn = cs * tt;
#pragma unroll
                for (int r = 0; r < 3; r++) {
                    const double wp = W[r * 3 + p], wq = W[r * 3 + q];
                    W[r * 3 + p] = cs * wp - sn * wq;
                    W[r * 3 + q] = sn * wp + cs * wq;
                    const double vp = V[r * 3 + p], vq = V[r * 3 + q];
                    V[r * 3 + p] = cs * vp - sn * vq;
                    V[r * 3 + q] = sn * vp + cs * vq;
                }
            }
        if (off < 1e-15) break;
    }
    double nrm[3];
    int ord[3] = {0, 1, 2};
    for (int c = 0; c < 3; c++) nrm[c] = sqrt(W[c] * W[c] + W[3 + c] * W[3 + c] + W[6 + c] * W[6 + c]);
    for (int a = 0; a < 2; a++)
        for (int b = a + 1; b < 3; b++)
            if (nrm[ord[b]] > nrm[ord[a]]) { int t = ord[a]; ord[a] = ord[b]; ord[b] = t; }
    double Vs[9];
    for (int c = 0; c < 3; c++) {
        const int o = ord[c];
        s[c] = nrm[o];
        for (int r = 0; r < 3; r++) {
            U[r * 3 + c] = nrm[o] > 0 ? W[r * 3 + o] / nrm[o] : 0.0;
            Vs[r * 3 + c] = V[r * 3 + o];
        }
    }
    for (int i = 0; i < 9; i++) V[i] = Vs[i];
    if (!(s[2] > 1e-300)) {  // rank-deficient: complete U with a cross product
        U[2] = U[3] * U[7] - U[6] * U[4];
        U[5] = U[6] * U[1] - U[0] * U[7];
        U[8] = U[0] * U[4] - U[3] * U[1];
    }
}

}  // namespace pcnn
