// pcnn_normals.cu -- surface normals by PCA over the nearest neighbours (sm_100a).
//
// Replaces the per-point loop of the reference's normal estimation
// (Homework1/YuF_KIT-第1章作业/pca_normal.py:85-101 and PCA(), :17-33): for every point a hybrid
// search (at most max_nn nearest neighbours within `radius`, the point itself included;
// open3d KDTreeFlann.search_hybrid_vector_3d, :93), and when at least 3 are found the PCA of
// those points -- centre = mean, XTX = sum (p - c)(p - c)^T, eigen-decomposition -- whose
// eigenvector of the SMALLEST eigenvalue is the normal (:99-100); fewer than 3 leave the
// normal at zero (:88).
// Here: the batched kNN of this library on the index's own points (k = max_nn, exact double
// distances; the radius bound then drops the tail of each sorted row), followed by one thread
// per point: mean and scatter matrix in double over the gathered neighbours, cyclic Jacobi
// with eigenvectors.  The sign of an eigenvector is arbitrary (LAPACK's in the reference, the
// Jacobi rotations' here): normals agree up to sign.
#include <cmath>
#include "pcnn_internal.cuh"

namespace pcnn {

// original point order -> xyz (the queries) and -> Morton position (to gather neighbours)
__global__ void normals_unsort_kernel(IndexView iv, float* __restrict__ xyz, int* __restrict__ pos_of) {
    const int pos = blockIdx.x * blockDim.x + threadIdx.x;
    if (pos >= iv.n) return;
    const float4 p = __ldg(iv.pts + pos);
    const int id = __float_as_int(p.w);
    xyz[3 * (size_t)id + 0] = p.x; xyz[3 * (size_t)id + 1] = p.y; xyz[3 * (size_t)id + 2] = p.z;
    pos_of[id] = pos;
}

// symmetric 3x3 (a00 a01 a02 a11 a12 a22): eigenvector of the smallest eigenvalue, unit length
__device__ void sym3_smallest_eigenvector(const double* m, double* nv) {
    double a[3][3] = {{m[0], m[1], m[2]}, {m[1], m[3], m[4]}, {m[2], m[4], m[5]}};
    double v[3][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
    for (int sweep = 0; sweep < 16; sweep++) {
        const double off = fabs(a[0][1]) + fabs(a[0][2]) + fabs(a[1][2]);
        if (off == 0.0) break;
        for (int p = 0; p < 2; p++)
            for (int q = p + 1; q < 3; q++) {
                const double apq = a[p][q];
                if (apq == 0.0) continue;
                const double theta = (a[q][q] - a[p][p]) / (2.0 * apq);
                const double t = (theta >= 0.0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
                const double c = 1.0 / sqrt(t * t + 1.0), s = t * c;
                const int r = 3 - p - q;
                const double arp = a[r][p], arq = a[r][q];
                a[p][p] -= t * apq;
                a[q][q] += t * apq;
                a[p][q] = a[q][p] = 0.0;
                a[r][p] = a[p][r] = c * arp - s * arq;
                a[r][q] = a[q][r] = s * arp + c * arq;
                for (int i = 0; i < 3; i++) {
                    const double vip = v[i][p], viq = v[i][q];
                    v[i][p] = c * vip - s * viq;
                    v[i][q] = s * vip + c * viq;
                }
            }
    }
    int k = 0;
    if (a[1][1] < a[k][k]) k = 1;
    if (a[2][2] < a[k][k]) k = 2;
    const double len = sqrt(v[0][k] * v[0][k] + v[1][k] * v[1][k] + v[2][k] * v[2][k]);
    nv[0] = v[0][k] / len; nv[1] = v[1][k] / len; nv[2] = v[2][k] / len;
}

__global__ void __launch_bounds__(128) normals_pca_kernel(IndexView iv, const int* __restrict__ pos_of,
                                                          const int32_t* __restrict__ nn_idx,
                                                          const double* __restrict__ nn_dist, int n, int k,
                                                          double radius, float* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int32_t* row = nn_idx + (size_t)i * k;
    const double* drow = nn_dist + (size_t)i * k;
    int m = 0;  // neighbours within the radius: a prefix of the ascending row
    while (m < k && row[m] >= 0 && drow[m] <= radius) m++;
    float nx = 0.f, ny = 0.f, nz = 0.f;
    if (m >= 3) {  // pca_normal.py:94
        double cx = 0.0, cy = 0.0, cz = 0.0;
        for (int j = 0; j < m; j++) {
            const float4 p = __ldg(iv.pts + pos_of[row[j]]);
            cx += (double)p.x; cy += (double)p.y; cz += (double)p.z;
        }
        cx /= m; cy /= m; cz /= m;  // :18
        double s[6] = {0, 0, 0, 0, 0, 0};
        for (int j = 0; j < m; j++) {
            const float4 p = __ldg(iv.pts + pos_of[row[j]]);
            const double dx = (double)p.x - cx, dy = (double)p.y - cy, dz = (double)p.z - cz;  // :19
            s[0] += dx * dx; s[1] += dx * dy; s[2] += dx * dz; s[3] += dy * dy; s[4] += dy * dz; s[5] += dz * dz;  // :20
        }
        double nv[3];
        sym3_smallest_eigenvector(s, nv);
        nx = (float)nv[0]; ny = (float)nv[1]; nz = (float)nv[2];
    }
    out[3 * (size_t)i + 0] = nx; out[3 * (size_t)i + 1] = ny; out[3 * (size_t)i + 2] = nz;
}

}  // namespace pcnn

using namespace pcnn;

extern "C" int pcnn_estimate_normals(const pcnn_index* index, int max_nn, double radius, float* out_normals_dev,
                                     pcnn_stream stream) {
    PCNN_CHECK_ARG(index, "pcnn_estimate_normals: index is NULL");
    PCNN_CHECK_ARG(max_nn >= 1, "pcnn_estimate_normals: max_nn must be >= 1");
    if (max_nn > PCNN_MAX_K) { set_error("pcnn_estimate_normals: max_nn %d exceeds PCNN_MAX_K", max_nn); return PCNN_ERR_UNSUPPORTED; }
    PCNN_CHECK_ARG(radius == radius, "pcnn_estimate_normals: radius is NaN");
    PCNN_CHECK_ARG(index->n == 0 || out_normals_dev, "pcnn_estimate_normals: out_normals is NULL");
    DeviceGuard g(index->device);
    if (!g.ok) { set_error("pcnn_estimate_normals: cannot select CUDA device %d (no CPU fallback exists)", index->device); return PCNN_ERR_CUDA; }
    cudaStream_t s = (cudaStream_t)stream;
    const int n = index->n;
    if (n == 0) return PCNN_OK;
    Scratch sc(s);
    float* xyz;
    int* pos_of;
    int32_t* nn_idx;
    double* nn_dist;
    PCNN_CUDA(sc.get(&xyz, (size_t)n * 3));
    PCNN_CUDA(sc.get(&pos_of, (size_t)n));
    PCNN_CUDA(sc.get(&nn_idx, (size_t)n * max_nn));
    PCNN_CUDA(sc.get(&nn_dist, (size_t)n * max_nn));
    normals_unsort_kernel<<<(n + 255) / 256, 256, 0, s>>>(index->view, xyz, pos_of);
    int rc = launch_knn(index, xyz, n, 3, max_nn, nn_idx, nullptr, nn_dist, s);
    if (rc != PCNN_OK) return rc;
    normals_pca_kernel<<<(n + 127) / 128, 128, 0, s>>>(index->view, pos_of, nn_idx, nn_dist, n, max_nn, radius,
                                                       out_normals_dev);
    PCNN_CUDA(cudaGetLastError());
    return PCNN_OK;
}

extern "C" int pcnn_estimate_normals_host(const pcnn_index* index, int max_nn, double radius, float* out_normals_host) {
    PCNN_CHECK_ARG(index, "pcnn_estimate_normals_host: index is NULL");
    PCNN_CHECK_ARG(index->n == 0 || out_normals_host, "pcnn_estimate_normals_host: out_normals is NULL");
    DeviceGuard g(index->device);
    if (!g.ok) { set_error("pcnn_estimate_normals_host: cannot select CUDA device %d (no CPU fallback exists)", index->device); return PCNN_ERR_CUDA; }
    if (index->n == 0) return PCNN_OK;
    cudaStream_t s = nullptr;
    PCNN_CUDA(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
    wait_build(index, s);
    int rc = PCNN_OK;
    {
        Scratch sc(s);
        float* out_dev = nullptr;
        if (sc.get(&out_dev, (size_t)index->n * 3) != cudaSuccess) { set_error("pcnn_estimate_normals_host: device allocation failed"); rc = PCNN_ERR_NOMEM; }
        if (rc == PCNN_OK) rc = pcnn_estimate_normals(index, max_nn, radius, out_dev, (pcnn_stream)s);
        if (rc == PCNN_OK && cudaMemcpyAsync(out_normals_host, out_dev, sizeof(float) * 3 * (size_t)index->n, cudaMemcpyDeviceToHost, s) != cudaSuccess) {
            set_error("pcnn_estimate_normals_host: copy back failed");
            rc = PCNN_ERR_CUDA;
        }
        cudaStreamSynchronize(s);
    }
    cudaStreamSynchronize(s);
    cudaStreamDestroy(s);
    return rc;
}
