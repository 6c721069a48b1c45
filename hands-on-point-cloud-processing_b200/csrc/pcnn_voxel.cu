// pcnn_voxel.cu -- voxel-grid downsample of a cloud with normals (sm_100a).
//
// Replaces the pcl::VoxelGrid<pcl::PointNormal> step of the reference:
//   readBinaryAndVoxelDown   Homework9/hw9/src/registration.cpp:8-68  (setDownsampleAllData(true),
//                            setLeafSize(v, v, v), filter)
//   VoxelGridSampling        registration.cpp:665-707                 (leaf 1.75)
// PCL is a third-party dependency that is not vendored in the reference (CMake takes whatever
// find_package(PCL) finds: 1.8 - 1.10 at the time); this follows the published algorithm of
// pcl/filters/impl/voxel_grid.hpp (applyFilter) and pcl/common/impl/accumulators.hpp:
//   1. bounding box of the cloud; min_b = floor(min * inv_leaf), max_b likewise (float
//      arithmetic, inv_leaf = 1 / leaf in float), div_b = max_b - min_b + 1;
//      if div_b.x * div_b.y * div_b.z exceeds INT_MAX the filter gives up and returns its input;
//   2. per point ijk = (int)(floor(p * inv_leaf) - (float)min_b), idx = ijk . (1, div.x, div.x*div.y);
//   3. points sorted by idx; one output point per occupied voxel, in ascending idx: xyz = mean
//      of the voxel's points, normal = sum of their normals, normalised (AccumulatorNormal).
// Here: keys by one kernel, the library's LSD radix sort (stable, so a voxel's points stay in
// input order), head flags + the tiled scan give every occupied voxel its output slot, and one
// warp per voxel forms the means: per-lane double partial sums over the lanes' strided share,
// folded in a fixed order (deterministic; PCL's own float sums depend on std::sort's order).
// HBM-bound integer/byte work: 28 B in and 4+4 B of sort traffic per pass and point, 24 B out
// per voxel.
#include <climits>
#include <cmath>
#include <vector>
#include "pcnn_internal.cuh"
#include "pcnn_sort.cuh"

namespace pcnn {

__global__ void __launch_bounds__(256) vox_bbox_kernel(const float* __restrict__ xyz, int n, int stride,
                                                       float* __restrict__ partial) {
    __shared__ float sm[6][8];
    const float INF = __int_as_float(0x7f800000);
    float lo[3] = {INF, INF, INF}, hi[3] = {-INF, -INF, -INF};
    for (long long i = blockIdx.x * 256ll + threadIdx.x; i < n; i += 256ll * gridDim.x) {
        const float* p = xyz + i * stride;
#pragma unroll
        for (int a = 0; a < 3; a++) { lo[a] = fminf(lo[a], p[a]); hi[a] = fmaxf(hi[a], p[a]); }
    }
#pragma unroll
    for (int a = 0; a < 3; a++)
        for (int o = 16; o > 0; o >>= 1) {
            lo[a] = fminf(lo[a], __shfl_xor_sync(0xffffffffu, lo[a], o));
            hi[a] = fmaxf(hi[a], __shfl_xor_sync(0xffffffffu, hi[a], o));
        }
    if ((threadIdx.x & 31) == 0)
        for (int a = 0; a < 3; a++) { sm[a][threadIdx.x >> 5] = lo[a]; sm[3 + a][threadIdx.x >> 5] = hi[a]; }
    __syncthreads();
    if (threadIdx.x < 6) {
        float v = sm[threadIdx.x][0];
        for (int w = 1; w < 8; w++) v = threadIdx.x < 3 ? fminf(v, sm[threadIdx.x][w]) : fmaxf(v, sm[threadIdx.x][w]);
        partial[blockIdx.x * 6 + threadIdx.x] = v;
    }
}

struct VoxGrid {
    float inv[3];
    int min_b[3];
    int mul[3];
};

__global__ void vox_key_kernel(const float* __restrict__ xyz, int n, int stride, VoxGrid g,
                               unsigned* __restrict__ keys, unsigned* __restrict__ vals) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float* p = xyz + (size_t)i * stride;
    int idx = 0;
#pragma unroll
    for (int a = 0; a < 3; a++) {
        // static_cast<int>(std::floor(p * inverse_leaf_size) - static_cast<float>(min_b)), voxel_grid.hpp
        const int ijk = (int)(__fsub_rn(floorf(__fmul_rn(p[a], g.inv[a])), (float)g.min_b[a]));
        idx += ijk * g.mul[a];
    }
    keys[i] = (unsigned)idx;
    vals[i] = (unsigned)i;
}

__global__ void vox_heads_kernel(const unsigned* __restrict__ keys, int n, long long* __restrict__ flag) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    flag[i] = (i == 0 || keys[i] != keys[i - 1]) ? 1 : 0;
}

__global__ void vox_starts_kernel(const long long* __restrict__ flag, const long long* __restrict__ offsets, int n,
                                  int* __restrict__ starts) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (flag[i]) starts[offsets[i]] = i;
    if (i == n - 1) starts[offsets[n]] = n;
}

// one warp per occupied voxel
__global__ void __launch_bounds__(256) vox_reduce_kernel(const float* __restrict__ xyz, const float* __restrict__ nrm,
                                                         int stride, int nstride, const unsigned* __restrict__ order,
                                                         const int* __restrict__ starts,
                                                         const long long* __restrict__ nvox_dev, long long capacity,
                                                         float* __restrict__ out_xyz, float* __restrict__ out_nrm) {
    const long long v = (blockIdx.x * 256ll + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    long long nvox = *nvox_dev;
    if (nvox > capacity) nvox = capacity;
    if (v >= nvox) return;
    const int b = starts[v], e = starts[v + 1];
    double s[6] = {0, 0, 0, 0, 0, 0};
    for (int i = b + lane; i < e; i += 32) {
        const size_t id = order[i];
        const float* p = xyz + id * stride;
        s[0] += (double)p[0]; s[1] += (double)p[1]; s[2] += (double)p[2];
        if (nrm) {
            const float* q = nrm + id * nstride;
            s[3] += (double)q[0]; s[4] += (double)q[1]; s[5] += (double)q[2];
        }
    }
#pragma unroll
    for (int a = 0; a < 6; a++)
        for (int o = 16; o > 0; o >>= 1) s[a] += __shfl_xor_sync(0xffffffffu, s[a], o);
    if (lane == 0) {
        const double cnt = (double)(e - b);
        out_xyz[3 * v + 0] = (float)(s[0] / cnt);
        out_xyz[3 * v + 1] = (float)(s[1] / cnt);
        out_xyz[3 * v + 2] = (float)(s[2] / cnt);
        if (nrm && out_nrm) {
            // AccumulatorNormal::get: the summed normal, normalised (a zero sum gives NaN, as Eigen's normalize())
            const double len = sqrt(s[3] * s[3] + s[4] * s[4] + s[5] * s[5]);
            out_nrm[3 * v + 0] = (float)(s[3] / len);
            out_nrm[3 * v + 1] = (float)(s[4] / len);
            out_nrm[3 * v + 2] = (float)(s[5] / len);
        }
    }
}

__global__ void vox_copy_kernel(const float* __restrict__ xyz, const float* __restrict__ nrm, int n, int stride,
                                int nstride, long long capacity, float* __restrict__ out_xyz,
                                float* __restrict__ out_nrm) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n || i >= capacity) return;
    for (int a = 0; a < 3; a++) {
        out_xyz[3 * (size_t)i + a] = xyz[(size_t)i * stride + a];
        if (nrm && out_nrm) out_nrm[3 * (size_t)i + a] = nrm[(size_t)i * nstride + a];
    }
}

}  // namespace pcnn

using namespace pcnn;

extern "C" int pcnn_voxel_downsample(const float* xyz_dev, const float* normals_dev, int64_t n64, int stride,
                                     int normal_stride, float leaf_x, float leaf_y, float leaf_z,
                                     float* out_xyz_dev, float* out_normals_dev, int64_t capacity,
                                     int64_t* out_count_host, int device, pcnn_stream stream) {
    PCNN_CHECK_ARG(out_count_host, "pcnn_voxel_downsample: out_count is NULL");
    PCNN_CHECK_ARG(n64 >= 0 && n64 < 0x7fffffff, "pcnn_voxel_downsample: n out of range");
    PCNN_CHECK_ARG(n64 == 0 || xyz_dev, "pcnn_voxel_downsample: xyz is NULL");
    PCNN_CHECK_ARG(stride >= 3 && (normals_dev == nullptr || normal_stride >= 3), "pcnn_voxel_downsample: stride must be >= 3");
    PCNN_CHECK_ARG(leaf_x > 0.f && leaf_y > 0.f && leaf_z > 0.f, "pcnn_voxel_downsample: leaf sizes must be positive");
    PCNN_CHECK_ARG(capacity >= 0 && (capacity == 0 || out_xyz_dev), "pcnn_voxel_downsample: out_xyz is NULL or capacity negative");
    DeviceGuard g(device);
    if (!g.ok) { set_error("pcnn_voxel_downsample: cannot select CUDA device %d (no CPU fallback exists)", device); return PCNN_ERR_CUDA; }
    cudaStream_t s = (cudaStream_t)stream;
    const int n = (int)n64;
    *out_count_host = 0;
    if (n == 0) return PCNN_OK;
    Scratch sc(s);
    // ---- 1. bounding box (one small read-back: the grid dimensions decide the key width)
    const int nb = 148 * 4 < (n + 255) / 256 ? 148 * 4 : (n + 255) / 256;
    float* partial;
    PCNN_CUDA(sc.get(&partial, (size_t)nb * 6));
    vox_bbox_kernel<<<nb, 256, 0, s>>>(xyz_dev, n, stride, partial);
    std::vector<float> hp((size_t)nb * 6);
    PCNN_CUDA(cudaMemcpyAsync(hp.data(), partial, sizeof(float) * hp.size(), cudaMemcpyDeviceToHost, s));
    PCNN_CUDA(cudaStreamSynchronize(s));
    float lo[3], hi[3];
    for (int a = 0; a < 3; a++) { lo[a] = INFINITY; hi[a] = -INFINITY; }
    for (int b = 0; b < nb; b++)
        for (int a = 0; a < 3; a++) { lo[a] = std::fmin(lo[a], hp[(size_t)b * 6 + a]); hi[a] = std::fmax(hi[a], hp[(size_t)b * 6 + 3 + a]); }
    PCNN_CHECK_ARG(std::isfinite(lo[0]) && std::isfinite(lo[1]) && std::isfinite(lo[2]) && std::isfinite(hi[0]) &&
                       std::isfinite(hi[1]) && std::isfinite(hi[2]),
                   "pcnn_voxel_downsample: the cloud holds non-finite coordinates");
    const float leaf[3] = {leaf_x, leaf_y, leaf_z};
    VoxGrid vg;
    long long div[3];
    long long cells = 1;
    bool too_fine = false;
    for (int a = 0; a < 3; a++) {
        vg.inv[a] = 1.0f / leaf[a];
        // the INT_MAX guard of voxel_grid.hpp, on (max - min) * inv + 1 per axis
        const long long d = (long long)((hi[a] - lo[a]) * vg.inv[a]) + 1;  // float, as PCL
        if (d > INT_MAX || cells * d > INT_MAX) too_fine = true; else cells *= d;
        const float fmin_b = std::floor(lo[a] * vg.inv[a]), fmax_b = std::floor(hi[a] * vg.inv[a]);
        if (!(std::fabs(fmin_b) < 2e9f && std::fabs(fmax_b) < 2e9f)) { too_fine = true; vg.min_b[a] = 0; div[a] = 1; continue; }
        vg.min_b[a] = (int)fmin_b;
        div[a] = (long long)(int)fmax_b - vg.min_b[a] + 1;
    }
    if (!too_fine && div[0] * div[1] * div[2] > INT_MAX) too_fine = true;
    if (too_fine) {
        // "Leaf size is too small for the input dataset. Integer indices would overflow.": PCL returns the input
        vox_copy_kernel<<<(n + 255) / 256, 256, 0, s>>>(xyz_dev, normals_dev, n, stride, normal_stride, capacity,
                                                         out_xyz_dev, out_normals_dev);
        PCNN_CUDA(cudaGetLastError());
        PCNN_CUDA(cudaStreamSynchronize(s));
        *out_count_host = n;
        return PCNN_OK;
    }
    vg.mul[0] = 1; vg.mul[1] = (int)div[0]; vg.mul[2] = (int)(div[0] * div[1]);
    int bits = 1;
    while (bits < 32 && (1ll << bits) < div[0] * div[1] * div[2]) bits++;
    // ---- 2. keys, stable sort
    unsigned *k0, *k1, *v0, *v1, *hist;
    long long *flag, *offsets;
    int* starts;
    PCNN_CUDA(sc.get(&k0, (size_t)n)); PCNN_CUDA(sc.get(&k1, (size_t)n));
    PCNN_CUDA(sc.get(&v0, (size_t)n)); PCNN_CUDA(sc.get(&v1, (size_t)n));
    PCNN_CUDA(sc.get(&hist, radix_hist_elems(n)));
    PCNN_CUDA(sc.get(&flag, (size_t)n)); PCNN_CUDA(sc.get(&offsets, (size_t)n + 1));
    PCNN_CUDA(sc.get(&starts, (size_t)n + 1));
    vox_key_kernel<<<(n + 255) / 256, 256, 0, s>>>(xyz_dev, n, stride, vg, k0, v0);
    const int where = radix_sort_pairs<unsigned>(k0, v0, k1, v1, hist, n, bits, s);
    const unsigned* keys = where ? k1 : k0;
    const unsigned* order = where ? v1 : v0;
    // ---- 3. one slot per occupied voxel, ascending key; means by one warp per voxel
    vox_heads_kernel<<<(n + 255) / 256, 256, 0, s>>>(keys, n, flag);
    if (int rc = scan_counts(flag, n, offsets, sc, s)) return rc;
    vox_starts_kernel<<<(n + 255) / 256, 256, 0, s>>>(flag, offsets, n, starts);
    const long long max_out = capacity < n ? capacity : n;
    if (max_out > 0) {
        const long long warps = max_out;
        vox_reduce_kernel<<<(unsigned)((warps * 32 + 255) / 256), 256, 0, s>>>(
            xyz_dev, normals_dev, stride, normal_stride, order, starts, offsets + n, capacity, out_xyz_dev,
            out_normals_dev);
    }
    PCNN_CUDA(cudaGetLastError());
    long long total = 0;
    PCNN_CUDA(cudaMemcpyAsync(&total, offsets + n, sizeof(long long), cudaMemcpyDeviceToHost, s));
    PCNN_CUDA(cudaStreamSynchronize(s));
    *out_count_host = total;
    return PCNN_OK;
}
