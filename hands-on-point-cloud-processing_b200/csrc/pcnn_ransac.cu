// pcnn_ransac.cu -- global initialisation before ICP: descriptor matching and RANSAC (sm_100a).
//
// Replaces (Homework9/hw9/src/registration.cpp):
//   findRANSACCorrespondencesInter  :437-527   mutual exact 1-NN in descriptor space, sorted by
//                                              distance, worst x% dropped
//   findRANSACCorrespondencesUnion  :530-611   both directions' 1-NN pairs, sorted, worst x% dropped
//   RANSAC                          :288-434   4-point hypotheses, coplanarity check, Kabsch by 3x3
//                                              SVD, consensus = #pairs within dist_threshold
// The 1-NN is nanoflann's in the reference (KDTreeEigenMatrixAdaptor, 33 dimensions, exact):
// float32 squared L2 accumulated dimension by dimension, `result += diff * diff`
// (nanoflann.hpp L2_Simple_Adaptor), strict `<` acceptance.  In 33 dimensions a tree prunes
// almost nothing; here it is a dense all-pairs sweep with exactly that arithmetic (explicitly
// rounded product and sum, no FMA), candidates staged through shared memory, lowest index
// winning an exact tie.  (A -2ab GEMM on tensor cores would change the rounding and with it the
// arg-min of near-ties; the sizes -- thousands of keypoints -- make this launch-bound anyway.)
// RANSAC: the reference seeds std::mt19937 from std::random_device, so its hypotheses are not
// reproducible; here hypothesis `it` draws from a counter-based generator keyed by (seed, it),
// one warp per hypothesis (lane 0 samples and solves, all lanes count the consensus), and the
// best one is the largest consensus, earliest iteration -- what the sequential loop's strict
// `>` keeps (:423).  All max_iter hypotheses run concurrently.
#include <cmath>
#include "pcnn_internal.cuh"
#include "pcnn_linalg.cuh"
#include "pcnn_sort.cuh"

namespace pcnn {

constexpr int DM_MAXDIM = 64;
constexpr int DM_TILE = 64;  // candidates per shared-memory stage

// nearest b for every a.  grid.x tiles the queries (one per thread), grid.y slices the candidates
// so that a few thousand queries still fill the machine; each slice folds its best into
// packed[i] = (float bits of d2) << 32 | index by a 64-bit atomicMin: d2 >= 0, so the bit
// patterns order like the values, and among equal distances the lowest index wins -- the same
// answer as one sequential scan with a strict `<`.  D = compile-time dimension (0: any <= 64).
template <int D>
__global__ void __launch_bounds__(128) desc_nn1_kernel(const float* __restrict__ a, int na, const float* __restrict__ b,
                                                       int nb, int dim_rt, int slice,
                                                       unsigned long long* __restrict__ packed) {
    constexpr int CAP = D > 0 ? D : DM_MAXDIM;
    __shared__ float sb[DM_TILE][CAP + 1];
    const int dim = D > 0 ? D : dim_rt;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    float q[CAP];
#pragma unroll
    for (int d = 0; d < CAP; d++) q[d] = (i < na && d < dim) ? a[(size_t)i * dim + d] : 0.f;
    float best = __int_as_float(0x7f800000);
    int best_j = -1;
    const int jbeg = blockIdx.y * slice, jend = min(nb, jbeg + slice);
    for (int j0 = jbeg; j0 < jend; j0 += DM_TILE) {
        const int cnt = min(DM_TILE, jend - j0);
        __syncthreads();
        for (int e = threadIdx.x; e < cnt * dim; e += blockDim.x) sb[e / dim][e % dim] = b[(size_t)j0 * dim + e];
        __syncthreads();
        for (int j = 0; j < cnt; j++) {
            float r = 0.f;
#pragma unroll
            for (int d = 0; d < CAP; d++) {
                if (D > 0 || d < dim) {
                    const float diff = __fsub_rn(q[d], sb[j][d]);  // a[i] - data(b_idx, i)
                    r = __fadd_rn(r, __fmul_rn(diff, diff));       // result += diff * diff
                }
            }
            if (r < best) { best = r; best_j = j0 + j; }  // strict: the first (lowest index) wins a tie
        }
    }
    if (i < na && best_j >= 0)
        atomicMin(packed + i, ((unsigned long long)__float_as_uint(best) << 32) | (unsigned)best_j);
}

__global__ void desc_unpack_kernel(const unsigned long long* __restrict__ packed, int na, int32_t* __restrict__ out_idx,
                                   float* __restrict__ out_d2) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= na) return;
    const unsigned long long p = packed[i];
    const bool none = p == ~0ull;
    out_idx[i] = none ? -1 : (int32_t)(p & 0xffffffffull);
    out_d2[i] = none ? __int_as_float(0x7f800000) : __uint_as_float((unsigned)(p >> 32));
}

// Inter: pair (i, s2t[i]) kept when t2s[s2t[i]] == i.  Union: every target's and every source's pair.
__global__ void match_flag_kernel(const int32_t* __restrict__ s2t, const int32_t* __restrict__ t2s, int ns, int nt, int mode,
                                  long long* __restrict__ flag) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    const int total = mode == 0 ? ns : ns + nt;
    if (e >= total) return;
    if (mode == 0) flag[e] = (s2t[e] >= 0 && t2s[s2t[e]] == e) ? 1 : 0;
    else flag[e] = (e < nt ? t2s[e] : s2t[e - nt]) >= 0 ? 1 : 0;
}

__global__ void match_emit_kernel(const int32_t* __restrict__ s2t, const float* __restrict__ d_s2t,
                                  const int32_t* __restrict__ t2s, const float* __restrict__ d_t2s, int ns, int nt,
                                  int mode, const long long* __restrict__ flag, const long long* __restrict__ offsets,
                                  unsigned* __restrict__ keys, unsigned* __restrict__ vals, int2* __restrict__ pairs) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    const int total = mode == 0 ? ns : ns + nt;
    if (e >= total || !flag[e]) return;
    const long long o = offsets[e];
    int src, tar;
    float d;
    if (mode == 0) { src = e; tar = s2t[e]; d = d_s2t[e]; }              // dist_source2target, :505
    else if (e < nt) { src = t2s[e]; tar = e; d = d_t2s[e]; }            // :566
    else { src = e - nt; tar = s2t[e - nt]; d = d_s2t[e - nt]; }         // :584
    pairs[o] = make_int2(src, tar);
    keys[o] = __float_as_uint(d);  // distances are >= 0: their bit patterns order like the floats
    vals[o] = (unsigned)o;
}

__global__ void match_gather_kernel(const int2* __restrict__ pairs, const unsigned* __restrict__ order, long long keep,
                                    long long capacity, int32_t* __restrict__ out_pairs) {
    const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i >= keep || i >= capacity) return;
    const int2 p = pairs[order[i]];
    out_pairs[2 * i] = p.x;
    out_pairs[2 * i + 1] = p.y;
}

// ---- RANSAC
__device__ __forceinline__ unsigned long long splitmix64(unsigned long long x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}

struct RansacHyp {
    float R[9], t[3];
    bool ok;
};

// hypothesis `it`: sampling (:321-360), Kabsch (:362-396).  Float arithmetic without FMA where the
// reference computes in float; the SVD in double.
__device__ void ransac_hypothesis(const float* __restrict__ src, const float* __restrict__ tar,
                                  const int32_t* __restrict__ pairs, long long np, unsigned long long seed,
                                  unsigned long long it, float coplanar_min, RansacHyp& h) {
    unsigned long long ctr = 0;
    const unsigned long long key = splitmix64(seed ^ splitmix64(it));
    auto draw = [&]() -> long long { return (long long)(splitmix64(key + (ctr++)) % (unsigned long long)np); };
    long long id[4];
    h.ok = false;
    for (int attempt = 0; attempt < 4096 && !h.ok; attempt++) {
        id[0] = draw();
        for (int i = 1; i < 4; i++) {
            id[i] = draw();
            for (int j = 0; j < i; j++)
                while (id[i] == id[j]) id[i] = draw();  // as written at :329-333 (earlier j are not re-checked)
        }
        const float* s0 = src + 3 * (size_t)pairs[2 * id[0]];
        float p[3][3];
        for (int k = 1; k < 4; k++) {
            const float* sk = src + 3 * (size_t)pairs[2 * id[k]];
            for (int a = 0; a < 3; a++) p[k - 1][a] = __fsub_rn(sk[a], s0[a]);
        }
        const float nx = __fsub_rn(__fmul_rn(p[0][1], p[1][2]), __fmul_rn(p[0][2], p[1][1]));
        const float ny = __fsub_rn(__fmul_rn(p[0][2], p[1][0]), __fmul_rn(p[0][0], p[1][2]));
        const float nz = __fsub_rn(__fmul_rn(p[0][0], p[1][1]), __fmul_rn(p[0][1], p[1][0]));
        const float len = __fsqrt_rn(__fadd_rn(__fadd_rn(__fmul_rn(nx, nx), __fmul_rn(ny, ny)), __fmul_rn(nz, nz)));
        const float dist = __fdiv_rn(__fadd_rn(__fadd_rn(__fmul_rn(nx, p[2][0]), __fmul_rn(ny, p[2][1])), __fmul_rn(nz, p[2][2])), len);
        if (dist > coplanar_min) h.ok = true;  // signed, as at :358
    }
    if (!h.ok) return;
    float S[3][4], T[3][4], sc[3], tc[3];
    for (int i = 0; i < 4; i++)
        for (int a = 0; a < 3; a++) {
            S[a][i] = src[3 * (size_t)pairs[2 * id[i]] + a];
            T[a][i] = tar[3 * (size_t)pairs[2 * id[i] + 1] + a];
        }
    for (int a = 0; a < 3; a++) {  // rowwise().mean()
        sc[a] = __fdiv_rn(__fadd_rn(__fadd_rn(__fadd_rn(S[a][0], S[a][1]), S[a][2]), S[a][3]), 4.f);
        tc[a] = __fdiv_rn(__fadd_rn(__fadd_rn(__fadd_rn(T[a][0], T[a][1]), T[a][2]), T[a][3]), 4.f);
        for (int i = 0; i < 4; i++) { S[a][i] = __fsub_rn(S[a][i], sc[a]); T[a][i] = __fsub_rn(T[a][i], tc[a]); }
    }
    double H[9], U[9], sv[3], V[9];
    for (int r = 0; r < 3; r++)
        for (int c = 0; c < 3; c++) {  // target_mat * source_mat^T, float
            float acc = 0.f;
            for (int i = 0; i < 4; i++) acc = __fadd_rn(acc, __fmul_rn(T[r][i], S[c][i]));
            H[r * 3 + c] = (double)acc;
        }
    svd3_dev(H, U, sv, V);
    double Rd[9];
    for (int r = 0; r < 3; r++)
        for (int c = 0; c < 3; c++) Rd[r * 3 + c] = U[r * 3 + 0] * V[c * 3 + 0] + U[r * 3 + 1] * V[c * 3 + 1] + U[r * 3 + 2] * V[c * 3 + 2];
    float Rf[9];
    for (int i = 0; i < 9; i++) Rf[i] = (float)Rd[i];
    for (int a = 0; a < 3; a++)  // t_ = target_center - R_ * source_center
        h.t[a] = __fsub_rn(tc[a], __fadd_rn(__fadd_rn(__fmul_rn(Rf[a * 3], sc[0]), __fmul_rn(Rf[a * 3 + 1], sc[1])), __fmul_rn(Rf[a * 3 + 2], sc[2])));
    const double det = Rd[0] * (Rd[4] * Rd[8] - Rd[5] * Rd[7]) - Rd[1] * (Rd[3] * Rd[8] - Rd[5] * Rd[6]) + Rd[2] * (Rd[3] * Rd[7] - Rd[4] * Rd[6]);
    if (det < 0) {  // R_ = V * diag(1, 1, det) * U^T, t_ kept (:391-396)
        for (int r = 0; r < 3; r++)
            for (int c = 0; c < 3; c++) Rf[r * 3 + c] = (float)(V[r * 3 + 0] * U[c * 3 + 0] + V[r * 3 + 1] * U[c * 3 + 1] + det * V[r * 3 + 2] * U[c * 3 + 2]);
    }
    for (int i = 0; i < 9; i++) h.R[i] = Rf[i];
}

__device__ __forceinline__ bool ransac_inlier(const float* R, const float* t, const float* s, const float* q, float thr) {
    // (target_point - (R_ * source_point + t_)).norm() <= threshold, float (:411, :424)
    float acc = 0.f;
#pragma unroll
    for (int a = 0; a < 3; a++) {
        const float m = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(R[a * 3], s[0]), __fmul_rn(R[a * 3 + 1], s[1])), __fmul_rn(R[a * 3 + 2], s[2])), t[a]);
        const float d = __fsub_rn(q[a], m);
        acc = __fadd_rn(acc, __fmul_rn(d, d));
    }
    return __fsqrt_rn(acc) <= thr;
}

// Two kernels: the hypotheses (sampling + SVD) one THREAD each -- on one lane of a warp they cost
// half of a fused kernel's issue slots -- then one warp per hypothesis counts its consensus;
// best = max over (consensus << 32 | ~iteration).
__global__ void __launch_bounds__(128) ransac_hyp_kernel(const float* __restrict__ src, const float* __restrict__ tar,
                                                         const int32_t* __restrict__ pairs, long long np,
                                                         unsigned long long seed, int max_iter, float coplanar_min,
                                                         float* __restrict__ hyp /* [max_iter][13] */) {
    const int it = blockIdx.x * blockDim.x + threadIdx.x;
    if (it >= max_iter) return;
    RansacHyp h;
    ransac_hypothesis(src, tar, pairs, np, seed, (unsigned long long)it, coplanar_min, h);
    float* o = hyp + 13 * (size_t)it;
    o[12] = h.ok ? 1.f : 0.f;
    if (h.ok) {
        for (int i = 0; i < 9; i++) o[i] = h.R[i];
        for (int i = 0; i < 3; i++) o[9 + i] = h.t[i];
    }
}

__global__ void __launch_bounds__(256) ransac_kernel(const float* __restrict__ src, const float* __restrict__ tar,
                                                     const int32_t* __restrict__ pairs, long long np, int max_iter,
                                                     float thr, const float* __restrict__ hyp,
                                                     unsigned long long* __restrict__ best) {
    const int lane = threadIdx.x & 31;
    const long long it = blockIdx.x * 8ll + (threadIdx.x >> 5);
    if (it >= max_iter) return;
    const float* hp = hyp + 13 * (size_t)it;
    if (hp[12] == 0.f) return;
    float R[9], t[3];
    for (int i = 0; i < 9; i++) R[i] = hp[i];
    for (int i = 0; i < 3; i++) t[i] = hp[9 + i];
    int cnt = 0;
    for (long long c = lane; c < np; c += 32)
        cnt += ransac_inlier(R, t, src + 3 * (size_t)pairs[2 * c], tar + 3 * (size_t)pairs[2 * c + 1], thr) ? 1 : 0;
    cnt = __reduce_add_sync(0xffffffffu, cnt);
    if (lane == 0 && cnt > 0)
        atomicMax(best, ((unsigned long long)cnt << 32) | (unsigned long long)(0xffffffffu - (unsigned)it));
}

__global__ void ransac_final_kernel(const float* __restrict__ src, const float* __restrict__ tar,
                                    const int32_t* __restrict__ pairs, long long np, unsigned long long seed,
                                    float coplanar_min, const unsigned long long* __restrict__ best,
                                    pcnn_ransac_result* __restrict__ out) {
    const unsigned long long b = *best;
    pcnn_ransac_result r;
    for (int i = 0; i < 9; i++) r.R[i] = (i % 4 == 0) ? 1.f : 0.f;
    for (int i = 0; i < 3; i++) r.t[i] = 0.f;
    r.consensus = (long long)(b >> 32);
    r.best_iter = -1;
    if (r.consensus > 0) {
        const unsigned long long it = 0xffffffffull - (b & 0xffffffffull);
        RansacHyp h;
        ransac_hypothesis(src, tar, pairs, np, seed, it, coplanar_min, h);
        for (int i = 0; i < 9; i++) r.R[i] = h.R[i];
        for (int i = 0; i < 3; i++) r.t[i] = h.t[i];
        r.best_iter = (long long)it;
    }
    *out = r;
}

}  // namespace pcnn

using namespace pcnn;

extern "C" int pcnn_desc_nn1(const float* a_dev, int64_t na, const float* b_dev, int64_t nb, int dim,
                             int32_t* out_idx_dev, float* out_d2_dev, int device, pcnn_stream stream) {
    PCNN_CHECK_ARG(na >= 0 && nb >= 0 && na < 0x7fffffff && nb < 0x7fffffff, "pcnn_desc_nn1: sizes out of range");
    PCNN_CHECK_ARG(dim >= 1, "pcnn_desc_nn1: dim must be >= 1");
    if (dim > DM_MAXDIM) { set_error("pcnn_desc_nn1: dim %d exceeds %d", dim, DM_MAXDIM); return PCNN_ERR_UNSUPPORTED; }
    PCNN_CHECK_ARG(na == 0 || (a_dev && out_idx_dev && out_d2_dev), "pcnn_desc_nn1: NULL argument");
    PCNN_CHECK_ARG(nb == 0 || b_dev, "pcnn_desc_nn1: b is NULL");
    DeviceGuard g(device);
    if (!g.ok) { set_error("pcnn_desc_nn1: cannot select CUDA device %d (no CPU fallback exists)", device); return PCNN_ERR_CUDA; }
    if (na == 0) return PCNN_OK;
    cudaStream_t s = (cudaStream_t)stream;
    Scratch sc(s);
    unsigned long long* packed;
    PCNN_CUDA(sc.get(&packed, (size_t)na));
    PCNN_CUDA(cudaMemsetAsync(packed, 0xff, sizeof(unsigned long long) * (size_t)na, s));
    if (nb > 0) {
        const int gx = (int)((na + 127) / 128);
        int gy = (148 * 4 + gx - 1) / gx;  // aim at four blocks per SM
        const int max_gy = (int)((nb + DM_TILE - 1) / DM_TILE);
        if (gy > max_gy) gy = max_gy;
        if (gy < 1) gy = 1;
        int slice = (int)((nb + gy - 1) / gy);
        slice = (slice + DM_TILE - 1) / DM_TILE * DM_TILE;
        gy = (int)((nb + slice - 1) / slice);
        const dim3 grid((unsigned)gx, (unsigned)gy);
        if (dim == 33)
            desc_nn1_kernel<33><<<grid, 128, 0, s>>>(a_dev, (int)na, b_dev, (int)nb, dim, slice, packed);
        else
            desc_nn1_kernel<0><<<grid, 128, 0, s>>>(a_dev, (int)na, b_dev, (int)nb, dim, slice, packed);
    }
    desc_unpack_kernel<<<(unsigned)((na + 255) / 256), 256, 0, s>>>(packed, (int)na, out_idx_dev, out_d2_dev);
    PCNN_CUDA(cudaGetLastError());
    return PCNN_OK;
}

extern "C" int pcnn_match_descriptors(const float* src_desc_dev, int64_t n_src, const float* tar_desc_dev, int64_t n_tar,
                                      int dim, int mode, float rejection_rate, int32_t* out_pairs_dev, int64_t capacity,
                                      int64_t* out_count_host, int device, pcnn_stream stream) {
    PCNN_CHECK_ARG(out_count_host, "pcnn_match_descriptors: out_count is NULL");
    PCNN_CHECK_ARG(mode == 0 || mode == 1, "pcnn_match_descriptors: mode must be 0 (mutual) or 1 (union)");
    PCNN_CHECK_ARG(n_src >= 0 && n_tar >= 0 && n_src + n_tar < 0x7fffffff, "pcnn_match_descriptors: sizes out of range");
    PCNN_CHECK_ARG(rejection_rate >= 0.f && rejection_rate <= 1.f, "pcnn_match_descriptors: rejection_rate must be in [0, 1]");
    PCNN_CHECK_ARG(capacity >= 0 && (capacity == 0 || out_pairs_dev), "pcnn_match_descriptors: out_pairs is NULL or capacity negative");
    *out_count_host = 0;
    if (n_src == 0 || n_tar == 0) return PCNN_OK;
    DeviceGuard g(device);
    if (!g.ok) { set_error("pcnn_match_descriptors: cannot select CUDA device %d (no CPU fallback exists)", device); return PCNN_ERR_CUDA; }
    cudaStream_t s = (cudaStream_t)stream;
    const int ns = (int)n_src, nt = (int)n_tar, total = mode == 0 ? ns : ns + nt;
    Scratch sc(s);
    int32_t *s2t, *t2s;
    float *d_s2t, *d_t2s;
    long long *flag, *offsets;
    unsigned *k0, *k1, *v0, *v1, *hist;
    int2* pairs;
    PCNN_CUDA(sc.get(&s2t, (size_t)ns)); PCNN_CUDA(sc.get(&d_s2t, (size_t)ns));
    PCNN_CUDA(sc.get(&t2s, (size_t)nt)); PCNN_CUDA(sc.get(&d_t2s, (size_t)nt));
    PCNN_CUDA(sc.get(&flag, (size_t)total)); PCNN_CUDA(sc.get(&offsets, (size_t)total + 1));
    PCNN_CUDA(sc.get(&k0, (size_t)total)); PCNN_CUDA(sc.get(&k1, (size_t)total));
    PCNN_CUDA(sc.get(&v0, (size_t)total)); PCNN_CUDA(sc.get(&v1, (size_t)total));
    PCNN_CUDA(sc.get(&hist, radix_hist_elems(total))); PCNN_CUDA(sc.get(&pairs, (size_t)total));
    int rc = pcnn_desc_nn1(tar_desc_dev, nt, src_desc_dev, ns, dim, t2s, d_t2s, device, stream);  // :458-473
    if (rc == PCNN_OK) rc = pcnn_desc_nn1(src_desc_dev, ns, tar_desc_dev, nt, dim, s2t, d_s2t, device, stream);  // :476-492
    if (rc != PCNN_OK) return rc;
    match_flag_kernel<<<(total + 255) / 256, 256, 0, s>>>(s2t, t2s, ns, nt, mode, flag);
    if (int rc2 = scan_counts(flag, total, offsets, sc, s)) return rc2;
    match_emit_kernel<<<(total + 255) / 256, 256, 0, s>>>(s2t, d_s2t, t2s, d_t2s, ns, nt, mode, flag, offsets, k0, v0, pairs);
    long long count = 0;
    PCNN_CUDA(cudaMemcpyAsync(&count, offsets + total, sizeof(long long), cudaMemcpyDeviceToHost, s));
    PCNN_CUDA(cudaStreamSynchronize(s));
    // std::sort by distance (:512-514, :597-601; stable here: equal distances keep emission order),
    // N_inliers = floor((1 - rate) * size) in float (:516, :603)
    const int where = radix_sort_pairs<unsigned>(k0, v0, k1, v1, hist, (int)count, 32, s);
    const long long keep = (long long)std::floor((1.f - rejection_rate) * (float)count);
    if (keep > 0 && capacity > 0)
        match_gather_kernel<<<(unsigned)((keep + 255) / 256), 256, 0, s>>>(pairs, where ? v1 : v0, keep, capacity, out_pairs_dev);
    PCNN_CUDA(cudaGetLastError());
    PCNN_CUDA(cudaStreamSynchronize(s));
    *out_count_host = keep;
    return PCNN_OK;
}

extern "C" int pcnn_ransac(const float* src_kp_dev, const float* tar_kp_dev, const int32_t* pairs_dev, int64_t n_pairs,
                           const pcnn_ransac_params* prm, pcnn_ransac_result* out_host, int device, pcnn_stream stream) {
    PCNN_CHECK_ARG(prm && out_host, "pcnn_ransac: NULL argument");
    PCNN_CHECK_ARG(n_pairs >= 4 && n_pairs < 0x7fffffff, "pcnn_ransac: at least 4 correspondences are needed");
    PCNN_CHECK_ARG(src_kp_dev && tar_kp_dev && pairs_dev, "pcnn_ransac: NULL argument");
    PCNN_CHECK_ARG(prm->max_iter >= 1, "pcnn_ransac: max_iter must be >= 1");
    DeviceGuard g(device);
    if (!g.ok) { set_error("pcnn_ransac: cannot select CUDA device %d (no CPU fallback exists)", device); return PCNN_ERR_CUDA; }
    cudaStream_t s = (cudaStream_t)stream;
    Scratch sc(s);
    unsigned long long* best;
    pcnn_ransac_result* out_dev;
    float* hyp;
    PCNN_CUDA(sc.get(&best, 1));
    PCNN_CUDA(sc.get(&out_dev, 1));
    PCNN_CUDA(sc.get(&hyp, (size_t)prm->max_iter * 13));
    PCNN_CUDA(cudaMemsetAsync(best, 0, sizeof(unsigned long long), s));
    ransac_hyp_kernel<<<(unsigned)((prm->max_iter + 127) / 128), 128, 0, s>>>(src_kp_dev, tar_kp_dev, pairs_dev, n_pairs,
                                                                             prm->seed, prm->max_iter, prm->coplanar_min, hyp);
    ransac_kernel<<<(unsigned)((prm->max_iter + 7) / 8), 256, 0, s>>>(src_kp_dev, tar_kp_dev, pairs_dev, n_pairs, prm->max_iter,
                                                                     prm->dist_threshold, hyp, best);
    ransac_final_kernel<<<1, 1, 0, s>>>(src_kp_dev, tar_kp_dev, pairs_dev, n_pairs, prm->seed, prm->coplanar_min, best, out_dev);
    PCNN_CUDA(cudaGetLastError());
    PCNN_CUDA(cudaMemcpyAsync(out_host, out_dev, sizeof(pcnn_ransac_result), cudaMemcpyDeviceToHost, s));
    PCNN_CUDA(cudaStreamSynchronize(s));
    return PCNN_OK;
}
