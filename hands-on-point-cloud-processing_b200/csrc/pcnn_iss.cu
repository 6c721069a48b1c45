// pcnn_iss.cu -- ISS keypoints on top of the batched radius search (sm_100a).
//
// Replaces ISSKeypoint::compute / getEigenvalues (Homework7/hw7/src/iss_detector.cpp:38-152):
//   1. radius neighbourhood of EVERY point of the cloud at local_radius (:48-56),
//   2. per point with >= 3 neighbours the scatter matrix about the point itself, each neighbour
//      weighted by 1 / (its own neighbour count) when weighted (:118-134), eigenvalues, the two
//      ratio tests and lambda3 > 0 (:69-82),
//   3. non-maximum suppression over a second radius search at non_max_radius, with a minimum
//      neighbour count (:87-108); keypoints come out in ascending point index.
// The reference materialises every neighbour list (rnn_idx); here nothing is written per
// neighbour: the three passes are three instantiations of ONE walk -- the dense stack walk of
// radius_stack_kernel (pcnn_radius_stack.cuh): one warp per point, queries are the index's own
// Morton-ordered points (perfectly coherent, no query ordering needed) -- with a per-lane
// visitor instead of CSR stores: a neighbour count, seven double sums, or a count plus "is any
// neighbour's lambda3 larger".  One warp reduction per point at the end of its walk.
//
// Inclusion follows Homework7's float KD-tree exactly (kdtree.cpp:347-355, resultSet.hpp:127-136):
//   key = f32(f32(f32(dx^2) + dy^2) + dz^2), every square and sum formed in double and rounded to
//   float (that is what `ElemType diff; diff += pow(db - query, 2)` does), dist = sqrtf(key),
//   hit iff dist <= radius.  sqrtf is monotone, so that is key <= t for the largest float t with
//   sqrtf(t) <= radius; a float32 fma chain decides all but the sliver within 2^-20 of t.
#include <cmath>
#include <cfloat>
#include "pcnn_internal.cuh"

namespace pcnn {

struct IssThr {
    float t;      // hit  <=>  key <= t;  < 0: nothing matches
    float t_in;   // fma-chain key <= t_in  => hit;  -1: no float pre-test
    float t_out;  // fma-chain key >  t_out => miss
};

static IssThr make_iss_thr(float r) {
    IssThr th;
    th.t = -1.f; th.t_in = -1.f; th.t_out = INFINITY;
    if (!(r >= 0.f)) return th;
    if (std::isinf(r)) { th.t = FLT_MAX; return th; }
    float t = r * r;
    if (std::isinf(t)) t = FLT_MAX;
    for (int i = 0; i < 8 && t < FLT_MAX && std::sqrt(t) <= r; i++) t = std::nextafterf(t, INFINITY);
    for (int i = 0; i < 16 && std::sqrt(t) > r; i++) t = std::nextafterf(t, -INFINITY);
    th.t = t;
    if (t > 1e-30f && t < 1e30f) {
        th.t_in = std::nextafterf((float)((double)t * (1.0 - 9.5367431640625e-07)), 0.f);
        th.t_out = std::nextafterf((float)((double)t * (1.0 + 9.5367431640625e-07)), INFINITY);
    }
    return th;
}

__device__ __forceinline__ bool iss_hit(const float4& p, float qx, float qy, float qz, const IssThr& th) {
    const float dx = p.x - qx, dy = p.y - qy, dz = p.z - qz;  // db - query, as the reference
    const float s = fmaf(dz, dz, fmaf(dy, dy, dx * dx));
    if (s <= th.t_in) return true;
    if (s > th.t_out) return false;
    float a = __double2float_rn(__dmul_rn((double)dx, (double)dx));
    a = __double2float_rn(__dadd_rn((double)a, __dmul_rn((double)dy, (double)dy)));
    a = __double2float_rn(__dadd_rn((double)a, __dmul_rn((double)dz, (double)dz)));
    return a <= th.t;
}

// 0 = out of reach, 1 = straddles the sphere, 2 = wholly inside the ball (conservative: the
// 2^-20 margins of t_in / t_out cover the rounding of these float bounds and of the key)
__device__ __forceinline__ int iss_classify(const float4& lo, const float4& hi, float qx, float qy, float qz,
                                            const IssThr& th) {
    const float ax = lo.x - qx, bx = qx - hi.x;
    const float ay = lo.y - qy, by = qy - hi.y;
    const float az = lo.z - qz, bz = qz - hi.z;
    const float gx = fmaxf(fmaxf(ax, bx), 0.f), gy = fmaxf(fmaxf(ay, by), 0.f), gz = fmaxf(fmaxf(az, bz), 0.f);
    if (fmaf(gx, gx, fmaf(gy, gy, gz * gz)) > th.t_out) return 0;
    const float fx = -fminf(ax, bx), fy = -fminf(ay, by), fz = -fminf(az, bz);
    return fmaf(fx, fx, fmaf(fy, fy, fz * fz)) <= th.t_in ? 2 : 1;
}

// Eigenvalues of a symmetric 3x3 matrix (a00 a01 a02 a11 a12 a22), ascending: cyclic Jacobi
// rotations in double with explicitly rounded products (no FMA contraction), so the oracle's
// restatement of the same iteration agrees to the last bits.  The reference uses
// Eigen::SelfAdjointEigenSolver<Matrix3f> (iss_detector.cpp:148-150).
__device__ void sym3_eigenvalues(const double* m, double* ev) {
    double a[3][3] = {{m[0], m[1], m[2]}, {m[1], m[3], m[4]}, {m[2], m[4], m[5]}};
    for (int sweep = 0; sweep < 12; sweep++) {
        const double off = dadd(dadd(fabs(a[0][1]), fabs(a[0][2])), fabs(a[1][2]));
        if (off == 0.0) break;
        for (int p = 0; p < 2; p++)
            for (int q = p + 1; q < 3; q++) {
                const double apq = a[p][q];
                if (apq == 0.0) continue;
                const double theta = dadd(a[q][q], -a[p][p]) / dmul(2.0, apq);
                const double t = (theta >= 0.0 ? 1.0 : -1.0) / dadd(fabs(theta), sqrt(dadd(dmul(theta, theta), 1.0)));
                const double c = 1.0 / sqrt(dadd(dmul(t, t), 1.0));
                const double s = dmul(t, c);
                const int r = 3 - p - q;
                const double arp = a[r][p], arq = a[r][q];
                a[p][p] = dadd(a[p][p], -dmul(t, apq));
                a[q][q] = dadd(a[q][q], dmul(t, apq));
                a[p][q] = a[q][p] = 0.0;
                a[r][p] = a[p][r] = dadd(dmul(c, arp), -dmul(s, arq));
                a[r][q] = a[q][r] = dadd(dmul(s, arp), dmul(c, arq));
            }
    }
    double e0 = a[0][0], e1 = a[1][1], e2 = a[2][2], tmp;
    if (e0 > e1) { tmp = e0; e0 = e1; e1 = tmp; }
    if (e1 > e2) { tmp = e1; e1 = e2; e2 = tmp; }
    if (e0 > e1) { tmp = e0; e0 = e1; e1 = tmp; }
    ev[0] = e0; ev[1] = e1; ev[2] = e2;
}

enum { ISS_COUNT = 0, ISS_COV = 1, ISS_NMS = 2 };
constexpr int IK_WARPS = 8;
constexpr int IK_STACK = 512;
constexpr int IK_TLIST = 96;
constexpr unsigned IK_NODE_MASK = 0x07ffffffu;

struct IssArgs {
    IssThr th;
    int* cnt_pos;      // [n] neighbours within local_radius, by Morton position
    float* wgt_pos;    // [n] 1 / cnt_pos in float (iss_detector.cpp:128), by Morton position
    float* lam_pos;    // [n] lambda3 or -1, by Morton position
    double* cov_pos;   // [n*6] scatter matrix (a00 a01 a02 a11 a12 a22), by Morton position; a00 = NaN: < 3 neighbours
    int* flag_pos;     // [n] keypoint flags, by Morton position
    float g21, g32;
    int min_neighbors;
    int weighted;
};

// One warp per point (Morton position `pos` = blockIdx.x * IK_WARPS + warp): the dense stack
// walk (see pcnn_radius_stack.cuh for the walk itself), every lane feeding its own visitor.
template <int PASS>
__global__ void __launch_bounds__(IK_WARPS * 32) iss_pass_kernel(IndexView iv, IssArgs A) {
    __shared__ unsigned s_stack[IK_WARPS][IK_STACK];
    __shared__ unsigned s_tl[IK_WARPS][IK_TLIST];
    __shared__ int s_loff[PCNN_MAX_LEVELS];
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (threadIdx.x < PCNN_MAX_LEVELS) s_loff[threadIdx.x] = threadIdx.x < iv.nlev ? 2 * iv.lev_off[threadIdx.x] : 0;
    __syncthreads();
    const int pos = blockIdx.x * IK_WARPS + w;
    if (pos >= iv.n) return;  // warp-uniform
    const float4 c = __ldg(iv.pts + pos);
    const IssThr th = A.th;
    const unsigned lt = (1u << lane) - 1u;

    // ---- per-lane visitor state
    int cnt = 0;
    double sw = 0.0, sxx = 0.0, sxy = 0.0, sxz = 0.0, syy = 0.0, syz = 0.0, szz = 0.0;
    bool beaten = false;
    float mine = 0.f;
    if (PASS == ISS_COV && A.cnt_pos[pos] < 3) {  // iss_detector.cpp:72
        if (lane == 0) A.cov_pos[6 * (size_t)pos] = __longlong_as_double(0x7ff8000000000000ll);
        return;
    }
    if (PASS == ISS_NMS) {
        mine = A.lam_pos[pos];
        if (mine == -1.f) {  // iss_detector.cpp:91
            if (lane == 0) A.flag_pos[pos] = 0;
            return;
        }
    }
    auto visit = [&](int j, const float4& p) {
        if (PASS == ISS_COUNT) {
            cnt++;
        } else if (PASS == ISS_COV) {
            // weight 1 / |N(j)| in float (iss_detector.cpp:128), differences in float (:131-133)
            const float wf = A.weighted ? __ldg(A.wgt_pos + j) : 1.f;
            const double wd = (double)wf;
            const double dx = (double)(p.x - c.x), dy = (double)(p.y - c.y), dz = (double)(p.z - c.z);
            const double wx = wd * dx, wy = wd * dy, wz = wd * dz;
            sw += wd;
            sxx += wx * dx; sxy += wx * dy; sxz += wx * dz;
            syy += wy * dy; syz += wy * dz; szz += wz * dz;
        } else {
            cnt++;
            beaten |= __ldg(A.lam_pos + j) > mine;  // iss_detector.cpp:99: lambda3[i] < lambda3[neighbour]
        }
    };
    auto level_nodes = [&](int L) -> int { return (int)(((unsigned)iv.n0 + (1u << L) - 1u) >> L); };
    int tcnt = 0;
    auto drain = [&]() {
        __syncwarp();
        const int j8 = lane & 7;
        for (int base = 0; base < tcnt; base += 4) {
            const int it = base + (lane >> 3);
            const unsigned ent = it < tcnt ? s_tl[w][it] : 0u;
            const int pidx = (int)((ent & 0x7fffffffu) << PCNN_GROUP_LOG) + j8;
            if (it < tcnt && pidx < iv.n) {
                const float4 p = __ldg(iv.pts + pidx);
                if ((ent >> 31) != 0u || iss_hit(p, c.x, c.y, c.z, th)) visit(pidx, p);
            }
        }
        tcnt = 0;
        __syncwarp();
    };

    if (th.t >= 0.f) {
        const int top = iv.nlev - 1;
        int sp = 0;
        if (top == 0) {
            if (lane == 0) s_tl[w][0] = 0u;
            tcnt = 1;
        } else {
            if (lane == 0) s_stack[w][0] = (unsigned)top << 27;
            sp = 1;
        }
        __syncwarp();
        while (sp > 0) {
            const int k = sp <= IK_STACK - 80 ? min(sp, 16) : 1;
            const int sh = k > 8 ? 1 : (k > 4 ? 2 : (k > 2 ? 3 : (k > 1 ? 4 : (sp <= IK_STACK - 80 ? 5 : 1))));
            const int e = lane >> sh, sub = lane & ((1 << sh) - 1);
            const unsigned ent = e < k ? s_stack[w][sp - k + e] : 0u;
            sp -= k;
            const int lvl = (int)(ent >> 27);
            const int she = min(sh, lvl);
            const int L = lvl - she;
            const int nd = ((int)(ent & IK_NODE_MASK) << she) + sub;
            int cls = 0;
            if (e < k && sub < (1 << she) && nd < level_nodes(L)) {
                const float4* b = iv.box + (size_t)s_loff[L] + 2 * (size_t)nd;
                cls = th.t_in >= 0.f ? iss_classify(__ldg(b), __ldg(b + 1), c.x, c.y, c.z, th) : 1;
            }
            __syncwarp();
            const bool to_stack = (cls == 1 && L > 0) || (PASS != ISS_COUNT && cls == 2 && L == 1);
            const unsigned ms = __ballot_sync(full, to_stack);
            if (to_stack) s_stack[w][sp + __popc(ms & lt)] = (unsigned)nd | ((unsigned)L << 27);
            sp += __popc(ms);
            const bool to_list = L == 0 && (cls == 1 || (PASS != ISS_COUNT && cls == 2));
            const unsigned mt = __ballot_sync(full, to_list);
            if (to_list) s_tl[w][tcnt + __popc(mt & lt)] = (unsigned)nd | (cls == 2 ? 0x80000000u : 0u);
            tcnt += __popc(mt);
            if (PASS == ISS_COUNT) {
                // a contained subtree only adds its size
                if (cls == 2) cnt += min(((nd + 1) << L) << PCNN_GROUP_LOG, iv.n) - ((nd << L) << PCNN_GROUP_LOG);
            } else {
                unsigned mb = __ballot_sync(full, cls == 2 && L >= 2);
                while (mb) {
                    const int b = __ffs((int)mb) - 1;
                    mb &= mb - 1;
                    const int cL = __shfl_sync(full, L, b), cn = __shfl_sync(full, nd, b);
                    const int p0 = (cn << cL) << PCNN_GROUP_LOG;
                    const int p1 = min(((cn + 1) << cL) << PCNN_GROUP_LOG, iv.n);
                    for (int pb = p0 + lane; pb < p1; pb += 32) visit(pb, __ldg(iv.pts + pb));
                }
            }
            if (tcnt > IK_TLIST - 32) drain();
            else __syncwarp();
        }
        drain();
    }

    // ---- one reduction per point
    if (PASS == ISS_COUNT) {
        cnt = __reduce_add_sync(full, cnt);
        if (lane == 0) A.cnt_pos[pos] = cnt;
    } else if (PASS == ISS_COV) {
        double v[7] = {sw, sxx, sxy, sxz, syy, syz, szz};
#pragma unroll
        for (int i = 0; i < 7; i++)
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) v[i] += __shfl_xor_sync(full, v[i], o);
        // the eigenvalues are left to iss_eigen_kernel, one THREAD per point: done here they would
        // run on one lane of the warp (ncu: 22 % of this kernel's instructions with 1 lane active)
        if (lane == 0) {
            // cov /= weight_sum (iss_detector.cpp:135)
            for (int i = 0; i < 6; i++) A.cov_pos[6 * (size_t)pos + i] = A.weighted ? v[1 + i] / v[0] : v[1 + i];
        }
    } else {
        cnt = __reduce_add_sync(full, cnt);
        const bool any_beaten = __any_sync(full, beaten);
        if (lane == 0) A.flag_pos[pos] = (cnt >= A.min_neighbors && !any_beaten) ? 1 : 0;
    }
}

// eigenvalues and the candidate tests (iss_detector.cpp:73-80), one thread per point
__global__ void __launch_bounds__(128) iss_eigen_kernel(IssArgs A, int n) {
    const int pos = blockIdx.x * blockDim.x + threadIdx.x;
    if (pos >= n) return;
    double m[6];
    for (int i = 0; i < 6; i++) m[i] = A.cov_pos[6 * (size_t)pos + i];
    float lam = -1.f;
    if (m[0] == m[0]) {
        double ev[3];
        sym3_eigenvalues(m, ev);
        // float, as Eigen::Vector3f in the reference
        const float l1 = (float)ev[2], l2 = (float)ev[1], l3 = (float)ev[0];
        if (l2 / l1 < A.g21 && l3 / l2 < A.g32 && l3 > 0.f) lam = l3;
    }
    A.lam_pos[pos] = lam;
}

__global__ void iss_weight_kernel(const int* __restrict__ cnt_pos, int n, float* __restrict__ wgt_pos) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) wgt_pos[i] = __fdiv_rn(1.f, (float)cnt_pos[i]);  // weight_nn = 1.f / rnn_idx[idx_neighbor].size()
}

// position order -> original point order
__global__ void iss_scatter_kernel(IndexView iv, const float* __restrict__ lam_pos, const int* __restrict__ flag_pos,
                                   float* __restrict__ lam_out, long long* __restrict__ flag_out) {
    const int pos = blockIdx.x * blockDim.x + threadIdx.x;
    if (pos >= iv.n) return;
    const int id = __float_as_int(__ldg(iv.pts + pos).w);
    if (lam_out) lam_out[id] = lam_pos[pos];
    flag_out[id] = flag_pos[pos];
}

__global__ void iss_compact_kernel(const long long* __restrict__ flag, const long long* __restrict__ offsets, int n,
                                   int32_t* __restrict__ out_idx, long long capacity) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (flag[i] && offsets[i] < capacity) out_idx[offsets[i]] = i;
}

int launch_iss(const pcnn_index* ix, const pcnn_iss_params* prm, int32_t* out_idx, int64_t capacity,
               int64_t* out_count_dev, float* out_lambda3, cudaStream_t s) {
    const int n = ix->n;
    Scratch sc(s);
    if (n == 0) {
        PCNN_CUDA(cudaMemsetAsync(out_count_dev, 0, sizeof(int64_t), s));
        return PCNN_OK;
    }
    IssArgs A;
    long long *flag, *offsets;
    PCNN_CUDA(sc.get(&A.cnt_pos, (size_t)n));
    PCNN_CUDA(sc.get(&A.wgt_pos, (size_t)n));
    PCNN_CUDA(sc.get(&A.lam_pos, (size_t)n));
    PCNN_CUDA(sc.get(&A.cov_pos, (size_t)n * 6));
    PCNN_CUDA(sc.get(&A.flag_pos, (size_t)n));
    PCNN_CUDA(sc.get(&flag, (size_t)n));
    PCNN_CUDA(sc.get(&offsets, (size_t)n + 1));
    A.g21 = prm->gamma21; A.g32 = prm->gamma32; A.min_neighbors = prm->min_neighbors; A.weighted = prm->weighted;
    const int grid = (n + IK_WARPS - 1) / IK_WARPS;
    A.th = make_iss_thr(prm->local_radius);
    prof_begin(ix, s);
    iss_pass_kernel<ISS_COUNT><<<grid, IK_WARPS * 32, 0, s>>>(ix->view, A);
    if (A.weighted) iss_weight_kernel<<<(n + 255) / 256, 256, 0, s>>>(A.cnt_pos, n, A.wgt_pos);
    iss_pass_kernel<ISS_COV><<<grid, IK_WARPS * 32, 0, s>>>(ix->view, A);
    iss_eigen_kernel<<<(n + 127) / 128, 128, 0, s>>>(A, n);
    A.th = make_iss_thr(prm->non_max_radius);
    iss_pass_kernel<ISS_NMS><<<grid, IK_WARPS * 32, 0, s>>>(ix->view, A);
    prof_end(ix, s);
    iss_scatter_kernel<<<(n + 255) / 256, 256, 0, s>>>(ix->view, A.lam_pos, A.flag_pos, out_lambda3, flag);
    if (int rc = scan_counts(flag, n, offsets, sc, s)) return rc;
    iss_compact_kernel<<<(n + 255) / 256, 256, 0, s>>>(flag, offsets, n, out_idx, (long long)capacity);
    PCNN_CUDA(cudaMemcpyAsync(out_count_dev, offsets + n, sizeof(int64_t), cudaMemcpyDeviceToDevice, s));
    PCNN_CUDA(cudaGetLastError());
    return PCNN_OK;
}

}  // namespace pcnn
