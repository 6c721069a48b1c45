// pcnn_radius.cu -- batched fixed-radius search (sm_100a), two passes with CSR output.
//
// Replaces, for a whole batch, `RadiusNNResultSet rs(r); KDTreeRadiusNNSearch(...)`
// (Homework2/hw2/include/kdtree.hpp:367-402) / `OctreeRadiusNNSearch(...)` (octree.hpp:189-237)
// and the append list RadiusNNResultSet::addPoint (resultSet.hpp:130-140, dist <= radius).
//
// One warp per query.  Per step every lane tests one child box of the current node: a lower
// bound prunes (as kdtree.hpp:391-401 prunes on the split plane), and an upper bound gives
// the octree's `contains` fast path (octree.hpp:133-139, :196-207): a subtree whose farthest
// corner is within the ball is emitted wholesale without per-point tests.  A leaf is 32
// points, one per lane; hits are compacted with a ballot so the index/distance writes of a
// row are contiguous.  The test is on d2 against the largest double whose sqrt is <= radius,
// which is exactly the reference's `sqrt(d2) <= radius`.
// Pass 1 (FILL = false) counts, an exclusive scan makes the CSR offsets, pass 2 repeats the
// identical traversal and writes.  This operation is output-bound: see DESIGN.md.
#include <cmath>
#include <cfloat>
#include "pcnn_internal.cuh"
#include "pcnn_warp.cuh"

namespace pcnn {

template <bool FILL>
__device__ __forceinline__ void emit_range(const IndexView& iv, int p0, int p1, int lane, double qx,
                                           double qy, double qz, long long& run, int32_t* out_idx,
                                           float* out_dist, double* out_dist64) {
    if (FILL) {
        for (int p = p0 + lane; p < p1; p += 32) {
            const float4 pt = __ldg(iv.pts + p);
            const long long o = run + (p - p0);
            out_idx[o] = __float_as_int(pt.w);
            if (out_dist || out_dist64) {
                const double d = sqrt(key64(pt.x, pt.y, pt.z, qx, qy, qz));
                if (out_dist) out_dist[o] = (float)d;
                if (out_dist64) out_dist64[o] = d;
            }
        }
    }
    run += (long long)(p1 - p0);
}

template <bool FILL>
__global__ void __launch_bounds__(WQ_WARPS * 32)
radius_warp_kernel(IndexView iv, WarpPlan wp, const float* __restrict__ q, int nq, int stride, double thr,
                   long long* __restrict__ counts, const long long* __restrict__ offsets,
                   int32_t* __restrict__ out_idx, float* __restrict__ out_dist,
                   double* __restrict__ out_dist64) {
    __shared__ unsigned s_mask[WQ_WARPS][WQ_MAXD];
    __shared__ int s_base[WQ_WARPS][WQ_MAXD];
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int qi = blockIdx.x * WQ_WARPS + w;
    if (qi >= nq) return;  // warp-uniform
    const float* qp = q + (size_t)qi * stride;
    const double qx = (double)qp[0], qy = (double)qp[1], qz = (double)qp[2];
    long long run = FILL ? offsets[qi] : 0;

    if (iv.n > 0 && thr >= 0.0) {
        int d = 0;
        unsigned mask;
        {
            const int l = wp.lvl[0];
            const int cnt = dev_level_count(iv.n0, l);
            bool pend = false, cont = false;
            if (lane < cnt) {
                const float4* b = iv.box + 2 * ((size_t)iv.lev_off[l] + lane);
                const float4 lo = __ldg(b), hi = __ldg(b + 1);
                pend = boxlb64(lo, hi, qx, qy, qz) <= thr;
                cont = pend && boxub64(lo, hi, qx, qy, qz) <= thr;
            }
            unsigned cm = __ballot_sync(full, cont);
            mask = __ballot_sync(full, pend && !cont);
            while (cm) {
                const int j = __ffs((int)cm) - 1;
                cm &= cm - 1;
                const long long p0 = (long long)j << (l + PCNN_GROUP_LOG);
                long long p1 = (long long)(j + 1) << (l + PCNN_GROUP_LOG);
                if (p1 > iv.n) p1 = iv.n;
                emit_range<FILL>(iv, (int)p0, (int)p1, lane, qx, qy, qz, run, out_idx, out_dist, out_dist64);
            }
            if (lane == 0) s_base[w][0] = 0;
            __syncwarp();
        }
        while (true) {
            if (mask == 0) {
                if (d == 0) break;
                d--;
                mask = s_mask[w][d];
                continue;
            }
            const int j = __ffs((int)mask) - 1;
            mask &= mask - 1;
            const int node = s_base[w][d] + j;
            if (d < wp.nd - 1) {
                const int l = wp.lvl[d + 1];
                const int sh = wp.lvl[d] - l;
                const int base = node << sh;
                const int cnt = dev_level_count(iv.n0, l);
                bool pend = false, cont = false;
                if (lane < (1 << sh) && base + lane < cnt) {
                    const float4* b = iv.box + 2 * ((size_t)iv.lev_off[l] + base + lane);
                    const float4 lo = __ldg(b), hi = __ldg(b + 1);
                    pend = boxlb64(lo, hi, qx, qy, qz) <= thr;
                    cont = pend && boxub64(lo, hi, qx, qy, qz) <= thr;
                }
                unsigned cm = __ballot_sync(full, cont);
                const unsigned nm = __ballot_sync(full, pend && !cont);
                while (cm) {
                    const int c = __ffs((int)cm) - 1;
                    cm &= cm - 1;
                    const long long p0 = (long long)(base + c) << (l + PCNN_GROUP_LOG);
                    long long p1 = (long long)(base + c + 1) << (l + PCNN_GROUP_LOG);
                    if (p1 > iv.n) p1 = iv.n;
                    emit_range<FILL>(iv, (int)p0, (int)p1, lane, qx, qy, qz, run, out_idx, out_dist, out_dist64);
                }
                if (lane == 0) { s_mask[w][d] = mask; s_base[w][d + 1] = base; }
                __syncwarp();
                d++;
                mask = nm;
            } else {
                const int pidx = (node << wp.leaf_pts_log) + lane;
                const bool valid = lane < (1 << wp.leaf_pts_log) && pidx < iv.n;
                double dk = 0.0;
                int id = 0;
                if (valid) {
                    const float4 p = __ldg(iv.pts + pidx);
                    dk = key64(p.x, p.y, p.z, qx, qy, qz);
                    id = __float_as_int(p.w);
                }
                const bool hit = valid && dk <= thr;
                const unsigned m = __ballot_sync(full, hit);
                if (FILL && hit) {
                    const long long o = run + __popc(m & ((1u << lane) - 1u));
                    out_idx[o] = id;
                    if (out_dist || out_dist64) {
                        const double dist = sqrt(dk);
                        if (out_dist) out_dist[o] = (float)dist;
                        if (out_dist64) out_dist64[o] = dist;
                    }
                }
                run += __popc(m);
            }
        }
    }
    if (!FILL && lane == 0) counts[qi] = run;
}

// Exclusive scan of nq counts into offsets[nq+1] (one block; counts may alias offsets).
__global__ void __launch_bounds__(1024) scan_counts_kernel(const long long* __restrict__ counts, int nq,
                                                           long long* __restrict__ offsets) {
    __shared__ long long part[1024];
    const int t = threadIdx.x;
    const int chunk = (nq + 1023) / 1024;
    const int lo = min(t * chunk, nq), hi = min(lo + chunk, nq);
    long long s = 0;
    for (int i = lo; i < hi; i++) s += counts[i];
    part[t] = s;
    __syncthreads();
    for (int d = 1; d < 1024; d <<= 1) {
        long long v = (t >= d) ? part[t - d] : 0;
        __syncthreads();
        part[t] += v;
        __syncthreads();
    }
    long long run = part[t] - s;
    for (int i = lo; i < hi; i++) {
        long long c = counts[i];
        offsets[i] = run;
        run += c;
    }
    if (t == 1023) offsets[nq] = part[1023];
}

// Largest double t with sqrt(t) <= r under IEEE rounding; -1 if nothing can match.
static double radius_threshold(double r) {
    if (!(r >= 0.0)) return -1.0;
    if (std::isinf(r)) return DBL_MAX;
    double t = r * r;
    if (std::isinf(t)) return DBL_MAX;
    for (int i = 0; i < 8 && std::sqrt(t) <= r; i++) t = std::nextafter(t, INFINITY);
    for (int i = 0; i < 16 && std::sqrt(t) > r; i++) t = std::nextafter(t, -INFINITY);
    return t;
}

int launch_radius_count(const pcnn_index* ix, const float* q, int64_t nq64, int stride, double radius,
                        int64_t* out_offsets, cudaStream_t s) {
    const int nq = (int)nq64;
    static_assert(sizeof(long long) == sizeof(int64_t), "int64 layout");
    if (nq == 0) {
        PCNN_CUDA(cudaMemsetAsync(out_offsets, 0, sizeof(int64_t), s));
        return PCNN_OK;
    }
    const WarpPlan wp = make_warp_plan(ix->view);
    const double thr = radius_threshold(radius);
    // counts are written into offsets[1..nq] then scanned in place into offsets[0..nq]
    long long* counts = (long long*)out_offsets + 1;
    prof_begin(ix, s);
    radius_warp_kernel<false><<<(nq + WQ_WARPS - 1) / WQ_WARPS, WQ_WARPS * 32, 0, s>>>(
        ix->view, wp, q, nq, stride, thr, counts, nullptr, nullptr, nullptr, nullptr);
    prof_end(ix, s);
    Scratch sc(s);
    long long* tmp;
    PCNN_CUDA(sc.get(&tmp, (size_t)nq + 1));
    scan_counts_kernel<<<1, 1024, 0, s>>>(counts, nq, tmp);
    PCNN_CUDA(cudaMemcpyAsync(out_offsets, tmp, sizeof(int64_t) * ((size_t)nq + 1), cudaMemcpyDeviceToDevice, s));
    PCNN_CUDA(cudaGetLastError());
    return PCNN_OK;
}

int launch_radius_fill(const pcnn_index* ix, const float* q, int64_t nq64, int stride, double radius,
                       const int64_t* offsets, int32_t* out_idx, float* out_dist, double* out_dist64,
                       cudaStream_t s) {
    const int nq = (int)nq64;
    if (nq == 0) return PCNN_OK;
    const WarpPlan wp = make_warp_plan(ix->view);
    const double thr = radius_threshold(radius);
    prof_begin(ix, s);
    radius_warp_kernel<true><<<(nq + WQ_WARPS - 1) / WQ_WARPS, WQ_WARPS * 32, 0, s>>>(
        ix->view, wp, q, nq, stride, thr, nullptr, (const long long*)offsets, out_idx, out_dist, out_dist64);
    prof_end(ix, s);
    PCNN_CUDA(cudaGetLastError());
    return PCNN_OK;
}

}  // namespace pcnn
