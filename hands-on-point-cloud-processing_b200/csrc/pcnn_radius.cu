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
#include "pcnn_sort.cuh"

namespace pcnn {

// float32 pre-tests around the exact double test.  s = fma(dx,dx, fma(dy,dy, dz*dz)) with
// dx = p - q in float32 differs from the exact key by less than 2^-21 relative (three
// roundings of the differences, three of the accumulation), so
//     s <= t_in  = thr * (1 - 2^-20), rounded down  =>  key <= thr   (hit, no double needed)
//     s >  t_out = thr * (1 + 2^-20), rounded up    =>  key >  thr   (miss)
// and only the sliver in between pays for key64().  The same margins make the box tests
// conservative: prune on lower bound > t_out, "contains" on farthest corner <= t_in.
struct RadiusThr {
    double thr;   // largest double whose sqrt is <= radius; < 0: nothing matches
    float t_in;   // -1 when the float pre-tests are disabled (tiny / huge radii)
    float t_out;
};

__device__ __forceinline__ float sq3(float ax, float ay, float az) {
    return fmaf(ax, ax, fmaf(ay, ay, az * az));
}

}  // namespace pcnn
#include "pcnn_radius_lockstep.cuh"  // needs RadiusThr
#include "pcnn_radius_stack.cuh"     // needs RadiusLane
namespace pcnn {

template <bool FILL, bool D64>
__device__ __forceinline__ void emit_range(const IndexView& iv, int p0, int p1, int lane, float qxf, float qyf,
                                           float qzf, double qx, double qy, double qz, bool fast32,
                                           long long& run, int32_t* out_idx, float* out_dist,
                                           double* out_dist64) {
    if (FILL) {
        for (int p = p0 + lane; p < p1; p += 32) {
            const float4 pt = __ldg(iv.pts + p);
            const long long o = run + (p - p0);
            out_idx[o] = __float_as_int(pt.w);
            if (D64) {
                const double d = sqrt(key64(pt.x, pt.y, pt.z, qx, qy, qz));
                if (out_dist) out_dist[o] = (float)d;
                out_dist64[o] = d;
            } else if (out_dist) {
                // float32 distances: sqrt of the float32 key, within 3e-7 relative of the exact
                // value (the contract asks for 1e-6); the exact double path serves out_dist64
                out_dist[o] = fast32 ? sqrt_approx32(sq3(pt.x - qxf, pt.y - qyf, pt.z - qzf))
                                     : (float)sqrt(key64(pt.x, pt.y, pt.z, qx, qy, qz));
            }
        }
    }
    run += (long long)(p1 - p0);
}

template <bool FILL, bool D64>
__global__ void __launch_bounds__(WQ_WARPS * 32)
radius_warp_kernel(IndexView iv, WarpPlan wp, const float* __restrict__ q, int nq, int stride,
                   const unsigned* __restrict__ perm, RadiusThr rt, long long* __restrict__ counts,
                   const long long* __restrict__ offsets, int32_t* __restrict__ out_idx,
                   float* __restrict__ out_dist, double* __restrict__ out_dist64) {
    __shared__ unsigned s_mask[WQ_WARPS][WQ_MAXD];
    __shared__ int s_base[WQ_WARPS][WQ_MAXD];
    __shared__ int s_lvl[WQ_MAXD], s_cnt[WQ_MAXD], s_off[WQ_MAXD];  // per chain depth: level, nodes, box offset
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (threadIdx.x < WQ_MAXD) {
        const int l = wp.lvl[threadIdx.x];
        s_lvl[threadIdx.x] = l;
        s_cnt[threadIdx.x] = (int)(((unsigned)iv.n0 + (1u << l) - 1u) >> l);
        s_off[threadIdx.x] = 2 * iv.lev_off[l];
    }
    __syncthreads();
    const int nd = wp.nd, leaf_log = wp.leaf_pts_log;
    const int slot = blockIdx.x * WQ_WARPS + w;
    if (slot >= nq) return;  // warp-uniform
    const int qi = perm ? (int)perm[slot] : slot;
    const float* qp = q + (size_t)qi * stride;
    const float qxf = qp[0], qyf = qp[1], qzf = qp[2];
    const double qx = (double)qxf, qy = (double)qyf, qz = (double)qzf;
    const double thr = rt.thr;
    const bool pre = rt.t_in >= 0.f;
    long long run = FILL ? offsets[qi] : 0;

    // box classification: 0 = prune, 1 = descend, 2 = wholly inside the ball
    auto classify = [&](const float4& lo, const float4& hi) -> int {
        if (pre) {
            const float gx = fmaxf(fmaxf(lo.x - qxf, qxf - hi.x), 0.f);
            const float gy = fmaxf(fmaxf(lo.y - qyf, qyf - hi.y), 0.f);
            const float gz = fmaxf(fmaxf(lo.z - qzf, qzf - hi.z), 0.f);
            if (sq3(gx, gy, gz) > rt.t_out) return 0;
            const float fx = fmaxf(qxf - lo.x, hi.x - qxf);
            const float fy = fmaxf(qyf - lo.y, hi.y - qyf);
            const float fz = fmaxf(qzf - lo.z, hi.z - qzf);
            return sq3(fx, fy, fz) <= rt.t_in ? 2 : 1;
        }
        if (!(boxlb64(lo, hi, qx, qy, qz) <= thr)) return 0;
        return boxub64(lo, hi, qx, qy, qz) <= thr ? 2 : 1;
    };

    if (iv.n > 0 && thr >= 0.0) {
        int d = 0;
        unsigned mask;
        {
            const int l = s_lvl[0];
            int cls = 0;
            if (lane < s_cnt[0]) {
                const float4* b = iv.box + (size_t)s_off[0] + 2 * lane;
                cls = classify(__ldg(b), __ldg(b + 1));
            }
            unsigned cm = __ballot_sync(full, cls == 2);
            mask = __ballot_sync(full, cls == 1);
            while (cm) {
                const int j = __ffs((int)cm) - 1;
                cm &= cm - 1;
                const long long p0 = (long long)j << (l + PCNN_GROUP_LOG);
                long long p1 = (long long)(j + 1) << (l + PCNN_GROUP_LOG);
                if (p1 > iv.n) p1 = iv.n;
                emit_range<FILL, D64>(iv, (int)p0, (int)p1, lane, qxf, qyf, qzf, qx, qy, qz, pre, run, out_idx,
                                      out_dist, out_dist64);
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
            if (d < nd - 1) {
                const int l = s_lvl[d + 1];
                const int sh = s_lvl[d] - l;
                const int base = node << sh;
                int cls = 0;
                if (lane < (1 << sh) && base + lane < s_cnt[d + 1]) {
                    const float4* b = iv.box + (size_t)s_off[d + 1] + 2 * (size_t)(base + lane);
                    cls = classify(__ldg(b), __ldg(b + 1));
                }
                unsigned cm = __ballot_sync(full, cls == 2);
                const unsigned nm = __ballot_sync(full, cls == 1);
                while (cm) {
                    const int c = __ffs((int)cm) - 1;
                    cm &= cm - 1;
                    const long long p0 = (long long)(base + c) << (l + PCNN_GROUP_LOG);
                    long long p1 = (long long)(base + c + 1) << (l + PCNN_GROUP_LOG);
                    if (p1 > iv.n) p1 = iv.n;
                    emit_range<FILL, D64>(iv, (int)p0, (int)p1, lane, qxf, qyf, qzf, qx, qy, qz, pre, run,
                                          out_idx, out_dist, out_dist64);
                }
                if (lane == 0) { s_mask[w][d] = mask; s_base[w][d + 1] = base; }
                __syncwarp();
                d++;
                mask = nm;
            } else {
                // a leaf is 2^leaf_log points: chunks of 32, one point per lane
                const int pbase0 = node << leaf_log;
                const int pend = min(pbase0 + (1 << leaf_log), iv.n);
                for (int pbase = pbase0; pbase < pend; pbase += 32) {
                    const int pidx = pbase + lane;
                    const bool valid = pidx < pend;
                    bool hit = false;
                    float4 p = make_float4(0.f, 0.f, 0.f, 0.f);
                    double dk = 0.0;
                    bool have_dk = false;
                    float s = 0.f;
                    if (valid) {
                        p = __ldg(iv.pts + pidx);
                        if (pre) {
                            s = sq3(p.x - qxf, p.y - qyf, p.z - qzf);
                            if (s <= rt.t_in) hit = true;
                            else if (!(s > rt.t_out)) { dk = key64(p.x, p.y, p.z, qx, qy, qz); have_dk = true; hit = dk <= thr; }
                        } else {
                            dk = key64(p.x, p.y, p.z, qx, qy, qz);
                            have_dk = true;
                            hit = dk <= thr;
                        }
                    }
                    const unsigned m = __ballot_sync(full, hit);
                    if (FILL && hit) {
                        const long long o = run + __popc(m & ((1u << lane) - 1u));
                        out_idx[o] = __float_as_int(p.w);
                        if (!D64 && pre) {
                            if (out_dist) out_dist[o] = sqrt_approx32(s);
                        } else if (D64 || out_dist) {
                            if (!have_dk) dk = key64(p.x, p.y, p.z, qx, qy, qz);
                            const double dist = sqrt(dk);
                            if (out_dist) out_dist[o] = (float)dist;
                            if (D64) out_dist64[o] = dist;
                        }
                    }
                    run += __popc(m);
                }
            }
        }
    }
    if (!FILL && lane == 0) counts[qi] = run;
}

// ------------------------------------------------------------------ batched form
// Neighbouring queries have nearly the same neighbourhood: with r = 0.5 m and queries a few
// centimetres apart the balls of consecutive queries overlap almost completely.  One warp owns
// 32 consecutive queries of the finely Morton-ordered batch (lane j keeps query j and its
// running output position) and cuts them into RUNS whose bounding box stays within
// radius / run_div per axis.  A run is walked ONCE, boxes against the run's bounding box:
//   out of reach of every query -> pruned;  wholly inside EVERY query's ball (the octree's
//   `contains`, octree.hpp:133-139, for the whole run) -> all its points go to every row
//   without a test;  otherwise, at a 32-point leaf, each lane loads one point and tests it
//   against every query of the run in turn (the query broadcast by shuffle): a dense
//   points x queries tile with the point in registers, six float32 operations per pair, one
//   ballot per (leaf, query) to compact that query's hits into its CSR row.
// Per query the walk and the point loads are paid once per run, not once per query.
template <bool FILL, bool D64>
__global__ void __launch_bounds__(WQ_WARPS * 32)
radius_batch_kernel(IndexView iv, WarpPlan wp, const float* __restrict__ q, int nq, int stride,
                    const unsigned* __restrict__ perm, RadiusThr rt, float run_extent,
                    long long* __restrict__ counts, const long long* __restrict__ offsets,
                    int32_t* __restrict__ out_idx, float* __restrict__ out_dist,
                    double* __restrict__ out_dist64) {
    __shared__ unsigned s_mask[WQ_WARPS][WQ_MAXD];
    __shared__ int s_base[WQ_WARPS][WQ_MAXD];
    __shared__ int s_lvl[WQ_MAXD], s_cnt[WQ_MAXD], s_off[WQ_MAXD];
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (threadIdx.x < WQ_MAXD) {
        const int l = wp.lvl[threadIdx.x];
        s_lvl[threadIdx.x] = l;
        s_cnt[threadIdx.x] = (int)(((unsigned)iv.n0 + (1u << l) - 1u) >> l);
        s_off[threadIdx.x] = 2 * iv.lev_off[l];
    }
    __syncthreads();
    const int slot0 = (blockIdx.x * WQ_WARPS + w) * 32;
    if (slot0 >= nq) return;  // warp-uniform
    const int nb = min(32, nq - slot0);  // queries of this warp
    const bool haveq = lane < nb;
    const int qi = perm ? (int)perm[slot0 + (haveq ? lane : 0)] : slot0 + (haveq ? lane : 0);
    const float* qp = q + (size_t)qi * stride;
    const float qxf = qp[0], qyf = qp[1], qzf = qp[2];
    const double thr = rt.thr;
    const bool pre = rt.t_in >= 0.f;
    const int nd = wp.nd, leaf_log = wp.leaf_pts_log;
    const unsigned lt = (1u << lane) - 1u;
    long long run = (FILL && haveq) ? offsets[qi] : 0;

    if (iv.n > 0 && thr >= 0.0) {
        int s0 = 0;
        while (s0 < nb) {
            // ---- next run [s0, e): extend while the bounding box stays within run_extent per axis
            float lox = __shfl_sync(full, qxf, s0), loy = __shfl_sync(full, qyf, s0), loz = __shfl_sync(full, qzf, s0);
            float hix = lox, hiy = loy, hiz = loz;
            int e = s0 + 1;
            while (e < nb) {
                const float x = __shfl_sync(full, qxf, e), y = __shfl_sync(full, qyf, e), z = __shfl_sync(full, qzf, e);
                const float nlx = fminf(lox, x), nly = fminf(loy, y), nlz = fminf(loz, z);
                const float nhx = fmaxf(hix, x), nhy = fmaxf(hiy, y), nhz = fmaxf(hiz, z);
                if (!(nhx - nlx <= run_extent && nhy - nly <= run_extent && nhz - nlz <= run_extent)) break;
                lox = nlx; loy = nly; loz = nlz; hix = nhx; hiy = nhy; hiz = nhz;
                e++;
            }
            // 0 = out of reach of every query, 2 = inside every query's ball, 1 = look closer
            auto classify = [&](const float4& lo, const float4& hi) -> int {
                if (pre) {
                    const float gx = fmaxf(fmaxf(lo.x - hix, lox - hi.x), 0.f);
                    const float gy = fmaxf(fmaxf(lo.y - hiy, loy - hi.y), 0.f);
                    const float gz = fmaxf(fmaxf(lo.z - hiz, loz - hi.z), 0.f);
                    if (sq3(gx, gy, gz) > rt.t_out) return 0;
                    const float fx = fmaxf(hi.x - lox, hix - lo.x);
                    const float fy = fmaxf(hi.y - loy, hiy - lo.y);
                    const float fz = fmaxf(hi.z - loz, hiz - lo.z);
                    return sq3(fx, fy, fz) <= rt.t_in ? 2 : 1;
                }
                // exact path (tiny / huge radii): distance between the two boxes, in double
                const double gx = dmax0((double)lo.x - (double)hix, (double)lox - (double)hi.x);
                const double gy = dmax0((double)lo.y - (double)hiy, (double)loy - (double)hi.y);
                const double gz = dmax0((double)lo.z - (double)hiz, (double)loz - (double)hi.z);
                return dadd(dadd(dmul(gx, gx), dmul(gy, gy)), dmul(gz, gz)) <= thr ? 1 : 0;
            };
            // every point of [p0, p1) is a hit of every query of the run
            auto emit_all = [&](int p0, int p1) {
                if (FILL) {
                    for (int pb = p0; pb < p1; pb += 32) {
                        const int pidx = pb + lane;
                        const bool valid = pidx < p1;
                        float4 p = make_float4(0.f, 0.f, 0.f, 0.f);
                        if (valid) p = __ldg(iv.pts + pidx);
                        for (int c = s0; c < e; c++) {
                            const long long ro = __shfl_sync(full, run, c) + (pb - p0);
                            const float cx = __shfl_sync(full, qxf, c), cy = __shfl_sync(full, qyf, c),
                                        cz = __shfl_sync(full, qzf, c);
                            if (valid) {
                                out_idx[ro + lane] = __float_as_int(p.w);
                                if (!D64) {
                                    if (out_dist) out_dist[ro + lane] = sqrt_approx32(sq3(p.x - cx, p.y - cy, p.z - cz));
                                } else {
                                    const double dist = sqrt(key64(p.x, p.y, p.z, (double)cx, (double)cy, (double)cz));
                                    if (out_dist) out_dist[ro + lane] = (float)dist;
                                    out_dist64[ro + lane] = dist;
                                }
                            }
                        }
                    }
                }
                if (lane >= s0 && lane < e) run += (long long)(p1 - p0);
            };
            int d = 0;
            unsigned mask;
            {
                int cls = 0;
                const int l = s_lvl[0];
                if (lane < s_cnt[0]) {
                    const float4* b = iv.box + (size_t)s_off[0] + 2 * lane;
                    cls = classify(__ldg(b), __ldg(b + 1));
                }
                unsigned cm = __ballot_sync(full, cls == 2);
                mask = __ballot_sync(full, cls == 1);
                while (cm) {
                    const int c = __ffs((int)cm) - 1;
                    cm &= cm - 1;
                    const long long p0 = (long long)c << (l + PCNN_GROUP_LOG);
                    long long p1 = (long long)(c + 1) << (l + PCNN_GROUP_LOG);
                    if (p1 > iv.n) p1 = iv.n;
                    emit_all((int)p0, (int)p1);
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
                if (d < nd - 1) {
                    const int l = s_lvl[d + 1];
                    const int sh = s_lvl[d] - l;
                    const int base = node << sh;
                    int cls = 0;
                    if (lane < (1 << sh) && base + lane < s_cnt[d + 1]) {
                        const float4* b = iv.box + (size_t)s_off[d + 1] + 2 * (size_t)(base + lane);
                        cls = classify(__ldg(b), __ldg(b + 1));
                    }
                    unsigned cm = __ballot_sync(full, cls == 2);
                    const unsigned nm = __ballot_sync(full, cls == 1);
                    while (cm) {
                        const int c = __ffs((int)cm) - 1;
                        cm &= cm - 1;
                        const long long p0 = (long long)(base + c) << (l + PCNN_GROUP_LOG);
                        long long p1 = (long long)(base + c + 1) << (l + PCNN_GROUP_LOG);
                        if (p1 > iv.n) p1 = iv.n;
                        emit_all((int)p0, (int)p1);
                    }
                    if (lane == 0) { s_mask[w][d] = mask; s_base[w][d + 1] = base; }
                    __syncwarp();
                    d++;
                    mask = nm;
                } else {
                    // ---- leaf: one point per lane against every query of the run
                    const int pidx = (node << leaf_log) + lane;
                    const bool valid = lane < (1 << leaf_log) && pidx < iv.n;
                    float4 p = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (valid) p = __ldg(iv.pts + pidx);
                    const int id = __float_as_int(p.w);
                    for (int c = s0; c < e; c++) {
                        const float cx = __shfl_sync(full, qxf, c), cy = __shfl_sync(full, qyf, c),
                                    cz = __shfl_sync(full, qzf, c);
                        bool hit = false;
                        float s = 0.f;
                        double dk = 0.0;
                        bool have_dk = false;
                        if (valid) {
                            if (pre) {
                                s = sq3(p.x - cx, p.y - cy, p.z - cz);
                                hit = s <= rt.t_in;
                                if (!hit && !(s > rt.t_out)) {
                                    dk = key64(p.x, p.y, p.z, (double)cx, (double)cy, (double)cz);
                                    have_dk = true;
                                    hit = dk <= thr;
                                }
                            } else {
                                dk = key64(p.x, p.y, p.z, (double)cx, (double)cy, (double)cz);
                                have_dk = true;
                                hit = dk <= thr;
                            }
                        }
                        const unsigned m = __ballot_sync(full, hit);
                        if (m == 0) continue;
                        if (FILL) {
                            const long long ro = __shfl_sync(full, run, c);
                            if (hit) {
                                const long long o = ro + __popc(m & lt);
                                out_idx[o] = id;
                                if (!D64 && pre) {
                                    if (out_dist) out_dist[o] = sqrt_approx32(s);
                                } else if (D64 || out_dist) {
                                    if (!have_dk) dk = key64(p.x, p.y, p.z, (double)cx, (double)cy, (double)cz);
                                    const double dist = sqrt(dk);
                                    if (out_dist) out_dist[o] = (float)dist;
                                    if (D64) out_dist64[o] = dist;
                                }
                            }
                        }
                        if (lane == c) run += __popc(m);
                    }
                }
            }
            s0 = e;
        }
    }
    if (!FILL && haveq) counts[qi] = run;
}

// Exclusive scan of nq counts into offsets[nq+1]: tiles of 4096 counts, one block per tile, four
// consecutive counts per thread and warp-shuffle scans inside the tile (scan_tiles_kernel writes
// tile-local offsets and the tile's sum); one block scans the tile sums (<= 2^19 tiles for 2^31
// queries, looped); scan_add_kernel adds each tile's base and writes offsets[nq].
constexpr int SCAN_TILE = 4096;

__device__ __forceinline__ long long block_exclusive_1024(long long s, long long* wsum, long long* total) {
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    long long inc = s;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const long long u = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += u;
    }
    if (lane == 31) wsum[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        const long long wv = wsum[lane];
        long long winc = wv;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const long long u = __shfl_up_sync(0xffffffffu, winc, o);
            if (lane >= o) winc += u;
        }
        wsum[lane] = winc - wv;
        if (lane == 31) *total = winc;
    }
    __syncthreads();
    return wsum[warp] + inc - s;
}

__global__ void __launch_bounds__(1024) scan_tiles_kernel(const long long* __restrict__ counts, int nq,
                                                          long long* __restrict__ offsets,
                                                          long long* __restrict__ tile_sums) {
    __shared__ long long wsum[32];
    __shared__ long long total;
    const int base = blockIdx.x * SCAN_TILE + 4 * threadIdx.x;
    long long v[4];
#pragma unroll
    for (int j = 0; j < 4; j++) v[j] = base + j < nq ? counts[base + j] : 0;
    long long run = block_exclusive_1024((v[0] + v[1]) + (v[2] + v[3]), wsum, &total);
#pragma unroll
    for (int j = 0; j < 4; j++) {
        if (base + j < nq) offsets[base + j] = run;
        run += v[j];
    }
    if (threadIdx.x == 0) tile_sums[blockIdx.x] = total;
}

__global__ void __launch_bounds__(1024) scan_sums_kernel(long long* __restrict__ tile_sums, int ntiles) {
    __shared__ long long wsum[32];
    __shared__ long long total;
    long long carry = 0;
    for (int base = 0; base < ntiles; base += 1024) {
        const int i = base + threadIdx.x;
        const long long v = i < ntiles ? tile_sums[i] : 0;
        const long long ex = block_exclusive_1024(v, wsum, &total);
        if (i < ntiles) tile_sums[i] = carry + ex;
        carry += total;
        __syncthreads();
    }
    if (threadIdx.x == 0) tile_sums[ntiles] = carry;
}

__global__ void __launch_bounds__(1024) scan_add_kernel(long long* __restrict__ offsets, int nq,
                                                        const long long* __restrict__ tile_sums, int ntiles) {
    const long long add = tile_sums[blockIdx.x];
    const int base = blockIdx.x * SCAN_TILE + 4 * threadIdx.x;
#pragma unroll
    for (int j = 0; j < 4; j++)
        if (base + j < nq) offsets[base + j] += add;
    if (blockIdx.x == 0 && threadIdx.x == 0) offsets[nq] = tile_sums[ntiles];
}

int scan_counts(const long long* counts, int nq, long long* offsets, Scratch& sc, cudaStream_t s) {
    const int ntiles = (nq + SCAN_TILE - 1) / SCAN_TILE;
    long long* tile_sums;
    PCNN_CUDA(sc.get(&tile_sums, (size_t)ntiles + 1));
    scan_tiles_kernel<<<ntiles, 1024, 0, s>>>(counts, nq, offsets, tile_sums);
    scan_sums_kernel<<<1, 1024, 0, s>>>(tile_sums, ntiles);
    scan_add_kernel<<<ntiles, 1024, 0, s>>>(offsets, nq, tile_sums, ntiles);
    return PCNN_OK;
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

static RadiusThr make_thr(double radius) {
    RadiusThr rt;
    rt.thr = radius_threshold(radius);
    rt.t_in = -1.f;
    rt.t_out = 0.f;
    if (rt.thr > 1e-30 && rt.thr < 1e30) {
        float a = (float)(rt.thr * (1.0 - 9.5367431640625e-07));
        if ((double)a > rt.thr * (1.0 - 9.5367431640625e-07)) a = std::nextafterf(a, 0.f);
        float b = (float)(rt.thr * (1.0 + 9.5367431640625e-07));
        if ((double)b < rt.thr * (1.0 + 9.5367431640625e-07)) b = std::nextafterf(b, INFINITY);
        rt.t_in = a;
        rt.t_out = b;
    }
    return rt;
}

// Which kernel serves a call (option radius_batch; the two passes decide independently, every
// kernel finds the same set):
//   0 = one warp per query, depth first (radius_warp_kernel), 2 = runs of neighbouring queries
//   share one depth-first walk (radius_batch_kernel), 3 = one query per lane, warp-wide writes
//   (radius_lockstep_kernel), 4 = one warp per query, dense stack walk (radius_stack_kernel),
//   5 = dense stack walk shared by a run of queries (radius_stack_batch_kernel),
//   1 = automatic.  Measured on the 1M aggregate (count + fill kernels, ms):
//       1M queries r=0.5:  0: 3.76+5.34   2: 2.05+4.23   3: 2.86+6.47   4: 3.34+4.90   5: 1.10+3.30
//       1M queries r=1.0:  0: 7.27+12.7   2: 3.86+16.7                  4: 5.91+11.5   5: 2.01+10.3
//       200k queries r=0.5: 0: 0.75+1.04  2: 0.82+1.54  3: 0.80+2.06    4: 0.65+0.94   5: 0.60+1.27
//   so: a batch about as dense as the cloud shares walks (5); sparse batches use 4.
enum { RM_WARP = 0, RM_BATCH = 2, RM_LOCKSTEP = 3, RM_STACK = 4, RM_STACK_BATCH = 5 };
static int radius_method(const pcnn_index* ix, int nq, double radius, bool fill, double hits_per_query) {
    int m = ix->opt_radius_batch;
#ifndef PCNN_STUDY
    if (m == RM_BATCH || m == RM_LOCKSTEP) m = 1;  // study variants: not in the shipped library
#endif
    if (m == 1) {
        const bool dense = ix->opt_sort_queries && nq > 1024 && 2 * (int64_t)nq >= ix->n;
        m = dense ? RM_STACK_BATCH : RM_STACK;
    }
    if (m == RM_BATCH && !(radius > 1e-12 && radius < 1e12)) m = RM_WARP;
    if (m == RM_STACK_BATCH && !(radius > 1e-12 && radius < 1e12)) m = RM_STACK;
    if ((m == RM_LOCKSTEP || m == RM_STACK || m == RM_STACK_BATCH) && ix->view.n0 >= RL_MAX_GROUPS) m = RM_WARP;  // 27-bit node fields
    return m;
}
// run extent handed to the run-batched kernels: radius / radius_run_div when the option is set,
// else negative = "that many mean group extents" (pcnn_radius_stack.cuh)
static float run_extent(const pcnn_index* ix, double radius, bool fill, bool stack) {
    if (ix->opt_radius_run_div > 0) return (float)(radius / ix->opt_radius_run_div);
    if (!stack) return (float)(radius / 4);
    return -0.01f * (float)(fill ? ix->opt_radius_run_mult_fill : ix->opt_radius_run_mult_count);
}

// The fill pass reuses the query ordering of the count pass before it (same stream, same query
// buffer, same kind of ordering).  Any permutation of the queries is correct -- the ordering
// only buys locality -- but it has to BE a permutation, so the one buffer is guarded: the
// handle's mutex is held from the key check to the launch that reads (or the copy that
// writes) it, and an event orders a writer on one stream after the last reader on another.
static void remember_perm(const pcnn_index* ix, const unsigned* perm, const float* q, int nq, int stride, int kind,
                          double radius, cudaStream_t s) {
    std::lock_guard<std::mutex> lk(ix->rperm_mu);
    ix->rperm_kind = -1;
    ix->rperm_hits_per_query = -1.0;
    if (!perm) return;
    if (!ix->rperm_ev && cudaEventCreateWithFlags(&ix->rperm_ev, cudaEventDisableTiming) != cudaSuccess) {
        ix->rperm_ev = nullptr;
        cudaGetLastError();
        return;
    }
    if ((size_t)nq > ix->rperm_cap) {
        if (ix->rperm) cudaFree(ix->rperm);  // waits for the device: no reader is left
        ix->rperm = nullptr;
        ix->rperm_cap = 0;
        if (cudaMalloc(&ix->rperm, (size_t)nq * sizeof(unsigned)) != cudaSuccess) { cudaGetLastError(); return; }
        ix->rperm_cap = (size_t)nq;
    } else if (ix->rperm_stream != s) {
        cudaStreamWaitEvent(s, ix->rperm_ev, 0);  // a no-op until the event has been recorded once
    }
    if (cudaMemcpyAsync(ix->rperm, perm, (size_t)nq * sizeof(unsigned), cudaMemcpyDeviceToDevice, s) != cudaSuccess) return;
    cudaEventRecord(ix->rperm_ev, s);
    ix->rperm_q = q; ix->rperm_nq = nq; ix->rperm_stride = stride; ix->rperm_kind = kind;
    ix->rperm_stream = s; ix->rperm_radius = radius;
}
// Returns the cached ordering with `lk` HELD when it matches (the caller launches its kernel, calls
// perm_used() and lets go), or nullptr with `lk` released.
static const unsigned* recall_perm(const pcnn_index* ix, const float* q, int nq, int stride, int kind, cudaStream_t s,
                                   std::unique_lock<std::mutex>& lk) {
    lk = std::unique_lock<std::mutex>(ix->rperm_mu);
    if (ix->rperm && ix->rperm_kind == kind && ix->rperm_q == q && ix->rperm_nq == nq && ix->rperm_stride == stride &&
        ix->rperm_stream == s)
        return ix->rperm;
    lk.unlock();
    return nullptr;
}
static void perm_used(const pcnn_index* ix, cudaStream_t s, std::unique_lock<std::mutex>& lk) {
    if (lk.owns_lock()) {
        if (ix->rperm_ev) cudaEventRecord(ix->rperm_ev, s);
        lk.unlock();
    }
}
void note_radius_total(const pcnn_index* ix, const float* q, int64_t nq, double radius, int64_t total) {
    std::lock_guard<std::mutex> lk(ix->rperm_mu);
    if (ix->rperm_q == q && ix->rperm_nq == (int)nq && ix->rperm_radius == radius && nq > 0)
        ix->rperm_hits_per_query = (double)total / (double)nq;
}
static double recall_hits(const pcnn_index* ix, const float* q, int nq, double radius) {
    std::lock_guard<std::mutex> lk(ix->rperm_mu);
    return (ix->rperm_q == q && ix->rperm_nq == nq && ix->rperm_radius == radius) ? ix->rperm_hits_per_query : -1.0;
}

int launch_radius_count(const pcnn_index* ix, const float* q, int64_t nq64, int stride, double radius,
                        int64_t* out_offsets, cudaStream_t s) {
    const int nq = (int)nq64;
    static_assert(sizeof(long long) == sizeof(int64_t), "int64 layout");
    if (nq == 0) {
        PCNN_CUDA(cudaMemsetAsync(out_offsets, 0, sizeof(int64_t), s));
        return PCNN_OK;
    }
    NvtxRange nvtx_(ix, "pcnn.radius.count");
    const WarpPlan wp = make_warp_plan(ix->view, ix->opt_radius_leaf_level);
    const RadiusThr rt = make_thr(radius);
    Scratch sc(s);
    unsigned* perm = nullptr;
    const int method = radius_method(ix, nq, radius, false, -1.0);
    const bool batch = method == RM_BATCH || method == RM_STACK_BATCH;
    if (ix->opt_sort_queries && nq > 1024 && ix->n > 0) {
        int rc = batch ? sort_queries_fine(ix, q, nq, stride, ix->opt_radius_sort_bits, &perm, sc, s)
                       : sort_queries(ix, q, nq, stride, &perm, sc, s);
        if (rc != PCNN_OK) return rc;
    }
    remember_perm(ix, perm, q, nq, stride, batch ? 1 : 0, radius, s);
    long long* counts;
    PCNN_CUDA(sc.get(&counts, (size_t)nq + 1));
    prof_begin(ix, s);
    if (method == RM_STACK_BATCH) {
        const int nwarps = (nq + 31) / 32;
        radius_stack_batch_kernel<false, false><<<(nwarps + RK_WARPS - 1) / RK_WARPS, RK_WARPS * 32, 0, s>>>(
            ix->view, q, nq, stride, perm, rt, run_extent(ix, radius, false, true), (float)radius, counts, nullptr,
            nullptr, nullptr, nullptr);
    } else if (method == RM_STACK) {
        radius_stack_kernel<false, false><<<(nq + RK_WARPS - 1) / RK_WARPS, RK_WARPS * 32, 0, s>>>(
            ix->view, q, nq, stride, perm, rt, counts, nullptr, nullptr, nullptr, nullptr);
#ifdef PCNN_STUDY
    } else if (method == RM_LOCKSTEP) {
        radius_lockstep_kernel<false, false><<<(nq + RL_THREADS - 1) / RL_THREADS, RL_THREADS, 0, s>>>(
            ix->view, q, nq, stride, perm, rt, counts, nullptr, nullptr, nullptr, nullptr);
    } else if (batch) {
        const int nwarps = (nq + 31) / 32;
        radius_batch_kernel<false, false><<<(nwarps + WQ_WARPS - 1) / WQ_WARPS, WQ_WARPS * 32, 0, s>>>(
            ix->view, wp, q, nq, stride, perm, rt, run_extent(ix, radius, false, false), counts, nullptr, nullptr,
            nullptr, nullptr);
#endif
    } else {
        radius_warp_kernel<false, false><<<(nq + WQ_WARPS - 1) / WQ_WARPS, WQ_WARPS * 32, 0, s>>>(
            ix->view, wp, q, nq, stride, perm, rt, counts, nullptr, nullptr, nullptr, nullptr);
    }
    prof_end(ix, s);
    NvtxRange nvtx_scan(ix, "pcnn.radius.scan");
    if (int rc = scan_counts(counts, nq, (long long*)out_offsets, sc, s)) return rc;
    PCNN_CUDA(cudaGetLastError());
    return PCNN_OK;
}

int launch_radius_fill(const pcnn_index* ix, const float* q, int64_t nq64, int stride, double radius,
                       const int64_t* offsets, int32_t* out_idx, float* out_dist, double* out_dist64,
                       cudaStream_t s) {
    const int nq = (int)nq64;
    if (nq == 0) return PCNN_OK;
    NvtxRange nvtx_(ix, "pcnn.radius.fill");
    const WarpPlan wp = make_warp_plan(ix->view, ix->opt_radius_leaf_level);
    const RadiusThr rt = make_thr(radius);
    Scratch sc(s);
    unsigned* perm = nullptr;
    const int method = radius_method(ix, nq, radius, true, recall_hits(ix, q, nq, radius));
    const bool batch = method == RM_BATCH || method == RM_STACK_BATCH;
    std::unique_lock<std::mutex> perm_lock;
    if (ix->opt_sort_queries && nq > 1024 && ix->n > 0) {
        perm = const_cast<unsigned*>(recall_perm(ix, q, nq, stride, batch ? 1 : 0, s, perm_lock));
        if (!perm) {
            int rc = batch ? sort_queries_fine(ix, q, nq, stride, ix->opt_radius_sort_bits, &perm, sc, s)
                           : sort_queries(ix, q, nq, stride, &perm, sc, s);
            if (rc != PCNN_OK) return rc;
        }
    }
    prof_begin(ix, s);
    if (method == RM_STACK_BATCH) {
        const int nwarps = (nq + 31) / 32;
        const int bgrid = (nwarps + RK_WARPS - 1) / RK_WARPS;
        const float ext = run_extent(ix, radius, true, true);
        if (out_dist64)
            radius_stack_batch_kernel<true, true><<<bgrid, RK_WARPS * 32, 0, s>>>(
                ix->view, q, nq, stride, perm, rt, ext, (float)radius, nullptr, (const long long*)offsets, out_idx, out_dist,
                out_dist64);
        else
            radius_stack_batch_kernel<true, false><<<bgrid, RK_WARPS * 32, 0, s>>>(
                ix->view, q, nq, stride, perm, rt, ext, (float)radius, nullptr, (const long long*)offsets, out_idx, out_dist, nullptr);
        prof_end(ix, s);
        perm_used(ix, s, perm_lock);
        PCNN_CUDA(cudaGetLastError());
        return PCNN_OK;
    }
    if (method == RM_STACK) {
        const int sgrid = (nq + RK_WARPS - 1) / RK_WARPS;
        if (out_dist64)
            radius_stack_kernel<true, true><<<sgrid, RK_WARPS * 32, 0, s>>>(
                ix->view, q, nq, stride, perm, rt, nullptr, (const long long*)offsets, out_idx, out_dist, out_dist64);
        else
            radius_stack_kernel<true, false><<<sgrid, RK_WARPS * 32, 0, s>>>(
                ix->view, q, nq, stride, perm, rt, nullptr, (const long long*)offsets, out_idx, out_dist, nullptr);
        prof_end(ix, s);
        perm_used(ix, s, perm_lock);
        PCNN_CUDA(cudaGetLastError());
        return PCNN_OK;
    }
#ifdef PCNN_STUDY
    if (method == RM_LOCKSTEP) {
        const int lgrid = (nq + RL_THREADS - 1) / RL_THREADS;
        if (out_dist64)
            radius_lockstep_kernel<true, true><<<lgrid, RL_THREADS, 0, s>>>(
                ix->view, q, nq, stride, perm, rt, nullptr, (const long long*)offsets, out_idx, out_dist, out_dist64);
        else
            radius_lockstep_kernel<true, false><<<lgrid, RL_THREADS, 0, s>>>(
                ix->view, q, nq, stride, perm, rt, nullptr, (const long long*)offsets, out_idx, out_dist, nullptr);
        prof_end(ix, s);
        perm_used(ix, s, perm_lock);
        PCNN_CUDA(cudaGetLastError());
        return PCNN_OK;
    }
    if (batch) {
        const int nwarps = (nq + 31) / 32;
        const int bgrid = (nwarps + WQ_WARPS - 1) / WQ_WARPS;
        const float ext = run_extent(ix, radius, true, false);
        if (out_dist64)
            radius_batch_kernel<true, true><<<bgrid, WQ_WARPS * 32, 0, s>>>(
                ix->view, wp, q, nq, stride, perm, rt, ext, nullptr, (const long long*)offsets, out_idx, out_dist, out_dist64);
        else
            radius_batch_kernel<true, false><<<bgrid, WQ_WARPS * 32, 0, s>>>(
                ix->view, wp, q, nq, stride, perm, rt, ext, nullptr, (const long long*)offsets, out_idx, out_dist, nullptr);
        prof_end(ix, s);
        perm_used(ix, s, perm_lock);
        PCNN_CUDA(cudaGetLastError());
        return PCNN_OK;
    }
#endif  // PCNN_STUDY
    const int grid = (nq + WQ_WARPS - 1) / WQ_WARPS;
    if (out_dist64)
        radius_warp_kernel<true, true><<<grid, WQ_WARPS * 32, 0, s>>>(
            ix->view, wp, q, nq, stride, perm, rt, nullptr, (const long long*)offsets, out_idx, out_dist, out_dist64);
    else
        radius_warp_kernel<true, false><<<grid, WQ_WARPS * 32, 0, s>>>(
            ix->view, wp, q, nq, stride, perm, rt, nullptr, (const long long*)offsets, out_idx, out_dist, nullptr);
    prof_end(ix, s);
        perm_used(ix, s, perm_lock);
    PCNN_CUDA(cudaGetLastError());
    return PCNN_OK;
}

}  // namespace pcnn

// ------------------------------------------------------------------ allocation-free count pass
// pcnn_radius_count_noalloc (include/pcnn.h): the count pass of the radius search for callers that
// cannot allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).  Same
// kernels as pcnn_radius_count on a sparse batch in the caller's query order (radius_stack_kernel,
// then the tiled scan); what differs is where the scratch lives: the per-query counts and the
// scan's tile sums sit in ONE file-scope device array instead of the stream-ordered pool.  Every
// GPU that loads the library holds the array until exit.  All calls share it and it keeps the last
// call's contents; every word this call reads it has written first (counts[0, nq) by the count
// kernel, the tile sums by the scan), so nothing needs clearing -- but two calls that overlap on
// the GPU (different streams, no ordering between them) race on it.
namespace pcnn {
constexpr size_t NOALLOC_COUNT_BYTES = ((size_t)PCNN_NOALLOC_MAX_QUERIES + 1) * sizeof(long long);
constexpr size_t NOALLOC_TILES = ((size_t)PCNN_NOALLOC_MAX_QUERIES + SCAN_TILE - 1) / SCAN_TILE + 1;
constexpr size_t NOALLOC_BYTES = ((NOALLOC_COUNT_BYTES + 255) / 256) * 256 + NOALLOC_TILES * sizeof(long long);
static_assert(NOALLOC_BYTES <= (size_t)PCNN_NOALLOC_SCRATCH_BYTES, "pcnn.h states the size of the scratch array");
}  // namespace pcnn
static __device__ __align__(256) unsigned char g_noalloc_scratch[PCNN_NOALLOC_SCRATCH_BYTES];

namespace {
// In host code `g_noalloc_scratch` is a host-side stand-in, not a device address: the address a
// kernel may use comes from cudaGetSymbolAddress, once per device, with that device current.
constexpr int NOALLOC_MAX_DEVICES = 64;
std::mutex g_noalloc_mu;
void* g_noalloc_ptr[NOALLOC_MAX_DEVICES] = {};

int noalloc_scratch(int device, unsigned char** out) {
    if (device < 0 || device >= NOALLOC_MAX_DEVICES) {
        pcnn::set_error("pcnn_radius_count_noalloc: device %d out of range", device);
        return PCNN_ERR_INVALID;
    }
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (!g_noalloc_ptr[device]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);  // the caller made `device` current
        if (e != cudaSuccess || !p) {
            pcnn::set_error("pcnn_radius_count_noalloc: cudaGetSymbolAddress failed on device %d: %s", device,
                            cudaGetErrorString(e));
            return PCNN_ERR_CUDA;
        }
        g_noalloc_ptr[device] = p;
    }
    *out = static_cast<unsigned char*>(g_noalloc_ptr[device]);
    return PCNN_OK;
}
}  // namespace

extern "C" int pcnn_radius_count_noalloc(const pcnn_index* index, const float* q_dev, int64_t nq, int stride,
                                         double radius, int64_t* out_offsets_dev, pcnn_stream stream) {
    using namespace pcnn;
    PCNN_CHECK_ARG(index != nullptr, "pcnn_radius_count_noalloc: index is NULL");
    PCNN_CHECK_ARG(nq >= 0, "pcnn_radius_count_noalloc: nq=%lld is negative", (long long)nq);
    PCNN_CHECK_ARG(stride >= 3, "pcnn_radius_count_noalloc: stride must be >= 3 floats (got %d)", stride);
    PCNN_CHECK_ARG(nq == 0 || q_dev != nullptr, "pcnn_radius_count_noalloc: query pointer is NULL");
    PCNN_CHECK_ARG(out_offsets_dev != nullptr, "pcnn_radius_count_noalloc: out_offsets is NULL");
    PCNN_CHECK_ARG(radius == radius, "pcnn_radius_count_noalloc: radius is NaN");
    if (nq > (int64_t)PCNN_NOALLOC_MAX_QUERIES) {
        set_error("pcnn_radius_count_noalloc: nq=%lld exceeds PCNN_NOALLOC_MAX_QUERIES=%d (the scratch is a fixed "
                  "%d-byte device array); split the batch or use pcnn_radius_count",
                  (long long)nq, PCNN_NOALLOC_MAX_QUERIES, PCNN_NOALLOC_SCRATCH_BYTES);
        return PCNN_ERR_UNSUPPORTED;
    }
    if (index->view.n0 >= RL_MAX_GROUPS) {
        set_error("pcnn_radius_count_noalloc: clouds of 2^30 points and more are not served by this entry point");
        return PCNN_ERR_UNSUPPORTED;
    }
    DeviceGuard g(index->device);
    if (!g.ok) {
        set_error("pcnn_radius_count_noalloc: cannot select CUDA device %d (no CPU fallback exists)", index->device);
        return PCNN_ERR_CUDA;
    }
    cudaStream_t s = (cudaStream_t)stream;
    if (nq == 0) {
        PCNN_CUDA(cudaMemsetAsync(out_offsets_dev, 0, sizeof(int64_t), s));
        return PCNN_OK;
    }
    unsigned char* base = nullptr;
    if (int rc = noalloc_scratch(index->device, &base)) return rc;
    long long* counts = reinterpret_cast<long long*>(base);
    long long* tile_sums = reinterpret_cast<long long*>(base + ((NOALLOC_COUNT_BYTES + 255) / 256) * 256);
    const int n = (int)nq;
    const RadiusThr rt = make_thr(radius);
    radius_stack_kernel<false, false><<<(n + RK_WARPS - 1) / RK_WARPS, RK_WARPS * 32, 0, s>>>(
        index->view, q_dev, n, stride, nullptr, rt, counts, nullptr, nullptr, nullptr, nullptr);
    const int ntiles = (n + SCAN_TILE - 1) / SCAN_TILE;
    scan_tiles_kernel<<<ntiles, 1024, 0, s>>>(counts, n, (long long*)out_offsets_dev, tile_sums);
    scan_sums_kernel<<<1, 1024, 0, s>>>(tile_sums, ntiles);
    scan_add_kernel<<<ntiles, 1024, 0, s>>>((long long*)out_offsets_dev, n, tile_sums, ntiles);
    PCNN_CUDA(cudaGetLastError());
    return PCNN_OK;
}

