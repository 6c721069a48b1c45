// pcnn_radius_lockstep.cuh -- fixed-radius search, one query per lane for the WALK and the
// TESTS, the whole warp per query for the WRITES (device only; included by pcnn_radius.cu).
//
// Same contract as the other radius kernels (kdtree.hpp:367-402, octree.hpp:189-237,
// resultSet.hpp:130-140): every point with sqrt(d2) <= radius, float32 pre-tests around the
// exact double test (RadiusThr), the octree's `contains` fast path for boxes wholly inside the
// ball.  What differs is who does what:
//
//   walk   Each lane walks the hierarchy for ITS query, left to right, without a stack: a step
//          loads the two children of a straddling node and classifies both (prune / contained /
//          straddling); the right child, when the left one has to be entered first, is
//          remembered in two 32-bit masks by level (owed a visit as straddling / as contained),
//          so no box is ever loaded twice.  The 32 walks of a warp run in lock step, rounds of
//          RL_WALK predicated steps, as in knn_dfs_kernel.
//   tests  A straddling 8-point group is queued and the queues are drained together: every lane
//          tests the eight points of its own next group -- dense, no idle lanes, no ballots.
//   items  The walk does not write hits.  It appends ITEMS to a per-lane list in shared memory:
//          (group, 8-bit hit mask) for a tested or contained group, (first group, level) for a
//          contained subtree.  Classification is per query and per 8-point group, so a query
//          tests ~500 points for ~900 hits at r = 0.5 m on the 1M aggregate (the warp-per-query
//          kernels, with 32-point leaves or run-wide boxes, test 1100-4300).
//   write  When no lane can go on (all lists full or all walks done) the warp turns to one
//          query at a time: lane i takes item i, a warp scan of the item sizes gives every
//          item its position in the CSR row, tested groups are packed four to a step (lane ->
//          item lane/8, point lane%8) and contained subtrees go out in 32-point chunks.  Row
//          writes are contiguous; positions come from the scan, so rows stay in the index's
//          Morton order.
//
// The count pass runs the same walk and tests without lists: it only adds up sizes.
#pragma once
#include "pcnn_internal.cuh"

namespace pcnn {

constexpr int RL_WARPS = 4;
constexpr int RL_THREADS = RL_WARPS * 32;
constexpr int RL_CAP = 32;    // items per lane between two write phases (one item per lane when read back)
constexpr int RL_WALK = 8;    // steps per round
constexpr unsigned RL_GID_MASK = 0x07ffffffu;  // item word: first group (27 bits) | (level + 1) << 27, 0 = single group
constexpr int RL_MAX_GROUPS = 1 << 27;

__device__ __forceinline__ float sqrt_approx32(float x) {
    float r;
    asm("sqrt.approx.f32 %0, %1;" : "=f"(r) : "f"(x));  // relative error <= 2^-23, subnormals kept
    return r;
}

struct RadiusLane {
    float qxf, qyf, qzf;
    double thr;
    float t_in, t_out;
    bool pre;

    // 0 = out of reach, 1 = straddles the sphere, 2 = wholly inside the ball
    __device__ __forceinline__ int classify(const float4& lo, const float4& hi) const {
        if (pre) {
            const float ax = lo.x - qxf, bx = qxf - hi.x;
            const float ay = lo.y - qyf, by = qyf - hi.y;
            const float az = lo.z - qzf, bz = qzf - hi.z;
            const float gx = fmaxf(fmaxf(ax, bx), 0.f), gy = fmaxf(fmaxf(ay, by), 0.f), gz = fmaxf(fmaxf(az, bz), 0.f);
            if (fmaf(gx, gx, fmaf(gy, gy, gz * gz)) > t_out) return 0;
            const float fx = -fminf(ax, bx), fy = -fminf(ay, by), fz = -fminf(az, bz);
            return fmaf(fx, fx, fmaf(fy, fy, fz * fz)) <= t_in ? 2 : 1;
        }
        const double qx = (double)qxf, qy = (double)qyf, qz = (double)qzf;
        if (!(boxlb64(lo, hi, qx, qy, qz) <= thr)) return 0;
        return boxub64(lo, hi, qx, qy, qz) <= thr ? 2 : 1;
    }

    __device__ __forceinline__ bool hit(const float4& p) const {
        if (pre) {
            const float dx = p.x - qxf, dy = p.y - qyf, dz = p.z - qzf;
            const float s = fmaf(dx, dx, fmaf(dy, dy, dz * dz));
            if (s <= t_in) return true;
            if (s > t_out) return false;
        }
        return key64(p.x, p.y, p.z, (double)qxf, (double)qyf, (double)qzf) <= thr;
    }

    // hit mask of the (up to) eight points of group g.  No branch per point: eight float32 keys,
    // two compare masks, and only the sliver between t_in and t_out goes to the double test.
    __device__ __forceinline__ unsigned test_group(const IndexView& iv, int g) const {
        const int first = g << PCNN_GROUP_LOG;
        const float4* pp = iv.pts + first;
        if (pre && first + PCNN_GROUP <= iv.n) {
            float4 p[PCNN_GROUP];
#pragma unroll
            for (int j = 0; j < PCNN_GROUP; j++) p[j] = __ldg(pp + j);
            unsigned in = 0, amb = 0;
#pragma unroll
            for (int j = 0; j < PCNN_GROUP; j++) {
                const float dx = p[j].x - qxf, dy = p[j].y - qyf, dz = p[j].z - qzf;
                const float s = fmaf(dx, dx, fmaf(dy, dy, dz * dz));
                in |= (s <= t_in ? 1u : 0u) << j;
                amb |= (s <= t_out ? 1u : 0u) << j;
            }
            amb &= ~in;
            while (amb) {  // rare
                const int j = __ffs((int)amb) - 1;
                amb &= amb - 1;
                const float4 pj = __ldg(pp + j);
                if (key64(pj.x, pj.y, pj.z, (double)qxf, (double)qyf, (double)qzf) <= thr) in |= 1u << j;
            }
            return in;
        }
        unsigned m = 0;
        for (int j = 0; j < PCNN_GROUP && first + j < iv.n; j++) m |= (hit(__ldg(pp + j)) ? 1u : 0u) << j;
        return m;
    }
};

template <bool FILL, bool D64>
__global__ void __launch_bounds__(RL_THREADS)
radius_lockstep_kernel(IndexView iv, const float* __restrict__ q, int nq, int stride,
                       const unsigned* __restrict__ perm, RadiusThr rt, long long* __restrict__ counts,
                       const long long* __restrict__ offsets, int32_t* __restrict__ out_idx,
                       float* __restrict__ out_dist, double* __restrict__ out_dist64) {
    constexpr int LW = FILL ? RL_WARPS : 1;  // list storage only exists in the fill pass
    __shared__ unsigned s_item[LW][32][RL_CAP + 1];
    __shared__ unsigned char s_imask[LW][32][RL_CAP + 4];
    __shared__ int s_sc_g[LW][32], s_sc_o[LW][32];
    __shared__ unsigned char s_sc_m[LW][32];
    __shared__ int s_tq[FILL ? 1 : RL_CAP][FILL ? 1 : RL_THREADS];
    __shared__ int s_loff[PCNN_MAX_LEVELS];
    const unsigned full = 0xffffffffu;
    const int tid = threadIdx.x, lane = tid & 31, w = FILL ? (tid >> 5) : 0;
    if (tid < PCNN_MAX_LEVELS) s_loff[tid] = tid < iv.nlev ? 2 * iv.lev_off[tid] : 0;
    __syncthreads();
    const int slot0 = (blockIdx.x * RL_WARPS + (tid >> 5)) * 32;
    if (slot0 >= nq) return;  // warp-uniform
    const bool haveq = slot0 + lane < nq;
    const int qs = haveq ? slot0 + lane : slot0;
    const int qi = perm ? (int)perm[qs] : qs;
    const float* qp = q + (size_t)qi * stride;
    RadiusLane rl;
    rl.qxf = qp[0]; rl.qyf = qp[1]; rl.qzf = qp[2];
    rl.thr = rt.thr; rl.t_in = rt.t_in; rl.t_out = rt.t_out; rl.pre = rt.t_in >= 0.f;
    const unsigned lt = (1u << lane) - 1u;
    long long run = (FILL && haveq) ? offsets[qi] : 0;  // fill: row cursor; count: the count

    const int top = iv.nlev - 1;
    int l = top, n = 0;
    unsigned pendP = 0, pendC = 0;  // by level: right sibling owed a visit as straddling / contained
    bool need_pop = false;
    bool done = !haveq || iv.n <= 0 || !(rt.thr >= 0.0);
    int cnt = 0;            // fill: items in my list; count: queued groups
    unsigned pendmask = 0;  // fill: list slots whose hit mask is still to be computed

    auto level_nodes = [&](int L) -> int { return (int)(((unsigned)iv.n0 + (1u << L) - 1u) >> L); };
    auto add_contained = [&](int L, int node) {
        const int g0 = node << L;
        if constexpr (FILL) {
            if (L == 0) {
                const int left = iv.n - (g0 << PCNN_GROUP_LOG);
                s_item[w][lane][cnt] = (unsigned)g0;
                s_imask[w][lane][cnt] = (unsigned char)(left >= PCNN_GROUP ? 0xffu : ((1u << left) - 1u));
            } else {
                s_item[w][lane][cnt] = (unsigned)g0 | ((unsigned)(L + 1) << 27);
            }
            cnt++;
        } else {
            const int p0 = g0 << PCNN_GROUP_LOG;
            const int p1 = min((g0 + (1 << L)) << PCNN_GROUP_LOG, iv.n);
            run += p1 - p0;
        }
    };
    auto queue_group = [&](int g) {
        if constexpr (FILL) {
            s_item[w][lane][cnt] = (unsigned)g;
            pendmask |= 1u << cnt;
            cnt++;
        } else {
            s_tq[cnt][tid] = g;
            cnt++;
        }
    };
    if (!done && top == 0) {  // a single group: no hierarchy to walk
        queue_group(0);
        done = true;
    }

    while (true) {
        // ---------------- walk: RL_WALK lock-step steps
#pragma unroll 1
        for (int s = 0; s < RL_WALK; s++) {
            const bool can = !done && cnt <= RL_CAP - 2;
            if (!__any_sync(full, can)) break;
            if (can) {
                bool examine = true;
                if (need_pop) {
                    const unsigned pend = pendP | pendC;
                    if (pend == 0) {
                        done = true;
                        examine = false;
                    } else {
                        const int lp = __ffs((int)pend) - 1;
                        const unsigned bit = 1u << lp;
                        n = (n >> (lp - l)) | 1;
                        l = lp;
                        if (pendC & bit) {
                            pendC &= ~bit;
                            add_contained(l, n);
                            examine = false;
                        } else {
                            pendP &= ~bit;
                            need_pop = false;
                        }
                    }
                }
                if (examine) {
                    const int L = l - 1, A = 2 * n;
                    const float4* b = iv.box + (size_t)s_loff[L] + 2 * (size_t)A;
                    const bool hasB = A + 1 < level_nodes(L);
                    const float4 alo = __ldg(b), ahi = __ldg(b + 1);
                    float4 blo = alo, bhi = ahi;
                    if (hasB) { blo = __ldg(b + 2); bhi = __ldg(b + 3); }
                    const int cA = rl.classify(alo, ahi);
                    const int cB = hasB ? rl.classify(blo, bhi) : 0;
                    if (cA == 2) add_contained(L, A);
                    else if (cA == 1 && L == 0) queue_group(A);
                    if (cA == 1 && L > 0) {
                        if (cB == 2) pendC |= 1u << L;
                        else if (cB == 1) pendP |= 1u << L;
                        l = L;
                        n = A;
                    } else {
                        if (cB == 2) add_contained(L, A + 1);
                        else if (cB == 1 && L == 0) queue_group(A + 1);
                        if (cB == 1 && L > 0) { l = L; n = A + 1; }
                        else need_pop = true;
                    }
                }
            }
        }
        // no lane can take another step (lists full or walks over): tests, then writes
        if (__any_sync(full, !done && cnt <= RL_CAP - 2)) continue;
        // ---------------- tests: every lane scans its own queued groups
        if constexpr (FILL) {
            const int maxp = __reduce_max_sync(full, __popc(pendmask));
            for (int r = 0; r < maxp; r++) {
                if (pendmask) {
                    const int slot = __ffs((int)pendmask) - 1;
                    pendmask &= pendmask - 1;
                    s_imask[w][lane][slot] = (unsigned char)rl.test_group(iv, (int)s_item[w][lane][slot]);
                }
            }
        } else {
            const int maxq = __reduce_max_sync(full, cnt);
            for (int r = 0; r < maxq; r++)
                if (r < cnt) run += __popc(rl.test_group(iv, s_tq[r][tid]));
            cnt = 0;
        }
        const bool more = __any_sync(full, !done);
        if constexpr (FILL) {
            // ---------------- write
            {
                __syncwarp();
                unsigned have = __ballot_sync(full, cnt > 0);
                while (have) {
                    const int c = __ffs((int)have) - 1;
                    have &= have - 1;
                    const int nc = __shfl_sync(full, cnt, c);
                    const long long ro = __shfl_sync(full, run, c);
                    const float cx = __shfl_sync(full, rl.qxf, c), cy = __shfl_sync(full, rl.qyf, c),
                                cz = __shfl_sync(full, rl.qzf, c);
                    const bool valid = lane < nc;
                    const unsigned word = valid ? s_item[w][c][lane] : 0u;
                    const unsigned m8 = valid ? (unsigned)s_imask[w][c][lane] : 0u;
                    const int kind = (int)(word >> 27);
                    const int g0 = (int)(word & RL_GID_MASK);
                    int size = 0;
                    if (valid) {
                        if (kind == 0) size = __popc(m8);
                        else size = min((g0 + (1 << (kind - 1))) << PCNN_GROUP_LOG, iv.n) - (g0 << PCNN_GROUP_LOG);
                    }
                    int incl = size;
#pragma unroll
                    for (int o = 1; o < 32; o <<= 1) {
                        const int u = __shfl_up_sync(full, incl, o);
                        if (lane >= o) incl += u;
                    }
                    const int excl = incl - size;
                    const int total = __shfl_sync(full, incl, 31);
                    auto put = [&](long long dst, const float4& p) {
                        out_idx[dst] = __float_as_int(p.w);
                        if constexpr (D64) {
                            const double dist = sqrt(key64(p.x, p.y, p.z, (double)cx, (double)cy, (double)cz));
                            if (out_dist) out_dist[dst] = (float)dist;
                            out_dist64[dst] = dist;
                        } else if (out_dist) {
                            // float32 distances: within 4e-7 relative of the exact value (contract: 1e-6)
                            if (rl.pre) {
                                const float dx = p.x - cx, dy = p.y - cy, dz = p.z - cz;
                                out_dist[dst] = sqrt_approx32(fmaf(dx, dx, fmaf(dy, dy, dz * dz)));
                            } else {
                                out_dist[dst] = (float)sqrt(key64(p.x, p.y, p.z, (double)cx, (double)cy, (double)cz));
                            }
                        }
                    };
                    // single groups, packed four to a step
                    const bool small = valid && kind == 0 && m8 != 0;
                    const unsigned sm = __ballot_sync(full, small);
                    if (small) {
                        const int pos = __popc(sm & lt);
                        s_sc_g[w][pos] = g0;
                        s_sc_o[w][pos] = excl;
                        s_sc_m[w][pos] = (unsigned char)m8;
                    }
                    __syncwarp();
                    const int ns = __popc(sm);
                    const int j = lane & 7;
                    for (int s0 = 0; s0 < ns; s0 += 4) {
                        const int it = s0 + (lane >> 3);
                        if (it < ns) {
                            const unsigned mb = s_sc_m[w][it];
                            if ((mb >> j) & 1u) {
                                const float4 p = __ldg(iv.pts + ((size_t)s_sc_g[w][it] << PCNN_GROUP_LOG) + j);
                                put(ro + s_sc_o[w][it] + __popc(mb & ((1u << j) - 1u)), p);
                            }
                        }
                    }
                    // contained subtrees, 32 points at a time
                    unsigned bm = __ballot_sync(full, valid && kind != 0);
                    while (bm) {
                        const int bi = __ffs((int)bm) - 1;
                        bm &= bm - 1;
                        const int p0 = __shfl_sync(full, g0, bi) << PCNN_GROUP_LOG;
                        const int len = __shfl_sync(full, size, bi);
                        const long long o = ro + __shfl_sync(full, excl, bi);
                        for (int pb = lane; pb < len; pb += 32) put(o + pb, __ldg(iv.pts + p0 + pb));
                    }
                    if (lane == c) run += total;
                    __syncwarp();
                }
                cnt = 0;
            }
        }
        if (!more) break;
    }
    if (!FILL && haveq) counts[qi] = run;
}

}  // namespace pcnn
