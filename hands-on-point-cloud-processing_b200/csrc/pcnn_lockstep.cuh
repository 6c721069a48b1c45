// pcnn_lockstep.cuh -- exact float32 1-NN for one query per lane, the 32 walks of a warp run
// in lock step (device only).
//
// Replaces nanoflann's k = 1 search as the ICP loop calls it
// (Homework9/hw9/src/registration.cpp:772-785, :925-941; nanoflann.hpp:1347-1410): same key
// (float32, ((dx*dx)+dy*dy)+dz*dz with dx = q - p, nanoflann.hpp:383-408), lowest original
// index wins an exact tie.  Same walk as knn_dfs_kernel (pcnn_knn.cu): nearest child first,
// the farther siblings still owed a visit kept in one 32-bit mask, no stack, no recursion;
// rounds of predicated steps instead of data-dependent loops, so a warp runs at the pace of
// its longest walk with all its loads in flight together.  A caller that iterates (ICP) passes
// the group its previous answer was found in: the first bound is then already (nearly) final;
// and a caller that discards neighbours beyond a gate passes the gate as the initial bound.
//
// Every lane of the warp must call this together (tail lanes shadow a valid query).
#pragma once
#include "pcnn_search.cuh"

namespace pcnn {

constexpr int LS_QDEPTH = 4;  // queued groups per lane
constexpr int LS_WALK = 8;    // steps per round

template <class BEST>
__device__ __forceinline__ void nn1_scan_group(const IndexView& iv, int g, float qx, float qy, float qz,
                                               BEST& best) {
    const int first = g << PCNN_GROUP_LOG;
    const int cnt = (iv.n - first) < PCNN_GROUP ? (iv.n - first) : PCNN_GROUP;
#pragma unroll
    for (int j = 0; j < PCNN_GROUP; j++) {
        if (j < cnt) {
            const float4 p = __ldg(iv.pts + first + j);
            const float d = key32(p.x, p.y, p.z, qx, qy, qz);
            const int id = __float_as_int(p.w);
            best.offer(d, id, first + j);
        }
    }
}

// s_qg / s_ql: shared-memory queues [LS_QDEPTH][THREADS], column `tid` belongs to this lane.
template <int THREADS, class BEST>
__device__ __forceinline__ void nn1_lockstep(const IndexView& iv, float qx, float qy, float qz, int seed_group,
                                             float bound, BEST& best, int (*s_qg)[THREADS],
                                             float (*s_ql)[THREADS], int tid) {
    const unsigned full = 0xffffffffu;
    const float INFf = __int_as_float(0x7f800000);
    // `bound`: the caller only wants a neighbour with key < bound (ICP's gate on the squared
    // distance, registration.cpp:783 / :936); farther ones may be reported as "none" (pos = -1)
    best.start(bound);
    if (iv.n <= 0) return;
    const unsigned long long qcode = morton63(qx, qy, qz, iv.cube, iv.cube[3]);
    if (seed_group < 0 || seed_group >= iv.n0) seed_group = home_group(iv, qcode);
    nn1_scan_group(iv, seed_group, qx, qy, qz, best);
    const int top = iv.nlev - 1;
    // start as low in the tree as the seed bound allows; with a warm start that is a few levels.
    // (Measured alternative: start AT the seed group with the sibling subtrees of its ancestors tested
    // up front by independent loads and only climb.  Slower both with many searches in flight, 570 vs
    // 500 us per 1M-point ICP iteration, and with a few hundred, 148 vs 122 us: every sibling that
    // passes costs a pop and a re-test, where the descent prunes most of them on the way down.)
    int l = walk_start_level(iv, seed_group, qx, qy, qz, best.bound());
    int n = seed_group >> l;
    unsigned pending = 0;
    int mode = 0;  // 0 = examine the children of (l, n), 1 = backtrack
    int cnt = 0;
    bool done = top == 0;
    while (true) {
#pragma unroll 1
        for (int s = 0; s < LS_WALK; s++) {
            const bool can = !done && cnt <= LS_QDEPTH - 2;
            if (!__any_sync(full, can)) break;
            if (can) {
                int L, P, far = 0;
                if (mode == 0) {
                    L = l - 1;
                    P = 2 * n;
                } else if (pending == 0) {
                    done = true;
                    L = 0;
                    P = 0;
                } else {
                    const int lp = __ffs((int)pending) - 1;
                    pending &= pending - 1;
                    far = (n >> (lp - l)) ^ 1;
                    l = lp;
                    n = far;
                    L = lp;
                    P = far & ~1;
                }
                if (!done) {
                    const int cl = (int)(((unsigned)iv.n0 + (1u << L) - 1u) >> L);
                    const float4* b = iv.box + 2 * ((size_t)iv.lev_off[L] + P);
                    const float4 alo = __ldg(b), ahi = __ldg(b + 1);
                    const float lbA = boxlb32(alo, ahi, qx, qy, qz);
                    float lbB = INFf;
                    const bool hasB = P + 1 < cl;
                    if (hasB) {
                        const float4 blo = __ldg(b + 2), bhi = __ldg(b + 3);
                        lbB = boxlb32(blo, bhi, qx, qy, qz);
                    }
                    if (mode == 0) {
                        bool first1 = lbB < lbA;
                        if (hasB && lbB == lbA) first1 = qcode >= __ldg(iv.gcode + ((size_t)(P + 1) << L));
                        const float lbn = first1 ? lbB : lbA, lbf = first1 ? lbA : lbB;
                        const int near = P + (first1 ? 1 : 0);
                        // strict test: a box at exactly the best key may hold a lower index
                        const bool passN = !(lbn > best.bound());
                        const bool passF = !(lbf > best.bound()) && lbf < INFf;
                        if (L == 0) {
                            if (passN && near != seed_group) { s_qg[cnt][tid] = near; s_ql[cnt][tid] = lbn; cnt++; }
                            if (passF && (near ^ 1) != seed_group) { s_qg[cnt][tid] = near ^ 1; s_ql[cnt][tid] = lbf; cnt++; }
                            mode = 1;
                        } else if (passN) {
                            if (passF) pending |= 1u << L;
                            l = L;
                            n = near;
                        } else {
                            mode = 1;
                        }
                    } else {
                        const float lbv = (far & 1) ? lbB : lbA;
                        if (!(lbv > best.bound())) mode = 0;
                    }
                }
            }
        }
        {
            const int maxc = __reduce_max_sync(full, cnt);
            for (int r = 0; r < maxc; r++)
                if (r < cnt && !(s_ql[r][tid] > best.bound())) nn1_scan_group(iv, s_qg[r][tid], qx, qy, qz, best);
            cnt = 0;
        }
        if (!__any_sync(full, !done)) break;
    }
}

}  // namespace pcnn
