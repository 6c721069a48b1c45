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

#ifndef PCNN_LS_QDEPTH
#define PCNN_LS_QDEPTH 4
#endif
#ifndef PCNN_LS_WALK
#define PCNN_LS_WALK 16  /* 6 / 8 / 12 / 16 steps: 1M p2plane 178 / 175 / 174 / 173 us per iteration */
#endif
constexpr int LS_QDEPTH = PCNN_LS_QDEPTH;  // queued groups per lane
constexpr int LS_WALK = PCNN_LS_WALK;    // steps per round

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
    // (Also measured and dropped here: knn_dfs_kernel's remembered bounds of deferred siblings, which turn
    // the backtracking step into a pop-and-compare -- standalone 1-NN 0.501 -> 0.487 ms per 1M queries, but
    // the ICP loops lose: 1M p2plane 177 -> 182 us per iteration, 100k 30.8 -> 32.3, with 8 KB more shared
    // memory per block and a data-dependent loop inside the step.)
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

// ------------------------------------------------------------------ one query, the whole warp
// nn1_coop: the same exact float32 1-NN for ONE query with all 32 lanes working on it.  The lock-step
// walk above runs at the pace of its longest walk; a warp that holds a single real query (what a
// nearly converged ICP iteration is made of: a few uncertified points, one per block) then spends
// 30-60 us walking alone, with the kernel waiting for it (ncu, converged 1M-point iteration: the SMs
// are active 21 us of the kernel's 49).  Here the warp keeps a stack of nodes in shared memory, pops
// up to 16 and tests their 32 children at once (as radius_stack_kernel), and scans 16 queued groups
// (128 points, four independent loads per lane) per step.  Every lane keeps a PARTIAL result over the
// points it happened to test and prunes with the warp-wide minimum of the partial pruning keys, which
// is never below the pruning key of the merged result (min of reach(key) is reach of the min; a lane's
// third-smallest key is at least the third-smallest of the union): nothing is pruned that the merged
// search would have kept, and every pruned box exceeds the final floor.  Partial results are merged
// by a butterfly at the end; all lanes return the merged result.
constexpr int CO_STACK = 256;  // stack entries per warp (growth bound: see radius_stack_kernel)
constexpr int CO_GLIST = 64;   // queued groups per warp
constexpr unsigned CO_NODE_MASK = 0x07ffffffu;

__device__ __forceinline__ float warp_min_nonneg(float v) {  // v >= 0 (or +inf): float order = bit order
    return __uint_as_float(__reduce_min_sync(0xffffffffu, __float_as_uint(v)));
}

__device__ __forceinline__ void best_merge_xor(Best1f& a, int o) {
    const unsigned full = 0xffffffffu;
    const float k = __shfl_xor_sync(full, a.key, o);
    const int i = __shfl_xor_sync(full, a.idx, o), p = __shfl_xor_sync(full, a.pos, o);
    if (kless32(k, i, a.key, a.idx)) { a.key = k; a.idx = i; a.pos = p; }
}
// Two partial Best2f over DISJOINT point sets: the better best wins; the runner-up is the smallest
// key among the loser's best and the two old runners-up (the loser's runner-up cannot beat the loser's
// best); whatever is left joins `lim`.  Entries that stand for "nothing found" carry the start bound
// and position -1, exactly as after start().
__device__ __forceinline__ void best_merge_xor(Best2f& a, int o) {
    const unsigned full = 0xffffffffu;
    const float bk = __shfl_xor_sync(full, a.key, o), bk2 = __shfl_xor_sync(full, a.key2, o), bl = __shfl_xor_sync(full, a.lim, o);
    const int bi = __shfl_xor_sync(full, a.idx, o), bp = __shfl_xor_sync(full, a.pos, o), bp2 = __shfl_xor_sync(full, a.pos2, o);
    const bool bw = kless32(bk, bi, a.key, a.idx);
    const float wk = bw ? bk : a.key, lk = bw ? a.key : bk;            // winner's / loser's best
    const int wi = bw ? bi : a.idx, wp = bw ? bp : a.pos, lp = bw ? a.pos : bp;
    const float wk2 = bw ? bk2 : a.key2, lk2 = bw ? a.key2 : bk2;      // their runners-up
    const int wp2 = bw ? bp2 : a.pos2;
    float rest;
    if (wk2 < lk) { a.key2 = wk2; a.pos2 = wp2; rest = lk; }
    else { a.key2 = lk; a.pos2 = lp; rest = fminf(wk2, lk2); }
    a.key = wk; a.idx = wi; a.pos = wp;
    a.lim = fminf(fminf(a.lim, bl), rest);
    a.prune = fminf(a.lim, a.reach(wk));
}

// s_stack [CO_STACK], s_gl [CO_GLIST]: this warp's shared memory.  qx, qy, qz, seed_group and bound are
// warp-uniform; `best` comes in with its parameters set (Best2f::gap) and returns the merged result in
// every lane.  Every lane of the warp must call this together.
template <class BEST>
__device__ __forceinline__ void nn1_coop(const IndexView& iv, float qx, float qy, float qz, int seed_group,
                                         float bound, BEST& best, unsigned* s_stack, unsigned* s_gl) {
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const unsigned lt = (1u << lane) - 1u;
    best.start(bound);
    if (iv.n <= 0) return;
    if (seed_group < 0 || seed_group >= iv.n0) seed_group = home_group(iv, morton63(qx, qy, qz, iv.cube, iv.cube[3]));
    {
        const int pos = (seed_group << PCNN_GROUP_LOG) + (lane & 7);
        if (lane < PCNN_GROUP && pos < iv.n) {
            const float4 p = __ldg(iv.pts + pos);
            best.offer(key32(p.x, p.y, p.z, qx, qy, qz), __float_as_int(p.w), pos);
        }
    }
    float wb = warp_min_nonneg(best.bound());
    const int top = iv.nlev - 1;
    int sp = 0, gc = 0;
    if (top > 0) {
        const int l = walk_start_level(iv, seed_group, qx, qy, qz, wb);
        if (lane == 0) s_stack[0] = (unsigned)(seed_group >> l) | ((unsigned)l << 27);
        sp = 1;
    }
    __syncwarp();
    while (sp > 0 || gc > 0) {
        if (gc >= 16 || sp == 0) {
            // scan up to 16 queued groups (taken from the end of the list): four points per lane
            const int take = gc < 16 ? gc : 16, base = gc - take;
            float4 p[4];
            int ps[4];
            bool v[4];
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const int it = base + 4 * u + (lane >> 3);
                v[u] = it < gc;
                ps[u] = v[u] ? ((int)s_gl[it] << PCNN_GROUP_LOG) + (lane & 7) : 0;
                v[u] = v[u] && ps[u] < iv.n;
                if (v[u]) p[u] = __ldg(iv.pts + ps[u]);
            }
#pragma unroll
            for (int u = 0; u < 4; u++)
                if (v[u]) best.offer(key32(p[u].x, p[u].y, p[u].z, qx, qy, qz), __float_as_int(p[u].w), ps[u]);
            gc = base;
            wb = warp_min_nonneg(best.bound());
            __syncwarp();
            continue;
        }
        const int k = sp <= CO_STACK - 80 ? (sp < 16 ? sp : 16) : 1;
        const int sh = k > 8 ? 1 : (k > 4 ? 2 : (k > 2 ? 3 : (k > 1 ? 4 : (sp <= CO_STACK - 80 ? 5 : 1))));
        const int en = lane >> sh, sub = lane & ((1 << sh) - 1);
        const unsigned ent = en < k ? s_stack[sp - k + en] : 0u;
        sp -= k;
        const int lvl = (int)(ent >> 27);
        const int she = sh < lvl ? sh : lvl;
        const int L = lvl - she;
        const int nd = ((int)(ent & CO_NODE_MASK) << she) + sub;
        bool pass = false;
        if (en < k && sub < (1 << she) && nd < (int)(((unsigned)iv.n0 + (1u << L) - 1u) >> L)) {
            const float4* b = iv.box + 2 * ((size_t)iv.lev_off[L] + nd);
            // strict test: a box at exactly the best key may hold a lower index
            pass = !(boxlb32(__ldg(b), __ldg(b + 1), qx, qy, qz) > wb);
        }
        __syncwarp();  // every lane has read its entry before the pushes below reuse the slots
        const bool to_stack = pass && L > 0;
        const unsigned ms = __ballot_sync(full, to_stack);
        if (to_stack) s_stack[sp + __popc(ms & lt)] = (unsigned)nd | ((unsigned)L << 27);
        sp += __popc(ms);
        const bool to_list = pass && L == 0 && nd != seed_group;
        const unsigned mg = __ballot_sync(full, to_list);
        if (to_list) s_gl[gc + __popc(mg & lt)] = (unsigned)nd;
        gc += __popc(mg);
        __syncwarp();
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) best_merge_xor(best, o);
}

}  // namespace pcnn
