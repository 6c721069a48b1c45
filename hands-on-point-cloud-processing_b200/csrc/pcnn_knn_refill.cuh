// pcnn_knn_refill.cuh -- knn_dfs_kernel's lock-step nearest-first walks with LANE REFILL
// (included by pcnn_knn.cu after TopK / scan_own_group / write_row).
//
// Same search as knn_dfs_kernel, same results bit for bit (KDTreeKNNSearch's order, near child
// first, kdtree.hpp:347-363; the batch loop of benchmark.hpp:59-66).  What changes is who does the
// work.  In knn_dfs_kernel a warp takes 32 queries and runs until the longest of the 32 walks
// ends: measured on the 1M x 1M workload the mean lane walks 31 steps, the longest of a warp 64
// (tools/warp_times.py: 49 % of the lane-steps of a round do anything).  Here the grid is
// persistent (as many blocks as are resident), queries are handed out from one global counter,
// and a lane whose query is finished writes its row and waits; as soon as REFILL_MIN lanes of
// the warp wait, they fetch the next queries of the (Morton-ordered) batch together, seed them
// and join the rounds of the lanes still walking.  The idea is the dynamic fetch of persistent
// ray-traversal kernels; lock step is kept (rounds of predicated steps, queues drained together).
#pragma once

namespace pcnn {

#ifndef PCNN_REFILL_MIN
#define PCNN_REFILL_MIN 8
#endif

template <int K>
__global__ void __launch_bounds__(KT_THREADS, K <= 8 ? PCNN_DFS_MINBLOCKS : PCNN_DFS_MINBLOCKS16)
knn_dfs_refill_kernel(IndexView iv, const float* __restrict__ q, int nq, int stride,
                      const unsigned* __restrict__ perm, int k, int32_t* __restrict__ out_idx,
                      float* __restrict__ out_dist, double* __restrict__ out_dist64,
                      unsigned* __restrict__ next_query, int refill_min, int chunk) {
    __shared__ int s_qg[KQ_DEPTH][KT_THREADS];
    __shared__ float s_ql[KQ_DEPTH][KT_THREADS];
    const unsigned full = 0xffffffffu;
    const int tid = threadIdx.x, lane = tid & 31;
    const unsigned lt = (1u << lane) - 1u;
    const float INFf = __int_as_float(0x7f800000);
    const int top = iv.nlev - 1;
    // per-lane query state
    int qi = -1;              // caller's row of the query this lane works on (-1: none)
    float qxf = 0.f, qyf = 0.f, qzf = 0.f;
    unsigned long long qcode = 0;
    TopK<K> tk;
    tk.init();
    float wf = INFf;
    int cnt = 0, l = 0, n = 0, mode = 0, h = 0, h0 = 0, h1 = 0;
    unsigned pending = 0;
    bool done = true, fresh = false;
    bool drained = false;     // the global counter has run past the batch
    // chunk > 0: no global counter -- every warp owns `chunk` consecutive queries of the (Morton-ordered)
    // batch and its lanes refill from that chunk only, so a lane that joins late is still a neighbour of
    // the lanes that are walking
    const long long wfirst = (long long)(blockIdx.x * (KT_THREADS / 32) + (tid >> 5)) * chunk;
    const int wcount = chunk > 0 ? (int)((wfirst >= nq) ? 0 : ((nq - wfirst) < chunk ? (nq - wfirst) : chunk)) : 0;
    int wnext = 0;
    if (chunk > 0 && wcount == 0) return;  // warp-uniform
    int ins_dummy = 0;
    while (true) {
        // ---- (A) refill: lanes without a query fetch the next ones together
        const unsigned idle = __ballot_sync(full, qi < 0);
        if (!drained && (idle == full || __popc(idle) >= refill_min)) {
            unsigned base = 0;
            unsigned limit = (unsigned)nq;
            if (chunk > 0) {
                base = (unsigned)(wfirst + wnext);
                limit = (unsigned)(wfirst + wcount);
                wnext += __popc(idle);
                if (wnext >= wcount) drained = true;
            } else {
                if (lane == 0) base = atomicAdd(next_query, (unsigned)__popc(idle));
                base = __shfl_sync(full, base, 0);
                if (base + (unsigned)__popc(idle) >= (unsigned)nq) drained = true;
            }
            if (qi < 0) {
                const unsigned mine = base + (unsigned)__popc(idle & lt);
                if (mine < limit) {
                    qi = perm ? (int)perm[mine] : (int)mine;
                    const float* qp = q + (size_t)qi * stride;
                    qxf = qp[0]; qyf = qp[1]; qzf = qp[2];
                    tk.init();
                    wf = INFf;
                    pending = 0; mode = 0; done = false; fresh = true;
                    // seed: the group holding the query's Morton code and its two neighbours give a
                    // first bound, so the walk prunes from its first step
                    qcode = morton63(qxf, qyf, qzf, iv.cube, iv.cube[3]);
                    h = home_group(iv, qcode);
                    h0 = h > 0 ? h - 1 : 0;
                    h1 = (h + 1 < iv.n0) ? h + 1 : iv.n0 - 1;
                    cnt = 0;
                    for (int g = h0; g <= h1; g++) { s_qg[cnt][tid] = g; s_ql[cnt][tid] = 0.f; cnt++; }
                }
            }
        }
        if (__all_sync(full, qi < 0)) break;  // nothing in flight and nothing left to fetch
        const double qx = (double)qxf, qy = (double)qyf, qz = (double)qzf;
        // ---- (B) walk: rounds of predicated steps (fresh lanes wait for their seed scan)
#pragma unroll 1
        for (int s = 0; s < KL_WALK; s++) {
            const bool can = !done && !fresh && cnt <= KQ_DEPTH - 2 && !(cnt > 0 && wf > 3.0e38f);
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
                    const int cl = (int)(((long long)iv.n0 + (1LL << L) - 1) >> L);
                    const float4* b = iv.box + 2 * ((size_t)iv.lev_off[L] + P);
                    const float4 alo = __ldg(b), ahi = __ldg(b + 1);
                    const float lbA = lbsq32(boxgap_rd(alo.x, ahi.x, qxf), boxgap_rd(alo.y, ahi.y, qyf),
                                             boxgap_rd(alo.z, ahi.z, qzf));
                    float lbB = INFf;
                    const bool hasB = P + 1 < cl;
                    if (hasB) {
                        const float4 blo = __ldg(b + 2), bhi = __ldg(b + 3);
                        lbB = lbsq32(boxgap_rd(blo.x, bhi.x, qxf), boxgap_rd(blo.y, bhi.y, qyf),
                                     boxgap_rd(blo.z, bhi.z, qzf));
                    }
                    if (mode == 0) {
                        bool first1 = lbB < lbA;
                        if (hasB && lbB == lbA)  // tie (query inside both boxes): follow the curve
                            first1 = qcode >= __ldg(iv.gcode + ((size_t)(P + 1) << L));
                        const float lbn = first1 ? lbB : lbA, lbf2 = first1 ? lbA : lbB;
                        const int near = P + (first1 ? 1 : 0);
                        const bool passN = !(lbn > wf);
                        const bool passF = !(lbf2 > wf) && lbf2 < INFf;
                        if (L == 0) {
                            if (passN && (near < h0 || near > h1)) { s_qg[cnt][tid] = near; s_ql[cnt][tid] = lbn; cnt++; }
                            if (passF && ((near ^ 1) < h0 || (near ^ 1) > h1)) { s_qg[cnt][tid] = near ^ 1; s_ql[cnt][tid] = lbf2; cnt++; }
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
                        if (!(lbv > wf)) mode = 0;  // visit it: next step examines its children
                    }
                }
            }
        }
        // ---- (C) drain the queues together
        {
            const int maxc = __reduce_max_sync(full, cnt);
            for (int r = 0; r < maxc; r++) {
                const bool act = r < cnt && !(s_ql[r][tid] > wf);
                if (__any_sync(full, act))
                    scan_own_group<K>(iv, act ? s_qg[r][tid] : 0, act, qxf, qyf, qzf, qx, qy, qz, tk, wf, ins_dummy);
            }
            cnt = 0;
        }
        // ---- (D) lanes whose seed has just been scanned start as low in the tree as its bound allows
        if (__any_sync(full, fresh)) {
            if (fresh) {
                fresh = false;
                if (top == 0) {
                    done = true;  // a single group, already scanned
                } else {
                    l = walk_start_level(iv, h, qxf, qyf, qzf, wf);
                    n = h >> l;
                }
            }
        }
        // ---- (E) finished queries leave: write the row, the lane waits for the next refill
        if (qi >= 0 && done) {
            write_row<K>(tk, qi, k, out_idx, out_dist, out_dist64);
            qi = -1;
        }
    }
}

}  // namespace pcnn
