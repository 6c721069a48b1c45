// pcnn_radius_stack.cuh -- fixed-radius search, one warp per query, the walk kept DENSE
// (device only; included by pcnn_radius.cu after RadiusThr / RadiusLane are defined).
//
// Same contract as the other radius kernels (kdtree.hpp:367-402, octree.hpp:189-237,
// resultSet.hpp:130-140).  ncu on radius_warp_kernel showed ~60 loop iterations of ~90
// instructions per query, most of them with a handful of lanes busy: a depth-first walk hands the
// warp one node at a time.  Here the warp keeps a STACK of straddling nodes in shared memory
// and every step pops up to 16 of them at once and classifies their 32 children, one per lane
// (with fewer than 16 on the stack the step reaches 2..5 levels down instead, so the lanes are
// never idle near the root); straddling children go back on the stack with one ballot.  What
// falls out of the walk is sorted into
//   * contained subtrees of >= 32 points: written at once, 32 points per store;
//   * 8-point groups, straddling (to be tested) or contained: appended to a list and worked off
//     four groups (32 points) at a time -- lane -> group lane/8, point lane%8 -- so tests and
//     writes run with all lanes busy whatever the shape of the leaves.
// Classification is per query and per 8-point group (the 32-point leaves of radius_warp_kernel
// test 2.3x as many points).  The row cursor, the query and the thresholds are warp-uniform:
// no shuffles.  Rows come out in walk order, not in Morton order: deterministic, and the same
// in both passes, but callers must treat a row as a set (as they must for the reference,
// whose order is its own traversal order).
#pragma once
#include "pcnn_internal.cuh"

namespace pcnn {

constexpr int RK_WARPS = 8;
constexpr int RK_STACK = 512;  // stack entries per warp; see the growth bound at the pop
constexpr int RK_TLIST = 96;   // queued groups per warp
constexpr unsigned RK_NODE_MASK = 0x07ffffffu;  // entry: node (27 bits) | level << 27

template <bool FILL, bool D64>
__global__ void __launch_bounds__(RK_WARPS * 32)
radius_stack_kernel(IndexView iv, const float* __restrict__ q, int nq, int stride,
                    const unsigned* __restrict__ perm, RadiusThr rt, long long* __restrict__ counts,
                    const long long* __restrict__ offsets, int32_t* __restrict__ out_idx,
                    float* __restrict__ out_dist, double* __restrict__ out_dist64) {
    __shared__ unsigned s_stack[RK_WARPS][RK_STACK];
    __shared__ unsigned s_tl[RK_WARPS][RK_TLIST];
    __shared__ int s_loff[PCNN_MAX_LEVELS];
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (threadIdx.x < PCNN_MAX_LEVELS) s_loff[threadIdx.x] = threadIdx.x < iv.nlev ? 2 * iv.lev_off[threadIdx.x] : 0;
    __syncthreads();
    const int slot = blockIdx.x * RK_WARPS + w;
    if (slot >= nq) return;  // warp-uniform
    const int qi = perm ? (int)perm[slot] : slot;
    const float* qp = q + (size_t)qi * stride;
    RadiusLane rl;
    rl.qxf = qp[0]; rl.qyf = qp[1]; rl.qzf = qp[2];
    rl.thr = rt.thr; rl.t_in = rt.t_in; rl.t_out = rt.t_out; rl.pre = rt.t_in >= 0.f;
    const unsigned lt = (1u << lane) - 1u;
    long long run = FILL ? offsets[qi] : 0;  // fill: row cursor (warp-uniform)
    int acc = 0;                             // count: this lane's share of the count

    auto level_nodes = [&](int L) -> int { return (int)(((unsigned)iv.n0 + (1u << L) - 1u) >> L); };
    auto put = [&](long long dst, const float4& p, float s) {
        out_idx[dst] = __float_as_int(p.w);
        if constexpr (D64) {
            const double dist = sqrt(key64(p.x, p.y, p.z, (double)rl.qxf, (double)rl.qyf, (double)rl.qzf));
            if (out_dist) out_dist[dst] = (float)dist;
            out_dist64[dst] = dist;
        } else if (out_dist) {
            // float32 distances: within 4e-7 relative of the exact value (contract: 1e-6)
            out_dist[dst] = rl.pre ? sqrt_approx32(s)
                                   : (float)sqrt(key64(p.x, p.y, p.z, (double)rl.qxf, (double)rl.qyf, (double)rl.qzf));
        }
    };
    // work off the queued groups, four to a step
    int tcnt = 0;
    auto drain = [&]() {
        __syncwarp();
        const int j = lane & 7;
        for (int base = 0; base < tcnt; base += 4) {
            const int it = base + (lane >> 3);
            const unsigned ent = it < tcnt ? s_tl[w][it] : 0u;
            const int pidx = (int)((ent & 0x7fffffffu) << PCNN_GROUP_LOG) + j;
            const bool v = it < tcnt && pidx < iv.n;
            float4 p = make_float4(0.f, 0.f, 0.f, 0.f);
            if (v) p = __ldg(iv.pts + pidx);
            const float dx = p.x - rl.qxf, dy = p.y - rl.qyf, dz = p.z - rl.qzf;
            const float s = fmaf(dx, dx, fmaf(dy, dy, dz * dz));
            bool hit = v && ((ent >> 31) != 0u || (rl.pre && s <= rl.t_in));
            if (v && !hit && !(rl.pre && s > rl.t_out))  // the sliver (or a radius without float pre-tests)
                hit = key64(p.x, p.y, p.z, (double)rl.qxf, (double)rl.qyf, (double)rl.qzf) <= rl.thr;
            if constexpr (FILL) {
                const unsigned m = __ballot_sync(full, hit);
                if (hit) put(run + __popc(m & lt), p, s);
                run += __popc(m);
            } else {
                acc += hit ? 1 : 0;
            }
        }
        tcnt = 0;
        __syncwarp();
    };

    if (iv.n > 0 && rt.thr >= 0.0) {
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
            // Pop k nodes and look `sh` levels below them, k << sh <= 32.  Growth bound: a step
            // pushes at most 32 and pops k, and with one node per step (k = 1 near the limit) the
            // stack grows by less than the remaining depth (< 32) before it shrinks again.
            const int k = sp <= RK_STACK - 80 ? min(sp, 16) : 1;
            const int sh = k > 8 ? 1 : (k > 4 ? 2 : (k > 2 ? 3 : (k > 1 ? 4 : (sp <= RK_STACK - 80 ? 5 : 1))));
            const int e = lane >> sh, sub = lane & ((1 << sh) - 1);
            const unsigned ent = e < k ? s_stack[w][sp - k + e] : 0u;
            sp -= k;
            const int lvl = (int)(ent >> 27);
            const int she = min(sh, lvl);
            const int L = lvl - she;
            const int nd = ((int)(ent & RK_NODE_MASK) << she) + sub;
            int cls = 0;
            if (e < k && sub < (1 << she) && nd < level_nodes(L)) {
                const float4* b = iv.box + (size_t)s_loff[L] + 2 * (size_t)nd;
                cls = rl.classify(__ldg(b), __ldg(b + 1));
            }
            __syncwarp();  // every lane has read its entry before the pushes below reuse the slots
            // straddling inner nodes (and contained 16-point nodes, split next step) go back on the stack
            const bool to_stack = (cls == 1 && L > 0) || (cls == 2 && L == 1);
            const unsigned ms = __ballot_sync(full, to_stack);
            if (to_stack) s_stack[w][sp + __popc(ms & lt)] = (unsigned)nd | ((unsigned)L << 27);
            sp += __popc(ms);
            // groups: straddling -> tests, contained -> all eight points
            const bool to_list = cls != 0 && L == 0;
            const unsigned mt = __ballot_sync(full, to_list);
            if (to_list) s_tl[w][tcnt + __popc(mt & lt)] = (unsigned)nd | (cls == 2 ? 0x80000000u : 0u);
            tcnt += __popc(mt);
            // contained subtrees of 32 points and more
            const bool big = cls == 2 && L >= 2;
            if constexpr (FILL) {
                unsigned mb = __ballot_sync(full, big);
                while (mb) {
                    const int c = __ffs((int)mb) - 1;
                    mb &= mb - 1;
                    const int cL = __shfl_sync(full, L, c), cn = __shfl_sync(full, nd, c);
                    const int p0 = (cn << cL) << PCNN_GROUP_LOG;
                    const int p1 = min(((cn + 1) << cL) << PCNN_GROUP_LOG, iv.n);
                    for (int pb = p0 + lane; pb < p1; pb += 32) {
                        const float4 p = __ldg(iv.pts + pb);
                        const float dx = p.x - rl.qxf, dy = p.y - rl.qyf, dz = p.z - rl.qzf;
                        put(run + (pb - p0), p, fmaf(dx, dx, fmaf(dy, dy, dz * dz)));
                    }
                    run += p1 - p0;
                }
            } else if (big) {
                acc += min(((nd + 1) << L) << PCNN_GROUP_LOG, iv.n) - ((nd << L) << PCNN_GROUP_LOG);
            }
            if (tcnt > RK_TLIST - 32) drain();
            else __syncwarp();
        }
        drain();
    }
    if constexpr (!FILL) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(full, acc, o);
        if (lane == 0) counts[qi] = acc;
    }
}

// ------------------------------------------------------------------ runs of queries
// The same dense walk, shared by a RUN of neighbouring queries (as radius_batch_kernel shares its
// depth-first walk): one warp owns 32 consecutive queries of the finely ordered batch, lane c
// keeps query c and its row cursor, and the queries are cut into runs whose bounding box stays
// within run_extent per axis.  Boxes are classified against the run's box -- out of reach of
// every query / inside every query's ball / look closer -- so the walk is paid once per run.
// Queued groups are worked off four at a time with the point in registers and the queries of the
// run taken in turn from shared memory: a points x queries tile.  Float pre-tests only (the
// caller falls back to the per-query kernel for radii without them).
// run_extent > 0: the box of a run spans at most that per axis.  run_extent < 0: automatic,
// -run_extent times the mean extent of an 8-point group of this cloud (cube[8]), kept within
// [radius / 16, radius / 2].  Longer runs share the walk among more queries; shorter ones
// classify more sharply.  Per query the walk costs ~ r / e^2 and the looser classification
// ~ r * e, so the best e is a property of the sampling density, not of the radius (measured on
// the 1M aggregate, group extent ~18 cm: count pass best at e = 12-25 cm, fill pass at 6-10 cm, for
// r = 0.2 .. 1.0).
template <bool FILL, bool D64>
__global__ void __launch_bounds__(RK_WARPS * 32)
radius_stack_batch_kernel(IndexView iv, const float* __restrict__ q, int nq, int stride,
                          const unsigned* __restrict__ perm, RadiusThr rt, float run_extent, float radius_f,
                          long long* __restrict__ counts, const long long* __restrict__ offsets,
                          int32_t* __restrict__ out_idx, float* __restrict__ out_dist,
                          double* __restrict__ out_dist64) {
    __shared__ unsigned s_stack[RK_WARPS][RK_STACK];
    __shared__ unsigned s_tl[RK_WARPS][RK_TLIST];
    __shared__ float4 s_q[RK_WARPS][32];
    if (run_extent < 0.f)
        run_extent = fminf(fmaxf(-run_extent * __ldg(iv.cube + 8), radius_f * 0.0625f), radius_f * 0.5f);
    __shared__ int s_loff[PCNN_MAX_LEVELS];
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (threadIdx.x < PCNN_MAX_LEVELS) s_loff[threadIdx.x] = threadIdx.x < iv.nlev ? 2 * iv.lev_off[threadIdx.x] : 0;
    __syncthreads();
    // (The kNN kernels' interleaved launch order was measured here and dropped: count 1.106 -> 1.186 ms,
    // fill 3.295 -> 3.495 ms at r = 0.5 -- neighbouring runs share the lines they write and the nodes they walk.)
    const int slot0 = (blockIdx.x * RK_WARPS + w) * 32;
    if (slot0 >= nq) return;  // warp-uniform
    const int nb = min(32, nq - slot0);
    const bool haveq = lane < nb;
    const int qi = perm ? (int)perm[slot0 + (haveq ? lane : 0)] : slot0 + (haveq ? lane : 0);
    const float* qp = q + (size_t)qi * stride;
    const float qxf = qp[0], qyf = qp[1], qzf = qp[2];
    s_q[w][lane] = make_float4(qxf, qyf, qzf, 0.f);
    const float t_in = rt.t_in, t_out = rt.t_out;
    const double thr = rt.thr;
    const unsigned lt = (1u << lane) - 1u;
    long long run = (FILL && haveq) ? offsets[qi] : 0;  // fill: my query's row cursor; count: its count
    __syncwarp();

    auto level_nodes = [&](int L) -> int { return (int)(((unsigned)iv.n0 + (1u << L) - 1u) >> L); };
    auto put = [&](long long dst, const float4& p, const float4& c, float s) {
        out_idx[dst] = __float_as_int(p.w);
        if constexpr (D64) {
            const double dist = sqrt(key64(p.x, p.y, p.z, (double)c.x, (double)c.y, (double)c.z));
            if (out_dist) out_dist[dst] = (float)dist;
            out_dist64[dst] = dist;
        } else if (out_dist) {
            out_dist[dst] = sqrt_approx32(s);
        }
    };

    if (iv.n > 0 && thr >= 0.0) {
        const int top = iv.nlev - 1;
        int s0 = 0;
        while (s0 < nb) {
            // ---- next run [s0, e)
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
            const bool mine = lane >= s0 && lane < e;
            // 0 = out of reach of every query, 2 = inside every query's ball, 1 = look closer
            auto classify = [&](const float4& lo, const float4& hi) -> int {
                const float gx = fmaxf(fmaxf(lo.x - hix, lox - hi.x), 0.f);
                const float gy = fmaxf(fmaxf(lo.y - hiy, loy - hi.y), 0.f);
                const float gz = fmaxf(fmaxf(lo.z - hiz, loz - hi.z), 0.f);
                if (fmaf(gx, gx, fmaf(gy, gy, gz * gz)) > t_out) return 0;
                const float fx = fmaxf(hi.x - lox, hix - lo.x);
                const float fy = fmaxf(hi.y - loy, hiy - lo.y);
                const float fz = fmaxf(hi.z - loz, hiz - lo.z);
                return fmaf(fx, fx, fmaf(fy, fy, fz * fz)) <= t_in ? 2 : 1;
            };
            int tcnt = 0;
            auto drain = [&]() {
                __syncwarp();
                // eight groups (two points per lane) per step: the per-query overhead of the inner loop --
                // query load, cursor broadcast, loop control -- is paid once for 64 points.  A lane
                // without a point holds +inf coordinates: its key is +inf and fails every test.
                const int j = lane & 7;
                const float INFf = __int_as_float(0x7f800000);
                auto load = [&](int it, float4& p, bool& allhit) {
                    const unsigned ent = it < tcnt ? s_tl[w][it] : 0u;
                    const int pidx = (int)((ent & 0x7fffffffu) << PCNN_GROUP_LOG) + j;
                    const bool v = it < tcnt && pidx < iv.n;
                    allhit = v && (ent >> 31) != 0u;
                    p = make_float4(INFf, INFf, INFf, 0.f);
                    if (v) p = __ldg(iv.pts + pidx);
                };
                for (int base = 0; base < tcnt; base += 8) {
                    float4 p0, p1;
                    bool all0, all1;
                    load(base + (lane >> 3), p0, all0);
                    load(base + 4 + (lane >> 3), p1, all1);
                    for (int c = s0; c < e; c++) {
                        const float4 cq = s_q[w][c];
                        const float ax = p0.x - cq.x, ay = p0.y - cq.y, az = p0.z - cq.z;
                        const float bx = p1.x - cq.x, by = p1.y - cq.y, bz = p1.z - cq.z;
                        const float sa = fmaf(ax, ax, fmaf(ay, ay, az * az));
                        const float sb = fmaf(bx, bx, fmaf(by, by, bz * bz));
                        bool h0 = all0 || sa <= t_in, h1 = all1 || sb <= t_in;
                        const bool amb0 = !h0 && !(sa > t_out), amb1 = !h1 && !(sb > t_out);
                        if (__any_sync(full, amb0 || amb1)) {  // the sliver between the float thresholds: rare
                            if (amb0) h0 = key64(p0.x, p0.y, p0.z, (double)cq.x, (double)cq.y, (double)cq.z) <= thr;
                            if (amb1) h1 = key64(p1.x, p1.y, p1.z, (double)cq.x, (double)cq.y, (double)cq.z) <= thr;
                        }
                        const unsigned m0 = __ballot_sync(full, h0), m1 = __ballot_sync(full, h1);
                        const int n0 = __popc(m0), n1 = __popc(m1);
                        if constexpr (FILL) {
                            if ((m0 | m1) == 0u) continue;
                            const long long ro = __shfl_sync(full, run, c);
                            if (h0) put(ro + __popc(m0 & lt), p0, cq, sa);
                            if (h1) put(ro + n0 + __popc(m1 & lt), p1, cq, sb);
                        }
                        if (lane == c) run += n0 + n1;
                    }
                }
                tcnt = 0;
                __syncwarp();
            };
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
                const int k = sp <= RK_STACK - 80 ? min(sp, 16) : 1;
                const int sh = k > 8 ? 1 : (k > 4 ? 2 : (k > 2 ? 3 : (k > 1 ? 4 : (sp <= RK_STACK - 80 ? 5 : 1))));
                const int en = lane >> sh, sub = lane & ((1 << sh) - 1);
                const unsigned ent = en < k ? s_stack[w][sp - k + en] : 0u;
                sp -= k;
                const int lvl = (int)(ent >> 27);
                const int she = min(sh, lvl);
                const int L = lvl - she;
                const int nd = ((int)(ent & RK_NODE_MASK) << she) + sub;
                int cls = 0;
                if (en < k && sub < (1 << she) && nd < level_nodes(L)) {
                    const float4* b = iv.box + (size_t)s_loff[L] + 2 * (size_t)nd;
                    cls = classify(__ldg(b), __ldg(b + 1));
                }
                __syncwarp();
                const bool to_stack = (cls == 1 && L > 0) || (FILL && cls == 2 && L == 1);
                const unsigned ms = __ballot_sync(full, to_stack);
                if (to_stack) s_stack[w][sp + __popc(ms & lt)] = (unsigned)nd | ((unsigned)L << 27);
                sp += __popc(ms);
                const bool to_list = L == 0 && (cls == 1 || (FILL && cls == 2));
                const unsigned mt = __ballot_sync(full, to_list);
                if (to_list) s_tl[w][tcnt + __popc(mt & lt)] = (unsigned)nd | (cls == 2 ? 0x80000000u : 0u);
                tcnt += __popc(mt);
                if constexpr (FILL) {
                    unsigned mb = __ballot_sync(full, cls == 2 && L >= 2);
                    while (mb) {
                        const int b = __ffs((int)mb) - 1;
                        mb &= mb - 1;
                        const int cL = __shfl_sync(full, L, b), cn = __shfl_sync(full, nd, b);
                        const int p0 = (cn << cL) << PCNN_GROUP_LOG;
                        const int p1 = min(((cn + 1) << cL) << PCNN_GROUP_LOG, iv.n);
                        for (int pb = p0; pb < p1; pb += 64) {  // 64 points per step, two per lane
                            const bool va = pb + lane < p1, vb = pb + 32 + lane < p1;
                            float4 pa = make_float4(0.f, 0.f, 0.f, 0.f), pbv = pa;
                            if (va) pa = __ldg(iv.pts + pb + lane);
                            if (vb) pbv = __ldg(iv.pts + pb + 32 + lane);
                            for (int c = s0; c < e; c++) {
                                const float4 cq = s_q[w][c];
                                const long long ro = __shfl_sync(full, run, c) + (pb - p0) + lane;
                                const float ax = pa.x - cq.x, ay = pa.y - cq.y, az = pa.z - cq.z;
                                const float bx = pbv.x - cq.x, by = pbv.y - cq.y, bz = pbv.z - cq.z;
                                if (va) put(ro, pa, cq, fmaf(ax, ax, fmaf(ay, ay, az * az)));
                                if (vb) put(ro + 32, pbv, cq, fmaf(bx, bx, fmaf(by, by, bz * bz)));
                            }
                        }
                        if (mine) run += p1 - p0;
                    }
                } else {
                    // every point of a contained subtree counts for every query of the run
                    int sz = 0;
                    if (cls == 2) sz = min(((nd + 1) << L) << PCNN_GROUP_LOG, iv.n) - ((nd << L) << PCNN_GROUP_LOG);
                    const int tot = __reduce_add_sync(full, sz);
                    if (mine) run += tot;
                }
                if (tcnt > RK_TLIST - 32) drain();
                else __syncwarp();
            }
            drain();
            s0 = e;
        }
    }
    if (!FILL && haveq) counts[qi] = run;
}

}  // namespace pcnn
