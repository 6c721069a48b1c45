// pcnn_knn.cu -- batched kNN kernels (sm_100a) and their C ABI.
//
// Replaces, for a whole batch of queries, the serial loop of the reference
// (Homework2/hw2/include/benchmark.hpp:59-66): `KNNResultSet rs(k); KDTreeKNNSearch(...)`
// (kdtree.hpp:329-364) / `OctreeKNNSearch(...)` (octree.hpp:143-186).
//
// Three kernels, one result contract (k smallest under (d2 double, original index)):
//   knn_thread_kernel<K>  one query per thread, sorted top-K in registers, stackless sweep
//                         of the implicit AABB hierarchy; queries are first ordered along
//                         the Morton curve so a warp walks neighbouring nodes (k <= 16)
//   knn_warp_kernel       one query per warp, 32 candidates / 32 child boxes per step,
//                         nearest-first descent, top-k list distributed over the lanes and
//                         updated with ballots + shuffles (k <= 32; larger k in passes)
//   knn_brute_kernel<K>   brute-force tiles: the cloud streamed through shared memory with
//                         bulk async copies (TMA, cp.async.bulk + mbarrier), one query per
//                         thread; small clouds and the FP64-pipe-bound reference point
#include "pcnn_internal.cuh"
#include "pcnn_search.cuh"
#include "pcnn_sort.cuh"
#include "pcnn_warp.cuh"

namespace pcnn {

// ------------------------------------------------------------------ query ordering
__global__ void query_code_kernel(const float* __restrict__ q, int nq, int stride,
                                  const float* __restrict__ cube, unsigned* __restrict__ codes,
                                  unsigned* __restrict__ perm) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nq) return;
    const float* p = q + (size_t)i * stride;
    codes[i] = (unsigned)(morton63(p[0], p[1], p[2], cube, cube[3]) >> 33);  // top 30 bits
    perm[i] = (unsigned)i;
}

int sort_queries(const pcnn_index* ix, const float* q, int64_t nq, int stride, unsigned** perm_out,
                 Scratch& sc, cudaStream_t s) {
    const int n = (int)nq;
    unsigned *c0, *c1, *p0, *p1, *hist;
    PCNN_CUDA(sc.get(&c0, (size_t)n));
    PCNN_CUDA(sc.get(&c1, (size_t)n));
    PCNN_CUDA(sc.get(&p0, (size_t)n));
    PCNN_CUDA(sc.get(&p1, (size_t)n));
    PCNN_CUDA(sc.get(&hist, radix_hist_elems(n)));
    query_code_kernel<<<(n + 255) / 256, 256, 0, s>>>(q, n, stride, ix->cube, c0, p0);
    int where = radix_sort_pairs<unsigned>(c0, p0, c1, p1, hist, n, 30, s);
    *perm_out = where ? p1 : p0;
    PCNN_CUDA(cudaGetLastError());
    return PCNN_OK;
}

// ------------------------------------------------------------------ thread-per-query
constexpr int KT_THREADS = 128;

template <int K>
__device__ __forceinline__ void write_row(const TopK<K>& tk, int64_t row, int k, int32_t* out_idx,
                                          float* out_dist, double* out_dist64) {
#pragma unroll
    for (int j = 0; j < K; j++) {
        if (j < k) {
            const bool have = tk.idx[j] != 0x7fffffff;
            const double d = have ? sqrt(tk.key[j]) : tk.key[j];  // IEEE sqrt, as the reference stores it
            out_idx[row * k + j] = have ? tk.idx[j] : -1;
            if (out_dist) out_dist[row * k + j] = (float)d;
            if (out_dist64) out_dist64[row * k + j] = d;
        }
    }
}

template <int K>
__global__ void __launch_bounds__(KT_THREADS)
knn_thread_kernel(IndexView iv, const float* __restrict__ q, int nq, int stride,
                  const unsigned* __restrict__ perm, int k, int leaf_log, int32_t* __restrict__ out_idx,
                  float* __restrict__ out_dist, double* __restrict__ out_dist64) {
    const int t = blockIdx.x * KT_THREADS + threadIdx.x;
    if (t >= nq) return;
    const int qi = perm ? (int)perm[t] : t;
    const float* qp = q + (size_t)qi * stride;
    TopK<K> tk;
    knn_thread<K>(iv, qp[0], qp[1], qp[2], leaf_log, tk);
    write_row<K>(tk, qi, k, out_idx, out_dist, out_dist64);
}

// ------------------------------------------------------------------ warp-coherent traversal
// 32 Morton-consecutive queries share ONE traversal: control flow is warp-uniform (a subtree
// is skipped only when no lane needs it, decided by a vote), every box and every candidate
// point is fetched with a single broadcast load, and each lane filters the candidate against
// its own query and its own k-th distance.  Lanes that pass the float32 filter for some point
// of a group then run the exact double key + sorted insertion together, one candidate per
// lane per round, so the expensive part is not serialised lane by lane.  Same result
// contract as knn_thread(): a lane sees a superset of what its own traversal would visit.
constexpr int KC_SEED_SPAN = 16;  // widest union of home ranges (groups) scanned by all lanes

template <int K>
__device__ __forceinline__ void scan_group_shared(const IndexView& iv, int g, bool mine, float qxf, float qyf,
                                                  float qzf, double qx, double qy, double qz, TopK<K>& tk,
                                                  float& wf) {
    const unsigned full = 0xffffffffu;
    const int first = g << PCNN_GROUP_LOG;
    const int cnt = (iv.n - first) < PCNN_GROUP ? (iv.n - first) : PCNN_GROUP;  // warp-uniform
    unsigned pass = 0;
#pragma unroll
    for (int j = 0; j < PCNN_GROUP; j++) {
        if (j < cnt) {
            const float4 p = __ldg(iv.pts + first + j);  // same address in every lane: one broadcast
            const float lbf = lbsq32(absdiff_rz(p.x, qxf), absdiff_rz(p.y, qyf), absdiff_rz(p.z, qzf));
            if (!(lbf > wf)) pass |= 1u << j;
        }
    }
    if (!mine) pass = 0;
    while (__any_sync(full, pass != 0)) {
        if (pass) {
            const int j = __ffs((int)pass) - 1;
            pass &= pass - 1;
            const float4 p = __ldg(iv.pts + first + j);
            const double d = key64(p.x, p.y, p.z, qx, qy, qz);
            const int id = __float_as_int(p.w);
            if (d <= tk.worst() && tk.accepts(d, id)) {
                tk.insert(d, id);
                wf = ru32(tk.worst());
            }
        }
    }
}

template <int K>
__global__ void __launch_bounds__(KT_THREADS)
knn_coherent_kernel(IndexView iv, const float* __restrict__ q, int nq, int stride,
                    const unsigned* __restrict__ perm, int k, int leaf_log, int32_t* __restrict__ out_idx,
                    float* __restrict__ out_dist, double* __restrict__ out_dist64) {
    const unsigned full = 0xffffffffu;
    const int t = blockIdx.x * KT_THREADS + threadIdx.x;
    const bool live = t < nq;
    const int tt = live ? t : nq - 1;  // tail lanes shadow the last query so the warp stays whole
    const int qi = perm ? (int)perm[tt] : tt;
    const float* qp = q + (size_t)qi * stride;
    const float qxf = qp[0], qyf = qp[1], qzf = qp[2];
    const double qx = (double)qxf, qy = (double)qyf, qz = (double)qzf;
    TopK<K> tk;
    tk.init();
    if (iv.n > 0) {
        float wf = __int_as_float(0x7f800000);
        const int h = home_group(iv, morton63(qxf, qyf, qzf, iv.cube, iv.cube[3]));
        int h0 = h > 0 ? h - 1 : 0, h1 = (h + 1 < iv.n0) ? h + 1 : iv.n0 - 1;
        const int hmin = __reduce_min_sync(full, h0), hmax = __reduce_max_sync(full, h1);
        if (hmax - hmin < KC_SEED_SPAN) {
            // coherent warp: everyone scans the union of the home ranges
            h0 = hmin;
            h1 = hmax;
            for (int g = hmin; g <= hmax; g++)
                scan_group_shared<K>(iv, g, true, qxf, qyf, qzf, qx, qy, qz, tk, wf);
        } else {
            // the curve jumps inside this warp: every lane seeds from its own three groups
            for (int g = h0; g <= h1; g++) scan_group64<K>(iv, g, qxf, qyf, qzf, qx, qy, qz, tk, wf);
        }
        const int top = iv.nlev - 1;
        if (leaf_log > top) leaf_log = top;
        int g_lo = 0, g_hi = iv.n0 - 1;
        if (wf < 3.0e38f) {
            const float r = __fmul_ru(__fsqrt_ru(wf), 1.000001f);
            const float lx = __fsub_rd(qxf, r), ly = __fsub_rd(qyf, r), lz = __fsub_rd(qzf, r);
            const float hx = __fadd_ru(qxf, r), hy = __fadd_ru(qyf, r), hz = __fadd_ru(qzf, r);
            g_lo = group_before(iv, morton63(lx, ly, lz, iv.cube, iv.cube[3]));
            g_hi = home_group(iv, morton63(hx, hy, hz, iv.cube, iv.cube[3]));
        }
        g_lo = __reduce_min_sync(full, g_lo);
        g_hi = __reduce_max_sync(full, g_hi);
        int i = g_lo;
        int l = i == 0 ? top : min(__ffs(i) - 1, top);
        while (i <= g_hi) {
            const int node = i >> l;
            const float4* b = iv.box + 2 * ((size_t)iv.lev_off[l] + node);
            const float4 lo = __ldg(b), hi = __ldg(b + 1);
            const float lbf = lbsq32(boxgap_rd(lo.x, hi.x, qxf), boxgap_rd(lo.y, hi.y, qyf),
                                     boxgap_rd(lo.z, hi.z, qzf));
            const bool need = !(lbf > wf);
            if (!__any_sync(full, need)) {  // no lane can gain anything in this subtree
                i = (node + 1) << l;
                l = min(__ffs(i) - 1, top);
                continue;
            }
            if (l > leaf_log) {
                l -= 1;
                continue;
            }
            const int g0 = node << l;
            int g1 = (node + 1) << l;
            if (g1 > iv.n0) g1 = iv.n0;
            for (int g = g0; g < g1; g++) {
                const bool mine = need && (g < h0 || g > h1);
                if (__any_sync(full, mine)) scan_group_shared<K>(iv, g, mine, qxf, qyf, qzf, qx, qy, qz, tk, wf);
            }
            i = g1;
            if (i >= iv.n0) break;
            l = min(__ffs(i) - 1, top);
        }
    }
    if (live) write_row<K>(tk, qi, k, out_idx, out_dist, out_dist64);
}

// ------------------------------------------------------------------ brute-force tiles (TMA)
constexpr int BF_THREADS = 128;
constexpr int BF_TILE = 1024;  // points per stage: 16 KB

__device__ __forceinline__ unsigned smem_u32(const void* p) {
    return (unsigned)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(unsigned long long* bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, unsigned parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity)
        : "memory");
}
// 1-D bulk async copy global -> shared (TMA engine; SASS UBLKCP), completion on an mbarrier
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, unsigned bytes, unsigned long long* bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
        "l"(src), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}

template <int K>
__global__ void __launch_bounds__(BF_THREADS)
knn_brute_kernel(const float4* __restrict__ pts, int n, const float* __restrict__ q, int nq, int stride,
                 int k, int32_t* __restrict__ out_idx, float* __restrict__ out_dist,
                 double* __restrict__ out_dist64) {
    __shared__ __align__(128) float4 tile[2][BF_TILE];
    __shared__ __align__(8) unsigned long long bar[2];
    const int t = blockIdx.x * BF_THREADS + threadIdx.x;
    const int qi = t < nq ? t : nq - 1;
    const float* qp = q + (size_t)qi * stride;
    const double qx = (double)qp[0], qy = (double)qp[1], qz = (double)qp[2];
    TopK<K> tk;
    tk.init();
    const int ntiles = (n + BF_TILE - 1) / BF_TILE;
    if (threadIdx.x == 0) {
        mbar_init(&bar[0], 1);
        mbar_init(&bar[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int s = 0; s < 2 && s < ntiles; s++) {
            unsigned bytes = (unsigned)(min(BF_TILE, n - s * BF_TILE) * (int)sizeof(float4));
            mbar_expect_tx(&bar[s], bytes);
            bulk_g2s(&tile[s][0], pts + (size_t)s * BF_TILE, bytes, &bar[s]);
        }
    }
    for (int it = 0; it < ntiles; it++) {
        const int s = it & 1;
        mbar_wait(&bar[s], (unsigned)((it >> 1) & 1));
        const int cnt = min(BF_TILE, n - it * BF_TILE);
#pragma unroll 4
        for (int j = 0; j < cnt; j++) {
            const float4 p = tile[s][j];  // same address across the warp: shared-memory broadcast
            const double d = key64(p.x, p.y, p.z, qx, qy, qz);
            const int id = __float_as_int(p.w);
            if (d <= tk.worst() && tk.accepts(d, id)) tk.insert(d, id);
        }
        __syncthreads();  // everyone is done with stage s before it is refilled
        if (threadIdx.x == 0 && it + 2 < ntiles) {
            unsigned bytes = (unsigned)(min(BF_TILE, n - (it + 2) * BF_TILE) * (int)sizeof(float4));
            mbar_expect_tx(&bar[s], bytes);
            bulk_g2s(&tile[s][0], pts + (size_t)(it + 2) * BF_TILE, bytes, &bar[s]);
        }
    }
    if (t < nq) write_row<K>(tk, t, k, out_idx, out_dist, out_dist64);
}

// ------------------------------------------------------------------ warp-per-query
// One warp owns one query.  Lane j tests child box j of the current node (up to 32 children
// per step) and, in a leaf, evaluates candidate j of 32.  The next node is the pending child
// with the smallest lower bound, found with one REDUX over (float bound | lane).  The
// running top-k is a 32-entry sorted list, one entry per lane; a candidate is inserted by
// one ballot (its rank) and one shuffle-up (the shift).  (lo_key, lo_idx) is an exclusive
// lower bound used to fetch neighbours 33..64, 65..96, ... in further passes.
__global__ void __launch_bounds__(WQ_WARPS * 32)
knn_warp_kernel(IndexView iv, WarpPlan wp, const float* __restrict__ q, int nq, int stride,
                int k_total, int col0, int kk, double* __restrict__ lower_key,
                int* __restrict__ lower_idx, int32_t* __restrict__ out_idx,
                float* __restrict__ out_dist, double* __restrict__ out_dist64) {
    __shared__ float s_lb[WQ_WARPS][WQ_MAXD][32];
    __shared__ int s_base[WQ_WARPS][WQ_MAXD];
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int qi = blockIdx.x * WQ_WARPS + w;
    if (qi >= nq) return;  // warp-uniform
    const float* qp = q + (size_t)qi * stride;
    const double qx = (double)qp[0], qy = (double)qp[1], qz = (double)qp[2];
    const double INF = __longlong_as_double(0x7ff0000000000000LL);
    const float INFf = __int_as_float(0x7f800000);
    double lok = -1.0;
    int loi = -1;
    if (col0 > 0) { lok = lower_key[qi]; loi = lower_idx[qi]; }
    double lk = INF;       // list entry `lane`
    int li = 0x7fffffff;
    double wk = INF;       // entry kk-1: the pruning bound
    int wi = 0x7fffffff;

    if (iv.n > 0) {
        int d = 0;
        {   // root step: lane j <-> node j of the top chain level
            const int l = wp.lvl[0];
            const int cnt = dev_level_count(iv.n0, l);
            double lb = INF;
            if (lane < cnt) {
                const float4* b = iv.box + 2 * ((size_t)iv.lev_off[l] + lane);
                lb = boxlb64(__ldg(b), __ldg(b + 1), qx, qy, qz);
            }
            s_lb[w][0][lane] = (lane < cnt) ? __double2float_rd(lb) : INFf;
            if (lane == 0) s_base[w][0] = 0;
            __syncwarp();
        }
        while (true) {
            const float lbf = s_lb[w][d][lane];
            const unsigned sel = __reduce_min_sync(full, (__float_as_uint(lbf) & ~31u) | (unsigned)lane);
            const unsigned selb = sel & ~31u;
            if (selb >= 0x7f800000u || (double)__uint_as_float(selb) > wk) {
                if (d == 0) break;  // nothing left that could hold a neighbour
                d--;
                continue;
            }
            const int j = (int)(sel & 31u);
            const int node = s_base[w][d] + j;
            __syncwarp();
            if (lane == j) s_lb[w][d][lane] = INFf;  // consumed
            __syncwarp();
            if (d < wp.nd - 1) {
                const int l = wp.lvl[d + 1];
                const int sh = wp.lvl[d] - l;
                const int base = node << sh;
                const int cnt = dev_level_count(iv.n0, l);
                const bool valid = lane < (1 << sh) && base + lane < cnt;
                double lb = INF;
                if (valid) {
                    const float4* b = iv.box + 2 * ((size_t)iv.lev_off[l] + base + lane);
                    lb = boxlb64(__ldg(b), __ldg(b + 1), qx, qy, qz);
                }
                d++;
                s_lb[w][d][lane] = (valid && !(lb > wk)) ? __double2float_rd(lb) : INFf;
                if (lane == 0) s_base[w][d] = base;
                __syncwarp();
            } else {
                // leaf: one candidate per lane
                const int pidx = (node << wp.leaf_pts_log) + lane;
                const bool valid = lane < (1 << wp.leaf_pts_log) && pidx < iv.n;
                double dk = INF;
                int id = 0x7fffffff;
                if (valid) {
                    const float4 p = __ldg(iv.pts + pidx);
                    dk = key64(p.x, p.y, p.z, qx, qy, qz);
                    id = __float_as_int(p.w);
                }
                unsigned m = __ballot_sync(full, valid && kless(lok, loi, dk, id) && kless(dk, id, wk, wi));
                while (m) {
                    const int s = __ffs((int)m) - 1;
                    m &= m - 1;
                    const double ck = shfl_d(dk, s);
                    const int ci = __shfl_sync(full, id, s);
                    if (!kless(ck, ci, wk, wi)) continue;  // bound tightened meanwhile (uniform)
                    const int pos = __popc(__ballot_sync(full, kless(lk, li, ck, ci)));
                    const double uk = shfl_up_d(lk, 1);
                    const int ui = __shfl_up_sync(full, li, 1);
                    if (lane > pos) { lk = uk; li = ui; }
                    else if (lane == pos) { lk = ck; li = ci; }
                    wk = shfl_d(lk, kk - 1);
                    wi = __shfl_sync(full, li, kk - 1);
                }
            }
        }
    }
    if (lane < kk) {
        const bool have = li != 0x7fffffff;
        const double dist = have ? sqrt(lk) : INF;
        const size_t o = (size_t)qi * k_total + col0 + lane;
        out_idx[o] = have ? li : -1;
        if (out_dist) out_dist[o] = (float)dist;
        if (out_dist64) out_dist64[o] = dist;
    }
    if (lower_key && lane == kk - 1) { lower_key[qi] = lk; lower_idx[qi] = li; }
}

// ------------------------------------------------------------------ launcher
template <int K>
static void launch_thread_k(const pcnn_index* ix, const float* q, int nq, int stride, const unsigned* perm,
                            int k, int32_t* oi, float* od, double* od64, cudaStream_t s) {
    knn_thread_kernel<K><<<(nq + KT_THREADS - 1) / KT_THREADS, KT_THREADS, 0, s>>>(
        ix->view, q, nq, stride, perm, k, ix->opt_leaf_log, oi, od, od64);
}
template <int K>
static void launch_coherent_k(const pcnn_index* ix, const float* q, int nq, int stride, const unsigned* perm,
                              int k, int32_t* oi, float* od, double* od64, cudaStream_t s) {
    knn_coherent_kernel<K><<<(nq + KT_THREADS - 1) / KT_THREADS, KT_THREADS, 0, s>>>(
        ix->view, q, nq, stride, perm, k, ix->opt_coherent_leaf_log, oi, od, od64);
}
template <int K>
static void launch_brute_k(const pcnn_index* ix, const float* q, int nq, int stride, int k, int32_t* oi,
                           float* od, double* od64, cudaStream_t s) {
    knn_brute_kernel<K><<<(nq + BF_THREADS - 1) / BF_THREADS, BF_THREADS, 0, s>>>(
        ix->pts, (int)ix->n, q, nq, stride, k, oi, od, od64);
}

int launch_knn(const pcnn_index* ix, const float* q, int64_t nq64, int stride, int k, int32_t* out_idx,
               float* out_dist, double* out_dist64, cudaStream_t s) {
    if (nq64 == 0) return PCNN_OK;
    const int nq = (int)nq64;
    int method = ix->opt_method;
    if (method == 0) method = (k <= 16) ? 4 : 2;
    if ((method == 1 || method == 3 || method == 4) && k > 16) method = 2;
    if (method == 3 && ix->n == 0) method = 1;
    if (method == 4) {
        Scratch sc(s);
        unsigned* perm = nullptr;
        if (ix->opt_sort_queries && nq > 32 && ix->n > 0) {
            int rc = sort_queries(ix, q, nq, stride, &perm, sc, s);
            if (rc != PCNN_OK) return rc;
        }
        prof_begin(ix, s);
        if (k <= 1) launch_coherent_k<1>(ix, q, nq, stride, perm, k, out_idx, out_dist, out_dist64, s);
        else if (k <= 2) launch_coherent_k<2>(ix, q, nq, stride, perm, k, out_idx, out_dist, out_dist64, s);
        else if (k <= 4) launch_coherent_k<4>(ix, q, nq, stride, perm, k, out_idx, out_dist, out_dist64, s);
        else if (k <= 8) launch_coherent_k<8>(ix, q, nq, stride, perm, k, out_idx, out_dist, out_dist64, s);
        else launch_coherent_k<16>(ix, q, nq, stride, perm, k, out_idx, out_dist, out_dist64, s);
        prof_end(ix, s);
    } else if (method == 1) {
        Scratch sc(s);
        unsigned* perm = nullptr;
        if (ix->opt_sort_queries && nq > 1024 && ix->n > 0) {
            int rc = sort_queries(ix, q, nq, stride, &perm, sc, s);
            if (rc != PCNN_OK) return rc;
        }
        prof_begin(ix, s);
        if (k <= 1) launch_thread_k<1>(ix, q, nq, stride, perm, k, out_idx, out_dist, out_dist64, s);
        else if (k <= 2) launch_thread_k<2>(ix, q, nq, stride, perm, k, out_idx, out_dist, out_dist64, s);
        else if (k <= 4) launch_thread_k<4>(ix, q, nq, stride, perm, k, out_idx, out_dist, out_dist64, s);
        else if (k <= 8) launch_thread_k<8>(ix, q, nq, stride, perm, k, out_idx, out_dist, out_dist64, s);
        else launch_thread_k<16>(ix, q, nq, stride, perm, k, out_idx, out_dist, out_dist64, s);
        prof_end(ix, s);
    } else if (method == 3) {
        prof_begin(ix, s);
        if (k <= 1) launch_brute_k<1>(ix, q, nq, stride, k, out_idx, out_dist, out_dist64, s);
        else if (k <= 2) launch_brute_k<2>(ix, q, nq, stride, k, out_idx, out_dist, out_dist64, s);
        else if (k <= 4) launch_brute_k<4>(ix, q, nq, stride, k, out_idx, out_dist, out_dist64, s);
        else if (k <= 8) launch_brute_k<8>(ix, q, nq, stride, k, out_idx, out_dist, out_dist64, s);
        else launch_brute_k<16>(ix, q, nq, stride, k, out_idx, out_dist, out_dist64, s);
        prof_end(ix, s);
    } else {
        Scratch sc(s);
        double* lower_key = nullptr;
        int* lower_idx = nullptr;
        if (k > 32) {
            PCNN_CUDA(sc.get(&lower_key, (size_t)nq));
            PCNN_CUDA(sc.get(&lower_idx, (size_t)nq));
        }
        const WarpPlan wp = make_warp_plan(ix->view);
        prof_begin(ix, s);
        for (int col0 = 0; col0 < k; col0 += 32) {
            const int kk = (k - col0) < 32 ? (k - col0) : 32;
            knn_warp_kernel<<<(nq + WQ_WARPS - 1) / WQ_WARPS, WQ_WARPS * 32, 0, s>>>(
                ix->view, wp, q, nq, stride, k, col0, kk, lower_key, lower_idx, out_idx, out_dist, out_dist64);
        }
        prof_end(ix, s);
    }
    PCNN_CUDA(cudaGetLastError());
    return PCNN_OK;
}

}  // namespace pcnn
