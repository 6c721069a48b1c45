// pcnn_knn.cu -- batched kNN kernels (sm_100a) and their C ABI.
//
// Replaces, for a whole batch of queries, the serial loop of the reference
// (Homework2/hw2/include/benchmark.hpp:59-66): `KNNResultSet rs(k); KDTreeKNNSearch(...)`
// (kdtree.hpp:329-364) / `OctreeKNNSearch(...)` (octree.hpp:143-186).
//
// One result contract -- the k smallest under (squared distance in double, original index), rows in that
// order (resultSet.hpp:65-91) -- and four kernels in the shipped library:
//   knn_dfs_kernel<K>     k <= 16 (method 7, the headline kernel): one query per lane, 32 queries of the
//                         Morton-ordered batch per warp, lock-step nearest-first walks of the implicit AABB
//                         hierarchy in rounds of 12 box steps, queued groups drained together; the k best
//                         as (float32 key rounded up, position) ordered exactly (TopKf); deferred siblings
//                         remembered with their bounds; launch order interleaved over the curve; a warp's
//                         last walks finished cooperatively
//   knn_dfs_heap_kernel   16 < k <= 32 (method 9): the same walk, the k best in a per-lane max-heap in shared
//                         memory
//   knn_warp_kernel       k > 32 (method 2): one query per warp, 32 child boxes / 32 candidates per step, the
//                         list distributed over the lanes; larger k in passes of 32
//   knn_brute_kernel<K>   method 3: the cloud streamed through shared memory by 1-D bulk async copies
//                         (cp.async.bulk + mbarrier), one query per thread; tiny clouds and the FP64-bound
//                         reference point
// The measured-slower variants (methods 1, 4, 5, 6, 8, 10) exist in the -DPCNN_STUDY build only.
#include "pcnn_internal.cuh"
#include "pcnn_search.cuh"
#include "pcnn_sort.cuh"
#include "pcnn_warp.cuh"

namespace pcnn {

// ------------------------------------------------------------------ query ordering
__global__ void query_code_kernel(const float* __restrict__ q, int nq, int stride,
                                  const float* __restrict__ cube, int bits, unsigned* __restrict__ codes,
                                  unsigned* __restrict__ perm) {
    pdl_prologue();
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nq) return;
    const float* p = q + (size_t)i * stride;
    codes[i] = (unsigned)(morton63(p[0], p[1], p[2], cube, cube[3]) >> (63 - bits));  // top `bits` bits
    perm[i] = (unsigned)i;
}

int sort_queries_fine(const pcnn_index* ix, const float* q, int64_t nq, int stride, int bits, unsigned** perm_out,
                      Scratch& sc, cudaStream_t s);
int sort_queries(const pcnn_index* ix, const float* q, int64_t nq, int stride, unsigned** perm_out,
                 Scratch& sc, cudaStream_t s) {
    if (ix->opt_sort_bits > 32) return sort_queries_fine(ix, q, nq, stride, ix->opt_sort_bits, perm_out, sc, s);  // 64-bit keys
    NvtxRange nvtx_(ix, "pcnn.query_order");
    const int n = (int)nq;
    unsigned *c0, *c1, *p0, *p1, *hist;
    PCNN_CUDA(sc.get(&c0, (size_t)n));
    PCNN_CUDA(sc.get(&c1, (size_t)n));
    PCNN_CUDA(sc.get(&p0, (size_t)n));
    PCNN_CUDA(sc.get(&p1, (size_t)n));
    PCNN_CUDA(sc.get(&hist, radix_hist_elems(n)));
    const int bits = ix->opt_sort_bits;
    const bool pdl = ix->opt_pdl != 0;
    launch_pdl(query_code_kernel, dim3((n + 255) / 256), dim3(256), s, pdl, q, n, stride, (const float*)ix->cube, bits, c0, p0);
    int where = radix_sort_pairs<unsigned>(c0, p0, c1, p1, hist, n, bits, s, pdl);
    *perm_out = where ? p1 : p0;
    PCNN_CUDA(cudaGetLastError());
    return PCNN_OK;
}

// Measured and dropped (round 2): ordering by a counting sort over the leading 12-18 Morton bits (count +
// last-block scan, scatter by atomics: two kernels instead of thirteen).  On the bench clouds the 64^3
// cells of the bounding cube that hold points are few: the atomics pile onto hot counters (ordering 0.47 ms
// per 1M queries against 0.12 for the radix sort) and a cell is too coarse to give a warp neighbours
// (search 1.24 ms against 1.05 sorted, 1.27 shuffled).
// Finer ordering (up to all 63 bits, 64-bit keys, one radix pass per 8 bits) for callers whose
// work per query is large enough to pay for it: the batched radius search wants consecutive
// queries centimetres apart, not merely in the same 15 cm cell.
__global__ void query_code64_kernel(const float* __restrict__ q, int nq, int stride,
                                    const float* __restrict__ cube, int bits,
                                    unsigned long long* __restrict__ codes, unsigned* __restrict__ perm) {
    pdl_prologue();
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nq) return;
    const float* p = q + (size_t)i * stride;
    codes[i] = morton63(p[0], p[1], p[2], cube, cube[3]) >> (63 - bits);
    perm[i] = (unsigned)i;
}

int sort_queries_fine(const pcnn_index* ix, const float* q, int64_t nq, int stride, int bits, unsigned** perm_out,
                      Scratch& sc, cudaStream_t s) {
    NvtxRange nvtx_(ix, "pcnn.query_order");
    const int n = (int)nq;
    unsigned long long *c0, *c1;
    unsigned *p0, *p1, *hist;
    PCNN_CUDA(sc.get(&c0, (size_t)n));
    PCNN_CUDA(sc.get(&c1, (size_t)n));
    PCNN_CUDA(sc.get(&p0, (size_t)n));
    PCNN_CUDA(sc.get(&p1, (size_t)n));
    PCNN_CUDA(sc.get(&hist, radix_hist_elems(n)));
    const bool pdl = ix->opt_pdl != 0;
    launch_pdl(query_code64_kernel, dim3((n + 255) / 256), dim3(256), s, pdl, q, n, stride, (const float*)ix->cube, bits, c0, p0);
    int where = radix_sort_pairs<unsigned long long>(c0, p0, c1, p1, hist, n, bits, s, pdl);
    *perm_out = where ? p1 : p0;
    PCNN_CUDA(cudaGetLastError());
    return PCNN_OK;
}

// ------------------------------------------------------------------ thread-per-query
[[maybe_unused]] constexpr int KT_THREADS = 128;
#ifndef PCNN_DFS_MINBLOCKS
#define PCNN_DFS_MINBLOCKS 8  // resident blocks per SM asked of ptxas for knn_dfs_kernel (64 registers; measured)
#endif
#ifndef PCNN_DFS_MINBLOCKS16
#define PCNN_DFS_MINBLOCKS16 6  // the same for the K = 16 instance (80 registers): 1M x 1M, k = 10: 8 -> 2.33 ms, 6 -> 2.23, 5 -> 2.35, 4 -> 2.53
#endif

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
#ifdef PCNN_STUDY  // measured-slower variant kept for study: not in the shipped library

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
                wf = fminf(wf, ru32(tk.worst()));  // wf may already be tighter (box upper bounds)
            }
        }
    }
}

template <int K>
__global__ void __launch_bounds__(KT_THREADS)
knn_coherent_kernel(IndexView iv, const float* __restrict__ q, int nq, int stride,
                    const unsigned* __restrict__ perm, int k, int leaf_log, int32_t* __restrict__ out_idx,
                    float* __restrict__ out_dist, double* __restrict__ out_dist64, int budget,
                    int budget2, int* __restrict__ qlist, int* __restrict__ qcount,
                    unsigned long long* __restrict__ dbg) {
    const unsigned full = 0xffffffffu;
    const int t = blockIdx.x * KT_THREADS + threadIdx.x;
    bool handed_over = false;
    unsigned long long dbg_t0 = 0, dbg_c0 = 0;
    if (dbg) {  // diagnostics: per-warp start time (ns) and duration (cycles)
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(dbg_t0));
        dbg_c0 = clock64();
    }
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
        int steps = 0;  // box tests + group scans of this warp (uniform)
        bool reseeded = false;
        while (i <= g_hi) {
            if (steps > budget) {
                if (!reseeded) {
                    // The sweep drags on: some lane's seed bound is poor (its Morton neighbours are
                    // not its spatial neighbours, e.g. a query next to a coarse cell boundary of
                    // the curve).  Tighten the bounds WITHOUT scanning points: walk the rest of
                    // the window once on boxes only; a full leaf box holds >= K points, all of
                    // them within its farthest corner, so that distance bounds the k-th neighbour.
                    reseeded = true;
                    steps = 0;
                    budget = budget2;
                    constexpr int LD = K <= 8 ? 0 : 1;  // 8 << LD points per leaf box >= K
                    int i2 = i, l2 = l;
                    while (i2 <= g_hi) {
                        const int node = i2 >> l2;
                        const float4* bb = iv.box + 2 * ((size_t)iv.lev_off[l2] + node);
                        const float4 lo = __ldg(bb), hi = __ldg(bb + 1);
                        const float lbf = lbsq32(boxgap_rd(lo.x, hi.x, qxf), boxgap_rd(lo.y, hi.y, qyf),
                                                 boxgap_rd(lo.z, hi.z, qzf));
                        const bool need = !(lbf > wf);
                        if (!__any_sync(full, need)) {
                            i2 = (node + 1) << l2;
                            if (i2 >= iv.n0) break;
                            l2 = min(__ffs(i2) - 1, top);
                            continue;
                        }
                        if (l2 > LD) {
                            l2 -= 1;
                            continue;
                        }
                        if (((long long)(node + 1) << (l2 + PCNN_GROUP_LOG)) <= (long long)iv.n) {
                            // farthest corner, rounded up, with the same 2^-20 margin as the lower bound
                            const float ax = fmaxf(fabsf(__fsub_ru(qxf, lo.x)), fabsf(__fsub_ru(hi.x, qxf)));
                            const float ay = fmaxf(fabsf(__fsub_ru(qyf, lo.y)), fabsf(__fsub_ru(hi.y, qyf)));
                            const float az = fmaxf(fabsf(__fsub_ru(qzf, lo.z)), fabsf(__fsub_ru(hi.z, qzf)));
                            const float ub = __fmul_ru(
                                __fadd_ru(__fadd_ru(__fmul_ru(ax, ax), __fmul_ru(ay, ay)), __fmul_ru(az, az)),
                                1.000001f);
                            wf = fminf(wf, ub);
                        }
                        i2 = (node + 1) << l2;
                        if (i2 >= iv.n0) break;
                        l2 = min(__ffs(i2) - 1, top);
                    }
                    int nh = iv.n0 - 1;
                    if (wf < 3.0e38f) {
                        const float r = __fmul_ru(__fsqrt_ru(wf), 1.000001f);
                        const float hx = __fadd_ru(qxf, r), hy = __fadd_ru(qyf, r), hz = __fadd_ru(qzf, r);
                        nh = home_group(iv, morton63(hx, hy, hz, iv.cube, iv.cube[3]));
                    }
                    g_hi = __reduce_max_sync(full, nh);
                    continue;
                }
                // still dragging on: hand the warp's queries to the nearest-first kernel
                const unsigned m = __ballot_sync(full, live);
                int base = 0;
                if ((threadIdx.x & 31) == 0) base = atomicAdd(qcount, __popc(m));
                base = __shfl_sync(full, base, 0);
                if (live) qlist[base + __popc(m & ((1u << (threadIdx.x & 31)) - 1u))] = qi;
                handed_over = true;
                break;
            }
            steps++;
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
                if (__any_sync(full, mine)) {
                    scan_group_shared<K>(iv, g, mine, qxf, qyf, qzf, qx, qy, qz, tk, wf);
                    steps++;
                }
            }
            i = g1;
            if (i >= iv.n0) break;
            l = min(__ffs(i) - 1, top);
        }
    }
    if (live && !handed_over) write_row<K>(tk, qi, k, out_idx, out_dist, out_dist64);
    if (dbg && (threadIdx.x & 31) == 0) {
        dbg[2 * (size_t)(t >> 5)] = dbg_t0;
        dbg[2 * (size_t)(t >> 5) + 1] = (unsigned long long)(clock64() - dbg_c0);
    }
}

#endif  // PCNN_STUDY
// ------------------------------------------------------------------ shared walk, per-lane scan queues
// Same shared, warp-uniform walk over the boxes as knn_coherent_kernel, but a lane that needs a
// group does not make the whole warp scan it: it pushes the group on its own small queue (shared
// memory).  When some lane's queue is full all lanes drain their queues together, every lane
// scanning ITS OWN next group in lock step -- in a coherent warp those are the same addresses
// (broadcast loads), in a warp that straddles a jump of the curve the lanes work on different
// places at the same time.  The warp then pays for the largest per-lane need, not for the sum.
#ifndef PCNN_KQ_DEPTH
#define PCNN_KQ_DEPTH 8
#endif
constexpr int KQ_DEPTH = PCNN_KQ_DEPTH;  // queue entries per lane

template <int K>
__device__ __forceinline__ void scan_own_group(const IndexView& iv, int g, bool active, float qxf, float qyf,
                                               float qzf, double qx, double qy, double qz, TopK<K>& tk,
                                               float& wf, int& ins_rounds) {
    const unsigned full = 0xffffffffu;
    const int first = g << PCNN_GROUP_LOG;
    unsigned pass = 0;
    if (active) {
        const int cnt = (iv.n - first) < PCNN_GROUP ? (iv.n - first) : PCNN_GROUP;
#pragma unroll
        for (int j = 0; j < PCNN_GROUP; j++) {
            if (j < cnt) {
                const float4 p = __ldg(iv.pts + first + j);
                const float lbf = lbsq32(absdiff_rz(p.x, qxf), absdiff_rz(p.y, qyf), absdiff_rz(p.z, qzf));
                if (!(lbf > wf)) pass |= 1u << j;
            }
        }
    }
    while (__any_sync(full, pass != 0)) {
        ins_rounds++;
        if (pass) {
            const int j = __ffs((int)pass) - 1;
            pass &= pass - 1;
            const float4 p = __ldg(iv.pts + first + j);
            const double d = key64(p.x, p.y, p.z, qx, qy, qz);
            const int id = __float_as_int(p.w);
            if (d <= tk.worst() && tk.accepts(d, id)) {
                tk.insert(d, id);
                wf = fminf(wf, ru32(tk.worst()));
            }
        }
    }
}

// ------------------------------------------------------------------ the k best as (float key, position)
// knn_dfs_kernel's list.  The order is still the reference's -- ascending exact double key, ties by original
// index (resultSet.hpp:65-91 via kless) -- but an entry is two registers, not three, and a comparison is one
// FSETP: an entry carries f = ru32(key), the double key rounded UP to float32, and the point's POSITION in
// the Morton-ordered array.  Rounding is monotone, so f_a < f_b implies key_a < key_b; only when two floats
// are EQUAL (keys within one float32 ulp of each other, or true ties) are the double keys and the original
// indices recomputed from the two points and compared exactly (cold path).  The row's indices and distances
// are recomputed from the positions when the walk has ended.  f of the k-th entry is at once the float32
// pruning bound the walk needs (an upper bound of the k-th key).
template <int K>
struct TopKf {
    float key[K];
    int pos[K];
    __device__ __forceinline__ void init() {
#pragma unroll
        for (int j = 0; j < K; j++) { key[j] = __int_as_float(0x7f800000); pos[j] = -1; }
    }
};

// exact order of two entries whose float keys are equal (an empty slot is greater than everything)
__device__ __noinline__ bool entry_less_exact(const float4* __restrict__ pts, int pa, int pb, double qx, double qy,
                                              double qz) {
    if (pa < 0) return false;
    if (pb < 0) return true;
    const float4 a = __ldg(pts + pa), b = __ldg(pts + pb);
    return kless(key64(a.x, a.y, a.z, qx, qy, qz), __float_as_int(a.w), key64(b.x, b.y, b.z, qx, qy, qz),
                 __float_as_int(b.w));
}

// candidate (f, ppos) against the list; returns true if it went in
template <int K>
__device__ __forceinline__ bool topkf_offer(TopKf<K>& tk, float f, int ppos, const float4* __restrict__ pts,
                                            double qx, double qy, double qz) {
    if (f > tk.key[K - 1]) return false;
    bool tie = false;
#pragma unroll
    for (int m = 0; m < K; m++) tie |= (f == tk.key[m]);
    if (!tie) {
        // f < key[K-1]: it goes in.  c(m) = "new < entry m" is monotone in m; entry j becomes entry j-1 where
        // c(j-1) holds, the new one where c(j) holds and c(j-1) does not, and stays otherwise.
#pragma unroll
        for (int j = K - 1; j > 0; --j) {
            const bool cj = f < tk.key[j], cjm = f < tk.key[j - 1];
            tk.key[j] = cjm ? tk.key[j - 1] : (cj ? f : tk.key[j]);
            tk.pos[j] = cjm ? tk.pos[j - 1] : (cj ? ppos : tk.pos[j]);
        }
        if (f < tk.key[0]) { tk.key[0] = f; tk.pos[0] = ppos; }
        return true;
    }
    // cold: some float key equals f -- bubble the candidate down with exact comparisons where floats tie
    if (!(f < tk.key[K - 1] || entry_less_exact(pts, ppos, tk.pos[K - 1], qx, qy, qz))) return false;
    tk.key[K - 1] = f;
    tk.pos[K - 1] = ppos;
#pragma unroll
    for (int j = K - 1; j > 0; --j) {
        const float ka = tk.key[j - 1], kb = tk.key[j];
        const int pa = tk.pos[j - 1], pb = tk.pos[j];
        bool sw = kb < ka;
        if (kb == ka) sw = entry_less_exact(pts, pb, pa, qx, qy, qz);
        tk.key[j - 1] = sw ? kb : ka;
        tk.pos[j - 1] = sw ? pb : pa;
        tk.key[j] = sw ? ka : kb;
        tk.pos[j] = sw ? pa : pb;
    }
    return true;
}

template <int K>
__device__ __forceinline__ void scan_own_group_f(const IndexView& iv, int g, bool active, float qxf, float qyf,
                                                 float qzf, double qx, double qy, double qz, TopKf<K>& tk,
                                                 float& wf, int& ins_rounds) {
    const unsigned full = 0xffffffffu;
    const int first = g << PCNN_GROUP_LOG;
    unsigned pass = 0;
    if (active) {
        const int cnt = (iv.n - first) < PCNN_GROUP ? (iv.n - first) : PCNN_GROUP;
#pragma unroll
        for (int j = 0; j < PCNN_GROUP; j++) {
            if (j < cnt) {
                const float4 p = __ldg(iv.pts + first + j);
                const float lbf = lbsq32x(absdiff_rz(p.x, qxf), absdiff_rz(p.y, qyf), absdiff_rz(p.z, qzf));
                if (!(lbf > wf)) pass |= 1u << j;
            }
        }
    }
    while (__any_sync(full, pass != 0)) {
        ins_rounds++;
        if (pass) {
            const int j = __ffs((int)pass) - 1;
            pass &= pass - 1;
            const float4 p = __ldg(iv.pts + first + j);
            const float f = ru32(key64(p.x, p.y, p.z, qx, qy, qz));
            if (topkf_offer<K>(tk, f, first + j, iv.pts, qx, qy, qz)) wf = fminf(wf, tk.key[K - 1]);
        }
    }
}

// Two queued groups per call: both are filtered (16 points) before the insertion rounds start, so the
// rounds -- as many as the lane with the most survivors needs -- are shared by twice as many candidates.
// (Three groups per call: the same 0.777 ms per 1M queries at k = 8; four: 0.937, the lists spill.)
template <int K>
__device__ __forceinline__ void scan_own_groups2_f(const IndexView& iv, int g0, bool act0, int g1, bool act1,
                                                   float qxf, float qyf, float qzf, double qx, double qy, double qz,
                                                   TopKf<K>& tk, float& wf, int& ins_rounds) {
    const unsigned full = 0xffffffffu;
    unsigned pass = 0;
#pragma unroll
    for (int h = 0; h < 2; h++) {
        const int g = h ? g1 : g0;
        if (h ? act1 : act0) {
            const int first = g << PCNN_GROUP_LOG;
            const int cnt = (iv.n - first) < PCNN_GROUP ? (iv.n - first) : PCNN_GROUP;
#pragma unroll
            for (int j = 0; j < PCNN_GROUP; j++) {
                if (j < cnt) {
                    const float4 p = __ldg(iv.pts + first + j);
                    const float lbf = lbsq32x(absdiff_rz(p.x, qxf), absdiff_rz(p.y, qyf), absdiff_rz(p.z, qzf));
                    if (!(lbf > wf)) pass |= 1u << (h * PCNN_GROUP + j);
                }
            }
        }
    }
    while (__any_sync(full, pass != 0)) {
        ins_rounds++;
        if (pass) {
            const int j = __ffs((int)pass) - 1;
            pass &= pass - 1;
            const int pi = ((j < PCNN_GROUP ? g0 : g1) << PCNN_GROUP_LOG) + (j & (PCNN_GROUP - 1));
            const float4 p = __ldg(iv.pts + pi);
            const float f = ru32(key64(p.x, p.y, p.z, qx, qy, qz));
            if (topkf_offer<K>(tk, f, pi, iv.pts, qx, qy, qz)) wf = fminf(wf, tk.key[K - 1]);
        }
    }
}

// The first group of a query goes into an EMPTY list: its eight points are keyed and put in order by a
// 19-comparator network instead of eight insertions.  A lane whose sorted keys tie as floats (or a group
// that is not full) sends the whole warp through the general path -- rare, and warp-uniform.
template <int K>
__device__ __forceinline__ void seed_first_group_f(const IndexView& iv, int g, float qxf, float qyf, float qzf,
                                                   double qx, double qy, double qz, TopKf<K>& tk, float& wf,
                                                   int& ins_rounds) {
    const unsigned full = 0xffffffffu;
    const int first = g << PCNN_GROUP_LOG;
    float f[PCNN_GROUP];
    int ps[PCNN_GROUP];
    bool odd = iv.n - first < PCNN_GROUP;
    if (!odd) {
#pragma unroll
        for (int j = 0; j < PCNN_GROUP; j++) {
            const float4 p = __ldg(iv.pts + first + j);
            f[j] = ru32(key64(p.x, p.y, p.z, qx, qy, qz));
            ps[j] = first + j;
        }
#define KS_CAS(a, b)                                                \
    {                                                               \
        const bool sw = f[b] < f[a];                                \
        const float fa = f[a], fb = f[b];                           \
        const int pa = ps[a], pb = ps[b];                           \
        f[a] = sw ? fb : fa; f[b] = sw ? fa : fb;                   \
        ps[a] = sw ? pb : pa; ps[b] = sw ? pa : pb;                 \
    }
        KS_CAS(0, 1) KS_CAS(2, 3) KS_CAS(4, 5) KS_CAS(6, 7)
        KS_CAS(0, 2) KS_CAS(1, 3) KS_CAS(4, 6) KS_CAS(5, 7)
        KS_CAS(1, 2) KS_CAS(5, 6) KS_CAS(0, 4) KS_CAS(3, 7)
        KS_CAS(1, 5) KS_CAS(2, 6)
        KS_CAS(1, 4) KS_CAS(3, 6)
        KS_CAS(2, 4) KS_CAS(3, 5)
        KS_CAS(3, 4)
#undef KS_CAS
#pragma unroll
        for (int j = 0; j + 1 < PCNN_GROUP; j++) odd |= f[j] == f[j + 1];
    }
    if (__any_sync(full, odd)) {
        scan_own_group_f<K>(iv, g, true, qxf, qyf, qzf, qx, qy, qz, tk, wf, ins_rounds);
        return;
    }
#pragma unroll
    for (int j = 0; j < K && j < PCNN_GROUP; j++) { tk.key[j] = f[j]; tk.pos[j] = ps[j]; }
    if (K <= PCNN_GROUP) wf = fminf(wf, tk.key[K - 1]);
}

template <int K>
__device__ __forceinline__ void write_row_f(const TopKf<K>& tk, const float4* __restrict__ pts, double qx, double qy,
                                            double qz, int64_t row, int k, int32_t* out_idx, float* out_dist,
                                            double* out_dist64) {
#pragma unroll
    for (int j = 0; j < K; j++) {
        if (j < k) {
            const bool have = tk.pos[j] >= 0;
            double d = PCNN_INF_D;
            int id = -1;
            if (have) {
                const float4 p = __ldg(pts + tk.pos[j]);
                d = sqrt(key64(p.x, p.y, p.z, qx, qy, qz));  // IEEE sqrt, as the reference stores it
                id = __float_as_int(p.w);
            }
            out_idx[row * k + j] = id;
            if (out_dist) out_dist[row * k + j] = (float)d;
            if (out_dist64) out_dist64[row * k + j] = d;
        }
    }
}

#ifdef PCNN_STUDY  // measured-slower variant kept for study: not in the shipped library
template <int K>
__global__ void __launch_bounds__(KT_THREADS)
knn_queue_kernel(IndexView iv, const float* __restrict__ q, int nq, int stride,
                 const unsigned* __restrict__ perm, int k, int32_t* __restrict__ out_idx,
                 float* __restrict__ out_dist, double* __restrict__ out_dist64, int budget,
                 unsigned long long* __restrict__ dbg) {
    __shared__ int s_qg[KQ_DEPTH][KT_THREADS];
    __shared__ float s_ql[KQ_DEPTH][KT_THREADS];
    const unsigned full = 0xffffffffu;
    const int tid = threadIdx.x;
    const int t = blockIdx.x * KT_THREADS + tid;
    unsigned long long dbg_t0 = 0, dbg_c0 = 0;
    if (dbg) {
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(dbg_t0));
        dbg_c0 = clock64();
    }
    const bool live = t < nq;
    const int tt = live ? t : nq - 1;
    const int qi = perm ? (int)perm[tt] : tt;
    const float* qp = q + (size_t)qi * stride;
    const float qxf = qp[0], qyf = qp[1], qzf = qp[2];
    const double qx = (double)qxf, qy = (double)qyf, qz = (double)qzf;
    TopK<K> tk;
    tk.init();
    int c_box = 0, c_scan = 0, c_ins = 0, c_rbox = 0, c_drain = 0, g_lo_dbg = 0, g_hi_dbg = 0;  // diagnostics
    if (iv.n > 0) {
        float wf = __int_as_float(0x7f800000);
        int cnt = 0;  // entries on this lane's queue
        int steps = 0;
#define KQ_DRAIN()                                                                                   \
        do {                                                                                         \
            const int maxc = __reduce_max_sync(full, cnt);                                           \
            c_drain++;                                                                               \
            for (int r = 0; r < maxc; r++) {                                                         \
                const bool act = r < cnt && !(s_ql[r][tid] > wf);                                    \
                if (__any_sync(full, act)) {                                                         \
                    scan_own_group<K>(iv, act ? s_qg[r][tid] : 0, act, qxf, qyf, qzf, qx, qy, qz, tk, wf, c_ins); \
                    steps++;                                                                         \
                    c_scan++;                                                                        \
                }                                                                                    \
            }                                                                                        \
            cnt = 0;                                                                                 \
        } while (0)

        // seed: own home group and its two neighbours
        const int h = home_group(iv, morton63(qxf, qyf, qzf, iv.cube, iv.cube[3]));
        const int h0 = h > 0 ? h - 1 : 0, h1 = (h + 1 < iv.n0) ? h + 1 : iv.n0 - 1;
        for (int g = h0; g <= h1; g++) { s_qg[cnt][tid] = g; s_ql[cnt][tid] = 0.f; cnt++; }
        KQ_DRAIN();
        const int top = iv.nlev - 1;
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
        g_lo_dbg = g_lo; g_hi_dbg = g_hi;
        int i = g_lo;
        int l = i == 0 ? top : min(__ffs(i) - 1, top);
        bool reseeded = false;
        steps = 0;
        while (i <= g_hi) {
            if (steps > budget && !reseeded) {
                // Poor seed somewhere in the warp: tighten the bounds on boxes alone over the rest
                // of the window (a full group box holds 8 >= K points... or 16 at level 1, all
                // within its farthest corner), then go on.
                reseeded = true;
                KQ_DRAIN();
                constexpr int LD = K <= 8 ? 0 : 1;
                int i2 = i, l2 = l;
                while (i2 <= g_hi) {
                    c_rbox++;
                    const int node = i2 >> l2;
                    const float4* bb = iv.box + 2 * ((size_t)iv.lev_off[l2] + node);
                    const float4 lo = __ldg(bb), hi = __ldg(bb + 1);
                    const float lbf = lbsq32(boxgap_rd(lo.x, hi.x, qxf), boxgap_rd(lo.y, hi.y, qyf),
                                             boxgap_rd(lo.z, hi.z, qzf));
                    const bool need = !(lbf > wf);
                    if (!__any_sync(full, need)) {
                        i2 = (node + 1) << l2;
                        if (i2 >= iv.n0) break;
                        l2 = min(__ffs(i2) - 1, top);
                        continue;
                    }
                    if (l2 > LD) {
                        l2 -= 1;
                        continue;
                    }
                    if (((long long)(node + 1) << (l2 + PCNN_GROUP_LOG)) <= (long long)iv.n) {
                        const float ax = fmaxf(fabsf(__fsub_ru(qxf, lo.x)), fabsf(__fsub_ru(hi.x, qxf)));
                        const float ay = fmaxf(fabsf(__fsub_ru(qyf, lo.y)), fabsf(__fsub_ru(hi.y, qyf)));
                        const float az = fmaxf(fabsf(__fsub_ru(qzf, lo.z)), fabsf(__fsub_ru(hi.z, qzf)));
                        const float ub = __fmul_ru(
                            __fadd_ru(__fadd_ru(__fmul_ru(ax, ax), __fmul_ru(ay, ay)), __fmul_ru(az, az)),
                            1.000001f);
                        wf = fminf(wf, ub);
                    }
                    i2 = (node + 1) << l2;
                    if (i2 >= iv.n0) break;
                    l2 = min(__ffs(i2) - 1, top);
                }
                int nh = iv.n0 - 1;
                if (wf < 3.0e38f) {
                    const float r = __fmul_ru(__fsqrt_ru(wf), 1.000001f);
                    const float hx = __fadd_ru(qxf, r), hy = __fadd_ru(qyf, r), hz = __fadd_ru(qzf, r);
                    nh = home_group(iv, morton63(hx, hy, hz, iv.cube, iv.cube[3]));
                }
                g_hi = __reduce_max_sync(full, nh);
                continue;
            }
            steps++;
            c_box++;
            const int node = i >> l;
            const float4* b = iv.box + 2 * ((size_t)iv.lev_off[l] + node);
            const float4 lo = __ldg(b), hi = __ldg(b + 1);
            const float lbf = lbsq32(boxgap_rd(lo.x, hi.x, qxf), boxgap_rd(lo.y, hi.y, qyf),
                                     boxgap_rd(lo.z, hi.z, qzf));
            const bool need = !(lbf > wf);
            if (!__any_sync(full, need)) {
                i = (node + 1) << l;
                if (i >= iv.n0) break;
                l = min(__ffs(i) - 1, top);
                continue;
            }
            if (l > 0) {
                l -= 1;
                continue;
            }
            if (need && (node < h0 || node > h1)) {
                s_qg[cnt][tid] = node;
                s_ql[cnt][tid] = lbf;
                cnt++;
            }
            if (__any_sync(full, cnt >= KQ_DEPTH)) KQ_DRAIN();
            i = node + 1;
            if (i >= iv.n0) break;
            l = min(__ffs(i) - 1, top);
        }
        KQ_DRAIN();
#undef KQ_DRAIN
    }
    if (live) write_row<K>(tk, qi, k, out_idx, out_dist, out_dist64);
    if (dbg && (tid & 31) == 0) {
        dbg[2 * (size_t)(t >> 5)] = dbg_t0;
        dbg[2 * (size_t)(t >> 5) + 1] = (unsigned long long)(clock64() - dbg_c0);
        unsigned long long* c = dbg + 2 * (size_t)((nq + 31) / 32) + 8 * (size_t)(t >> 5);
        c[0] = c_box; c[1] = c_scan; c[2] = c_ins; c[3] = c_rbox; c[4] = c_drain;
        c[5] = (unsigned long long)(g_hi_dbg - g_lo_dbg);
    }
}

#endif  // PCNN_STUDY
// ------------------------------------------------------------------ lock-step per-lane walks
// Every lane walks the hierarchy for its OWN query (own position, own window, own bound), but
// the warp executes the walks in lock step: a round is a fixed sequence of predicated box
// steps (no data-dependent inner loops for the hardware to serialise), groups a lane needs go
// to its queue, and queues are drained together with synchronised exact-insert rounds.  A warp
// costs the LONGEST lane walk, and its 32 box loads per step are in flight at once -- against
// the shared walk above, whose cost is the UNION of the lanes' walks, one dependent load at a
// time.  A lane whose walk drags on (poor Morton seed) switches to a boxes-only pass over the
// rest of its window that tightens its bound from the far corners of full leaf boxes, then
// resumes.
#ifndef PCNN_KL_WALK
#define PCNN_KL_WALK 12
#endif
constexpr int KL_WALK = PCNN_KL_WALK;  // box steps per round

#ifdef PCNN_STUDY  // measured-slower variant kept for study: not in the shipped library
template <int K>
__global__ void __launch_bounds__(KT_THREADS)
knn_lockstep_kernel(IndexView iv, const float* __restrict__ q, int nq, int stride,
                    const unsigned* __restrict__ perm, int k, int32_t* __restrict__ out_idx,
                    float* __restrict__ out_dist, double* __restrict__ out_dist64, int budget,
                    unsigned long long* __restrict__ dbg) {
    __shared__ int s_qg[KQ_DEPTH][KT_THREADS];
    __shared__ float s_ql[KQ_DEPTH][KT_THREADS];
    const unsigned full = 0xffffffffu;
    const int tid = threadIdx.x;
    const int t = blockIdx.x * KT_THREADS + tid;
    unsigned long long dbg_t0 = 0, dbg_c0 = 0;
    if (dbg) {
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(dbg_t0));
        dbg_c0 = clock64();
    }
    const bool live = t < nq;
    const int tt = live ? t : nq - 1;
    const int qi = perm ? (int)perm[tt] : tt;
    const float* qp = q + (size_t)qi * stride;
    const float qxf = qp[0], qyf = qp[1], qzf = qp[2];
    const double qx = (double)qxf, qy = (double)qyf, qz = (double)qzf;
    TopK<K> tk;
    tk.init();
    int c_round = 0, c_scan = 0, c_ins = 0, c_steps = 0, c_tight = 0;  // diagnostics
    if (iv.n > 0) {
        float wf = __int_as_float(0x7f800000);
        int cnt = 0;
#define KL_DRAIN()                                                                                   \
        do {                                                                                         \
            const int maxc = __reduce_max_sync(full, cnt);                                           \
            for (int r = 0; r < maxc; r++) {                                                         \
                const bool act = r < cnt && !(s_ql[r][tid] > wf);                                    \
                if (__any_sync(full, act)) {                                                         \
                    scan_own_group<K>(iv, act ? s_qg[r][tid] : 0, act, qxf, qyf, qzf, qx, qy, qz, tk, wf, c_ins); \
                    c_scan++;                                                                        \
                }                                                                                    \
            }                                                                                        \
            cnt = 0;                                                                                 \
        } while (0)

        const int h = home_group(iv, morton63(qxf, qyf, qzf, iv.cube, iv.cube[3]));
        const int h0 = h > 0 ? h - 1 : 0, h1 = (h + 1 < iv.n0) ? h + 1 : iv.n0 - 1;
        for (int g = h0; g <= h1; g++) { s_qg[cnt][tid] = g; s_ql[cnt][tid] = 0.f; cnt++; }
        KL_DRAIN();
        const int top = iv.nlev - 1;
        int g_lo = 0, g_hi = iv.n0 - 1;
        if (wf < 3.0e38f) {
            const float r = __fmul_ru(__fsqrt_ru(wf), 1.000001f);
            const float lx = __fsub_rd(qxf, r), ly = __fsub_rd(qyf, r), lz = __fsub_rd(qzf, r);
            const float hx = __fadd_ru(qxf, r), hy = __fadd_ru(qyf, r), hz = __fadd_ru(qzf, r);
            g_lo = group_before(iv, morton63(lx, ly, lz, iv.cube, iv.cube[3]));
            g_hi = home_group(iv, morton63(hx, hy, hz, iv.cube, iv.cube[3]));
        }
        // walk state of this lane
        int i = g_lo;
        int l = i == 0 ? top : min(__ffs(i) - 1, top);
        bool done = i > g_hi;
        int mode = 0;          // 0 = collect groups, 1 = boxes-only bound tightening
        int si = 0, sl = 0;    // saved position while in mode 1
        int mysteps = 0;
        bool tightened = false;
        constexpr int LD = K <= 8 ? 0 : 1;  // 8 << LD points per box >= K
        while (true) {
            c_round++;
#pragma unroll 1
            for (int s = 0; s < KL_WALK; s++) {
                const bool can = !done && cnt < KQ_DEPTH;
                if (!__any_sync(full, can)) break;
                if (can) {
                    mysteps++;
                    const int node = i >> l;
                    const float4* b = iv.box + 2 * ((size_t)iv.lev_off[l] + node);
                    const float4 lo = __ldg(b), hi = __ldg(b + 1);
                    const float lbf = lbsq32(boxgap_rd(lo.x, hi.x, qxf), boxgap_rd(lo.y, hi.y, qyf),
                                             boxgap_rd(lo.z, hi.z, qzf));
                    const bool need = !(lbf > wf);
                    const int floor_l = mode ? LD : 0;
                    if (need && l > floor_l) {
                        l -= 1;
                    } else {
                        if (need) {
                            if (mode == 0) {
                                if (node < h0 || node > h1) {
                                    s_qg[cnt][tid] = node;
                                    s_ql[cnt][tid] = lbf;
                                    cnt++;
                                }
                            } else if (((long long)(node + 1) << (l + PCNN_GROUP_LOG)) <= (long long)iv.n) {
                                const float ax = fmaxf(fabsf(__fsub_ru(qxf, lo.x)), fabsf(__fsub_ru(hi.x, qxf)));
                                const float ay = fmaxf(fabsf(__fsub_ru(qyf, lo.y)), fabsf(__fsub_ru(hi.y, qyf)));
                                const float az = fmaxf(fabsf(__fsub_ru(qzf, lo.z)), fabsf(__fsub_ru(hi.z, qzf)));
                                const float ub = __fmul_ru(
                                    __fadd_ru(__fadd_ru(__fmul_ru(ax, ax), __fmul_ru(ay, ay)), __fmul_ru(az, az)),
                                    1.000001f);
                                wf = fminf(wf, ub);
                            }
                        }
                        i = (node + 1) << l;
                        l = min(__ffs(i) - 1, top);
                        if (i > g_hi || i >= iv.n0) {
                            if (mode == 1) {  // tightening pass finished: resume collecting
                                mode = 0;
                                i = si;
                                l = sl;
                            } else {
                                done = true;
                            }
                        }
                    }
                    if (mode == 0 && !tightened && !done && mysteps > budget) {
                        tightened = true;
                        c_tight++;
                        mode = 1;
                        si = i;
                        sl = l;
                    }
                }
            }
            KL_DRAIN();
            if (!__any_sync(full, !done)) break;
        }
#undef KL_DRAIN
        c_steps = mysteps;
    }
    if (live) write_row<K>(tk, qi, k, out_idx, out_dist, out_dist64);
    if (dbg) {
        const int maxsteps = __reduce_max_sync(full, c_steps);
        const int sumsteps = __reduce_add_sync(full, c_steps);
        const int ntight = __reduce_add_sync(full, c_tight);
        if ((tid & 31) == 0) {
            dbg[2 * (size_t)(t >> 5)] = dbg_t0;
            dbg[2 * (size_t)(t >> 5) + 1] = (unsigned long long)(clock64() - dbg_c0);
            unsigned long long* c = dbg + 2 * (size_t)((nq + 31) / 32) + 8 * (size_t)(t >> 5);
            c[0] = maxsteps; c[1] = c_scan; c[2] = c_ins; c[3] = sumsteps; c[4] = c_round; c[5] = ntight;
        }
    }
}

#endif  // PCNN_STUDY
// ------------------------------------------------------------------ lock-step nearest-first walks
// The order of the reference's own search (near child first, kdtree.hpp:347-363), without its
// recursion: every lane runs a depth-first walk of the binary hierarchy for its own query,
// examining the two children of a node per step (one 64-byte fetch), descending into the
// nearer one and remembering in ONE 32-bit mask at which levels the farther sibling is still
// owed a visit -- the tree is complete, so that sibling is index arithmetic on the current
// node, no stack.  Nearest-first means the first leaf reached already gives a bound close to
// the final one wherever the query sits, including next to a coarse cell boundary of the
// curve where the Morton neighbours are not the spatial ones.  The warp runs the 32 walks in
// lock step (fixed rounds of predicated steps), groups a lane needs go to its queue and the
// queues are drained together, as in knn_lockstep_kernel.
// ------------------------------------------------------------------ cooperative finish of a warp's last walks
// A warp's time is that of its longest walk: on the 1M x 1M workload the mean warp takes 100 us, one in a
// hundred 156 us and more, and the kernel's last 130 us belong to a few dozen such warps, each stepping one
// or two lanes through a long walk (query in a sparse region) at the latency of one dependent load per step
// (tools/warp_times.py).  When few lanes are left after KD_COOP_ROUNDS rounds, the warp stops stepping them
// and finishes them one at a time with all 32 lanes: what is left of a lane's walk -- the node it stands on
// and the deferred siblings whose remembered bound still beats the k-th key -- goes on a stack of nodes;
// a node is replaced by its up to 32 descendants five levels down (one box per lane), a node of <= 32 points
// is scanned one point per lane, survivors are keyed exactly in parallel and handed to the owner lane, which
// inserts them.  The same points are offered as the lock-step walk would offer (the remaining subtrees, minus
// the seed groups), so the list -- exact keys, ties by index -- ends up the same, whatever the order.
// The stack lives in the warp's columns of the group queues, which are empty at a round boundary.
#ifndef PCNN_KD_COOP_LANES
#define PCNN_KD_COOP_LANES 4
#endif
#ifndef PCNN_KD_COOP_ROUNDS
#define PCNN_KD_COOP_ROUNDS 8
#endif
#ifndef PCNN_KD_SCAN2
#define PCNN_KD_SCAN2 1
#endif
#ifndef PCNN_KD_SEEDNET
#define PCNN_KD_SEEDNET 1
#endif
constexpr int KD_COOP_LANES = PCNN_KD_COOP_LANES, KD_COOP_ROUNDS = PCNN_KD_COOP_ROUNDS;

template <int K, int THREADS, int QDEPTH, int PLEVELS>
__device__ __forceinline__ void knn_coop_finish(const IndexView& iv, unsigned todo, TopKf<K>& tk, float& wf, float qxf,
                                                float qyf, float qzf, double qx, double qy, double qz, int l, int n,
                                                unsigned pending, int mode, int h0, int h1,
                                                int (*s_qg)[THREADS], float (*s_ql)[THREADS],
                                                float (*s_pl)[THREADS]) {
    static_assert(QDEPTH >= 4, "the warp's queue columns hold the node stack");
    constexpr int CAP = (QDEPTH - 1) * 32;  // stack entries; the last row holds the candidates of a leaf
    const unsigned full = 0xffffffffu;
    const int tid = threadIdx.x, lane = tid & 31, wb = tid & ~31;
    const unsigned lt = (1u << lane) - 1u;
#define KC_NODE(i) s_qg[(i) >> 5][wb + ((i) & 31)]
#define KC_LB(i) s_ql[(i) >> 5][wb + ((i) & 31)]
    while (todo) {
        const int src = __ffs((int)todo) - 1;
        todo &= todo - 1;
        const float ox = __shfl_sync(full, qxf, src), oy = __shfl_sync(full, qyf, src), oz = __shfl_sync(full, qzf, src);
        const int oh0 = __shfl_sync(full, h0, src), oh1 = __shfl_sync(full, h1, src);
        float owf = __shfl_sync(full, wf, src);
        int sp = 0;
        if (lane == src) {
            // deferred siblings, highest level first, so that the nearest one (lowest) is popped first
            unsigned pm = pending;
            while (pm) {
                const int lp = 31 - __clz((int)pm);
                pm &= ~(1u << lp);
                const float lbp = lp < PLEVELS ? s_pl[lp][tid] : 0.f;
                if (!(lbp > wf)) {
                    KC_NODE(sp) = (int)(((unsigned)lp << 27) | (unsigned)((n >> (lp - l)) ^ 1));
                    KC_LB(sp) = lbp;
                    sp++;
                }
            }
            if (mode == 0) {  // the node whose children were to be examined next
                KC_NODE(sp) = (int)(((unsigned)l << 27) | (unsigned)n);
                KC_LB(sp) = 0.f;
                sp++;
            }
        }
        sp = __shfl_sync(full, sp, src);
        __syncwarp();
        while (sp > 0) {
            sp--;
            const unsigned e = (unsigned)KC_NODE(sp);
            const float elb = KC_LB(sp);
            __syncwarp();
            if (elb > owf) continue;  // the bound has fallen since the node was pushed
            const int L = (int)(e >> 27), N = (int)(e & 0x7ffffffu);
            if (L <= 2) {
                // at most 32 points: one per lane
                const int pi = ((N << L) << PCNN_GROUP_LOG) + lane;
                const int g = pi >> PCNN_GROUP_LOG;
                bool hit = false;
                float f = 0.f;
                if (lane < (PCNN_GROUP << L) && pi < iv.n && (g < oh0 || g > oh1)) {
                    const float4 p = __ldg(iv.pts + pi);
                    const float lbf = lbsq32x(absdiff_rz(p.x, ox), absdiff_rz(p.y, oy), absdiff_rz(p.z, oz));
                    if (!(lbf > owf)) {
                        f = ru32(key64(p.x, p.y, p.z, (double)ox, (double)oy, (double)oz));
                        hit = !(f > owf);
                    }
                }
                const unsigned hm = __ballot_sync(full, hit);
                if (hm) {
                    if (hit) {
                        const int slot = CAP + __popc(hm & lt);
                        KC_LB(slot) = f;
                        KC_NODE(slot) = pi;
                    }
                    __syncwarp();
                    if (lane == src) {
                        const int nc = __popc(hm);
                        for (int c = 0; c < nc; c++)
                            if (topkf_offer<K>(tk, KC_LB(CAP + c), KC_NODE(CAP + c), iv.pts, qx, qy, qz))
                                wf = fminf(wf, tk.key[K - 1]);
                    }
                    owf = __shfl_sync(full, wf, src);
                }
            } else {
                // up to 32 descendants, one box per lane: five levels down while the stack has room for them
                // (a binary step otherwise: depth-first, the stack then grows by at most one node per level)
                const int d = (sp <= CAP - 32 - PCNN_MAX_LEVELS) ? (L - 5 >= 2 ? 5 : L - 2) : 1;
                const int Lc = L - d;
                const int idx = (N << d) + lane;
                const int cl = (int)(((unsigned)iv.n0 + (1u << Lc) - 1u) >> Lc);
                bool pass = false;
                float lb = 0.f;
                if (lane < (1 << d) && idx < cl) {
                    const float4* b = iv.box + 2 * ((size_t)iv.lev_off[Lc] + idx);
                    const float4 lo = __ldg(b), hi = __ldg(b + 1);
                    lb = lbsq32x(boxgap_rd(lo.x, hi.x, ox), boxgap_rd(lo.y, hi.y, oy), boxgap_rd(lo.z, hi.z, oz));
                    pass = !(lb > owf);
                }
                const unsigned pm = __ballot_sync(full, pass);
                if (pass) {
                    const int slot = sp + __popc(pm & lt);
                    KC_NODE(slot) = (int)(((unsigned)Lc << 27) | (unsigned)idx);
                    KC_LB(slot) = lb;
                }
                sp += __popc(pm);
            }
            __syncwarp();
        }
    }
#undef KC_NODE
#undef KC_LB
}

// levels whose deferred siblings have their lower bound remembered (per lane, shared memory); a sibling
// deferred higher up (clouds beyond 2^27 points) is simply visited
constexpr int KW_PLEVELS = 24;
#ifndef PCNN_KD_THREADS
#define PCNN_KD_THREADS 128
#endif
constexpr int KD_THREADS = PCNN_KD_THREADS;  // threads per block of knn_dfs_kernel
#ifndef PCNN_KD_STRATA
#define PCNN_KD_STRATA 24  /* measured 0 (curve order) / 4 / 8 / 12 / 16 / 24 / 32 / 37 / 64 / 148 / 1184: k=8 0.855 / 0.838 / 0.822 / 0.802 / 0.787 / 0.785 / 0.779 / 0.793 / 0.884 / 0.920 / 0.804 ms, k=16 1.565 / 1.549 / 1.527 / 1.500 / 1.498 / 1.483 / 1.641 / 1.491 / 1.647 / 1.719 / 1.539 */
#endif
#ifndef PCNN_KD_STRATA_ORDERED_ONLY
#define PCNN_KD_STRATA_ORDERED_ONLY 1  // batches searched in the caller's order keep the caller's launch order
#endif
constexpr int KD_STRATA = PCNN_KD_STRATA;   // launch-order interleave (0 = curve order)

template <int K>
__global__ void __launch_bounds__(KD_THREADS, (K <= 8 ? PCNN_DFS_MINBLOCKS : PCNN_DFS_MINBLOCKS16) * (128 / KD_THREADS))
knn_dfs_kernel(IndexView iv, const float* __restrict__ q, int nq, int stride,
               const unsigned* __restrict__ perm, int k, int32_t* __restrict__ out_idx,
               float* __restrict__ out_dist, double* __restrict__ out_dist64,
               unsigned long long* __restrict__ dbg) {
    __shared__ int s_qg[KQ_DEPTH][KD_THREADS];
    __shared__ float s_ql[KQ_DEPTH][KD_THREADS];
    __shared__ float s_pl[KW_PLEVELS][KD_THREADS];
    pdl_prologue();
    const unsigned full = 0xffffffffu;
    const int tid = threadIdx.x;
    // Launch order: block b of the grid takes slice (b % S) * per + b / S of the ordered batch, so blocks
    // launched one after another come from S different places on the curve.  Hard regions are contiguous
    // runs of the curve (tools/warp_times.py); in curve order a hard run at the end is the kernel's tail.
    int bid = blockIdx.x;
    if (KD_STRATA > 1 && (perm != nullptr || !PCNN_KD_STRATA_ORDERED_ONLY)) {
        const int per = (int)gridDim.x / KD_STRATA;
        bid = ((int)blockIdx.x % KD_STRATA) * per + (int)blockIdx.x / KD_STRATA;
        if ((long long)bid * KD_THREADS >= nq) return;
    }
    const int t = bid * KD_THREADS + tid;
    unsigned long long dbg_t0 = 0, dbg_c0 = 0;
    if (dbg) {
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(dbg_t0));
        dbg_c0 = clock64();
    }
    const bool live = t < nq;
    const int tt = live ? t : nq - 1;
    const int qi = perm ? (int)perm[tt] : tt;
    const float* qp = q + (size_t)qi * stride;
    const float qxf = qp[0], qyf = qp[1], qzf = qp[2];
    const double qx = (double)qxf, qy = (double)qyf, qz = (double)qzf;
    TopKf<K> tk;
    tk.init();
    int c_round = 0, c_scan = 0, c_ins = 0, c_steps = 0;  // diagnostics
    if (iv.n > 0) {
        const float INFf = __int_as_float(0x7f800000);
        float wf = INFf;
        int cnt = 0;
        const int top = iv.nlev - 1;
        const unsigned long long qcode = morton63(qxf, qyf, qzf, iv.cube, iv.cube[3]);
        int l = top, n = 0;      // current node
        unsigned pending = 0;    // bit L: the sibling of our ancestor at level L is still to visit
        int mode = 0;            // 0 = examine the children of (l, n); 1 = backtrack
        bool done = false;
        // seed: the group holding the query's Morton code and its two neighbours give a first
        // bound, so the walk prunes from its first step
        const int h = home_group(iv, qcode);
        const int h0 = h > 0 ? h - 1 : 0, h1 = (h + 1 < iv.n0) ? h + 1 : iv.n0 - 1;
        // (the home group first: its bound spares its neighbours most of their insertions)
#if PCNN_KD_SEEDNET
        seed_first_group_f<K>(iv, h, qxf, qyf, qzf, qx, qy, qz, tk, wf, c_ins);
#else
        scan_own_group_f<K>(iv, h, true, qxf, qyf, qzf, qx, qy, qz, tk, wf, c_ins);
#endif
        if constexpr (PCNN_KD_SCAN2 && K <= 8) {
            scan_own_groups2_f<K>(iv, h0, h0 < h, h1, h1 > h, qxf, qyf, qzf, qx, qy, qz, tk, wf, c_ins);
        } else {
            scan_own_group_f<K>(iv, h0, h0 < h, qxf, qyf, qzf, qx, qy, qz, tk, wf, c_ins);
            scan_own_group_f<K>(iv, h1, h1 > h, qxf, qyf, qzf, qx, qy, qz, tk, wf, c_ins);
        }
        if (top == 0) done = true;  // a single group, already scanned
        // start as low in the tree as the seed bound allows (usually 5-8 levels up, not 17)
        l = walk_start_level(iv, h, qxf, qyf, qzf, wf);
        n = h >> l;
        while (true) {
            c_round++;
#pragma unroll 1
            for (int s = 0; s < KL_WALK; s++) {
                // a lane waits for a drain when its queue may overflow or it has no bound yet
                const bool can = !done && cnt <= KQ_DEPTH - 2 && !(cnt > 0 && wf > 3.0e38f);
                if (!__any_sync(full, can)) break;
                if (can) {
                    if (mode != 0) {
                        // backtrack: the nearest deferred sibling whose remembered bound still beats the k-th
                        // key (no load, no box test: the bound was computed when the sibling was deferred)
                        mode = 2;
                        while (pending != 0) {
                            const int lp = __ffs((int)pending) - 1;
                            pending &= pending - 1;
                            const float lbp = lp < KW_PLEVELS ? s_pl[lp][tid] : 0.f;
                            if (!(lbp > wf)) {
                                n = (n >> (lp - l)) ^ 1;
                                l = lp;
                                mode = 0;
                                break;
                            }
                        }
                        if (mode != 0) done = true;
                    }
                    if (!done) {
                        // examine the two children of (l, n)
                        c_steps++;
                        const int L = l - 1, P = 2 * n;
                        // both boxes of the pair, unconditionally: a node without a sibling is followed by an
                        // empty box (pcnn_index.cu), whose lower bound is +inf
                        const float4* b = iv.box + 2 * ((size_t)iv.lev_off[L] + P);
                        const float4 alo = __ldg(b), ahi = __ldg(b + 1), blo = __ldg(b + 2), bhi = __ldg(b + 3);
                        const float lbA = lbsq32x(boxgap_rd(alo.x, ahi.x, qxf), boxgap_rd(alo.y, ahi.y, qyf),
                                                  boxgap_rd(alo.z, ahi.z, qzf));
                        const float lbB = lbsq32x(boxgap_rd(blo.x, bhi.x, qxf), boxgap_rd(blo.y, bhi.y, qyf),
                                                  boxgap_rd(blo.z, bhi.z, qzf));
                        bool first1 = lbB < lbA;
                        if (lbB == lbA)  // tie (query inside both boxes): follow the curve
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
                            if (passF) {
                                pending |= 1u << L;
                                if (L < KW_PLEVELS) s_pl[L][tid] = lbf2;
                            }
                            l = L;
                            n = near;
                        } else {
                            mode = 1;
                        }
                    }
                }
            }
            {
                const int maxc = __reduce_max_sync(full, cnt);
                // (K <= 8: 0.905 -> 0.890 ms per 1M queries at k = 8; the wider lists lose, 1.63 -> 1.71 at k = 16: spills)
                if constexpr (PCNN_KD_SCAN2 && K <= 8) {
                for (int r = 0; r < maxc; r += 2) {
                    const bool act0 = r < cnt && !(s_ql[r][tid] > wf);
                    const bool act1 = r + 1 < cnt && !(s_ql[r + 1][tid] > wf);
                    if (__any_sync(full, act0 || act1)) {
                        scan_own_groups2_f<K>(iv, act0 ? s_qg[r][tid] : 0, act0, act1 ? s_qg[r + 1][tid] : 0, act1, qxf, qyf,
                                              qzf, qx, qy, qz, tk, wf, c_ins);
                        c_scan++;
                    }
                }
                } else {
                for (int r = 0; r < maxc; r++) {
                    const bool act = r < cnt && !(s_ql[r][tid] > wf);
                    if (__any_sync(full, act)) {
                        scan_own_group_f<K>(iv, act ? s_qg[r][tid] : 0, act, qxf, qyf, qzf, qx, qy, qz, tk, wf, c_ins);
                        c_scan++;
                    }
                }
                }
                cnt = 0;
            }
            const unsigned alive = __ballot_sync(full, !done);
            if (alive == 0) break;
            if (__popc(alive) <= KD_COOP_LANES && c_round >= KD_COOP_ROUNDS) {
                knn_coop_finish<K, KD_THREADS, KQ_DEPTH, KW_PLEVELS>(iv, alive, tk, wf, qxf, qyf, qzf, qx, qy, qz, l, n, pending,
                                                                     mode, h0, h1, s_qg, s_ql, s_pl);
                break;
            }
        }
    }
    if (live) write_row_f<K>(tk, iv.pts, qx, qy, qz, qi, k, out_idx, out_dist, out_dist64);
    if (dbg) {
        const int maxsteps = __reduce_max_sync(full, c_steps);
        const int sumsteps = __reduce_add_sync(full, c_steps);
        if ((tid & 31) == 0) {
            dbg[2 * (size_t)(t >> 5)] = dbg_t0;
            dbg[2 * (size_t)(t >> 5) + 1] = (unsigned long long)(clock64() - dbg_c0);
            unsigned long long* c = dbg + 2 * (size_t)((nq + 31) / 32) + 8 * (size_t)(t >> 5);
            c[0] = maxsteps; c[1] = c_scan; c[2] = c_ins; c[3] = sumsteps; c[4] = c_round; c[5] = 0;
        }
    }
}

#ifdef PCNN_STUDY  // measured-slower variant kept for study: not in the shipped library
}  // namespace pcnn
#include "pcnn_knn_refill.cuh"
namespace pcnn {

// ------------------------------------------------------------------ hybrid: Morton sweep, nearest-first rescue
// knn_lockstep_kernel's per-lane Morton-window sweep does the least box work when the Morton
// seed is good (almost always); a lane whose sweep exceeds `budget` steps has a poor seed, and
// finishes with knn_dfs_kernel's nearest-first walk from the root instead (groups the sweep
// already covered, [g_lo, i), and the seed groups are skipped there).
template <int K>
__global__ void __launch_bounds__(KT_THREADS)
knn_hybrid_kernel(IndexView iv, const float* __restrict__ q, int nq, int stride,
                  const unsigned* __restrict__ perm, int k, int32_t* __restrict__ out_idx,
                  float* __restrict__ out_dist, double* __restrict__ out_dist64, int budget,
                  unsigned long long* __restrict__ dbg) {
    __shared__ int s_qg[KQ_DEPTH][KT_THREADS];
    __shared__ float s_ql[KQ_DEPTH][KT_THREADS];
    const unsigned full = 0xffffffffu;
    const int tid = threadIdx.x;
    const int t = blockIdx.x * KT_THREADS + tid;
    unsigned long long dbg_t0 = 0, dbg_c0 = 0;
    if (dbg) {
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(dbg_t0));
        dbg_c0 = clock64();
    }
    const bool live = t < nq;
    const int tt = live ? t : nq - 1;
    const int qi = perm ? (int)perm[tt] : tt;
    const float* qp = q + (size_t)qi * stride;
    const float qxf = qp[0], qyf = qp[1], qzf = qp[2];
    const double qx = (double)qxf, qy = (double)qyf, qz = (double)qzf;
    TopK<K> tk;
    tk.init();
    int c_round = 0, c_scan = 0, c_ins = 0, c_steps = 0, c_resc = 0;  // diagnostics
    if (iv.n > 0) {
        const float INFf = __int_as_float(0x7f800000);
        float wf = INFf;
        int cnt = 0;
#define KH_DRAIN()                                                                                   \
        do {                                                                                         \
            const int maxc = __reduce_max_sync(full, cnt);                                           \
            for (int r = 0; r < maxc; r++) {                                                         \
                const bool act = r < cnt && !(s_ql[r][tid] > wf);                                    \
                if (__any_sync(full, act)) {                                                         \
                    scan_own_group<K>(iv, act ? s_qg[r][tid] : 0, act, qxf, qyf, qzf, qx, qy, qz, tk, wf, c_ins); \
                    c_scan++;                                                                        \
                }                                                                                    \
            }                                                                                        \
            cnt = 0;                                                                                 \
        } while (0)

        const unsigned long long qcode = morton63(qxf, qyf, qzf, iv.cube, iv.cube[3]);
        const int h = home_group(iv, qcode);
        const int h0 = h > 0 ? h - 1 : 0, h1 = (h + 1 < iv.n0) ? h + 1 : iv.n0 - 1;
        for (int g = h0; g <= h1; g++) { s_qg[cnt][tid] = g; s_ql[cnt][tid] = 0.f; cnt++; }
        KH_DRAIN();
        const int top = iv.nlev - 1;
        int g_lo = 0, g_hi = iv.n0 - 1;
        if (wf < 3.0e38f) {
            const float r = __fmul_ru(__fsqrt_ru(wf), 1.000001f);
            const float lx = __fsub_rd(qxf, r), ly = __fsub_rd(qyf, r), lz = __fsub_rd(qzf, r);
            const float hx = __fadd_ru(qxf, r), hy = __fadd_ru(qyf, r), hz = __fadd_ru(qzf, r);
            g_lo = group_before(iv, morton63(lx, ly, lz, iv.cube, iv.cube[3]));
            g_hi = home_group(iv, morton63(hx, hy, hz, iv.cube, iv.cube[3]));
        }
        int i = g_lo;                                   // sweep position (mode 0)
        int l = i == 0 ? top : min(__ffs(i) - 1, top);  // level of the current node (all modes)
        int n = 0;                                      // current node (modes 1, 2)
        unsigned pending = 0;
        int mode = 0;  // 0 = Morton sweep, 1 = nearest-first: examine children, 2 = nearest-first: backtrack
        bool done = i > g_hi;
        int mysteps = 0;
        while (true) {
            c_round++;
#pragma unroll 1
            for (int s = 0; s < KL_WALK; s++) {
                const bool can = !done && cnt <= KQ_DEPTH - 2;
                if (!__any_sync(full, can)) break;
                if (can) {
                    mysteps++;
                    if (mode == 0) {
                        const int node = i >> l;
                        const float4* b = iv.box + 2 * ((size_t)iv.lev_off[l] + node);
                        const float4 lo = __ldg(b), hi = __ldg(b + 1);
                        const float lbf = lbsq32(boxgap_rd(lo.x, hi.x, qxf), boxgap_rd(lo.y, hi.y, qyf),
                                                 boxgap_rd(lo.z, hi.z, qzf));
                        const bool need = !(lbf > wf);
                        if (need && l > 0) {
                            l -= 1;
                        } else {
                            if (need && (node < h0 || node > h1)) {
                                s_qg[cnt][tid] = node;
                                s_ql[cnt][tid] = lbf;
                                cnt++;
                            }
                            i = (node + 1) << l;
                            l = min(__ffs(i) - 1, top);
                            if (i > g_hi || i >= iv.n0) done = true;
                        }
                        if (!done && mysteps > budget && top > 0) {  // poor seed: nearest-first from the root
                            c_resc++;
                            mode = 1;
                            l = top;
                            n = 0;
                        }
                    } else {
                        int L, P, far = 0;
                        if (mode == 1) {
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
                            if (mode == 1) {
                                bool first1 = lbB < lbA;
                                if (hasB && lbB == lbA)
                                    first1 = qcode >= __ldg(iv.gcode + ((size_t)(P + 1) << L));
                                const float lbn = first1 ? lbB : lbA, lbf2 = first1 ? lbA : lbB;
                                const int near = P + (first1 ? 1 : 0);
                                const bool passN = !(lbn > wf);
                                const bool passF = !(lbf2 > wf) && lbf2 < INFf;
                                if (L == 0) {
                                    const int fg = near ^ 1;
                                    // skip what the seed and the sweep already covered
                                    if (passN && (near < h0 || near > h1) && (near < g_lo || near >= i)) {
                                        s_qg[cnt][tid] = near; s_ql[cnt][tid] = lbn; cnt++;
                                    }
                                    if (passF && (fg < h0 || fg > h1) && (fg < g_lo || fg >= i)) {
                                        s_qg[cnt][tid] = fg; s_ql[cnt][tid] = lbf2; cnt++;
                                    }
                                    mode = 2;
                                } else if (passN) {
                                    if (passF) pending |= 1u << L;
                                    l = L;
                                    n = near;
                                } else {
                                    mode = 2;
                                }
                            } else {
                                const float lbv = (far & 1) ? lbB : lbA;
                                if (!(lbv > wf)) mode = 1;
                            }
                        }
                    }
                }
            }
            KH_DRAIN();
            if (!__any_sync(full, !done)) break;
        }
#undef KH_DRAIN
        c_steps = mysteps;
    }
    if (live) write_row<K>(tk, qi, k, out_idx, out_dist, out_dist64);
    if (dbg) {
        const int maxsteps = __reduce_max_sync(full, c_steps);
        const int sumsteps = __reduce_add_sync(full, c_steps);
        const int nresc = __reduce_add_sync(full, c_resc);
        if ((tid & 31) == 0) {
            dbg[2 * (size_t)(t >> 5)] = dbg_t0;
            dbg[2 * (size_t)(t >> 5) + 1] = (unsigned long long)(clock64() - dbg_c0);
            unsigned long long* c = dbg + 2 * (size_t)((nq + 31) / 32) + 8 * (size_t)(t >> 5);
            c[0] = maxsteps; c[1] = c_scan; c[2] = c_ins; c[3] = sumsteps; c[4] = c_round; c[5] = nresc;
        }
    }
}

#endif  // PCNN_STUDY
// ------------------------------------------------------------------ nearest-first walks, 16 < k <= 32
// Same lock-step nearest-first walk as knn_dfs_kernel; the k best of a lane do not fit its
// registers, so they live in a binary MAX-heap in shared memory (slot-major, one column per
// lane: conflict-free), keyed by (d2, index): the root is the k-th best = the pruning bound,
// an accepted candidate replaces the root and sifts down (<= 5 levels for k = 32) instead of
// shifting a sorted list, and the heap is sorted once at the end (KNNResultSet's ascending
// order, resultSet.hpp:65-91).
constexpr int KH_THREADS = 64;
constexpr int KH_CAP = 32;
#ifndef PCNN_KH_QDEPTH
#define PCNN_KH_QDEPTH 4
#endif
#ifndef PCNN_KH_WALK
#define PCNN_KH_WALK 20  /* 12 / 16 / 20 steps: 292 / 297 / 299 M queries/s at k = 32 on 1M x 1M */
#endif
#ifndef PCNN_KH_STRATA
#define PCNN_KH_STRATA 24
#endif
#ifndef PCNN_KH_SCAN2
#define PCNN_KH_SCAN2 1
#endif
#ifndef PCNN_KH_FILL
#define PCNN_KH_FILL 1
#endif
constexpr int KH_STRATA = PCNN_KH_STRATA;
constexpr int KH_WALK = PCNN_KH_WALK;  // box steps per round of the heap kernel
constexpr int KH_QDEPTH = PCNN_KH_QDEPTH;  // queue entries per lane in the heap kernel (shared memory decides its occupancy)

#ifndef PCNN_KH_FILL
#define PCNN_KH_FILL 1
#endif
// Entries are (float32 key rounded up, position in the Morton-ordered array) as in TopKf: 8 bytes instead
// of 12, one FSETP per comparison, exact double keys + original indices only where two floats are equal.
struct HeapRef {
    float* key;  // [KH_CAP][KH_THREADS], this lane's column
    int* pos;
    int k;
    const float4* pts;
    double qx, qy, qz;
    __device__ __forceinline__ float& K(int i) const { return key[i * KH_THREADS]; }
    __device__ __forceinline__ int& I(int i) const { return pos[i * KH_THREADS]; }
    __device__ __forceinline__ void init() const {
        for (int i = 0; i < k; i++) { K(i) = __int_as_float(0x7f800000); I(i) = -1; }
    }
    __device__ __forceinline__ bool less(float fa, int pa, float fb, int pb) const {
        return fa < fb || (fa == fb && entry_less_exact(pts, pa, pb, qx, qy, qz));
    }
    __device__ __forceinline__ bool accepts(float f, int p) const { return less(f, p, K(0), I(0)); }
    // place (f, p) at position i of a heap of size n and sift it down
    __device__ __forceinline__ void sift(int i, int n, float f, int p) const {
        while (true) {
            int c = 2 * i + 1;
            if (c >= n) break;
            float ck = K(c);
            int ci = I(c);
            if (c + 1 < n) {
                const float rk = K(c + 1);
                const int ri = I(c + 1);
                if (less(ck, ci, rk, ri)) { ck = rk; ci = ri; c = c + 1; }
            }
            if (!less(f, p, ck, ci)) break;
            K(i) = ck;
            I(i) = ci;
            i = c;
        }
        K(i) = f;
        I(i) = p;
    }
    __device__ __forceinline__ void replace_root(float f, int p) const { sift(0, k, f, p); }
    // bottom-up construction over all k slots (slots never filled hold +inf / -1 and sink nowhere)
    __device__ __forceinline__ void heapify() const {
        for (int i = k / 2 - 1; i >= 0; i--) sift(i, k, K(i), I(i));
    }
    // One candidate.  The first k go into the slots as they come and the heap is built once when the
    // k-th has arrived (O(k) instead of k sift-downs through a heap of +inf); from then on a candidate
    // that beats the root replaces it.  Returns true when the root (the pruning bound) may have changed.
    __device__ __forceinline__ bool offer(float f, int p, int& filled) const {
#if PCNN_KH_FILL
        if (filled < k) {
            K(filled) = f;
            I(filled) = p;
            filled++;
            if (filled < k) return false;
            heapify();
            return true;
        }
#endif
        if (!accepts(f, p)) return false;
        replace_root(f, p);
        return true;
    }
    // heap -> ascending order in place
    __device__ __forceinline__ void sort() const {
        for (int n = k - 1; n > 0; n--) {
            const float lk = K(n);
            const int li = I(n);
            K(n) = K(0);
            I(n) = I(0);
            sift(0, n, lk, li);
        }
    }
};

__device__ __forceinline__ void scan_own_group_heap(const IndexView& iv, int g, bool active, float qxf, float qyf,
                                                    float qzf, double qx, double qy, double qz,
                                                    const HeapRef& hp, float& wf, int& filled) {
    const unsigned full = 0xffffffffu;
    const int first = g << PCNN_GROUP_LOG;
    unsigned pass = 0;
    if (active) {
        const int cnt = (iv.n - first) < PCNN_GROUP ? (iv.n - first) : PCNN_GROUP;
#pragma unroll
        for (int j = 0; j < PCNN_GROUP; j++) {
            if (j < cnt) {
                const float4 p = __ldg(iv.pts + first + j);
                const float lbf = lbsq32x(absdiff_rz(p.x, qxf), absdiff_rz(p.y, qyf), absdiff_rz(p.z, qzf));
                if (!(lbf > wf)) pass |= 1u << j;
            }
        }
    }
    while (__any_sync(full, pass != 0)) {
        if (pass) {
            const int j = __ffs((int)pass) - 1;
            pass &= pass - 1;
            const float4 p = __ldg(iv.pts + first + j);
            const float f = ru32(key64(p.x, p.y, p.z, qx, qy, qz));
            if (hp.offer(f, first + j, filled)) wf = fminf(wf, hp.K(0));
        }
    }
}

// two queued groups per call (as scan_own_groups2_f): the replacement rounds are shared by 16 points' survivors
__device__ __forceinline__ void scan_own_groups2_heap(const IndexView& iv, int g0, bool act0, int g1, bool act1,
                                                      float qxf, float qyf, float qzf, double qx, double qy, double qz,
                                                      const HeapRef& hp, float& wf, int& filled) {
    const unsigned full = 0xffffffffu;
    unsigned pass = 0;
#pragma unroll
    for (int h = 0; h < 2; h++) {
        if (h ? act1 : act0) {
            const int first = (h ? g1 : g0) << PCNN_GROUP_LOG;
            const int cnt = (iv.n - first) < PCNN_GROUP ? (iv.n - first) : PCNN_GROUP;
#pragma unroll
            for (int j = 0; j < PCNN_GROUP; j++) {
                if (j < cnt) {
                    const float4 p = __ldg(iv.pts + first + j);
                    const float lbf = lbsq32x(absdiff_rz(p.x, qxf), absdiff_rz(p.y, qyf), absdiff_rz(p.z, qzf));
                    if (!(lbf > wf)) pass |= 1u << (h * PCNN_GROUP + j);
                }
            }
        }
    }
    while (__any_sync(full, pass != 0)) {
        if (pass) {
            const int j = __ffs((int)pass) - 1;
            pass &= pass - 1;
            const int pi = ((j < PCNN_GROUP ? g0 : g1) << PCNN_GROUP_LOG) + (j & (PCNN_GROUP - 1));
            const float4 p = __ldg(iv.pts + pi);
            const float f = ru32(key64(p.x, p.y, p.z, qx, qy, qz));
            if (hp.offer(f, pi, filled)) wf = fminf(wf, hp.K(0));
        }
    }
}

__global__ void __launch_bounds__(KH_THREADS)
knn_dfs_heap_kernel(IndexView iv, const float* __restrict__ q, int nq, int stride,
                    const unsigned* __restrict__ perm, int k, int32_t* __restrict__ out_idx,
                    float* __restrict__ out_dist, double* __restrict__ out_dist64) {
    __shared__ float s_hk[KH_CAP][KH_THREADS];
    __shared__ int s_hi[KH_CAP][KH_THREADS];
    __shared__ int s_qg[KH_QDEPTH][KH_THREADS];
    __shared__ float s_ql[KH_QDEPTH][KH_THREADS];
    const unsigned full = 0xffffffffu;
    const int tid = threadIdx.x;
    int bid = blockIdx.x;  // launch order interleaved over the curve, as knn_dfs_kernel
    if (KH_STRATA > 1 && perm != nullptr) {  // only batches the library ordered itself (as knn_dfs_kernel)
        const int per = (int)gridDim.x / KH_STRATA;
        bid = ((int)blockIdx.x % KH_STRATA) * per + (int)blockIdx.x / KH_STRATA;
        if ((long long)bid * KH_THREADS >= nq) return;
    }
    const int t = bid * KH_THREADS + tid;
    const bool live = t < nq;
    const int tt = live ? t : nq - 1;
    const int qi = perm ? (int)perm[tt] : tt;
    const float* qp = q + (size_t)qi * stride;
    const float qxf = qp[0], qyf = qp[1], qzf = qp[2];
    const double qx = (double)qxf, qy = (double)qyf, qz = (double)qzf;
    HeapRef hp;
    hp.key = &s_hk[0][tid];
    hp.pos = &s_hi[0][tid];
    hp.k = k;
    hp.pts = iv.pts;
    hp.qx = qx; hp.qy = qy; hp.qz = qz;
    hp.init();
    int filled = 0;  // slots taken before the heap was built (HeapRef::offer)
    if (iv.n > 0) {
        const float INFf = __int_as_float(0x7f800000);
        float wf = INFf;
        int cnt = 0;
        const int top = iv.nlev - 1;
        const unsigned long long qcode = morton63(qxf, qyf, qzf, iv.cube, iv.cube[3]);
        int l = top, n = 0;
        unsigned pending = 0;
        int mode = 0;
        bool done = false;
        // seed: enough groups around the query's place on the curve to fill the heap
        const int h = home_group(iv, qcode);
        const int half = (k + 2 * PCNN_GROUP - 1) / (2 * PCNN_GROUP);  // groups on each side
        const int h0 = max(h - half, 0), h1 = min(h + half, iv.n0 - 1);
        const int nseed = __reduce_max_sync(full, (h1 - h0 + KH_QDEPTH) / KH_QDEPTH);  // same trip count in every lane
        for (int sr = 0; sr < nseed; sr++) {
            const int g0 = h0 + sr * KH_QDEPTH;
            for (int g = g0; g <= h1 && g < g0 + KH_QDEPTH; g++) { s_qg[cnt][tid] = g; s_ql[cnt][tid] = 0.f; cnt++; }
            const int maxc = __reduce_max_sync(full, cnt);
            for (int r = 0; r < maxc; r++) {
                const bool act = r < cnt;
                scan_own_group_heap(iv, act ? s_qg[r][tid] : 0, act, qxf, qyf, qzf, qx, qy, qz, hp, wf, filled);
            }
            cnt = 0;
        }
        if (top == 0) done = true;
        l = walk_start_level(iv, h, qxf, qyf, qzf, wf);
        n = h >> l;
        while (true) {
#pragma unroll 1
            for (int s = 0; s < KH_WALK; s++) {
                const bool can = !done && cnt <= KH_QDEPTH - 2 && !(cnt > 0 && wf > 3.0e38f);
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
                        const float lbA = lbsq32x(boxgap_rd(alo.x, ahi.x, qxf), boxgap_rd(alo.y, ahi.y, qyf),
                                                 boxgap_rd(alo.z, ahi.z, qzf));
                        float lbB = INFf;
                        const bool hasB = P + 1 < cl;
                        if (hasB) {
                            const float4 blo = __ldg(b + 2), bhi = __ldg(b + 3);
                            lbB = lbsq32x(boxgap_rd(blo.x, bhi.x, qxf), boxgap_rd(blo.y, bhi.y, qyf),
                                         boxgap_rd(blo.z, bhi.z, qzf));
                        }
                        if (mode == 0) {
                            bool first1 = lbB < lbA;
                            if (hasB && lbB == lbA) first1 = qcode >= __ldg(iv.gcode + ((size_t)(P + 1) << L));
                            const float lbn = first1 ? lbB : lbA, lbf2 = first1 ? lbA : lbB;
                            const int near = P + (first1 ? 1 : 0);
                            const bool passN = !(lbn > wf);
                            const bool passF = !(lbf2 > wf) && lbf2 < INFf;
                            if (L == 0) {
                                const int fg = near ^ 1;
                                if (passN && (near < h0 || near > h1)) { s_qg[cnt][tid] = near; s_ql[cnt][tid] = lbn; cnt++; }
                                if (passF && (fg < h0 || fg > h1)) { s_qg[cnt][tid] = fg; s_ql[cnt][tid] = lbf2; cnt++; }
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
                            if (!(lbv > wf)) mode = 0;
                        }
                    }
                }
            }
            {
                const int maxc = __reduce_max_sync(full, cnt);
#if PCNN_KH_SCAN2
                for (int r = 0; r < maxc; r += 2) {
                    const bool act0 = r < cnt && !(s_ql[r][tid] > wf);
                    const bool act1 = r + 1 < cnt && !(s_ql[r + 1][tid] > wf);
                    if (__any_sync(full, act0 || act1))
                        scan_own_groups2_heap(iv, act0 ? s_qg[r][tid] : 0, act0, act1 ? s_qg[r + 1][tid] : 0, act1, qxf, qyf, qzf,
                                              qx, qy, qz, hp, wf, filled);
                }
#else
                for (int r = 0; r < maxc; r++) {
                    const bool act = r < cnt && !(s_ql[r][tid] > wf);
                    if (__any_sync(full, act))
                        scan_own_group_heap(iv, act ? s_qg[r][tid] : 0, act, qxf, qyf, qzf, qx, qy, qz, hp, wf, filled);
                }
#endif
                cnt = 0;
            }
            if (!__any_sync(full, !done)) break;
        }
    }
    if (PCNN_KH_FILL && filled < k) hp.heapify();  // fewer than k candidates in all: the slots are not a heap yet
    hp.sort();
    if (live) {
        for (int j = 0; j < k; j++) {
            const int pj = hp.I(j);
            double d = PCNN_INF_D;
            int id = -1;
            if (pj >= 0) {
                const float4 p = __ldg(iv.pts + pj);
                d = sqrt(key64(p.x, p.y, p.z, qx, qy, qz));  // IEEE sqrt, as the reference stores it
                id = __float_as_int(p.w);
            }
            out_idx[(size_t)qi * k + j] = id;
            if (out_dist) out_dist[(size_t)qi * k + j] = (float)d;
            if (out_dist64) out_dist64[(size_t)qi * k + j] = d;
        }
    }
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
__device__ __forceinline__ void knn_warp_one(const IndexView& iv, const WarpPlan& wp, float (*s_lb)[32],
                                             int* s_base, const float* __restrict__ q, int qi, int stride,
                                             int k_total, int col0, int kk, double* __restrict__ lower_key,
                                             int* __restrict__ lower_idx, int32_t* __restrict__ out_idx,
                                             float* __restrict__ out_dist, double* __restrict__ out_dist64) {
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31;
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
            s_lb[0][lane] = (lane < cnt) ? __double2float_rd(lb) : INFf;
            if (lane == 0) s_base[0] = 0;
            __syncwarp();
        }
        while (true) {
            const float lbf = s_lb[d][lane];
            const unsigned sel = __reduce_min_sync(full, (__float_as_uint(lbf) & ~31u) | (unsigned)lane);
            const unsigned selb = sel & ~31u;
            if (selb >= 0x7f800000u || (double)__uint_as_float(selb) > wk) {
                if (d == 0) break;  // nothing left that could hold a neighbour
                d--;
                continue;
            }
            const int j = (int)(sel & 31u);
            const int node = s_base[d] + j;
            __syncwarp();
            if (lane == j) s_lb[d][lane] = INFf;  // consumed
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
                s_lb[d][lane] = (valid && !(lb > wk)) ? __double2float_rd(lb) : INFf;
                if (lane == 0) s_base[d] = base;
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

__global__ void __launch_bounds__(WQ_WARPS * 32)
knn_warp_kernel(IndexView iv, WarpPlan wp, const float* __restrict__ q, int nq, int stride,
                int k_total, int col0, int kk, double* __restrict__ lower_key,
                int* __restrict__ lower_idx, int32_t* __restrict__ out_idx,
                float* __restrict__ out_dist, double* __restrict__ out_dist64) {
    __shared__ float s_lb[WQ_WARPS][WQ_MAXD][32];
    __shared__ int s_base[WQ_WARPS][WQ_MAXD];
    const int w = threadIdx.x >> 5;
    const int qi = blockIdx.x * WQ_WARPS + w;
    if (qi >= nq) return;  // warp-uniform
    knn_warp_one(iv, wp, s_lb[w], s_base[w], q, qi, stride, k_total, col0, kk, lower_key, lower_idx, out_idx,
                 out_dist, out_dist64);
}

#ifdef PCNN_STUDY  // measured-slower variant kept for study: not in the shipped library
// Same search for the queries a budgeted kernel gave up on: qlist[0 .. *qcount) holds their
// ids; a fixed grid of warps walks the list.
__global__ void __launch_bounds__(WQ_WARPS * 32)
knn_warp_list_kernel(IndexView iv, WarpPlan wp, const float* __restrict__ q, int stride, int k,
                     const int* __restrict__ qlist, const int* __restrict__ qcount,
                     int32_t* __restrict__ out_idx, float* __restrict__ out_dist,
                     double* __restrict__ out_dist64) {
    __shared__ float s_lb[WQ_WARPS][WQ_MAXD][32];
    __shared__ int s_base[WQ_WARPS][WQ_MAXD];
    const int w = threadIdx.x >> 5;
    const int total = *qcount;
    for (int i = blockIdx.x * WQ_WARPS + w; i < total; i += gridDim.x * WQ_WARPS) {
        knn_warp_one(iv, wp, s_lb[w], s_base[w], q, qlist[i], stride, k, 0, k, nullptr, nullptr, out_idx,
                     out_dist, out_dist64);
        __syncwarp();
    }
}

#endif  // PCNN_STUDY
// ------------------------------------------------------------------ launcher
#ifdef PCNN_STUDY
template <int K>
static void launch_thread_k(const pcnn_index* ix, const float* q, int nq, int stride, const unsigned* perm,
                            int k, int32_t* oi, float* od, double* od64, cudaStream_t s) {
    knn_thread_kernel<K><<<(nq + KT_THREADS - 1) / KT_THREADS, KT_THREADS, 0, s>>>(
        ix->view, q, nq, stride, perm, k, ix->opt_leaf_log, oi, od, od64);
}
template <int K>
static void launch_coherent_k(const pcnn_index* ix, const float* q, int nq, int stride, const unsigned* perm,
                              int k, int32_t* oi, float* od, double* od64, int* qlist, int* qcount,
                              cudaStream_t s) {
    knn_coherent_kernel<K><<<(nq + KT_THREADS - 1) / KT_THREADS, KT_THREADS, 0, s>>>(
        ix->view, q, nq, stride, perm, k, ix->opt_coherent_leaf_log, oi, od, od64,
        ix->opt_coherent_budget > 0 ? ix->opt_coherent_budget : 0x7fffffff,
        qlist ? ix->opt_coherent_budget2 : 0x7fffffff, qlist, qcount, ix->dbg_warp_times);
}
template <int K>
static void launch_queue_k(const pcnn_index* ix, const float* q, int nq, int stride, const unsigned* perm,
                           int k, int32_t* oi, float* od, double* od64, cudaStream_t s) {
    knn_queue_kernel<K><<<(nq + KT_THREADS - 1) / KT_THREADS, KT_THREADS, 0, s>>>(
        ix->view, q, nq, stride, perm, k, oi, od, od64,
        ix->opt_coherent_budget > 0 ? ix->opt_coherent_budget : 0x7fffffff, ix->dbg_warp_times);
}
template <int K>
static void launch_lockstep_k(const pcnn_index* ix, const float* q, int nq, int stride, const unsigned* perm,
                              int k, int32_t* oi, float* od, double* od64, cudaStream_t s) {
    knn_lockstep_kernel<K><<<(nq + KT_THREADS - 1) / KT_THREADS, KT_THREADS, 0, s>>>(
        ix->view, q, nq, stride, perm, k, oi, od, od64,
        ix->opt_coherent_budget > 0 ? ix->opt_coherent_budget : 0x7fffffff, ix->dbg_warp_times);
}
#endif  // PCNN_STUDY
template <int K>
static void launch_dfs_k(const pcnn_index* ix, const float* q, int nq, int stride, const unsigned* perm,
                         int k, int32_t* oi, float* od, double* od64, cudaStream_t s) {
    int nb = (nq + KD_THREADS - 1) / KD_THREADS;
    if (KD_STRATA > 1 && (perm != nullptr || !PCNN_KD_STRATA_ORDERED_ONLY)) nb = ((nb + KD_STRATA - 1) / KD_STRATA) * KD_STRATA;
    launch_pdl(knn_dfs_kernel<K>, dim3(nb), dim3(KD_THREADS), s, ix->opt_pdl != 0 && perm != nullptr,
               ix->view, q, nq, stride, perm, k, oi, od, od64, ix->dbg_warp_times);
}
#ifdef PCNN_STUDY
template <int K>
static void launch_refill_k(const pcnn_index* ix, const float* q, int nq, int stride, const unsigned* perm,
                            int k, int32_t* oi, float* od, double* od64, unsigned* counter, cudaStream_t s) {
    // persistent grid: what is resident at once; queries come from the global counter
    static int blocks = 0;
    if (blocks == 0) {
        int per_sm = 0, sms = 148;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, (const void*)knn_dfs_refill_kernel<K>, KT_THREADS, 0) != cudaSuccess || per_sm < 1)
            per_sm = 4;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, ix->device);
        blocks = per_sm * sms;
    }
    static int refill_min = getenv("PCNN_REFILL_MIN") ? atoi(getenv("PCNN_REFILL_MIN")) : PCNN_REFILL_MIN;  // study build only
    static int chunk = getenv("PCNN_REFILL_CHUNK") ? atoi(getenv("PCNN_REFILL_CHUNK")) : 0;
    const int need = chunk > 0 ? (nq + chunk * (KT_THREADS / 32) - 1) / (chunk * (KT_THREADS / 32)) : (nq + KT_THREADS - 1) / KT_THREADS;
    knn_dfs_refill_kernel<K><<<(chunk > 0 || need < blocks) ? need : blocks, KT_THREADS, 0, s>>>(
        ix->view, q, nq, stride, perm, k, oi, od, od64, counter, refill_min, chunk);
}
template <int K>
static void launch_hybrid_k(const pcnn_index* ix, const float* q, int nq, int stride, const unsigned* perm,
                            int k, int32_t* oi, float* od, double* od64, cudaStream_t s) {
    knn_hybrid_kernel<K><<<(nq + KT_THREADS - 1) / KT_THREADS, KT_THREADS, 0, s>>>(
        ix->view, q, nq, stride, perm, k, oi, od, od64,
        ix->opt_coherent_budget > 0 ? ix->opt_coherent_budget : 0x7fffffff, ix->dbg_warp_times);
}
#endif  // PCNN_STUDY
template <int K>
static void launch_brute_k(const pcnn_index* ix, const float* q, int nq, int stride, int k, int32_t* oi,
                           float* od, double* od64, cudaStream_t s) {
    knn_brute_kernel<K><<<(nq + BF_THREADS - 1) / BF_THREADS, BF_THREADS, 0, s>>>(
        ix->pts, (int)ix->n, q, nq, stride, k, oi, od, od64);
}

int launch_knn(const pcnn_index* ix, const float* q, int64_t nq64, int stride, int k, int32_t* out_idx,
               float* out_dist, double* out_dist64, cudaStream_t s, int sort_override) {
    if (nq64 == 0) return PCNN_OK;
    const int sortq = sort_override >= 0 ? sort_override : ix->opt_sort_queries;
    NvtxRange nvtx_(ix, "pcnn.knn");
    const int nq = (int)nq64;
    int method = ix->opt_method;
#ifndef PCNN_STUDY
    // the shipped library holds methods 2, 3, 7 and 9 only (pcnn_set_option refuses the others)
    if (method != 2 && method != 3 && method != 7 && method != 9) method = 0;
#endif
    if (method == 0) method = (k <= 16) ? 7 : (k <= KH_CAP ? 9 : 2);
    if (method == 9 && (k > KH_CAP || ix->n == 0)) method = 2;
    if ((method == 1 || method == 3 || method == 4 || method == 5 || method == 6 || method == 7 || method == 8 || method == 10) && k > 16) method = 2;
    if (method == 10 && ix->n == 0) method = 7;
    if (method == 3 && ix->n == 0) method = 1;
    if (method == 9) {
        Scratch sc(s);
        unsigned* perm = nullptr;
        if (sortq && nq > 32 && ix->n > 0) {
            int rc = sort_queries(ix, q, nq, stride, &perm, sc, s);
            if (rc != PCNN_OK) return rc;
        }
        prof_begin(ix, s);
        int nbh = (nq + KH_THREADS - 1) / KH_THREADS;
        if (KH_STRATA > 1 && perm != nullptr) nbh = ((nbh + KH_STRATA - 1) / KH_STRATA) * KH_STRATA;
        knn_dfs_heap_kernel<<<nbh, KH_THREADS, 0, s>>>(
            ix->view, q, nq, stride, perm, k, out_idx, out_dist, out_dist64);
        prof_end(ix, s);
#ifdef PCNN_STUDY
    } else if (method == 8) {
        Scratch sc(s);
        unsigned* perm = nullptr;
        if (sortq && nq > 32 && ix->n > 0) {
            int rc = sort_queries(ix, q, nq, stride, &perm, sc, s);
            if (rc != PCNN_OK) return rc;
        }
        prof_begin(ix, s);
        if (k <= 1) launch_hybrid_k<1>(ix, q, nq, stride, perm, k, out_idx, out_dist, out_dist64, s);
        else if (k <= 2) launch_hybrid_k<2>(ix, q, nq, stride, perm, k, out_idx, out_dist, out_dist64, s);
        else if (k <= 4) launch_hybrid_k<4>(ix, q, nq, stride, perm, k, out_idx, out_dist, out_dist64, s);
        else if (k <= 8) launch_hybrid_k<8>(ix, q, nq, stride, perm, k, out_idx, out_dist, out_dist64, s);
        else launch_hybrid_k<16>(ix, q, nq, stride, perm, k, out_idx, out_dist, out_dist64, s);
        prof_end(ix, s);
    } else if (method == 10) {
        Scratch sc(s);
        unsigned* perm = nullptr;
        if (sortq && nq > 32 && ix->n > 0) {
            int rc = sort_queries(ix, q, nq, stride, &perm, sc, s);
            if (rc != PCNN_OK) return rc;
        }
        unsigned* counter = nullptr;
        if (sc.get(&counter, 1) != cudaSuccess) { set_error("pcnn_knn: scratch allocation failed"); return PCNN_ERR_NOMEM; }
        PCNN_CUDA(cudaMemsetAsync(counter, 0, sizeof(unsigned), s));
        prof_begin(ix, s);
        if (k <= 1) launch_refill_k<1>(ix, q, nq, stride, perm, k, out_idx, out_dist, out_dist64, counter, s);
        else if (k <= 2) launch_refill_k<2>(ix, q, nq, stride, perm, k, out_idx, out_dist, out_dist64, counter, s);
        else if (k <= 4) launch_refill_k<4>(ix, q, nq, stride, perm, k, out_idx, out_dist, out_dist64, counter, s);
        else if (k <= 8) launch_refill_k<8>(ix, q, nq, stride, perm, k, out_idx, out_dist, out_dist64, counter, s);
        else launch_refill_k<16>(ix, q, nq, stride, perm, k, out_idx, out_dist, out_dist64, counter, s);
        prof_end(ix, s);
#endif  // PCNN_STUDY
    } else if (method == 7) {
        Scratch sc(s);
        unsigned* perm = nullptr;
        if (sortq && nq > 32 && ix->n > 0) {
            int rc = sort_queries(ix, q, nq, stride, &perm, sc, s);
            if (rc != PCNN_OK) return rc;
        }
        prof_begin(ix, s);
        if (k <= 1) launch_dfs_k<1>(ix, q, nq, stride, perm, k, out_idx, out_dist, out_dist64, s);
        else if (k <= 2) launch_dfs_k<2>(ix, q, nq, stride, perm, k, out_idx, out_dist, out_dist64, s);
        else if (k <= 4) launch_dfs_k<4>(ix, q, nq, stride, perm, k, out_idx, out_dist, out_dist64, s);
        else if (k <= 8) launch_dfs_k<8>(ix, q, nq, stride, perm, k, out_idx, out_dist, out_dist64, s);
        else if (k <= 12) launch_dfs_k<12>(ix, q, nq, stride, perm, k, out_idx, out_dist, out_dist64, s);  // 9..12: pca_normal.py's max_nn = 10
        else launch_dfs_k<16>(ix, q, nq, stride, perm, k, out_idx, out_dist, out_dist64, s);
        prof_end(ix, s);
#ifdef PCNN_STUDY
    } else if (method == 6) {
        Scratch sc(s);
        unsigned* perm = nullptr;
        if (sortq && nq > 32 && ix->n > 0) {
            int rc = sort_queries(ix, q, nq, stride, &perm, sc, s);
            if (rc != PCNN_OK) return rc;
        }
        prof_begin(ix, s);
        if (k <= 1) launch_lockstep_k<1>(ix, q, nq, stride, perm, k, out_idx, out_dist, out_dist64, s);
        else if (k <= 2) launch_lockstep_k<2>(ix, q, nq, stride, perm, k, out_idx, out_dist, out_dist64, s);
        else if (k <= 4) launch_lockstep_k<4>(ix, q, nq, stride, perm, k, out_idx, out_dist, out_dist64, s);
        else if (k <= 8) launch_lockstep_k<8>(ix, q, nq, stride, perm, k, out_idx, out_dist, out_dist64, s);
        else launch_lockstep_k<16>(ix, q, nq, stride, perm, k, out_idx, out_dist, out_dist64, s);
        prof_end(ix, s);
    } else if (method == 5) {
        Scratch sc(s);
        unsigned* perm = nullptr;
        if (sortq && nq > 32 && ix->n > 0) {
            int rc = sort_queries(ix, q, nq, stride, &perm, sc, s);
            if (rc != PCNN_OK) return rc;
        }
        prof_begin(ix, s);
        if (k <= 1) launch_queue_k<1>(ix, q, nq, stride, perm, k, out_idx, out_dist, out_dist64, s);
        else if (k <= 2) launch_queue_k<2>(ix, q, nq, stride, perm, k, out_idx, out_dist, out_dist64, s);
        else if (k <= 4) launch_queue_k<4>(ix, q, nq, stride, perm, k, out_idx, out_dist, out_dist64, s);
        else if (k <= 8) launch_queue_k<8>(ix, q, nq, stride, perm, k, out_idx, out_dist, out_dist64, s);
        else launch_queue_k<16>(ix, q, nq, stride, perm, k, out_idx, out_dist, out_dist64, s);
        prof_end(ix, s);
    } else if (method == 4) {
        Scratch sc(s);
        unsigned* perm = nullptr;
        if (sortq && nq > 32 && ix->n > 0) {
            int rc = sort_queries(ix, q, nq, stride, &perm, sc, s);
            if (rc != PCNN_OK) return rc;
        }
        // queries whose warp exceeds the sweep budget are finished by the nearest-first kernel
        int *qlist = nullptr, *qcount = nullptr;
        if (ix->opt_coherent_budget > 0 && ix->opt_coherent_budget2 > 0) {
            PCNN_CUDA(sc.get(&qlist, (size_t)nq + 32));
            PCNN_CUDA(sc.get(&qcount, 4));
            PCNN_CUDA(cudaMemsetAsync(qcount, 0, sizeof(int), s));
        }
        prof_begin(ix, s);
        if (k <= 1) launch_coherent_k<1>(ix, q, nq, stride, perm, k, out_idx, out_dist, out_dist64, qlist, qcount, s);
        else if (k <= 2) launch_coherent_k<2>(ix, q, nq, stride, perm, k, out_idx, out_dist, out_dist64, qlist, qcount, s);
        else if (k <= 4) launch_coherent_k<4>(ix, q, nq, stride, perm, k, out_idx, out_dist, out_dist64, qlist, qcount, s);
        else if (k <= 8) launch_coherent_k<8>(ix, q, nq, stride, perm, k, out_idx, out_dist, out_dist64, qlist, qcount, s);
        else launch_coherent_k<16>(ix, q, nq, stride, perm, k, out_idx, out_dist, out_dist64, qlist, qcount, s);
        if (qlist) {
            const WarpPlan wp = make_warp_plan(ix->view);
            int nblk = (nq + WQ_WARPS - 1) / WQ_WARPS;
            if (nblk > 148 * 4) nblk = 148 * 4;
            knn_warp_list_kernel<<<nblk, WQ_WARPS * 32, 0, s>>>(ix->view, wp, q, stride, k, qlist, qcount, out_idx,
                                                              out_dist, out_dist64);
        }
        prof_end(ix, s);
    } else if (method == 1) {
        Scratch sc(s);
        unsigned* perm = nullptr;
        if (sortq && nq > 1024 && ix->n > 0) {
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
#endif  // PCNN_STUDY
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
