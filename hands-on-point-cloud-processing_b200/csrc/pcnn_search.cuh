// pcnn_search.cuh -- per-query traversal of the implicit AABB hierarchy (thread-per-query
// form).  Compiles for host and device so the exact same logic can be exercised on the CPU
// build box (tests/hostsim) before it meets a GPU.
//
// Replaces the recursive searches of the reference:
//   KDTreeKNNSearch  Homework2/hw2/include/kdtree.hpp:329-364
//   OctreeKNNSearch  Homework2/hw2/include/octree.hpp:143-186
//   KNNResultSet::addPoint (sorted insertion)  resultSet.hpp:65-91
//   nanoflann searchLevel, k = 1, float32      Homework9/hw9/include/nanoflann.hpp:1347-1410
//
// Traversal is stackless: the hierarchy is a complete binary tree over 8-point groups in
// Morton order, so "skip this subtree" is index arithmetic (i = (node+1) << level) and the
// next level to test is ctz(i).  A home-group scan (binary search of the query's Morton
// code) seeds the k-th distance bound before the ordered sweep starts.
#pragma once
#include "pcnn_common.cuh"

namespace pcnn {

// host-only work counters for tests/hostsim (algorithmic cost studies); no-ops in the product
#if defined(PCNN_HOST_COUNTERS) && !defined(__CUDACC__)
extern thread_local long g_cnt_box, g_cnt_pt;
#define PCNN_COUNT_BOX() (++pcnn::g_cnt_box)
#define PCNN_COUNT_PT() (++pcnn::g_cnt_pt)
#else
#define PCNN_COUNT_BOX() ((void)0)
#define PCNN_COUNT_PT() ((void)0)
#endif

#define PCNN_INF_D (__longlong_as_double_compat(0x7ff0000000000000LL))

PCNN_HD double __longlong_as_double_compat(long long v) {
#if defined(__CUDA_ARCH__)
    return __longlong_as_double(v);
#else
    union { long long i; double d; } u; u.i = v; return u.d;
#endif
}

PCNN_HD int ctz32(unsigned v) {
#if defined(__CUDA_ARCH__)
    return __ffs((int)v) - 1;
#else
    return __builtin_ctz(v);
#endif
}

PCNN_HD float4 ldpt(const float4* p) {
#if defined(__CUDA_ARCH__)
    return __ldg(p);
#else
    return *p;
#endif
}

// Sorted list of the K best (key, index) pairs, ascending; lives in registers.
template <int K>
struct TopK {
    double key[K];
    int idx[K];
    PCNN_HD void init() {
#pragma unroll
        for (int j = 0; j < K; j++) { key[j] = PCNN_INF_D; idx[j] = 0x7fffffff; }
    }
    PCNN_HD double worst() const { return key[K - 1]; }
    PCNN_HD bool accepts(double d, int i) const { return kless(d, i, key[K - 1], idx[K - 1]); }
    // precondition: accepts(d, i)
    PCNN_HD void insert(double d, int i) {
        key[K - 1] = d;
        idx[K - 1] = i;
#pragma unroll
        for (int j = K - 1; j > 0; --j) {
            bool sw = kless(key[j], idx[j], key[j - 1], idx[j - 1]);
            double ka = key[j - 1], kb = key[j];
            int ia = idx[j - 1], ib = idx[j];
            key[j - 1] = sw ? kb : ka;
            idx[j - 1] = sw ? ib : ia;
            key[j] = sw ? ka : kb;
            idx[j] = sw ? ia : ib;
        }
    }
};

// float32 single-best variant (nanoflann k = 1 as ICP uses it)
struct Best1f {
    float key;
    int idx;  // original index
    int pos;  // position in Morton order (lets callers fetch per-point attributes)
    PCNN_HD void init() { key = i2f(0x7f800000); idx = 0x7fffffff; pos = -1; }
    PCNN_HD void start(float b) { init(); key = b; }
    PCNN_HD float bound() const { return key; }
    PCNN_HD void offer(float d, int id, int p) {
        if (kless32(d, id, key, idx)) { key = d; idx = id; pos = p; }
    }
};
// The nearest point, ONE other point remembered by position, and a FLOOR under everybody else:
// once the search has ended, every point other than `pos` and `pos2` has key >= floor() (examined
// ones by comparison, pruned ones because their box bound exceeded the pruning key, which only
// falls).  ICP turns that into a certificate that lets later iterations keep the correspondence
// without searching (pcnn_icp.cu).  The floor is only chased up to (sqrt(key) + gap)^2: a caller
// that expects its query to move by less than `gap` needs no more, and the walk then costs little
// more than a plain 1-NN walk (gap < 0: exactly that, prune == key, no certificate).  Remembering
// the runner-up by position is what makes near-ties certifiable: two target points within 1e-5 m
// of equidistant leave no room between key and key2, but plenty between key2 and the third.
struct Best2f {
    float key;
    int idx;
    int pos;
    float key2;   // runner-up (smallest key seen among the others; ties broken arbitrarily)
    int pos2;
    float lim;    // smallest key seen outside the two above
    float prune;  // pruning key = min(lim, (sqrt(key)(1 + 1e-4) + gap)^2)
    float gap;
    PCNN_HD void start(float b) { key = b; idx = 0x7fffffff; pos = -1; key2 = b; pos2 = -1; lim = b; prune = b; }
    PCNN_HD float bound() const { return prune; }
    PCNN_HD float floor_key() const { return prune; }
#if defined(__CUDACC__)
    __device__ __forceinline__ float reach(float k) const {
        if (gap < 0.f) return k;
        const float r = __fadd_ru(__fmul_ru(__fsqrt_ru(k), 1.0001f), gap);
        return __fmul_ru(r, r);
    }
    __device__ __forceinline__ void offer(float d, int id, int p) {
        if (kless32(d, id, key, idx)) {
            lim = fminf(lim, key2);  // the old runner-up joins the rest
            key2 = key; pos2 = pos;
            key = d; idx = id; pos = p;
            prune = fminf(lim, reach(d));
        } else if (d < key2) {
            lim = fminf(lim, key2);
            key2 = d; pos2 = p;
            prune = fminf(prune, lim);
        } else if (d < lim) {
            lim = d;
            prune = fminf(prune, d);
        }
    }
#endif
};

// index of the base group whose code range holds qc: last g with gcode[g] <= qc (or 0)
PCNN_HD int home_group(const IndexView& iv, unsigned long long qc) {
    int lo = 0, hi = iv.n0;
    while (lo < hi) {
        int mid = (lo + hi) >> 1;
#if defined(__CUDA_ARCH__)
        unsigned long long c = __ldg(iv.gcode + mid);
#else
        unsigned long long c = iv.gcode[mid];
#endif
        if (c <= qc) lo = mid + 1; else hi = mid;
    }
    return lo > 0 ? lo - 1 : 0;
}

// first g with gcode[g] >= qc, minus one (clamped to 0): every group before it holds only
// codes < qc
PCNN_HD int group_before(const IndexView& iv, unsigned long long qc) {
    int lo = 0, hi = iv.n0;
    while (lo < hi) {
        int mid = (lo + hi) >> 1;
#if defined(__CUDA_ARCH__)
        unsigned long long c = __ldg(iv.gcode + mid);
#else
        unsigned long long c = iv.gcode[mid];
#endif
        if (c < qc) lo = mid + 1; else hi = mid;
    }
    return lo > 0 ? lo - 1 : 0;
}

// Lowest level at which a walk for (q, bound) may start instead of the root.  Every point within
// sqrt(bound_f) of q lies componentwise inside [q - r, q + r]; the Morton code is monotone in
// each coordinate, so its code lies in [code(q - r), code(q + r)].  The ancestor of group h at
// level L holds the groups [a, a + 2^L): if every point before it has a smaller code than the
// window and every point after it a larger one, nothing outside that subtree can matter, and
// neither the descent from the root down to it nor its ancestors' siblings need a visit.
// (Loads are independent across levels: one latency, not a chain.)  Returns `top` when the
// window straddles a coarse cell boundary, or when there is no bound yet.
#if defined(__CUDACC__)
__device__ __forceinline__ int walk_start_level(const IndexView& iv, int h, float qxf, float qyf, float qzf,
                                                float bound_f) {
    const int top = iv.nlev - 1;
    if (!(bound_f < 3.0e38f)) return top;
    const float r = __fmul_ru(__fsqrt_ru(bound_f), 1.000001f);
    const unsigned long long wlo = morton63(__fsub_rd(qxf, r), __fsub_rd(qyf, r), __fsub_rd(qzf, r), iv.cube, iv.cube[3]);
    const unsigned long long whi = morton63(__fadd_ru(qxf, r), __fadd_ru(qyf, r), __fadd_ru(qzf, r), iv.cube, iv.cube[3]);
    int Ls = top;
    for (int L = top - 1; L >= 1; L--) {  // keeps the LOWEST level that qualifies
        const int a = (h >> L) << L;
        const long long b = (long long)a + (1LL << L);
        const bool ok_lo = a == 0 || __ldg(iv.gcode + a) < wlo;
        const bool ok_hi = b >= (long long)iv.n0 || __ldg(iv.gcode + b) > whi;
        if (ok_lo && ok_hi) Ls = L;
    }
    return Ls;
}
#endif

// Scan one 8-point group.  A float32 lower bound of d2 rejects most candidates on the FP32
// pipe; only survivors pay for the exact double key.  wf = ru32(tk.worst()) is kept current.
template <int K>
PCNN_HD void scan_group64(const IndexView& iv, int g, float qxf, float qyf, float qzf, double qx,
                          double qy, double qz, TopK<K>& tk, float& wf) {
    const int first = g << PCNN_GROUP_LOG;
    const int cnt = (iv.n - first) < PCNN_GROUP ? (iv.n - first) : PCNN_GROUP;
#pragma unroll
    for (int j = 0; j < PCNN_GROUP; j++) {
        if (j < cnt) {
            const float4 p = ldpt(iv.pts + first + j);
            PCNN_COUNT_PT();
            const float lbf = lbsq32(absdiff_rz(p.x, qxf), absdiff_rz(p.y, qyf), absdiff_rz(p.z, qzf));
            if (!(lbf > wf)) {
                const double d = key64(p.x, p.y, p.z, qx, qy, qz);
                const int id = f2i(p.w);
                if (d <= tk.worst() && tk.accepts(d, id)) {
                    tk.insert(d, id);
                    wf = ru32(tk.worst());
                }
            }
        }
    }
}

// kNN of one query; leaf_log = log2(groups per leaf), 0 <= leaf_log < nlev.
//  1. home scan: the group holding the query's Morton code and its two neighbours seed the
//     k-th distance bound r;
//  2. Morton window: every point within r of q lies componentwise inside [q-r, q+r], and the
//     Morton code is monotone in each coordinate, so its code lies in [code(q-r), code(q+r)]:
//     two binary searches turn that into a group range [g_lo, g_hi] and the sweep never
//     looks outside it (no walk down from the root);
//  3. stackless sweep of that range: float32 conservative box tests prune subtrees
//     (i = (node+1) << l), leaves are scanned with the exact double key.
template <int K>
PCNN_HD void knn_thread(const IndexView& iv, float qxf, float qyf, float qzf, int leaf_log,
                        TopK<K>& tk) {
    tk.init();
    if (iv.n <= 0) return;
    const double qx = (double)qxf, qy = (double)qyf, qz = (double)qzf;
    const int h = home_group(iv, morton63(qxf, qyf, qzf, iv.cube, iv.cube[3]));
    const int h0 = h > 0 ? h - 1 : 0, h1 = (h + 1 < iv.n0) ? h + 1 : iv.n0 - 1;
    float wf = i2f(0x7f800000);
    for (int g = h0; g <= h1; g++) scan_group64<K>(iv, g, qxf, qyf, qzf, qx, qy, qz, tk, wf);

    const int top = iv.nlev - 1;
    if (leaf_log > top) leaf_log = top;
    int g_lo = 0, g_hi = iv.n0 - 1;
    if (wf < 3.0e38f) {
        // r >= sqrt(worst): float sqrt of an upper bound, nudged up
#if defined(__CUDA_ARCH__)
        const float r = __fmul_ru(__fsqrt_ru(wf), 1.000001f);
        const float lx = __fsub_rd(qxf, r), ly = __fsub_rd(qyf, r), lz = __fsub_rd(qzf, r);
        const float hx = __fadd_ru(qxf, r), hy = __fadd_ru(qyf, r), hz = __fadd_ru(qzf, r);
#else
        const float r = (float)(std::sqrt((double)wf) * 1.000002);
        const float lx = i2f_host_prev(qxf - r), ly = i2f_host_prev(qyf - r), lz = i2f_host_prev(qzf - r);
        const float hx = i2f_host_next(qxf + r), hy = i2f_host_next(qyf + r), hz = i2f_host_next(qzf + r);
#endif
        g_lo = group_before(iv, morton63(lx, ly, lz, iv.cube, iv.cube[3]));
        g_hi = home_group(iv, morton63(hx, hy, hz, iv.cube, iv.cube[3]));
    }
    int i = g_lo;
    int l = i == 0 ? top : ctz32((unsigned)i);
    if (l > top) l = top;
    while (i <= g_hi) {
        // phase A: walk box tests until a leaf must be scanned
        int leaf = -1;
        while (i <= g_hi) {
            const int node = i >> l;
            const float4* b = iv.box + 2 * ((size_t)iv.lev_off[l] + node);
            const float4 lo = ldpt(b), hi = ldpt(b + 1);
            PCNN_COUNT_BOX();
            const float lbf = lbsq32(boxgap_rd(lo.x, hi.x, qxf), boxgap_rd(lo.y, hi.y, qyf),
                                     boxgap_rd(lo.z, hi.z, qzf));
            if (lbf > wf) {                       // nothing inside can enter (ties included)
                i = (node + 1) << l;
                const int a = ctz32((unsigned)i); // highest level whose node starts at i
                l = a < top ? a : top;
            } else if (l > leaf_log) {
                l -= 1;                           // test the left child next (same i)
            } else {
                leaf = node;
                break;
            }
        }
        if (leaf < 0) break;
        // phase B: scan the leaf
        const int g0 = leaf << l;
        int g1 = (leaf + 1) << l;
        if (g1 > iv.n0) g1 = iv.n0;
        for (int g = g0; g < g1; g++)
            if (g < h0 || g > h1) scan_group64<K>(iv, g, qxf, qyf, qzf, qx, qy, qz, tk, wf);
        i = g1;
        if (i >= iv.n0) break;
        const int a = ctz32((unsigned)i);
        l = a < top ? a : top;
    }
}

// float32 1-NN (ICP correspondence); lowest original index wins an exact tie.
PCNN_HD void nn1_thread(const IndexView& iv, float qx, float qy, float qz, int leaf_log,
                        Best1f& best) {
    best.init();
    if (iv.n <= 0) return;
    const int h = home_group(iv, morton63(qx, qy, qz, iv.cube, iv.cube[3]));
    {
        const int first = h << PCNN_GROUP_LOG;
        const int cnt = (iv.n - first) < PCNN_GROUP ? (iv.n - first) : PCNN_GROUP;
        for (int j = 0; j < cnt; j++) {
            float4 p = ldpt(iv.pts + first + j);
            float d = key32(p.x, p.y, p.z, qx, qy, qz);
            int id = f2i(p.w);
            if (kless32(d, id, best.key, best.idx)) { best.key = d; best.idx = id; best.pos = first + j; }
        }
    }
    const int top = iv.nlev - 1;
    if (leaf_log > top) leaf_log = top;
    int i = 0, l = top;
    while (i < iv.n0) {
        const int node = i >> l;
        const float4* b = iv.box + 2 * ((size_t)iv.lev_off[l] + node);
        const float4 lo = ldpt(b), hi = ldpt(b + 1);
        const float lb = boxlb32(lo, hi, qx, qy, qz);
        if (lb > best.key) {
            i = (node + 1) << l;
        } else if (l > leaf_log) {
            l -= 1;
            continue;
        } else {
            const int g0 = node << l;
            int g1 = (node + 1) << l;
            if (g1 > iv.n0) g1 = iv.n0;
            for (int g = g0; g < g1; g++) {
                if (g == h) continue;
                const int first = g << PCNN_GROUP_LOG;
                const int cnt = (iv.n - first) < PCNN_GROUP ? (iv.n - first) : PCNN_GROUP;
#pragma unroll
                for (int j = 0; j < PCNN_GROUP; j++) {
                    if (j < cnt) {
                        float4 p = ldpt(iv.pts + first + j);
                        float d = key32(p.x, p.y, p.z, qx, qy, qz);
                        int id = f2i(p.w);
                        if (kless32(d, id, best.key, best.idx)) { best.key = d; best.idx = id; best.pos = first + j; }
                    }
                }
            }
            i = g1 >= iv.n0 ? iv.n0 : (node + 1) << l;
        }
        if (i >= iv.n0) break;
        const int a = ctz32((unsigned)i);
        l = a < top ? a : top;
    }
}

}  // namespace pcnn
