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

template <int K>
PCNN_HD void scan_group64(const IndexView& iv, int g, double qx, double qy, double qz, TopK<K>& tk) {
    const int first = g << PCNN_GROUP_LOG;
    const int cnt = (iv.n - first) < PCNN_GROUP ? (iv.n - first) : PCNN_GROUP;
#pragma unroll
    for (int j = 0; j < PCNN_GROUP; j++) {
        if (j < cnt) {
            float4 p = ldpt(iv.pts + first + j);
            double d = key64(p.x, p.y, p.z, qx, qy, qz);
            int id = f2i(p.w);
            if (d <= tk.worst() && tk.accepts(d, id)) tk.insert(d, id);
        }
    }
}

// kNN of one query; leaf_log = log2(groups per leaf), 0 <= leaf_log < nlev.
template <int K>
PCNN_HD void knn_thread(const IndexView& iv, float qxf, float qyf, float qzf, int leaf_log,
                        TopK<K>& tk) {
    tk.init();
    if (iv.n <= 0) return;
    const double qx = (double)qxf, qy = (double)qyf, qz = (double)qzf;
    const int h = home_group(iv, morton63(qxf, qyf, qzf, iv.cube, iv.cube[3]));
    const int h0 = h > 0 ? h - 1 : 0, h1 = (h + 1 < iv.n0) ? h + 1 : iv.n0 - 1;
    for (int g = h0; g <= h1; g++) scan_group64<K>(iv, g, qx, qy, qz, tk);

    const int top = iv.nlev - 1;
    if (leaf_log > top) leaf_log = top;
    int i = 0, l = top;
    while (i < iv.n0) {
        const int node = i >> l;
        const float4* b = iv.box + 2 * ((size_t)iv.lev_off[l] + node);
        const float4 lo = ldpt(b), hi = ldpt(b + 1);
        const double lb = boxlb64(lo, hi, qx, qy, qz);
        if (lb > tk.worst()) {                // nothing inside can enter (ties included)
            i = (node + 1) << l;
        } else if (l > leaf_log) {
            l -= 1;                           // test the left child next (same i)
            continue;
        } else {
            const int g0 = node << l;
            int g1 = (node + 1) << l;
            if (g1 > iv.n0) g1 = iv.n0;
            for (int g = g0; g < g1; g++)
                if (g < h0 || g > h1) scan_group64<K>(iv, g, qx, qy, qz, tk);
            i = g1 >= iv.n0 ? iv.n0 : (node + 1) << l;
        }
        if (i >= iv.n0) break;
        const int a = ctz32((unsigned)i);     // highest level whose node starts at i
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
