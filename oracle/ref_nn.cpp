// oracle/ref_nn.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT.
//
// Thin C-ABI driver around the *unmodified* reference headers, included in place from
// /root/reference/Homework2/hw2/include (never copied into this repo).  The build recipe
// (oracle/Makefile, target _ref) compiles this file to oracle/_ref/libpcnn_ref.so.  Only
// tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
// load that library; the product (libpcnn.so) never links or calls it.
//
// What it drives (reference file:line):
//   KDTreeConstruction        Homework2/hw2/include/kdtree.hpp:419-429
//   KDTreeKNNSearch           kdtree.hpp:329-364
//   KDTreeRadiusNNSearch      kdtree.hpp:367-402
//   OctreeConstruction        octree.hpp:242-271
//   OctreeKNNSearch           octree.hpp:143-186
//   OctreeRadiusNNSearch      octree.hpp:189-237
//   KNNResultSet / RadiusNNResultSet   resultSet.hpp:28-93 / :96-142
// The per-query loop is the one at benchmark.hpp:59-72 (construct a result set, search),
// optionally split over host threads (the tree is read-only while searching).
//
// Must be compiled as exactly one translation unit with -std=c++14 (the headers define
// non-inline functions and use the `register` keyword, kdtree.hpp:58).
#include <cstdint>
#include <cstring>
#include <thread>
#include <vector>
#include <algorithm>

#include "resultSet.hpp"
#include "kdtree.hpp"
#include "octree.hpp"

namespace {

typedef std::vector<std::vector<double>> Db;

struct RefTree {
    int kind;  // 0 = kd-tree, 1 = octree
    Db db;
    Node* kd_root = nullptr;
    Octant* oc_root = nullptr;
    // last radius-search result (CSR), kept so the caller can size its buffers
    std::vector<int64_t> offsets;
    std::vector<int32_t> hit_idx;
    std::vector<double> hit_dist;
};

Db widen(const float* xyz, int64_t n, int stride) {
    // float32 -> double is lossless; mirrors readBinary (test.hpp:24-29) dropping intensity.
    Db db((size_t)n);
    for (int64_t i = 0; i < n; i++) {
        const float* p = xyz + i * stride;
        db[(size_t)i] = std::vector<double>{(double)p[0], (double)p[1], (double)p[2]};
    }
    return db;
}

// The reference frees every tree ever built from a process-global vector and never clears
// it (kdtree.hpp:431-435).  This driver owns one tree per handle and releases it by walking
// the handle's own nodes, then forgets the global bookkeeping.
void free_kd(Node* n) {
    if (!n) return;
    free_kd(n->left);
    free_kd(n->right);
    delete n;
}
void free_oc(Octant* o) {
    if (!o) return;
    for (auto* c : o->children) free_oc(c);
    delete o;
}

template <class F>
void parallel_for(int64_t n, int nthreads, F f) {
    if (nthreads <= 1 || n < 2) {
        f(0, n);
        return;
    }
    std::vector<std::thread> th;
    int64_t chunk = (n + nthreads - 1) / nthreads;
    for (int t = 0; t < nthreads; t++) {
        int64_t lo = t * chunk, hi = std::min<int64_t>(n, lo + chunk);
        if (lo >= hi) break;
        th.emplace_back([=] { f(lo, hi); });
    }
    for (auto& t : th) t.join();
}

}  // namespace

extern "C" {

void* pcref_build(int kind, const float* xyz, int64_t n, int stride, int leaf_size,
                  double min_extent) {
    RefTree* t = new RefTree;
    t->kind = kind;
    t->db = widen(xyz, n, stride);
    if (kind == 0) {
        t->kd_root = KDTreeConstruction(t->db, leaf_size);
        Node::address_set.clear();
    } else {
        t->oc_root = OctreeConstruction(t->db, leaf_size, min_extent);
        Octant::address_set.clear();
    }
    return t;
}

void pcref_destroy(void* h) {
    RefTree* t = (RefTree*)h;
    if (!t) return;
    free_kd(t->kd_root);
    free_oc(t->oc_root);
    delete t;
}

// kNN for nq queries.  Rows come back exactly as the reference's KNNResultSet holds them
// (ascending distance, ties in visit order, unfilled slots = (1e10, 0)).
// out_cmp (optional) receives comparisionCount per query.
int pcref_knn(void* h, const float* q, int64_t nq, int stride, int k, int32_t* out_idx,
              double* out_dist, int32_t* out_cmp, int nthreads) {
    RefTree* t = (RefTree*)h;
    parallel_for(nq, nthreads, [&](int64_t lo, int64_t hi) {
        for (int64_t i = lo; i < hi; i++) {
            const float* qp = q + i * stride;
            std::vector<double> query{(double)qp[0], (double)qp[1], (double)qp[2]};
            KNNResultSet rs(k);
            if (t->kind == 0)
                KDTreeKNNSearch(t->kd_root, t->db, rs, query);
            else
                OctreeKNNSearch(t->oc_root, t->db, rs, query);
            for (int j = 0; j < k; j++) {
                out_idx[i * k + j] = rs.distIndexList[j].index;
                out_dist[i * k + j] = rs.distIndexList[j].distance;
            }
            if (out_cmp) out_cmp[i] = rs.comparisionCount;
        }
    });
    return 0;
}

// Radius search for nq queries; results are kept inside the handle in the reference's own
// (traversal) order.  Returns the total hit count and fills offsets[nq+1].
int64_t pcref_radius(void* h, const float* q, int64_t nq, int stride, double radius,
                     int64_t* out_offsets, int nthreads) {
    RefTree* t = (RefTree*)h;
    std::vector<std::vector<DistIndex>> rows((size_t)nq);
    parallel_for(nq, nthreads, [&](int64_t lo, int64_t hi) {
        for (int64_t i = lo; i < hi; i++) {
            const float* qp = q + i * stride;
            std::vector<double> query{(double)qp[0], (double)qp[1], (double)qp[2]};
            RadiusNNResultSet rs(radius);
            if (t->kind == 0)
                KDTreeRadiusNNSearch(t->kd_root, t->db, rs, query);
            else
                OctreeRadiusNNSearch(t->oc_root, t->db, rs, query);
            rows[(size_t)i].swap(rs.distIndexList);
        }
    });
    t->offsets.assign((size_t)nq + 1, 0);
    for (int64_t i = 0; i < nq; i++) t->offsets[i + 1] = t->offsets[i] + (int64_t)rows[i].size();
    int64_t total = t->offsets[nq];
    t->hit_idx.resize((size_t)total);
    t->hit_dist.resize((size_t)total);
    for (int64_t i = 0; i < nq; i++) {
        int64_t o = t->offsets[i];
        for (size_t j = 0; j < rows[i].size(); j++) {
            t->hit_idx[o + j] = rows[i][j].index;
            t->hit_dist[o + j] = rows[i][j].distance;
        }
    }
    if (out_offsets) std::memcpy(out_offsets, t->offsets.data(), sizeof(int64_t) * (nq + 1));
    return total;
}

void pcref_radius_fetch(void* h, int32_t* out_idx, double* out_dist) {
    RefTree* t = (RefTree*)h;
    if (out_idx) std::memcpy(out_idx, t->hit_idx.data(), sizeof(int32_t) * t->hit_idx.size());
    if (out_dist) std::memcpy(out_dist, t->hit_dist.data(), sizeof(double) * t->hit_dist.size());
}

// Timing-only variants for the CPU baseline: same loops, results discarded except a checksum
// (so the optimiser cannot drop the search).  Returns the sum of hit counts / indices.
int64_t pcref_knn_checksum(void* h, const float* q, int64_t nq, int stride, int k, int nthreads) {
    RefTree* t = (RefTree*)h;
    std::vector<int64_t> part((size_t)std::max(1, nthreads), 0);
    int64_t chunk = (nq + std::max(1, nthreads) - 1) / std::max(1, nthreads);
    parallel_for(nq, nthreads, [&](int64_t lo, int64_t hi) {
        int64_t s = 0;
        for (int64_t i = lo; i < hi; i++) {
            const float* qp = q + i * stride;
            std::vector<double> query{(double)qp[0], (double)qp[1], (double)qp[2]};
            KNNResultSet rs(k);
            if (t->kind == 0)
                KDTreeKNNSearch(t->kd_root, t->db, rs, query);
            else
                OctreeKNNSearch(t->oc_root, t->db, rs, query);
            for (int j = 0; j < k; j++) s += rs.distIndexList[j].index;
        }
        part[(size_t)(chunk ? lo / chunk : 0)] = s;
    });
    int64_t s = 0;
    for (auto v : part) s += v;
    return s;
}

int64_t pcref_radius_checksum(void* h, const float* q, int64_t nq, int stride, double radius,
                              int nthreads) {
    RefTree* t = (RefTree*)h;
    std::vector<int64_t> part((size_t)std::max(1, nthreads), 0);
    int64_t chunk = (nq + std::max(1, nthreads) - 1) / std::max(1, nthreads);
    parallel_for(nq, nthreads, [&](int64_t lo, int64_t hi) {
        int64_t s = 0;
        for (int64_t i = lo; i < hi; i++) {
            const float* qp = q + i * stride;
            std::vector<double> query{(double)qp[0], (double)qp[1], (double)qp[2]};
            RadiusNNResultSet rs(radius);
            if (t->kind == 0)
                KDTreeRadiusNNSearch(t->kd_root, t->db, rs, query);
            else
                OctreeRadiusNNSearch(t->oc_root, t->db, rs, query);
            s += rs.size();
        }
        part[(size_t)(chunk ? lo / chunk : 0)] = s;
    });
    int64_t s = 0;
    for (auto v : part) s += v;
    return s;
}

int pcref_hardware_threads() { return (int)std::thread::hardware_concurrency(); }

}  // extern "C"
