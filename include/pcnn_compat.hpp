// pcnn_compat.hpp -- the reference's own C++ names on top of the C ABI (include/pcnn.h).
//
// A caller written against the reference's Homework2 headers
//   resultSet.hpp:7-18 (DistIndex), :28-93 (KNNResultSet), :96-142 (RadiusNNResultSet)
//   kdtree.hpp:419-435 (KDTreeConstruction / KDTreeDestruction), :329-364, :367-402 (searches)
//   octree.hpp:242-277 (OctreeConstruction / OctreeDestruction), :143-186, :189-237 (searches)
// compiles against this header instead and links libpcnn.so: same function names, same
// argument lists, same result-set members.  `Node` / `Octant` are handles to the device index
// instead of pointer trees; `db` is still passed to every search, as the reference does
// (kdtree.hpp:329-332), and ignored there because the index owns a copy of the cloud.
//
// The one-query functions are correct and slow (one launch per query).  The *Batch functions
// hoist the loop of benchmark.hpp:59-72 into one call: that is the fast path.
//
// Differences a caller can observe, all documented in DESIGN.md section 2:
//  - exact-distance ties straddling the k-th place are broken by point index, not visit order;
//  - RadiusNNResultSet lists hits in the index's order, not the reference's traversal order;
//  - comparisionCount counts the points handed to addPoint (k, or the hit count), not the
//    reference's number of distance evaluations;
//  - coordinates must be float32-representable doubles (every cloud the reference reads is:
//    test.hpp:24-29); anything else throws std::invalid_argument instead of being rounded;
//  - errors throw std::runtime_error with pcnn_last_error() instead of exit(1).
#ifndef PCNN_COMPAT_HPP
#define PCNN_COMPAT_HPP

#include <cstdint>
#include <iostream>
#include <stdexcept>
#include <string>
#include <vector>

#include "pcnn.h"

struct DistIndex {
    double distance;
    int index;
    DistIndex(double d, int i) : distance(d), index(i) {}
    bool operator<(const DistIndex& o) const { return distance < o.distance; }
};
inline std::ostream& operator<<(std::ostream& os, const DistIndex& v) {
    return os << "Distance = " << v.distance << ", Index = " << v.index;
}

// Bounded ascending list of capacity k; empty slots hold (1e10, 0) as in the reference.
class KNNResultSet {
    int capacity_;
    double worst_ = 1e10;

public:
    explicit KNNResultSet(int k) : capacity_(k), distIndexList((size_t)(k > 0 ? k : 0), DistIndex(1e10, 0)) {}
    int count = 0;
    int comparisionCount = 0;
    std::vector<DistIndex> distIndexList;
    int size() const { return capacity_; }
    double getWorstDist() const { return worst_; }
    void list() const {
        std::cout << "Distance-Index list: " << std::endl;
        for (const auto& e : distIndexList) std::cout << e << std::endl;
    }
    void addPoint(double dist, int index) {
        comparisionCount++;
        if (dist > worst_ || capacity_ <= 0) return;
        if (count < capacity_) count++;
        int pos = count - 1;
        for (; pos > 0 && distIndexList[(size_t)pos - 1].distance > dist; --pos)
            distIndexList[(size_t)pos] = distIndexList[(size_t)pos - 1];
        distIndexList[(size_t)pos] = DistIndex(dist, index);
        worst_ = distIndexList[(size_t)capacity_ - 1].distance;
    }
};

// Unbounded list of the hits with dist <= radius (inclusive).
class RadiusNNResultSet {
    double radius_;

public:
    explicit RadiusNNResultSet(double r) : radius_(r) {}
    int count = 0;
    int comparisionCount = 0;
    std::vector<DistIndex> distIndexList;
    int size() const { return count; }
    double getWorstDist() const { return radius_; }
    void list() const {
        std::cout << "Distance-Index list: " << std::endl;
        for (const auto& e : distIndexList) std::cout << e << std::endl;
    }
    void addPoint(double dist, int index) {
        comparisionCount++;
        if (dist <= radius_) {
            distIndexList.emplace_back(dist, index);
            count++;
        }
    }
};

namespace pcnn_compat {

inline void check(int rc, const char* what) {
    if (rc != PCNN_OK) throw std::runtime_error(std::string(what) + ": " + pcnn_last_error());
}

// vector<vector<double>> -> packed float32 xyz; lossless or it throws
inline std::vector<float> pack(const std::vector<std::vector<double>>& pts) {
    std::vector<float> out(pts.size() * 3);
    for (size_t i = 0; i < pts.size(); i++) {
        if (pts[i].size() < 3) throw std::invalid_argument("pcnn_compat: a point needs 3 coordinates");
        for (int a = 0; a < 3; a++) {
            const float f = (float)pts[i][(size_t)a];
            if ((double)f != pts[i][(size_t)a])
                throw std::invalid_argument("pcnn_compat: coordinate is not float32-representable");
            out[i * 3 + (size_t)a] = f;
        }
    }
    return out;
}
inline void pack1(const std::vector<double>& p, float out[3]) {
    std::vector<std::vector<double>> one(1, p);
    const std::vector<float> f = pack(one);
    out[0] = f[0]; out[1] = f[1]; out[2] = f[2];
}

struct Handle {
    pcnn_index* ix = nullptr;
};
// the reference frees "every tree built so far" from a global list (kdtree.hpp:431-435)
inline std::vector<Handle*>& registry(int kind) {
    static std::vector<Handle*> reg[2];
    return reg[kind];
}
inline Handle* construct(int kind, const std::vector<std::vector<double>>& db, int leaf_size, double min_extent) {
    const std::vector<float> xyz = pack(db);
    Handle* h = new Handle;
    const int rc = pcnn_index_build_host(&h->ix, kind, xyz.data(), (int64_t)db.size(), 3, leaf_size, min_extent, 0);
    if (rc != PCNN_OK) { delete h; check(rc, "pcnn_index_build_host"); }
    registry(kind).push_back(h);
    return h;
}
inline void destruct_all(int kind) {
    for (Handle* h : registry(kind)) { pcnn_index_destroy(h->ix); delete h; }
    registry(kind).clear();
}
inline void knn1(Handle* h, KNNResultSet& rs, const std::vector<double>& query) {
    const int k = rs.size();
    if (k <= 0) return;
    float q[3];
    pack1(query, q);
    std::vector<int32_t> idx((size_t)k);
    std::vector<double> dist((size_t)k);
    check(pcnn_knn_host(h->ix, q, 1, 3, k, idx.data(), nullptr, dist.data()), "pcnn_knn_host");
    for (int j = 0; j < k; j++)
        if (idx[(size_t)j] >= 0) rs.addPoint(dist[(size_t)j], idx[(size_t)j]);
}
inline void radius1(Handle* h, RadiusNNResultSet& rs, const std::vector<double>& query) {
    float q[3];
    pack1(query, q);
    int64_t off[2] = {0, 0};
    int rc = pcnn_radius_host(h->ix, q, 1, 3, rs.getWorstDist(), off, nullptr, nullptr, nullptr, 0);
    if (rc != PCNN_OK && rc != PCNN_ERR_CAPACITY) check(rc, "pcnn_radius_host");
    const int64_t n = off[1];
    if (n == 0) return;
    std::vector<int32_t> idx((size_t)n);
    std::vector<double> dist((size_t)n);
    check(pcnn_radius_host(h->ix, q, 1, 3, rs.getWorstDist(), off, idx.data(), nullptr, dist.data(), n), "pcnn_radius_host");
    for (int64_t j = 0; j < n; j++) rs.addPoint(dist[(size_t)j], idx[(size_t)j]);
}
// the loop of benchmark.hpp:59-66 in one call: rows of k (index, distance), ascending
inline void knn_batch(Handle* h, const std::vector<std::vector<double>>& queries, int k,
                      std::vector<int32_t>& idx, std::vector<double>& dist) {
    const std::vector<float> q = pack(queries);
    idx.assign(queries.size() * (size_t)k, -1);
    dist.assign(queries.size() * (size_t)k, 0.0);
    check(pcnn_knn_host(h->ix, q.data(), (int64_t)queries.size(), 3, k, idx.data(), nullptr, dist.data()), "pcnn_knn_host");
}
// the loop of benchmark.hpp:68-71 in one call: CSR (offsets[nq+1], idx, dist)
inline void radius_batch(Handle* h, const std::vector<std::vector<double>>& queries, double radius,
                         std::vector<int64_t>& offsets, std::vector<int32_t>& idx, std::vector<double>& dist) {
    const std::vector<float> q = pack(queries);
    const int64_t nq = (int64_t)queries.size();
    offsets.assign((size_t)nq + 1, 0);
    int rc = pcnn_radius_host(h->ix, q.data(), nq, 3, radius, offsets.data(), nullptr, nullptr, nullptr, 0);
    if (rc != PCNN_OK && rc != PCNN_ERR_CAPACITY) check(rc, "pcnn_radius_host");
    const int64_t n = offsets[(size_t)nq];
    idx.assign((size_t)n, 0);
    dist.assign((size_t)n, 0.0);
    if (n > 0)
        check(pcnn_radius_host(h->ix, q.data(), nq, 3, radius, offsets.data(), idx.data(), nullptr, dist.data(), n),
              "pcnn_radius_host");
}

}  // namespace pcnn_compat

// ---- the reference's names ------------------------------------------------------------
struct Node : pcnn_compat::Handle {};    // kdtree.hpp:10-35
struct Octant : pcnn_compat::Handle {};  // octree.hpp:10-34

inline Node* KDTreeConstruction(std::vector<std::vector<double>>& db, int leaf_size) {
    return static_cast<Node*>(pcnn_compat::construct(PCNN_KDTREE, db, leaf_size, 0.0));
}
inline void KDTreeDestruction() { pcnn_compat::destruct_all(PCNN_KDTREE); }
inline void KDTreeKNNSearch(Node*& root, std::vector<std::vector<double>>& /*db*/, KNNResultSet& result_set,
                            std::vector<double>& query) {
    pcnn_compat::knn1(root, result_set, query);
}
inline void KDTreeRadiusNNSearch(Node*& root, std::vector<std::vector<double>>& /*db*/,
                                 RadiusNNResultSet& result_set, std::vector<double>& query) {
    pcnn_compat::radius1(root, result_set, query);
}
inline Octant* OctreeConstruction(std::vector<std::vector<double>>& db, int leaf_size, double min_extent) {
    return static_cast<Octant*>(pcnn_compat::construct(PCNN_OCTREE, db, leaf_size, min_extent));
}
inline void OctreeDestruction() { pcnn_compat::destruct_all(PCNN_OCTREE); }
// the reference returns its early-exit flag (octree.hpp:143); a finished search reports true
inline bool OctreeKNNSearch(Octant*& root, std::vector<std::vector<double>>& /*db*/, KNNResultSet& result_set,
                            std::vector<double>& query) {
    pcnn_compat::knn1(root, result_set, query);
    return true;
}
inline bool OctreeRadiusNNSearch(Octant*& root, std::vector<std::vector<double>>& /*db*/,
                                 RadiusNNResultSet& result_set, std::vector<double>& query) {
    pcnn_compat::radius1(root, result_set, query);
    return true;
}

// ---- batched forms (the fast path) ----------------------------------------------------
inline void KDTreeKNNSearchBatch(Node* root, const std::vector<std::vector<double>>& queries, int k,
                                 std::vector<int32_t>& idx, std::vector<double>& dist) {
    pcnn_compat::knn_batch(root, queries, k, idx, dist);
}
inline void OctreeKNNSearchBatch(Octant* root, const std::vector<std::vector<double>>& queries, int k,
                                 std::vector<int32_t>& idx, std::vector<double>& dist) {
    pcnn_compat::knn_batch(root, queries, k, idx, dist);
}
inline void KDTreeRadiusNNSearchBatch(Node* root, const std::vector<std::vector<double>>& queries, double radius,
                                      std::vector<int64_t>& offsets, std::vector<int32_t>& idx,
                                      std::vector<double>& dist) {
    pcnn_compat::radius_batch(root, queries, radius, offsets, idx, dist);
}
inline void OctreeRadiusNNSearchBatch(Octant* root, const std::vector<std::vector<double>>& queries, double radius,
                                      std::vector<int64_t>& offsets, std::vector<int32_t>& idx,
                                      std::vector<double>& dist) {
    pcnn_compat::radius_batch(root, queries, radius, offsets, idx, dist);
}

#endif  // PCNN_COMPAT_HPP
