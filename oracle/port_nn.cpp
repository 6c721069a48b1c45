// oracle/port_nn.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT.
//
// CPU restatement ("port") of the reference's Homework2 nearest-neighbour path, written
// from scratch on flat arrays.  It exists so the parity checker is available on machines
// where /root/reference is absent (the GPU box); it is pinned in tests/test_oracle_cpu.py
// against (a) the golden vector Homework2/hw2/result_cpp.txt:11-33 / :44-66 and (b) the
// reference's own headers compiled unmodified (oracle/_ref/libpcnn_ref.so), bit for bit
// including traversal order and tie behaviour.
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
// may load this library.  The product path never does.
//
// What each piece follows (reference file:line, relative to Homework2/hw2/include):
//   ResultK::add          resultSet.hpp:65-91   (KNNResultSet::addPoint)
//   ResultR::add          resultSet.hpp:130-140 (RadiusNNResultSet::addPoint)
//   KdPort::build         kdtree.hpp:208-261    (KDTreeBuildFastMedian) + :419-429
//   KdPort::knn/radius    kdtree.hpp:329-364 / :367-402
//   OcPort::build         octree.hpp:36-90, :242-271
//   ball_inside/overlaps/contains  octree.hpp:93-100 / :103-130 / :133-139
//   OcPort::knn/radius    octree.hpp:143-186 / :189-237
// Distance: sqrt(((dx*dx)+dy*dy)+dz*dz) in double, dx = point - query, operands widened
// from float32 (kdtree.hpp:341-346; GCC folds pow(x,2) to x*x; x86-64 baseline has no FMA,
// and this file is built with -ffp-contract=off so neither does the port).
#include <cmath>
#include <cstdint>
#include <cstring>
#include <algorithm>
#include <thread>
#include <vector>

namespace {

const double kNone = 1e10;  // kdtree.hpp:9 / resultSet.hpp:32

struct P3 { double x, y, z; };

inline double ref_dist(const P3& p, const P3& q) {
    double dx = p.x - q.x, dy = p.y - q.y, dz = p.z - q.z;
    double s = dx * dx;
    s = s + dy * dy;
    s = s + dz * dz;
    return std::sqrt(s);
}

// Bounded ascending list of capacity k, slots initialised to (1e10, 0).
struct ResultK {
    int k, count = 0, cmp = 0;
    double worst = kNone;
    std::vector<double> d;
    std::vector<int> id;
    explicit ResultK(int k_) : k(k_), d((size_t)k_, kNone), id((size_t)k_, 0) {}
    void add(double dist, int index) {
        cmp++;
        if (dist > worst) return;      // equality is accepted
        if (count < k) count++;
        int pos = count - 1;           // when full: the last slot is overwritten
        while (pos > 0 && d[pos - 1] > dist) {  // strictly greater entries move right
            d[pos] = d[pos - 1];
            id[pos] = id[pos - 1];
            pos--;
        }
        d[pos] = dist;
        id[pos] = index;
        worst = d[k - 1];
    }
};

struct ResultR {
    double radius;
    int cmp = 0;
    std::vector<double> d;
    std::vector<int> id;
    explicit ResultR(double r) : radius(r) {}
    double worst_dist() const { return radius; }
    void add(double dist, int index) {
        cmp++;
        if (dist <= radius) { d.push_back(dist); id.push_back(index); }
    }
};

// ---------------------------------------------------------------- KD-tree
struct KdPort {
    struct N { int axis; double value; int left, right; int begin, end; };
    std::vector<P3> pts;
    std::vector<N> nodes;
    std::vector<int> leaf_pool;  // concatenated leaf index lists, each in stable input order

    static double coord(const P3& p, int a) { return a == 0 ? p.x : (a == 1 ? p.y : p.z); }

    int build_rec(std::vector<int>& idx, int axis, int leaf_size) {
        int me = (int)nodes.size();
        nodes.push_back(N{axis, kNone, -1, -1, 0, 0});
        if ((int)idx.size() > leaf_size) {
            std::vector<double> vals(idx.size());
            for (size_t i = 0; i < idx.size(); i++) vals[i] = coord(pts[idx[i]], axis);
            // lower median element = order statistic (n-1)/2 (kdtree.hpp:64, :232)
            size_t m = (idx.size() - 1) / 2;
            std::nth_element(vals.begin(), vals.begin() + m, vals.end());
            double median = vals[m];
            std::vector<int> L, R;
            for (int i : idx) (coord(pts[i], axis) <= median ? L : R).push_back(i);
            bool all_same = R.empty();
            if (all_same) {
                const P3& f = pts[idx[0]];
                for (int i : idx)
                    if (pts[i].x != f.x || pts[i].y != f.y || pts[i].z != f.z) { all_same = false; break; }
            }
            if (all_same) {
                // > leaf_size coincident points: the reference recurses forever here
                // (SURVEY 3.1).  The port stops instead; never hit by tests.  (An empty right
                // side alone is NOT this case: the reference then makes an empty right leaf
                // and keeps splitting the left side on the next axis, and so does the port.)
                nodes[me].begin = (int)leaf_pool.size();
                leaf_pool.insert(leaf_pool.end(), idx.begin(), idx.end());
                nodes[me].end = (int)leaf_pool.size();
                return me;
            }
            std::vector<int>().swap(idx);
            int next = (axis == 2) ? 0 : axis + 1;  // updateAxis, kdtree.hpp:38-46
            nodes[me].value = median;
            int l = build_rec(L, next, leaf_size);
            int r = build_rec(R, next, leaf_size);
            nodes[me].left = l;
            nodes[me].right = r;
        } else {
            nodes[me].begin = (int)leaf_pool.size();
            leaf_pool.insert(leaf_pool.end(), idx.begin(), idx.end());
            nodes[me].end = (int)leaf_pool.size();
        }
        return me;
    }

    template <class RS>
    void search(int n, RS& rs, const P3& q, double (*worst)(RS&)) const {
        if (n < 0) return;
        const N& nd = nodes[n];
        if (nd.value == kNone) {  // isLeaf, kdtree.hpp:28-31
            for (int i = nd.begin; i < nd.end; i++) rs.add(ref_dist(pts[leaf_pool[i]], q), leaf_pool[i]);
            return;
        }
        double qa = coord(q, nd.axis);
        int near = (qa <= nd.value) ? nd.left : nd.right;
        int far = (qa <= nd.value) ? nd.right : nd.left;
        search(near, rs, q, worst);
        if (std::fabs(qa - nd.value) < worst(rs)) search(far, rs, q, worst);
    }
};

double worst_k(ResultK& r) { return r.worst; }
double worst_r(ResultR& r) { return r.radius; }

// ---------------------------------------------------------------- Octree
struct OcPort {
    struct O { int child[8]; double cx, cy, cz, extent; int begin, end; bool leaf; };
    std::vector<P3> pts;
    std::vector<O> oct;
    std::vector<int> pool;  // every octant keeps its full index list (octree.hpp:14, :28)

    int build_rec(const std::vector<int>& idx, double cx, double cy, double cz, double extent,
                  int leaf_size, double min_extent) {
        if (idx.empty()) return -1;
        int me = (int)oct.size();
        O o;
        for (int c = 0; c < 8; c++) o.child[c] = -1;
        o.cx = cx; o.cy = cy; o.cz = cz; o.extent = extent; o.leaf = true;
        o.begin = (int)pool.size();
        pool.insert(pool.end(), idx.begin(), idx.end());
        o.end = (int)pool.size();
        oct.push_back(o);
        if (!((int)idx.size() <= leaf_size || extent <= min_extent)) {
            oct[me].leaf = false;
            std::vector<int> ch[8];
            for (int i : idx) {
                int code = 0;
                if (pts[i].x > cx) code |= 1;
                if (pts[i].y > cy) code |= 2;
                if (pts[i].z > cz) code |= 4;
                ch[code].push_back(i);
            }
            for (int c = 0; c < 8; c++) {
                double ccx = cx + ((c & 1) ? 0.5 : -0.5) * extent;
                double ccy = cy + ((c & 2) ? 0.5 : -0.5) * extent;
                double ccz = cz + ((c & 4) ? 0.5 : -0.5) * extent;
                int id = build_rec(ch[c], ccx, ccy, ccz, 0.5 * extent, leaf_size, min_extent);
                oct[me].child[c] = id;
            }
        }
        return me;
    }

    void construct(int leaf_size, double min_extent) {
        // octree.hpp:242-271: extent = FULL longest side, centre = min + extent.
        P3 mn = pts[0], mx = pts[0];
        for (const P3& p : pts) {
            mn.x = p.x < mn.x ? p.x : mn.x; mx.x = p.x > mx.x ? p.x : mx.x;
            mn.y = p.y < mn.y ? p.y : mn.y; mx.y = p.y > mx.y ? p.y : mx.y;
            mn.z = p.z < mn.z ? p.z : mn.z; mx.z = p.z > mx.z ? p.z : mx.z;
        }
        double extent = std::max(mx.x - mn.x, std::max(mx.y - mn.y, mx.z - mn.z));
        std::vector<int> idx(pts.size());
        for (size_t i = 0; i < idx.size(); i++) idx[i] = (int)i;
        build_rec(idx, mn.x + extent, mn.y + extent, mn.z + extent, extent, leaf_size, min_extent);
    }

    bool ball_inside(const P3& q, double r, const O& o) const {
        return (std::fabs(q.x - o.cx) + r) < o.extent && (std::fabs(q.y - o.cy) + r) < o.extent &&
               (std::fabs(q.z - o.cz) + r) < o.extent;
    }
    bool ball_overlaps(const P3& q, double r, const O& o) const {
        double md = r + o.extent;
        double a[3] = {std::fabs(q.x - o.cx), std::fabs(q.y - o.cy), std::fabs(q.z - o.cz)};
        if (a[0] > md || a[1] > md || a[2] > md) return false;
        int contact = (a[0] < o.extent) + (a[1] < o.extent) + (a[2] < o.extent);
        if (contact >= 2) return true;
        double x = std::max(a[0] - o.extent, 0.), y = std::max(a[1] - o.extent, 0.),
               z = std::max(a[2] - o.extent, 0.);
        return x * x + y * y + z * z < r * r;
    }
    bool ball_contains(const P3& q, double r, const O& o) const {
        double x = std::fabs(q.x - o.cx) + o.extent, y = std::fabs(q.y - o.cy) + o.extent,
               z = std::fabs(q.z - o.cz) + o.extent;
        return x * x + y * y + z * z < r * r;
    }
    void scan(const O& o, const P3& q, ResultK& rs) const {
        for (int i = o.begin; i < o.end; i++) rs.add(ref_dist(pts[pool[i]], q), pool[i]);
    }
    void scan(const O& o, const P3& q, ResultR& rs) const {
        for (int i = o.begin; i < o.end; i++) rs.add(ref_dist(pts[pool[i]], q), pool[i]);
    }

    bool knn(int n, ResultK& rs, const P3& q) const {
        if (n < 0) return false;
        const O& o = oct[n];
        if (o.leaf && o.end > o.begin) {
            scan(o, q, rs);
            return ball_inside(q, rs.worst, o);
        }
        int code = 0;
        if (q.x > o.cx) code |= 1;
        if (q.y > o.cy) code |= 2;
        if (q.z > o.cz) code |= 4;
        if (knn(o.child[code], rs, q)) return true;
        for (int c = 0; c < 8; c++) {
            if (c == code || o.child[c] < 0) continue;
            if (!ball_overlaps(q, rs.worst, oct[o.child[c]])) continue;
            if (knn(o.child[c], rs, q)) return true;
        }
        return ball_inside(q, rs.worst, o);
    }

    bool radius(int n, ResultR& rs, const P3& q) const {
        if (n < 0) return false;
        const O& o = oct[n];
        if (ball_contains(q, rs.radius, o)) {  // bulk append, no descent, returns false
            scan(o, q, rs);
            return false;
        }
        if (o.leaf && o.end > o.begin) {
            scan(o, q, rs);
            return ball_inside(q, rs.radius, o);
        }
        for (int c = 0; c < 8; c++) {
            if (o.child[c] < 0) continue;
            if (!ball_overlaps(q, rs.radius, oct[o.child[c]])) continue;
            if (radius(o.child[c], rs, q)) return true;
        }
        return ball_inside(q, rs.radius, o);
    }
};

struct PortTree {
    int kind;
    KdPort kd;
    OcPort oc;
    std::vector<int64_t> offsets;
    std::vector<int32_t> hit_idx;
    std::vector<double> hit_dist;
};

template <class F>
void parallel_for(int64_t n, int nthreads, F f) {
    if (nthreads <= 1 || n < 2) { f(0, n); return; }
    std::vector<std::thread> th;
    int64_t chunk = (n + nthreads - 1) / nthreads;
    for (int t = 0; t < nthreads; t++) {
        int64_t lo = t * chunk, hi = std::min<int64_t>(n, lo + chunk);
        if (lo >= hi) break;
        th.emplace_back([=] { f(lo, hi); });
    }
    for (auto& t : th) t.join();
}

inline P3 widen(const float* p) { return P3{(double)p[0], (double)p[1], (double)p[2]}; }

}  // namespace

extern "C" {

void* pcport_build(int kind, const float* xyz, int64_t n, int stride, int leaf_size,
                   double min_extent) {
    PortTree* t = new PortTree;
    t->kind = kind;
    std::vector<P3> pts((size_t)n);
    for (int64_t i = 0; i < n; i++) pts[(size_t)i] = widen(xyz + i * stride);
    if (kind == 0) {
        t->kd.pts.swap(pts);
        std::vector<int> idx((size_t)n);
        for (int64_t i = 0; i < n; i++) idx[(size_t)i] = (int)i;
        if (n > 0) t->kd.build_rec(idx, 0, leaf_size);
    } else {
        t->oc.pts.swap(pts);
        if (n > 0) t->oc.construct(leaf_size, min_extent);
    }
    return t;
}

void pcport_destroy(void* h) { delete (PortTree*)h; }

int pcport_knn(void* h, const float* q, int64_t nq, int stride, int k, int32_t* out_idx,
               double* out_dist, int32_t* out_cmp, int nthreads) {
    PortTree* t = (PortTree*)h;
    parallel_for(nq, nthreads, [&](int64_t lo, int64_t hi) {
        for (int64_t i = lo; i < hi; i++) {
            P3 query = widen(q + i * stride);
            ResultK rs(k);
            if (t->kind == 0) {
                if (!t->kd.nodes.empty()) t->kd.search<ResultK>(0, rs, query, worst_k);
            } else {
                if (!t->oc.oct.empty()) t->oc.knn(0, rs, query);
            }
            for (int j = 0; j < k; j++) {
                out_idx[i * k + j] = rs.id[j];
                out_dist[i * k + j] = rs.d[j];
            }
            if (out_cmp) out_cmp[i] = rs.cmp;
        }
    });
    return 0;
}

int64_t pcport_radius(void* h, const float* q, int64_t nq, int stride, double radius,
                      int64_t* out_offsets, int nthreads) {
    PortTree* t = (PortTree*)h;
    std::vector<ResultR> rows;
    rows.reserve((size_t)nq);
    for (int64_t i = 0; i < nq; i++) rows.emplace_back(radius);
    parallel_for(nq, nthreads, [&](int64_t lo, int64_t hi) {
        for (int64_t i = lo; i < hi; i++) {
            P3 query = widen(q + i * stride);
            if (t->kind == 0) {
                if (!t->kd.nodes.empty()) t->kd.search<ResultR>(0, rows[(size_t)i], query, worst_r);
            } else {
                if (!t->oc.oct.empty()) t->oc.radius(0, rows[(size_t)i], query);
            }
        }
    });
    t->offsets.assign((size_t)nq + 1, 0);
    for (int64_t i = 0; i < nq; i++) t->offsets[i + 1] = t->offsets[i] + (int64_t)rows[i].id.size();
    int64_t total = t->offsets[nq];
    t->hit_idx.resize((size_t)total);
    t->hit_dist.resize((size_t)total);
    for (int64_t i = 0; i < nq; i++) {
        int64_t o = t->offsets[i];
        for (size_t j = 0; j < rows[i].id.size(); j++) {
            t->hit_idx[o + j] = rows[i].id[j];
            t->hit_dist[o + j] = rows[i].d[j];
        }
    }
    if (out_offsets) std::memcpy(out_offsets, t->offsets.data(), sizeof(int64_t) * (nq + 1));
    return total;
}

void pcport_radius_fetch(void* h, int32_t* out_idx, double* out_dist) {
    PortTree* t = (PortTree*)h;
    if (out_idx) std::memcpy(out_idx, t->hit_idx.data(), sizeof(int32_t) * t->hit_idx.size());
    if (out_dist) std::memcpy(out_dist, t->hit_dist.data(), sizeof(double) * t->hit_dist.size());
}

int64_t pcport_knn_checksum(void* h, const float* q, int64_t nq, int stride, int k, int nthreads) {
    PortTree* t = (PortTree*)h;
    int nt = std::max(1, nthreads);
    std::vector<int64_t> part((size_t)nt, 0);
    int64_t chunk = (nq + nt - 1) / nt;
    parallel_for(nq, nthreads, [&](int64_t lo, int64_t hi) {
        int64_t s = 0;
        for (int64_t i = lo; i < hi; i++) {
            P3 query = widen(q + i * stride);
            ResultK rs(k);
            if (t->kind == 0) t->kd.search<ResultK>(0, rs, query, worst_k);
            else t->oc.knn(0, rs, query);
            for (int j = 0; j < k; j++) s += rs.id[j];
        }
        part[(size_t)(chunk ? lo / chunk : 0)] = s;
    });
    int64_t s = 0;
    for (auto v : part) s += v;
    return s;
}

int64_t pcport_radius_checksum(void* h, const float* q, int64_t nq, int stride, double radius,
                               int nthreads) {
    PortTree* t = (PortTree*)h;
    int nt = std::max(1, nthreads);
    std::vector<int64_t> part((size_t)nt, 0);
    int64_t chunk = (nq + nt - 1) / nt;
    parallel_for(nq, nthreads, [&](int64_t lo, int64_t hi) {
        int64_t s = 0;
        for (int64_t i = lo; i < hi; i++) {
            P3 query = widen(q + i * stride);
            ResultR rs(radius);
            if (t->kind == 0) t->kd.search<ResultR>(0, rs, query, worst_r);
            else t->oc.radius(0, rs, query);
            s += (int64_t)rs.id.size();
        }
        part[(size_t)(chunk ? lo / chunk : 0)] = s;
    });
    int64_t s = 0;
    for (auto v : part) s += v;
    return s;
}

int pcport_hardware_threads() { return (int)std::thread::hardware_concurrency(); }

// Brute-force canonical answer: the k smallest under the key (d2 in double, index), where
// d2 = ((dx*dx)+dy*dy)+dz*dz.  This is the contract the CUDA path is held to;
// the reference's answer differs from it only inside an exact-distance tie group.
int pcport_brute_knn(const float* xyz, int64_t n, int stride, const float* q, int64_t nq,
                     int qstride, int k, int32_t* out_idx, double* out_d2, int nthreads) {
    parallel_for(nq, nthreads, [&](int64_t lo, int64_t hi) {
        std::vector<std::pair<double, int32_t>> best;
        for (int64_t i = lo; i < hi; i++) {
            P3 query = widen(q + i * qstride);
            best.clear();
            for (int64_t j = 0; j < n; j++) {
                P3 p = widen(xyz + j * stride);
                double dx = p.x - query.x, dy = p.y - query.y, dz = p.z - query.z;
                double s = dx * dx; s = s + dy * dy; s = s + dz * dz;
                std::pair<double, int32_t> c(s, (int32_t)j);
                if ((int)best.size() < k) {
                    best.push_back(c);
                    std::push_heap(best.begin(), best.end());
                } else if (c < best.front()) {
                    std::pop_heap(best.begin(), best.end());
                    best.back() = c;
                    std::push_heap(best.begin(), best.end());
                }
            }
            std::sort(best.begin(), best.end());
            for (int j = 0; j < k; j++) {
                bool have = j < (int)best.size();
                out_idx[i * k + j] = have ? best[j].second : -1;
                out_d2[i * k + j] = have ? best[j].first : INFINITY;
            }
        }
    });
    return 0;
}

}  // extern "C"
