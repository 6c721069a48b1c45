// tests/compat/compat_main.cpp -- a caller written in the reference's own style
// (Homework2/hw2/include/test.hpp:162-280, benchmark.hpp:30-72) compiled against
// include/pcnn_compat.hpp instead of the reference headers.  Usage:
//   compat_main <cloud.f32 (packed xyz float32)> <query index> <k> <radius> <batch size>
// Prints the kNN / radius result sets of both trees for one query (the golden vector of
// Homework2/hw2/result_cpp.txt uses query 5, k = 8, r = 1.66) and a checksum of the batched
// search over the first <batch size> points.
#include <cstdio>
#include <cstdlib>
#include <algorithm>
#include "pcnn_compat.hpp"

static std::vector<std::vector<double>> read_cloud(const char* path) {
    FILE* f = std::fopen(path, "rb");
    if (!f) { std::perror(path); std::exit(2); }
    std::vector<std::vector<double>> db;
    float p[3];
    while (std::fread(p, sizeof(float), 3, f) == 3) db.push_back({(double)p[0], (double)p[1], (double)p[2]});
    std::fclose(f);
    return db;
}

int main(int argc, char** argv) {
    if (argc < 6) { std::fprintf(stderr, "usage: %s cloud.f32 query k radius batch\n", argv[0]); return 2; }
    auto db = read_cloud(argv[1]);
    const int qi = std::atoi(argv[2]), k = std::atoi(argv[3]), nb = std::atoi(argv[5]);
    const double radius = std::atof(argv[4]);
    auto query = db[(size_t)qi];

    Node* root_kd = KDTreeConstruction(db, 32);
    KNNResultSet knn_kd(k);
    KDTreeKNNSearch(root_kd, db, knn_kd, query);
    std::printf("kd knn:");
    for (const auto& e : knn_kd.distIndexList) std::printf(" %d:%.6g", e.index, e.distance);
    RadiusNNResultSet rnn_kd(radius);
    KDTreeRadiusNNSearch(root_kd, db, rnn_kd, query);
    std::vector<int> hits;
    for (const auto& e : rnn_kd.distIndexList) hits.push_back(e.index);
    std::sort(hits.begin(), hits.end());
    std::printf("\nkd radius:");
    for (int h : hits) std::printf(" %d", h);

    Octant* root_oc = OctreeConstruction(db, 32, 0.0001);
    KNNResultSet knn_oc(k);
    OctreeKNNSearch(root_oc, db, knn_oc, query);
    std::printf("\noc knn:");
    for (const auto& e : knn_oc.distIndexList) std::printf(" %d:%.6g", e.index, e.distance);
    RadiusNNResultSet rnn_oc(radius);
    OctreeRadiusNNSearch(root_oc, db, rnn_oc, query);
    std::printf("\noc radius count: %d", rnn_oc.size());

    // the hoisted loop of benchmark.hpp:59-66
    std::vector<std::vector<double>> queries(db.begin(), db.begin() + std::min<size_t>((size_t)nb, db.size()));
    std::vector<int32_t> idx;
    std::vector<double> dist;
    KDTreeKNNSearchBatch(root_kd, queries, k, idx, dist);
    long long cs = 0;
    for (size_t i = 0; i < idx.size(); i++) cs += (long long)idx[i] * (long long)(i % 97 + 1);
    std::printf("\nbatch %zu x %d checksum %lld first-row-self %d\n", queries.size(), k, cs, idx.empty() ? -1 : idx[0]);
    KDTreeDestruction();
    OctreeDestruction();
    return 0;
}
