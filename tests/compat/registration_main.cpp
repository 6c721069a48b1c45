// tests/compat/registration_main.cpp -- a caller in the style of Homework9/hw9/main.cpp:75-101 and
// registration.cpp:1141-1151, compiled against include/pcnn_registration_compat.hpp with ten-line
// stand-ins for the Eigen / PCL types the reference uses (neither library is installed here).
// Usage: registration_main <src.f32> <tar.f32> <tar_normals.f32> <max_iter> <max_corres_dist>
// Prints R and t of both ICP variants, 9 significant digits (float32 round-trips exactly).
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "pcnn_registration_compat.hpp"

struct Mat3 {  // stands in for Eigen::Matrix3f
    float m[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
    float& operator()(int r, int c) { return m[3 * r + c]; }
    const float& operator()(int r, int c) const { return m[3 * r + c]; }
};
struct Vec3 {  // stands in for Eigen::Vector3f
    float v[3] = {0, 0, 0};
    float& operator()(int i) { return v[i]; }
    const float& operator()(int i) const { return v[i]; }
};
struct PointXYZ { float x, y, z; };                       // pcl::PointXYZ
struct Normal { float normal_x, normal_y, normal_z; };   // pcl::Normal
using PointCloud = std::vector<PointXYZ>;                 // pcl::PointCloud<PointT>: size(), operator[]
using NormalCloud = std::vector<Normal>;
using Registration = pcnn_compat::RegistrationT<Mat3, Vec3, PointCloud, NormalCloud>;

template <class T>
static std::vector<T> read3(const char* path) {
    FILE* f = std::fopen(path, "rb");
    if (!f) { std::perror(path); std::exit(2); }
    std::vector<T> out;
    float p[3];
    while (std::fread(p, sizeof(float), 3, f) == 3) out.push_back(T{p[0], p[1], p[2]});
    std::fclose(f);
    return out;
}

static void show(const char* name, const Mat3& R, const Vec3& t, const pcnn_icp_result& r) {
    std::printf("%s:", name);
    for (int i = 0; i < 9; i++) std::printf(" %.9g", R.m[i]);
    for (int i = 0; i < 3; i++) std::printf(" %.9g", t.v[i]);
    std::printf(" iters %d pairs %lld status %d\n", r.iters, (long long)r.last_pairs, r.status);
}

int main(int argc, char** argv) {
    if (argc < 6) { std::fprintf(stderr, "usage: %s src tar tar_normals max_iter max_corres_dist\n", argv[0]); return 2; }
    PointCloud src = read3<PointXYZ>(argv[1]), tar = read3<PointXYZ>(argv[2]);
    NormalCloud ntar = read3<Normal>(argv[3]), nsrc;
    Registration reg;
    reg.setICPparams(10, 4000, (float)std::atof(argv[5]), (size_t)std::atoi(argv[4]), 0.f);  // main.cpp:88-95 shape
    Mat3 init_R, R;
    Vec3 init_t, t;
    R = init_R; t = init_t;                                   // registration.cpp:1141-1142
    reg.ICPpoint2point(init_R, init_t, src, tar, nsrc, ntar, R, t);
    show("p2p", R, t, reg.lastResult());
    R = init_R; t = init_t;
    reg.ICPpoint2plane(init_R, init_t, src, tar, nsrc, ntar, R, t);
    show("p2plane", R, t, reg.lastResult());
    // the free function of registration.cpp:165-178 on the same stand-in types
    PointCloud moved = src;
    pcnn_compat::transformCloudInplace(moved, R, t);
    std::printf("moved0: %.9g %.9g %.9g\n", moved[0].x, moved[0].y, moved[0].z);
    return 0;
}
