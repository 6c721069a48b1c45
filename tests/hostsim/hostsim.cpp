// tests/hostsim/hostsim.cpp -- TEST-ONLY host harness (never shipped, never a fallback).
// Compiles the product's __host__ __device__ traversal (csrc/pcnn_search.cuh) with g++ and
// runs it on the CPU over an index laid out exactly as the device build lays it out
// (stable sort by 63-bit Morton code, 8-point groups, binary AABB hierarchy), so that the
// traversal/tie/bound logic can be checked against the oracle on a machine without a GPU.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <numeric>
#include <vector>
#include <cuda_runtime.h>  // float4 only
#include "pcnn_search.cuh"

using namespace pcnn;

namespace {
struct HostIndex {
    std::vector<float4> pts;
    std::vector<unsigned long long> gcode;
    std::vector<float4> box;
    float cube[8];
    IndexView v;
};

HostIndex* build(const float* xyz, int n, int stride) {
    HostIndex* h = new HostIndex;
    float mn[3] = {INFINITY, INFINITY, INFINITY}, mx[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (int i = 0; i < n; i++)
        for (int a = 0; a < 3; a++) {
            mn[a] = std::fmin(mn[a], xyz[(size_t)i * stride + a]);
            mx[a] = std::fmax(mx[a], xyz[(size_t)i * stride + a]);
        }
    float side = std::fmax(mx[0] - mn[0], std::fmax(mx[1] - mn[1], mx[2] - mn[2]));
    if (!(side > 0.f) || !(side < 3.0e38f)) side = 1.f;
    for (int a = 0; a < 3; a++) { h->cube[a] = mn[a]; h->cube[4 + a] = mx[a]; }
    h->cube[7] = side;
    h->cube[3] = 2097152.0f / side;
    std::vector<unsigned long long> code((size_t)n);
    for (int i = 0; i < n; i++)
        code[i] = morton63(xyz[(size_t)i * stride], xyz[(size_t)i * stride + 1], xyz[(size_t)i * stride + 2],
                           h->cube, h->cube[3]);
    std::vector<int> perm((size_t)n);
    std::iota(perm.begin(), perm.end(), 0);
    std::stable_sort(perm.begin(), perm.end(), [&](int a, int b) { return code[a] < code[b]; });
    const int n0 = (n + PCNN_GROUP - 1) / PCNN_GROUP;
    h->pts.resize((size_t)n0 * PCNN_GROUP + PCNN_GROUP);
    h->gcode.resize((size_t)n0 + 1);
    for (int i = 0; i < n; i++) {
        const float* p = xyz + (size_t)perm[i] * stride;
        float4 f; f.x = p[0]; f.y = p[1]; f.z = p[2]; f.w = i2f(perm[i]);
        h->pts[i] = f;
        if (i % PCNN_GROUP == 0) h->gcode[i / PCNN_GROUP] = code[perm[i]];
    }
    IndexView& v = h->v;
    v.n = n; v.n0 = n0; v.nlev = 0;
    size_t nb = 0;
    for (int c = n0; c >= 1; c = (c + 1) / 2) {
        v.lev_off[v.nlev++] = (int)nb;
        nb += (size_t)c;
        if (c == 1) break;
    }
    if (n0 == 0) { v.nlev = 1; v.lev_off[0] = 0; nb = 1; }
    h->box.resize(2 * nb);
    for (int g = 0; g < n0; g++) {
        float4 lo{INFINITY, INFINITY, INFINITY, 0}, hi{-INFINITY, -INFINITY, -INFINITY, 0};
        for (int j = 0; j < PCNN_GROUP && g * PCNN_GROUP + j < n; j++) {
            const float4& p = h->pts[g * PCNN_GROUP + j];
            lo.x = std::fmin(lo.x, p.x); lo.y = std::fmin(lo.y, p.y); lo.z = std::fmin(lo.z, p.z);
            hi.x = std::fmax(hi.x, p.x); hi.y = std::fmax(hi.y, p.y); hi.z = std::fmax(hi.z, p.z);
        }
        h->box[2 * (size_t)g] = lo; h->box[2 * (size_t)g + 1] = hi;
    }
    int cnt = n0;
    for (int l = 1; l < v.nlev; l++) {
        int pc = (cnt + 1) / 2;
        for (int i = 0; i < pc; i++) {
            float4 lo = h->box[2 * ((size_t)v.lev_off[l - 1] + 2 * i)], hi = h->box[2 * ((size_t)v.lev_off[l - 1] + 2 * i) + 1];
            if (2 * i + 1 < cnt) {
                const float4& lo2 = h->box[2 * ((size_t)v.lev_off[l - 1] + 2 * i + 1)];
                const float4& hi2 = h->box[2 * ((size_t)v.lev_off[l - 1] + 2 * i + 1) + 1];
                lo.x = std::fmin(lo.x, lo2.x); lo.y = std::fmin(lo.y, lo2.y); lo.z = std::fmin(lo.z, lo2.z);
                hi.x = std::fmax(hi.x, hi2.x); hi.y = std::fmax(hi.y, hi2.y); hi.z = std::fmax(hi.z, hi2.z);
            }
            h->box[2 * ((size_t)v.lev_off[l] + i)] = lo; h->box[2 * ((size_t)v.lev_off[l] + i) + 1] = hi;
        }
        cnt = pc;
    }
    v.pts = h->pts.data(); v.gcode = h->gcode.data(); v.box = h->box.data(); v.cube = h->cube;
    return h;
}

template <int K>
void run_knn(const HostIndex* h, const float* q, int64_t nq, int stride, int k, int leaf_log, int32_t* oi, double* od) {
    for (int64_t i = 0; i < nq; i++) {
        TopK<K> tk;
        knn_thread<K>(h->v, q[i * stride], q[i * stride + 1], q[i * stride + 2], leaf_log, tk);
        for (int j = 0; j < k; j++) {
            bool have = tk.idx[j] != 0x7fffffff;
            oi[i * k + j] = have ? tk.idx[j] : -1;
            od[i * k + j] = have ? std::sqrt(tk.key[j]) : INFINITY;
        }
    }
}
}  // namespace

extern "C" {
void* hostsim_build(const float* xyz, int n, int stride) { return build(xyz, n, stride); }
void hostsim_destroy(void* h) { delete (HostIndex*)h; }
int hostsim_knn(void* hp, const float* q, int64_t nq, int stride, int k, int leaf_log, int32_t* oi, double* od) {
    const HostIndex* h = (const HostIndex*)hp;
    if (k <= 1) run_knn<1>(h, q, nq, stride, k, leaf_log, oi, od);
    else if (k <= 2) run_knn<2>(h, q, nq, stride, k, leaf_log, oi, od);
    else if (k <= 4) run_knn<4>(h, q, nq, stride, k, leaf_log, oi, od);
    else if (k <= 8) run_knn<8>(h, q, nq, stride, k, leaf_log, oi, od);
    else if (k <= 16) run_knn<16>(h, q, nq, stride, k, leaf_log, oi, od);
    else return -1;
    return 0;
}
int hostsim_nn1(void* hp, const float* q, int64_t nq, int stride, int leaf_log, int32_t* oi, float* od2) {
    const HostIndex* h = (const HostIndex*)hp;
    for (int64_t i = 0; i < nq; i++) {
        Best1f b;
        nn1_thread(h->v, q[i * stride], q[i * stride + 1], q[i * stride + 2], leaf_log, b);
        oi[i] = b.pos >= 0 ? b.idx : -1;
        od2[i] = b.key;
    }
    return 0;
}
// Morton order the device build must reproduce: perm[i] = original index at sorted slot i
void hostsim_order(void* hp, int32_t* perm) {
    const HostIndex* h = (const HostIndex*)hp;
    for (int i = 0; i < h->v.n; i++) perm[i] = f2i(h->pts[i].w);
}
}
