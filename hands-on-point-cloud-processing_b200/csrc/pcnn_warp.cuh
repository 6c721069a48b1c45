// pcnn_warp.cuh -- warp-per-query traversal plan: the binary hierarchy is walked in steps
// of up to 5 levels, so a node has up to 32 children and every lane tests one child box
// (two coalesced 16-byte loads per lane), and a leaf is 32 points = one coalesced 512-byte
// read, one candidate per lane.
#pragma once
#include "pcnn_common.cuh"

namespace pcnn {

constexpr int WQ_WARPS = 8;   // warps (= queries in flight) per block
constexpr int WQ_MAXD = 8;    // chain length bound: (31 levels) / 5 + 2

struct WarpPlan {
    int nd;            // chain length; lvl[nd-1] is the leaf level
    int lvl[WQ_MAXD];  // hierarchy level visited at depth d, strictly descending
    int leaf_pts_log;  // log2(points per leaf) = lvl[nd-1] + 3
};

inline int level_count(int n0, int l) { return (int)(((long long)n0 + (1LL << l) - 1) >> l); }

inline WarpPlan make_warp_plan(const IndexView& v, int leaf_level = 2) {
    WarpPlan p;
    const int top = v.nlev - 1;
    const int leaf = top < leaf_level ? top : leaf_level;
    int chain[WQ_MAXD];
    int m = 0;
    int l = leaf;
    chain[m++] = l;
    while (level_count(v.n0, l) > 32 && m < WQ_MAXD) {
        l = (l + 5 < top) ? l + 5 : top;
        chain[m++] = l;
    }
    p.nd = m;
    for (int d = 0; d < m; d++) p.lvl[d] = chain[m - 1 - d];
    for (int d = m; d < WQ_MAXD; d++) p.lvl[d] = 0;
    p.leaf_pts_log = leaf + PCNN_GROUP_LOG;
    return p;
}

__device__ __forceinline__ int dev_level_count(int n0, int l) {
    return (int)(((long long)n0 + (1LL << l) - 1) >> l);
}

__device__ __forceinline__ double shfl_d(double v, int src) {
    return __shfl_sync(0xffffffffu, v, src);
}
__device__ __forceinline__ double shfl_up_d(double v, int delta) {
    return __shfl_up_sync(0xffffffffu, v, delta);
}

}  // namespace pcnn
