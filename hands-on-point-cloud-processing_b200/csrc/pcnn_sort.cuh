// pcnn_sort.cuh -- hand-written LSD radix sort (8-bit digits) of (key, uint32 value) pairs,
// used for the Morton-code sort of the index build and for ordering query batches.
// Replaces the reference's per-node median selection (Homework2/hw2/include/kdtree.hpp:57-109,
// :208-261) and 8-way recursive partition (octree.hpp:36-90) as the spatial ordering step.
//
// Per pass: rs_hist (per-tile digit histogram, digit-major [256][ntiles]) -> rs_scan_chunk
// (exclusive scan of that table, chunk by chunk) -> rs_scatter (stable rank inside the tile via
// __match_any_sync + per-warp digit counters, then scatter).  HBM traffic per pass:
// read keys twice + values once, write both once.
#pragma once
#include "pcnn_common.cuh"

namespace pcnn {

constexpr int RS_THREADS = 256;
constexpr int RS_ITEMS = 16;
constexpr int RS_TILE = RS_THREADS * RS_ITEMS;
constexpr int RS_WARPS = RS_THREADS / 32;

template <typename KeyT>
__global__ void __launch_bounds__(RS_THREADS)
rs_hist(const KeyT* __restrict__ keys, int n, int shift, unsigned* __restrict__ hist, int ntiles) {
    __shared__ unsigned h[256];
    pdl_prologue();
    h[threadIdx.x] = 0;
    __syncthreads();
    const int base = blockIdx.x * RS_TILE;
#pragma unroll
    for (int it = 0; it < RS_ITEMS; it++) {
        int i = base + it * RS_THREADS + threadIdx.x;
        if (i < n) atomicAdd(&h[(unsigned)(keys[i] >> shift) & 255u], 1u);
    }
    __syncthreads();
    hist[(size_t)threadIdx.x * ntiles + blockIdx.x] = h[threadIdx.x];
}

// The digit table is scanned in chunks of 4096 counters by as many blocks (exclusive prefix
// inside the chunk + the chunk's total); rs_scatter adds the totals of the chunks before its own.
// (One block walking the whole table took 10 us per pass at 1M keys and 193 us at 10M.)
static __global__ void __launch_bounds__(1024) rs_scan_chunk(unsigned* __restrict__ a, int total,
                                                            unsigned* __restrict__ chunk_sum) {
    __shared__ unsigned wsum[32];
    pdl_prologue();
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int base = blockIdx.x * 4096;
    unsigned v[4];
#pragma unroll
    for (int j = 0; j < 4; j++) {
        const int i = base + 4 * t + j;
        v[j] = i < total ? a[i] : 0u;
    }
    const unsigned s = v[0] + v[1] + v[2] + v[3];
    unsigned inc = s;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned u = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += u;
    }
    if (lane == 31) wsum[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        unsigned w = wsum[lane], winc = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned u = __shfl_up_sync(0xffffffffu, winc, o);
            if (lane >= o) winc += u;
        }
        wsum[lane] = winc - w;
        if (lane == 31) chunk_sum[blockIdx.x] = winc;
    }
    __syncthreads();
    unsigned run = wsum[warp] + inc - s;
#pragma unroll
    for (int j = 0; j < 4; j++) {
        const int i = base + 4 * t + j;
        if (i < total) a[i] = run;
        run += v[j];
    }
}
template <typename KeyT>
__global__ void __launch_bounds__(RS_THREADS)
rs_scatter(const KeyT* __restrict__ kin, const unsigned* __restrict__ vin, KeyT* __restrict__ kout,
           unsigned* __restrict__ vout, int n, int shift, const unsigned* __restrict__ offs,
           const unsigned* __restrict__ chunk_sum, int ntiles) {
    __shared__ unsigned wcount[RS_WARPS][256];
    __shared__ unsigned dbase[256];
    pdl_prologue();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < RS_WARPS * 256; i += RS_THREADS) (&wcount[0][0])[i] = 0;
    __syncthreads();

    // warp-striped ownership keeps the in-tile order = (warp, item, lane) = global index order
    const int wbase = blockIdx.x * RS_TILE + warp * (32 * RS_ITEMS);
    KeyT key[RS_ITEMS];
    unsigned local[RS_ITEMS];
#pragma unroll
    for (int it = 0; it < RS_ITEMS; it++) {
        int i = wbase + it * 32 + lane;
        key[it] = (i < n) ? kin[i] : (KeyT)~(KeyT)0;  // padding sorts last, never written
    }
#pragma unroll
    for (int it = 0; it < RS_ITEMS; it++) {
        unsigned d = (unsigned)(key[it] >> shift) & 255u;
        unsigned peers = __match_any_sync(0xffffffffu, d);
        int leader = __ffs(peers) - 1;
        unsigned rank = __popc(peers & ((1u << lane) - 1u));
        unsigned b = 0;
        if (lane == leader) {
            b = wcount[warp][d];
            wcount[warp][d] = b + __popc(peers);
        }
        b = __shfl_sync(0xffffffffu, b, leader);
        local[it] = b + rank;
        __syncwarp();
    }
    __syncthreads();
    {   // digit = threadIdx.x: exclusive prefix over the warps + global base of (digit, tile)
        unsigned run = 0;
#pragma unroll
        for (int w = 0; w < RS_WARPS; w++) {
            unsigned c = wcount[w][threadIdx.x];
            wcount[w][threadIdx.x] = run;
            run += c;
        }
        // offs holds exclusive prefixes inside chunks of 4096 counters (rs_scan_chunk); the base of
        // this counter's chunk is the sum of the totals of the chunks before it (a handful of
        // independent loads: 16 chunks at 1M keys, 153 at 10M)
        const size_t at = (size_t)threadIdx.x * ntiles + blockIdx.x;
        const int myc = (int)(at >> 12);
        unsigned cb = 0;
#pragma unroll 8
        for (int c = 0; c < myc; c++) cb += __ldg(chunk_sum + c);
        dbase[threadIdx.x] = offs[at] + cb;
    }
    __syncthreads();
#pragma unroll
    for (int it = 0; it < RS_ITEMS; it++) {
        int i = wbase + it * 32 + lane;
        if (i < n) {
            unsigned d = (unsigned)(key[it] >> shift) & 255u;
            unsigned pos = dbase[d] + wcount[warp][d] + local[it];
            kout[pos] = key[it];
            vout[pos] = vin[i];
        }
    }
}

// Sort (keys, vals) of length n by the low `bits` bits of the key.  k0/v0 hold the input;
// k1/v1 are scratch of the same size; hist is scratch of 256*ceil(n/RS_TILE) unsigned.
// Returns 0 if the result is in (k0, v0), 1 if it is in (k1, v1).
// pdl: launch the passes with programmatic stream serialization (pcnn_common.cuh); for chains of small
// sorts, where launch latency is most of the cost (the query ordering of a 1M-query kNN call: 1.183 ->
// 1.166 ms for ordering + search).  Measured and dropped: 10-bit digits (three passes instead of four
// for the 30-bit query keys): 1.167 ms without PDL, 1.172 with -- fewer but slower kernels.
template <typename KeyT>
inline int radix_sort_pairs(KeyT* k0, unsigned* v0, KeyT* k1, unsigned* v1, unsigned* hist, int n,
                            int bits, cudaStream_t s, bool pdl = false) {
    if (n <= 0) return 0;
    const int ntiles = (n + RS_TILE - 1) / RS_TILE;
    int cur = 0;
    for (int shift = 0; shift < bits; shift += 8) {
        KeyT* ki = cur ? k1 : k0;
        KeyT* ko = cur ? k0 : k1;
        unsigned* vi = cur ? v1 : v0;
        unsigned* vo = cur ? v0 : v1;
        launch_pdl(rs_hist<KeyT>, dim3(ntiles), dim3(RS_THREADS), s, pdl, ki, n, shift, hist, ntiles);
        const int total = 256 * ntiles;
        const int nchunks = (total + 4095) / 4096;
        unsigned* chunk = hist + total;  // radix_hist_elems leaves room for the chunk totals
        launch_pdl(rs_scan_chunk, dim3(nchunks), dim3(1024), s, pdl, hist, total, chunk);
        launch_pdl(rs_scatter<KeyT>, dim3(ntiles), dim3(RS_THREADS), s, pdl, ki, vi, ko, vo, n, shift, hist, chunk, ntiles);
        cur ^= 1;
    }
    return cur;
}

inline size_t radix_hist_elems(int n) {
    const size_t total = (size_t)256 * ((n + RS_TILE - 1) / RS_TILE);
    return total + (total + 4095) / 4096;  // + chunk totals of the chunked scan
}

}  // namespace pcnn
