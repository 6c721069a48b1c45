// pcnn_internal.cuh -- host-side handle layout and helpers shared by the translation units.
#pragma once
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <new>
#include "../../include/pcnn.h"
#include "pcnn_common.cuh"
#include <nvtx3/nvToolsExt.h>

struct pcnn_index {
    int kind = 0;
    int device = 0;
    int64_t n = 0;
    int leaf_size = 32;
    double min_extent = 0;
    float4* pts = nullptr;
    unsigned long long* gcode = nullptr;
    float4* box = nullptr;
    float* cube = nullptr;  // device [8]: min.xyz, inv_cell, max.xyz, side
    pcnn::IndexView view{};
    size_t bytes = 0;
    float build_ms = 0.f;
    cudaEvent_t ev_begin = nullptr, ev_end = nullptr;  // bracket the last build
    // query-side options (pcnn_set_option)
    int opt_sort_queries = 1;
    int opt_radius_batch = 1;       // radius kernel: 0 warp per query, 1 auto, 2 batched runs, 3 lock-step (pcnn_radius.cu)
    int opt_radius_run_div = 0;     // a run spans at most radius / this per axis; 0 = 4 for the count pass, 8 for the fill pass
    int opt_radius_run_mult_count = 100;  // automatic run extent, percent of the mean 8-point-group extent
    int opt_radius_run_mult_fill = 40;
    int opt_radius_sort_bits = 45;  // leading Morton bits the batched radius search orders queries by
    int opt_radius_leaf_level = 2;  // radius search: a leaf is 8 << this points (32-point chunks)
    int opt_sort_bits = 32;  // leading Morton bits the query batch is ordered by (<= 32: 32-bit keys; measured 24 / 30 / 32 / 33 bits: 1.047 / 1.011 / 0.996 / 1.037 ms per 1M-query call)
    int opt_method = 0;
    int opt_leaf_log = 1;       // thread-per-query leaf = 2^leaf_log groups
    int opt_warp_leaf_log = 2;  // warp-per-query leaf = 32 points
    int opt_coherent_leaf_log = 1;  // warp-coherent traversal: leaf = 2^this groups
    int opt_coherent_budget = 150;   // box tests + group scans a warp may spend before it re-seeds; 0 = never
    int opt_coherent_budget2 = 600;  // ... and after the re-seed before it hands its queries over; 0 = never
    unsigned long long* dbg_warp_times = nullptr;  // diagnostics (pcnn_debug_warp_times)
    int opt_profile = 0;
    int opt_nvtx = 0;           // 1: host-side NVTX ranges around the phases of every call on this handle (pcnn.knn.order, ...)
    int opt_pdl = 1;            // query ordering + kNN search launched with programmatic stream serialization
    int opt_icp_pdl = 1;        // ICP: programmatic dependent launch between the kernels of an iteration
    int opt_icp_cert = 1;       // ICP correspondence certificates (pcnn_icp.cu); 0 = search every point every iteration
    cudaEvent_t prof0 = nullptr, prof1 = nullptr;  // bracket the last main search kernel
    // staging state of the host-pointer entry points (streams + device buffers), created on
    // first use and kept for the life of the handle so a call costs no allocation
    struct HostPipe* pipe = nullptr;
    std::mutex pipe_mu;
    // radius search: the query ordering of the last count pass, kept for the fill pass that follows
    // it on the same stream with the same query buffer (any permutation is correct; a stale one
    // only costs locality), and that count's hits per query (pcnn_radius.cu)
    mutable std::mutex rperm_mu;
    mutable unsigned* rperm = nullptr;
    mutable size_t rperm_cap = 0;  // entries
    mutable const float* rperm_q = nullptr;
    mutable int rperm_nq = 0, rperm_stride = 0, rperm_kind = -1;
    mutable cudaStream_t rperm_stream = nullptr;
    mutable double rperm_radius = -1.0;
    mutable double rperm_hits_per_query = -1.0;
    mutable cudaEvent_t rperm_ev = nullptr;  // last device-side use of rperm: a writer on another stream waits for it
};

namespace pcnn {

void set_error(const char* fmt, ...);

#define PCNN_CUDA(call)                                                                     \
    do {                                                                                    \
        cudaError_t e__ = (call);                                                           \
        if (e__ != cudaSuccess) {                                                           \
            pcnn::set_error("%s:%d: %s failed: %s", __FILE__, __LINE__, #call,              \
                            cudaGetErrorString(e__));                                       \
            return e__ == cudaErrorMemoryAllocation ? PCNN_ERR_NOMEM : PCNN_ERR_CUDA;       \
        }                                                                                   \
    } while (0)

#define PCNN_CHECK_ARG(cond, ...)              \
    do {                                       \
        if (!(cond)) {                         \
            pcnn::set_error(__VA_ARGS__);      \
            return PCNN_ERR_INVALID;           \
        }                                      \
    } while (0)

// RAII device switch
struct DeviceGuard {
    int prev = -1;
    bool ok = true;
    explicit DeviceGuard(int dev) {
        if (cudaGetDevice(&prev) != cudaSuccess) { ok = false; return; }
        if (prev != dev && cudaSetDevice(dev) != cudaSuccess) ok = false;
    }
    ~DeviceGuard() {
        if (prev >= 0) cudaSetDevice(prev);
    }
};

// The library's own stream-ordered memory pool on the current device.  Its release threshold
// is unlimited, so scratch freed by one call is reused by the next instead of going back to
// the driver at every synchronisation (pcnn_release_cached() trims it).
cudaMemPool_t lib_pool();

// Stream-ordered scratch allocation that is returned to the pool on scope exit.
struct Scratch {
    cudaStream_t s;
    void* p[24];
    int np = 0;
    explicit Scratch(cudaStream_t s_) : s(s_) {}
    template <typename T>
    cudaError_t get(T** out, size_t count) {
        void* q = nullptr;
        *out = nullptr;
        if (np >= (int)(sizeof(p) / sizeof(p[0]))) return cudaErrorMemoryAllocation;  // table full: refuse, never overrun
        cudaMemPool_t pool = lib_pool();
        cudaError_t e = pool ? cudaMallocFromPoolAsync(&q, count * sizeof(T) + 16, pool, s)
                             : cudaMallocAsync(&q, count * sizeof(T) + 16, s);
        if (e == cudaSuccess) p[np++] = q;
        *out = (T*)q;
        return e;
    }
    ~Scratch() {
        for (int i = 0; i < np; i++) cudaFreeAsync(p[i], s);
    }
};

// Grow-only device buffer.
struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
    cudaError_t ensure(size_t bytes) {
        if (bytes <= cap) return cudaSuccess;
        if (p) { cudaFree(p); p = nullptr; cap = 0; }
        cudaError_t e = cudaMalloc(&p, bytes);
        if (e == cudaSuccess) cap = bytes;
        return e;
    }
    void release() {
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
    }
};

void free_host_pipe(HostPipe* hp);

// Ordering contract of the handle (pcnn.h): pcnn_index_build only ENQUEUES the build on the caller's
// stream.  Work enqueued later on that same stream is ordered by the stream; every entry point that
// runs on a stream of its own (the *_host forms) makes that stream wait for the build first.
inline cudaError_t wait_build(const pcnn_index* ix, cudaStream_t s) {
    return ix->ev_end ? cudaStreamWaitEvent(s, ix->ev_end, 0) : cudaSuccess;
}

// NVTX range over the host-side enqueue of one phase (a profiler correlates the kernels launched inside
// it); one predictable branch when the option is off.
struct NvtxRange {
    bool on;
    NvtxRange(const pcnn_index* ix, const char* name) : on(ix && ix->opt_nvtx) {
        if (on) nvtxRangePushA(name);
    }
    ~NvtxRange() {
        if (on) nvtxRangePop();
    }
    NvtxRange(const NvtxRange&) = delete;
    NvtxRange& operator=(const NvtxRange&) = delete;
};

// profile brackets: no-ops unless option "profile" is set
inline void prof_begin(const pcnn_index* ix, cudaStream_t s) {
    if (ix->opt_profile && ix->prof0) cudaEventRecord(ix->prof0, s);
}
inline void prof_end(const pcnn_index* ix, cudaStream_t s) {
    if (ix->opt_profile && ix->prof1) cudaEventRecord(ix->prof1, s);
}

// kernels / launchers implemented across the .cu files
// sort_override: -1 = the handle's sort_queries option, 0 / 1 = this call only (the first, small stages
// of the host pipeline go unsorted so that result bytes start to flow early)
int launch_knn(const pcnn_index* ix, const float* q, int64_t nq, int stride, int k, int32_t* out_idx,
               float* out_dist, double* out_dist64, cudaStream_t s, int sort_override = -1);
int launch_radius_count(const pcnn_index* ix, const float* q, int64_t nq, int stride, double radius,
                        int64_t* out_offsets, cudaStream_t s);
// exclusive scan of nq int64 counts into offsets[nq + 1] (pcnn_radius.cu)
int scan_counts(const long long* counts, int nq, long long* offsets, Scratch& sc, cudaStream_t s);
int launch_iss(const pcnn_index* ix, const pcnn_iss_params* prm, int32_t* out_idx, int64_t capacity,
               int64_t* out_count_dev, float* out_lambda3, cudaStream_t s);
void note_radius_total(const pcnn_index* ix, const float* q, int64_t nq, double radius, int64_t total);
int launch_radius_fill(const pcnn_index* ix, const float* q, int64_t nq, int stride, double radius,
                       const int64_t* offsets, int32_t* out_idx, float* out_dist, double* out_dist64,
                       cudaStream_t s);
int launch_nn1(const pcnn_index* ix, const float* q, int64_t nq, int stride, int32_t* out_idx,
               float* out_d2, cudaStream_t s);
// Morton order of a query batch: perm[i] = original id of the i-th query along the curve.
int sort_queries(const pcnn_index* ix, const float* q, int64_t nq, int stride, unsigned** perm_out,
                 Scratch& scratch, cudaStream_t s);

int sort_queries_fine(const pcnn_index* ix, const float* q, int64_t nq, int stride, int bits, unsigned** perm_out,
                      Scratch& scratch, cudaStream_t s);

}  // namespace pcnn
