// pcnn_capi.cu -- C ABI entry points for kNN / radius search (declared in include/pcnn.h).
// Device-pointer variants validate and enqueue; host-pointer variants stage through HBM in
// stages on up to eight streams so that the PCIe copies of one chunk overlap the kernels of the
// next, each stage searched in the caller's order (one launch per stage) (the reference has no device boundary: Homework2/hw2/include/benchmark.hpp:59-72).
#include <algorithm>
#include <cstdlib>
#include <vector>
#include "pcnn_internal.cuh"
#include <chrono>
#include <vector>

using namespace pcnn;

// Streams and device staging buffers of the host-pointer entry points; one per index handle,
// created on first use, grown on demand, freed with the handle.
constexpr int HOST_STREAMS = 8;
struct HostPipe {
    cudaStream_t s[HOST_STREAMS] = {};
    pcnn::DevBuf q[HOST_STREAMS], idx[HOST_STREAMS], d32[HOST_STREAMS], d64[HOST_STREAMS];
    pcnn::DevBuf off;  // radius: CSR offsets
    cudaEvent_t ev[HOST_STREAMS] = {};
};

namespace pcnn {
void free_host_pipe(HostPipe* hp) {
    if (!hp) return;
    for (int i = 0; i < HOST_STREAMS; i++) {
        if (hp->s[i]) { cudaStreamSynchronize(hp->s[i]); cudaStreamDestroy(hp->s[i]); }
        if (hp->ev[i]) cudaEventDestroy(hp->ev[i]);
        hp->q[i].release(); hp->idx[i].release(); hp->d32[i].release(); hp->d64[i].release();
    }
    hp->off.release();
    delete hp;
}
}  // namespace pcnn

namespace {

// pipeline stage size: enough stages to hide the PCIe copies behind the kernels.  Measured on 1M queries,
// k = 8 (tools/e2e_sweep.sh, B200): stages of 65k / 125k / 131k / 200k queries -> 1.55 / 1.55 / 1.56 /
// 1.58 ms per call; round 1's pipeline, which ordered every stage's queries, took 1.82 ms.
int64_t host_chunk(int64_t nq) {
#ifdef PCNN_STUDY
    if (const char* e = std::getenv("PCNN_HOST_CHUNK")) {  // tuning aid
        const long long v = std::atoll(e);
        if (v > 0) return std::min<int64_t>(v, nq);
    }
#endif
    int64_t c = (nq + 7) / 8;
    c = std::max<int64_t>(c, 1 << 16);
    c = std::min<int64_t>(c, 1 << 18);
    return std::min(c, nq);
}

// caller holds index->pipe_mu
int get_pipe(pcnn_index* ix, HostPipe** out, const char* fn) {
    if (!ix->pipe) {
        HostPipe* hp = new (std::nothrow) HostPipe;
        if (!hp) { set_error("%s: host allocation failed", fn); return PCNN_ERR_NOMEM; }
        for (int i = 0; i < HOST_STREAMS; i++) {
            if (cudaStreamCreateWithFlags(&hp->s[i], cudaStreamNonBlocking) != cudaSuccess ||
                cudaEventCreateWithFlags(&hp->ev[i], cudaEventDisableTiming) != cudaSuccess) {
                set_error("%s: stream creation failed: %s", fn, cudaGetErrorString(cudaGetLastError()));
                free_host_pipe(hp);
                return PCNN_ERR_CUDA;
            }
        }
        ix->pipe = hp;
    }
    *out = ix->pipe;
    return PCNN_OK;
}

int check_query_args(const char* fn, const pcnn_index* index, const float* q, int64_t nq, int stride) {
    PCNN_CHECK_ARG(index != nullptr, "%s: index is NULL", fn);
    PCNN_CHECK_ARG(nq >= 0 && nq < (int64_t)0x7fffffff, "%s: nq=%lld out of range", fn, (long long)nq);
    PCNN_CHECK_ARG(stride >= 3, "%s: stride must be >= 3 floats (got %d)", fn, stride);
    PCNN_CHECK_ARG(nq == 0 || q != nullptr, "%s: query pointer is NULL", fn);
    return PCNN_OK;
}

}  // namespace

extern "C" {

int pcnn_knn(const pcnn_index* index, const float* q_dev, int64_t nq, int stride, int k,
             int32_t* out_idx_dev, float* out_dist_dev, double* out_dist64_dev, pcnn_stream stream) {
    int rc = check_query_args("pcnn_knn", index, q_dev, nq, stride);
    if (rc != PCNN_OK) return rc;
    PCNN_CHECK_ARG(k >= 1, "pcnn_knn: k must be >= 1 (got %d)", k);
    if (k > PCNN_MAX_K) { set_error("pcnn_knn: k=%d exceeds PCNN_MAX_K=%d", k, PCNN_MAX_K); return PCNN_ERR_UNSUPPORTED; }
    PCNN_CHECK_ARG(nq == 0 || out_idx_dev != nullptr, "pcnn_knn: out_idx is NULL");
    PCNN_CHECK_ARG((int64_t)nq * k < (int64_t)1 << 40, "pcnn_knn: nq*k too large");
    DeviceGuard g(index->device);
    if (!g.ok) { set_error("pcnn_knn: cannot select CUDA device %d (no CPU fallback exists)", index->device); return PCNN_ERR_CUDA; }
    return launch_knn(index, q_dev, nq, stride, k, out_idx_dev, out_dist_dev, out_dist64_dev, (cudaStream_t)stream);
}

int pcnn_knn_host(const pcnn_index* index, const float* q_host, int64_t nq, int stride, int k,
                  int32_t* out_idx_host, float* out_dist_host, double* out_dist64_host) {
    int rc = check_query_args("pcnn_knn_host", index, q_host, nq, stride);
    if (rc != PCNN_OK) return rc;
    PCNN_CHECK_ARG(k >= 1, "pcnn_knn_host: k must be >= 1 (got %d)", k);
    if (k > PCNN_MAX_K) { set_error("pcnn_knn_host: k=%d exceeds PCNN_MAX_K=%d", k, PCNN_MAX_K); return PCNN_ERR_UNSUPPORTED; }
    PCNN_CHECK_ARG(nq == 0 || out_idx_host != nullptr, "pcnn_knn_host: out_idx is NULL");
    if (nq == 0) return PCNN_OK;
    DeviceGuard g(index->device);
    if (!g.ok) { set_error("pcnn_knn_host: cannot select CUDA device %d (no CPU fallback exists)", index->device); return PCNN_ERR_CUDA; }
    pcnn_index* ix = const_cast<pcnn_index*>(index);
    std::lock_guard<std::mutex> lk(ix->pipe_mu);
    HostPipe* hp = nullptr;
    rc = get_pipe(ix, &hp, "pcnn_knn_host");
    if (rc != PCNN_OK) return rc;
    // Stage plan.  The link is the bottleneck of this call (64 MB of results at ~55 GB/s = 1.16 ms for 1M
    // queries, k = 8, against 1.05 ms of search), so what matters is how soon the first result byte
    // leaves and that the copy engine never starves afterwards.  Stages are searched in the CALLER'S
    // order: ordering a stage's queries along the curve is a chain of 13 small dependent kernels that
    // crawls (0.2-0.6 ms) while another stage's search fills the GPU; what it would buy is search time
    // (per million queries 0.77 ms ordered, 1.11 ms on LiDAR scan order, 1.18 ms on a random shuffle;
    // tools/knn_unsorted.py), and the search is not what this call waits for: from 0.25 ms on the D2H
    // engine is (stage timeline).  Rows do not depend on the order.  Two small stages lead so that
    // results are on the wire after ~0.2 ms.
    struct Stage { int64_t q0, m; int sort; };
    std::vector<Stage> plan;
    const int64_t chunk = host_chunk(nq);
    {
        int64_t q0 = 0;
        if (nq >= 4 * chunk) {
            plan.push_back({0, 32768, 0});
            plan.push_back({32768, 65536, 0});
            q0 = 32768 + 65536;
        }
        for (; q0 < nq; q0 += chunk) plan.push_back({q0, std::min<int64_t>(chunk, nq - q0), 0});
#ifdef PCNN_STUDY
        if (const char* ev = std::getenv("PCNN_HOST_LEAD_SORT"))  // tuning aid: order the queries of the first stages
            for (int i = 0; i < std::atoi(ev) && i < (int)plan.size(); i++) plan[i].sort = 1;
#endif
    }
    int max_streams = HOST_STREAMS;
#ifdef PCNN_STUDY
    if (const char* ev = std::getenv("PCNN_HOST_STREAMS")) max_streams = std::max(1, std::min(HOST_STREAMS, std::atoi(ev)));  // tuning aid
#endif
    const int nstreams = (int)std::min<size_t>((size_t)max_streams, plan.size());
    cudaError_t e = cudaSuccess;
    for (int i = 0; i < nstreams && e == cudaSuccess; i++) {
        e = hp->q[i].ensure(sizeof(float) * (size_t)chunk * stride);
        if (e == cudaSuccess) e = hp->idx[i].ensure(sizeof(int32_t) * (size_t)chunk * k);
        if (e == cudaSuccess && out_dist_host) e = hp->d32[i].ensure(sizeof(float) * (size_t)chunk * k);
        if (e == cudaSuccess && out_dist64_host) e = hp->d64[i].ensure(sizeof(double) * (size_t)chunk * k);
    }
    if (e != cudaSuccess) {
        set_error("pcnn_knn_host: staging allocation failed: %s", cudaGetErrorString(e));
        return PCNN_ERR_NOMEM;
    }
    // PCNN_PIPE_TRACE=1 (tuning aid, study build): per-stage event times of this call on stderr
#ifdef PCNN_STUDY
    static const bool trace = std::getenv("PCNN_PIPE_TRACE") != nullptr;
#else
    constexpr bool trace = false;
#endif
    std::vector<cudaEvent_t> tev;
    std::vector<double> thost;
    const auto t_host0 = std::chrono::steady_clock::now();
    auto mark = [&](cudaStream_t st) {
        if (!trace) return;
        cudaEvent_t ev;
        cudaEventCreate(&ev);
        cudaEventRecord(ev, st);
        tev.push_back(ev);
        thost.push_back(std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_host0).count());
    };
    for (int i = 0; i < nstreams; i++) wait_build(index, hp->s[i]);  // a device-side build may still be running on its own stream
    mark(hp->s[0]);
    for (size_t c = 0; c < plan.size() && rc == PCNN_OK; c++) {
        const int b = (int)(c % (size_t)nstreams);
        const int64_t q0 = plan[c].q0, m = plan[c].m;
        cudaStream_t s = hp->s[b];
        float* dq = (float*)hp->q[b].p;
        int32_t* di = (int32_t*)hp->idx[b].p;
        float* dd = out_dist_host ? (float*)hp->d32[b].p : nullptr;
        double* dd64 = out_dist64_host ? (double*)hp->d64[b].p : nullptr;
        NvtxRange nvtx_(index, "pcnn.knn_host.stage");
        e = cudaMemcpyAsync(dq, q_host + q0 * stride, sizeof(float) * (size_t)m * stride, cudaMemcpyHostToDevice, s);
        if (e != cudaSuccess) { set_error("pcnn_knn_host: H2D failed: %s", cudaGetErrorString(e)); rc = PCNN_ERR_CUDA; break; }
        mark(s);
        rc = launch_knn(index, dq, m, stride, k, di, dd, dd64, s, plan[c].sort);
        if (rc != PCNN_OK) break;
        mark(s);
        e = cudaMemcpyAsync(out_idx_host + q0 * k, di, sizeof(int32_t) * (size_t)m * k, cudaMemcpyDeviceToHost, s);
        if (e == cudaSuccess && out_dist_host)
            e = cudaMemcpyAsync(out_dist_host + q0 * k, dd, sizeof(float) * (size_t)m * k, cudaMemcpyDeviceToHost, s);
        if (e == cudaSuccess && out_dist64_host)
            e = cudaMemcpyAsync(out_dist64_host + q0 * k, dd64, sizeof(double) * (size_t)m * k, cudaMemcpyDeviceToHost, s);
        if (e != cudaSuccess) { set_error("pcnn_knn_host: D2H failed: %s", cudaGetErrorString(e)); rc = PCNN_ERR_CUDA; }
        mark(s);
    }
    for (int i = 0; i < nstreams; i++) {
        e = cudaStreamSynchronize(hp->s[i]);
        if (e != cudaSuccess && rc == PCNN_OK) {
            set_error("pcnn_knn_host: device work failed: %s", cudaGetErrorString(e));
            rc = PCNN_ERR_CUDA;
        }
    }
    if (trace) {
        const double t_end = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_host0).count();
        std::fprintf(stderr, "pcnn_knn_host trace (ms since call; device | host enqueue), call returned at %.3f\n", t_end);
        for (size_t i = 1; i < tev.size(); i++) {
            float ms = 0.f;
            cudaEventElapsedTime(&ms, tev[0], tev[i]);
            static const char* what[3] = {"H2D done", "search done", "D2H done"};
            std::fprintf(stderr, "  stage %zu %-11s %7.3f | %7.3f\n", (i - 1) / 3, what[(i - 1) % 3], ms, thost[i]);
        }
        for (cudaEvent_t ev : tev) cudaEventDestroy(ev);
    }
    return rc;
}

int pcnn_radius_count(const pcnn_index* index, const float* q_dev, int64_t nq, int stride, double radius,
                      int64_t* out_offsets_dev, int64_t* out_total_host, pcnn_stream stream) {
    int rc = check_query_args("pcnn_radius_count", index, q_dev, nq, stride);
    if (rc != PCNN_OK) return rc;
    PCNN_CHECK_ARG(out_offsets_dev != nullptr, "pcnn_radius_count: out_offsets is NULL");
    PCNN_CHECK_ARG(radius == radius, "pcnn_radius_count: radius is NaN");
    DeviceGuard g(index->device);
    if (!g.ok) { set_error("pcnn_radius_count: cannot select CUDA device %d (no CPU fallback exists)", index->device); return PCNN_ERR_CUDA; }
    cudaStream_t s = (cudaStream_t)stream;
    rc = launch_radius_count(index, q_dev, nq, stride, radius, out_offsets_dev, s);
    if (rc != PCNN_OK) return rc;
    if (out_total_host) {
        PCNN_CUDA(cudaMemcpyAsync(out_total_host, out_offsets_dev + nq, sizeof(int64_t), cudaMemcpyDeviceToHost, s));
        PCNN_CUDA(cudaStreamSynchronize(s));
        note_radius_total(index, q_dev, nq, radius, *out_total_host);
    }
    return PCNN_OK;
}

int pcnn_radius_fill(const pcnn_index* index, const float* q_dev, int64_t nq, int stride, double radius,
                     const int64_t* offsets_dev, int32_t* out_idx_dev, float* out_dist_dev,
                     double* out_dist64_dev, pcnn_stream stream) {
    int rc = check_query_args("pcnn_radius_fill", index, q_dev, nq, stride);
    if (rc != PCNN_OK) return rc;
    PCNN_CHECK_ARG(offsets_dev != nullptr, "pcnn_radius_fill: offsets is NULL");
    PCNN_CHECK_ARG(out_idx_dev != nullptr || nq == 0, "pcnn_radius_fill: out_idx is NULL");
    PCNN_CHECK_ARG(radius == radius, "pcnn_radius_fill: radius is NaN");
    DeviceGuard g(index->device);
    if (!g.ok) { set_error("pcnn_radius_fill: cannot select CUDA device %d (no CPU fallback exists)", index->device); return PCNN_ERR_CUDA; }
    return launch_radius_fill(index, q_dev, nq, stride, radius, offsets_dev, out_idx_dev, out_dist_dev,
                              out_dist64_dev, (cudaStream_t)stream);
}

int pcnn_radius_host(const pcnn_index* index, const float* q_host, int64_t nq, int stride, double radius,
                     int64_t* offsets_host, int32_t* out_idx_host, float* out_dist_host,
                     double* out_dist64_host, int64_t capacity) {
    int rc = check_query_args("pcnn_radius_host", index, q_host, nq, stride);
    if (rc != PCNN_OK) return rc;
    PCNN_CHECK_ARG(offsets_host != nullptr, "pcnn_radius_host: offsets is NULL");
    PCNN_CHECK_ARG(radius == radius, "pcnn_radius_host: radius is NaN");
    PCNN_CHECK_ARG(capacity >= 0, "pcnn_radius_host: negative capacity");
    DeviceGuard g(index->device);
    if (!g.ok) { set_error("pcnn_radius_host: cannot select CUDA device %d (no CPU fallback exists)", index->device); return PCNN_ERR_CUDA; }
    if (nq == 0) { offsets_host[0] = 0; return PCNN_OK; }
    pcnn_index* ix = const_cast<pcnn_index*>(index);
    std::lock_guard<std::mutex> lk(ix->pipe_mu);
    HostPipe* hp = nullptr;
    rc = get_pipe(ix, &hp, "pcnn_radius_host");
    if (rc != PCNN_OK) return rc;
    cudaStream_t s = hp->s[0], s1 = hp->s[1];
    wait_build(index, s);
    cudaError_t e = hp->q[0].ensure(sizeof(float) * (size_t)nq * stride);
    if (e == cudaSuccess) e = hp->off.ensure(sizeof(int64_t) * ((size_t)nq + 1));
    float* dq = (float*)hp->q[0].p;
    int64_t* doff = (int64_t*)hp->off.p;
    if (e == cudaSuccess) e = cudaMemcpyAsync(dq, q_host, sizeof(float) * (size_t)nq * stride, cudaMemcpyHostToDevice, s);
    if (e != cudaSuccess) { set_error("pcnn_radius_host: staging failed: %s", cudaGetErrorString(e)); return PCNN_ERR_NOMEM; }
    int64_t total = 0;
    rc = launch_radius_count(index, dq, nq, stride, radius, doff, s);
    if (rc == PCNN_OK) {
        e = cudaMemcpyAsync(offsets_host, doff, sizeof(int64_t) * ((size_t)nq + 1), cudaMemcpyDeviceToHost, s);
        if (e == cudaSuccess) e = cudaStreamSynchronize(s);
        if (e != cudaSuccess) { set_error("pcnn_radius_host: count pass failed: %s", cudaGetErrorString(e)); rc = PCNN_ERR_CUDA; }
        else total = offsets_host[nq];
    }
    if (rc == PCNN_OK && total > capacity) {
        set_error("pcnn_radius_host: %lld hits exceed capacity %lld", (long long)total, (long long)capacity);
        rc = PCNN_ERR_CAPACITY;
    }
    if (rc == PCNN_OK && total > 0) {
        PCNN_CHECK_ARG(out_idx_host != nullptr, "pcnn_radius_host: out_idx is NULL");
        e = hp->idx[0].ensure(sizeof(int32_t) * (size_t)total);
        if (e == cudaSuccess && out_dist_host) e = hp->d32[0].ensure(sizeof(float) * (size_t)total);
        if (e == cudaSuccess && out_dist64_host) e = hp->d64[0].ensure(sizeof(double) * (size_t)total);
        if (e != cudaSuccess) { set_error("pcnn_radius_host: output allocation failed: %s", cudaGetErrorString(e)); return PCNN_ERR_NOMEM; }
        int32_t* di = (int32_t*)hp->idx[0].p;
        float* dd = out_dist_host ? (float*)hp->d32[0].p : nullptr;
        double* dd64 = out_dist64_host ? (double*)hp->d64[0].p : nullptr;
        rc = launch_radius_fill(index, dq, nq, stride, radius, doff, di, dd, dd64, s);
        if (rc == PCNN_OK) {
            // the three result arrays leave on two streams so the copies share the link
            e = cudaEventRecord(hp->ev[0], s);
            if (e == cudaSuccess) e = cudaMemcpyAsync(out_idx_host, di, sizeof(int32_t) * (size_t)total, cudaMemcpyDeviceToHost, s);
            if (e == cudaSuccess && (out_dist_host || out_dist64_host)) e = cudaStreamWaitEvent(s1, hp->ev[0], 0);
            if (e == cudaSuccess && out_dist_host)
                e = cudaMemcpyAsync(out_dist_host, dd, sizeof(float) * (size_t)total, cudaMemcpyDeviceToHost, s1);
            if (e == cudaSuccess && out_dist64_host)
                e = cudaMemcpyAsync(out_dist64_host, dd64, sizeof(double) * (size_t)total, cudaMemcpyDeviceToHost, s1);
            if (e == cudaSuccess) e = cudaStreamSynchronize(s);
            if (e == cudaSuccess) e = cudaStreamSynchronize(s1);
            if (e != cudaSuccess) { set_error("pcnn_radius_host: fill pass failed: %s", cudaGetErrorString(e)); rc = PCNN_ERR_CUDA; }
        }
    }
    cudaStreamSynchronize(s);
    return rc;
}

int pcnn_icp_host(const pcnn_index* target, const float* tar_normals_host, const float* src_host,
                  int64_t n_src, int stride, const float* init_T_host, const pcnn_icp_params* params,
                  pcnn_icp_result* out_host) {
    PCNN_CHECK_ARG(target && params && out_host && init_T_host, "pcnn_icp_host: NULL argument");
    PCNN_CHECK_ARG(n_src >= 0 && stride >= 3, "pcnn_icp_host: bad n_src/stride");
    PCNN_CHECK_ARG(n_src == 0 || src_host, "pcnn_icp_host: src is NULL");
    PCNN_CHECK_ARG(params->mode == 0 || tar_normals_host || target->n == 0, "pcnn_icp_host: point-to-plane needs target normals");
    DeviceGuard g(target->device);
    if (!g.ok) { set_error("pcnn_icp_host: cannot select CUDA device %d (no CPU fallback exists)", target->device); return PCNN_ERR_CUDA; }
    cudaStream_t s = nullptr;
    PCNN_CUDA(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
    wait_build(target, s);
    float *dsrc = nullptr, *dn = nullptr;
    int rc = PCNN_OK;
    cudaError_t e = cudaSuccess;
    if (n_src > 0) {
        e = cudaMalloc(&dsrc, sizeof(float) * (size_t)n_src * stride);
        if (e == cudaSuccess) e = cudaMemcpyAsync(dsrc, src_host, sizeof(float) * (size_t)n_src * stride, cudaMemcpyHostToDevice, s);
    }
    if (e == cudaSuccess && params->mode == 1 && target->n > 0) {
        e = cudaMalloc(&dn, sizeof(float) * 3 * (size_t)target->n);
        if (e == cudaSuccess) e = cudaMemcpyAsync(dn, tar_normals_host, sizeof(float) * 3 * (size_t)target->n, cudaMemcpyHostToDevice, s);
    }
    if (e != cudaSuccess) { set_error("pcnn_icp_host: staging failed: %s", cudaGetErrorString(e)); rc = PCNN_ERR_NOMEM; }
    if (rc == PCNN_OK) rc = pcnn_icp(target, dn, dsrc, n_src, stride, init_T_host, params, out_host, s);
    cudaStreamSynchronize(s);
    cudaFree(dsrc); cudaFree(dn);
    cudaStreamDestroy(s);
    return rc;
}

int pcnn_iss_detect(const pcnn_index* index, const pcnn_iss_params* params, int32_t* out_idx_dev,
                    int64_t capacity, int64_t* out_count_host, float* out_lambda3_dev, pcnn_stream stream) {
    PCNN_CHECK_ARG(index && params && out_count_host, "pcnn_iss_detect: NULL argument");
    PCNN_CHECK_ARG(capacity >= 0 && (capacity == 0 || out_idx_dev), "pcnn_iss_detect: out_idx is NULL or capacity negative");
    PCNN_CHECK_ARG(params->local_radius == params->local_radius && params->non_max_radius == params->non_max_radius,
                   "pcnn_iss_detect: radius is NaN");
    DeviceGuard g(index->device);
    if (!g.ok) { set_error("pcnn_iss_detect: cannot select CUDA device %d (no CPU fallback exists)", index->device); return PCNN_ERR_CUDA; }
    cudaStream_t s = (cudaStream_t)stream;
    Scratch sc(s);
    int64_t* count_dev;
    PCNN_CUDA(sc.get(&count_dev, 1));
    int rc = launch_iss(index, params, out_idx_dev, capacity, count_dev, out_lambda3_dev, s);
    if (rc != PCNN_OK) return rc;
    PCNN_CUDA(cudaMemcpyAsync(out_count_host, count_dev, sizeof(int64_t), cudaMemcpyDeviceToHost, s));
    PCNN_CUDA(cudaStreamSynchronize(s));
    return PCNN_OK;
}

int pcnn_iss_detect_host(const pcnn_index* index, const pcnn_iss_params* params, int32_t* out_idx_host,
                         int64_t capacity, int64_t* out_count_host, float* out_lambda3_host) {
    PCNN_CHECK_ARG(index && params && out_count_host, "pcnn_iss_detect_host: NULL argument");
    PCNN_CHECK_ARG(capacity >= 0 && (capacity == 0 || out_idx_host), "pcnn_iss_detect_host: out_idx is NULL or capacity negative");
    DeviceGuard g(index->device);
    if (!g.ok) { set_error("pcnn_iss_detect_host: cannot select CUDA device %d (no CPU fallback exists)", index->device); return PCNN_ERR_CUDA; }
    cudaStream_t s = nullptr;
    PCNN_CUDA(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
    wait_build(index, s);
    int rc = PCNN_OK;
    {
        Scratch sc(s);
        int32_t* idx_dev = nullptr;
        float* lam_dev = nullptr;
        const int64_t cap = capacity < index->n ? capacity : index->n;  // never more keypoints than points
        cudaError_t e = sc.get(&idx_dev, (size_t)(cap > 0 ? cap : 1));
        if (e == cudaSuccess && out_lambda3_host) e = sc.get(&lam_dev, (size_t)(index->n > 0 ? index->n : 1));
        if (e != cudaSuccess) { set_error("pcnn_iss_detect_host: device allocation failed: %s", cudaGetErrorString(e)); rc = PCNN_ERR_NOMEM; }
        if (rc == PCNN_OK) rc = pcnn_iss_detect(index, params, idx_dev, cap, out_count_host, lam_dev, (pcnn_stream)s);
        if (rc == PCNN_OK) {
            const int64_t got = *out_count_host < cap ? *out_count_host : cap;
            if (got > 0 && cudaMemcpyAsync(out_idx_host, idx_dev, sizeof(int32_t) * (size_t)got, cudaMemcpyDeviceToHost, s) != cudaSuccess) rc = PCNN_ERR_CUDA;
            if (rc == PCNN_OK && lam_dev && index->n > 0 &&
                cudaMemcpyAsync(out_lambda3_host, lam_dev, sizeof(float) * (size_t)index->n, cudaMemcpyDeviceToHost, s) != cudaSuccess) rc = PCNN_ERR_CUDA;
            if (cudaStreamSynchronize(s) != cudaSuccess) rc = PCNN_ERR_CUDA;
            if (rc == PCNN_ERR_CUDA) set_error("pcnn_iss_detect_host: copy back failed: %s", cudaGetErrorString(cudaGetLastError()));
        }
    }
    cudaStreamSynchronize(s);
    cudaStreamDestroy(s);
    return rc;
}

}  // extern "C"
