// pcnn_index.cu -- spatial index build on device (sm_100a).
//
// Replaces KDTreeConstruction -> KDTreeBuildFastMedian (Homework2/hw2/include/kdtree.hpp:419-429,
// :208-261) and OctreeConstruction -> OctreeBuild (octree.hpp:242-271, :36-90).  The reference
// builds pointer trees by recursive median selection / 8-way partition on one core,
// O(N log N) with a full index-list copy per node.  Here:
//   1. bbox_partial/bbox_final  bounding box -> cubic root cell (as octree.hpp:252-264 takes
//                               the longest side), 2^21 cells per axis
//   2. morton_kernel            63-bit Morton code per point (the octree's child code
//                               x>cx | y>cy<<1 | z>cz<<2, octree.hpp:63-69, applied 21 levels deep)
//   3. radix_sort_pairs         hand-written LSD radix sort of (code, index)
//   4. gather_kernel            float4 cloud in Morton order, .w = original index
//   5. box_level0/box_up        AABBs of 8-point groups, then pairwise unions up to the root
// Everything is enqueued on one stream; no host round trip.
#include "pcnn_internal.cuh"
#include "pcnn_sort.cuh"

namespace pcnn {

static thread_local char g_err[512] = "";
void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

// one pool per device, created on first use
cudaMemPool_t lib_pool() {
    static std::mutex mu;
    static cudaMemPool_t pools[64] = {};
    static bool tried[64] = {};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
    std::lock_guard<std::mutex> lk(mu);
    if (!tried[dev]) {
        tried[dev] = true;
        cudaMemPoolProps props = {};
        props.allocType = cudaMemAllocationTypePinned;
        props.handleTypes = cudaMemHandleTypeNone;
        props.location.type = cudaMemLocationTypeDevice;
        props.location.id = dev;
        cudaMemPool_t pool = nullptr;
        if (cudaMemPoolCreate(&pool, &props) == cudaSuccess) {
            unsigned long long keep = ~0ULL;
            cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
            pools[dev] = pool;
        } else {
            cudaGetLastError();
        }
    }
    return pools[dev];
}

constexpr int BB_THREADS = 256;
constexpr int BB_BLOCKS = 592;  // 4 per SM on 148 SMs

__device__ __forceinline__ float warp_min(float v) {
    for (int o = 16; o > 0; o >>= 1) v = fminf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// partial[b*6 + 0..2] = min, +3..5 = max of block b's grid-stride slice
__global__ void __launch_bounds__(BB_THREADS)
bbox_partial(const float* __restrict__ xyz, int n, int stride, float* __restrict__ partial) {
    const float inf = __int_as_float(0x7f800000);
    float mn[3] = {inf, inf, inf}, mx[3] = {-inf, -inf, -inf};
    for (int i = blockIdx.x * BB_THREADS + threadIdx.x; i < n; i += gridDim.x * BB_THREADS) {
        const float* p = xyz + (size_t)i * stride;
#pragma unroll
        for (int a = 0; a < 3; a++) {
            float v = p[a];
            mn[a] = fminf(mn[a], v);
            mx[a] = fmaxf(mx[a], v);
        }
    }
    __shared__ float s[6][BB_THREADS / 32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int a = 0; a < 3; a++) {
        float lo = warp_min(mn[a]), hi = warp_max(mx[a]);
        if (lane == 0) { s[a][warp] = lo; s[3 + a][warp] = hi; }
    }
    __syncthreads();
    if (threadIdx.x < 6) {
        float v = s[threadIdx.x][0];
        for (int w = 1; w < BB_THREADS / 32; w++)
            v = threadIdx.x < 3 ? fminf(v, s[threadIdx.x][w]) : fmaxf(v, s[threadIdx.x][w]);
        partial[blockIdx.x * 6 + threadIdx.x] = v;
    }
}

// cube[0..2] = min, cube[3] = 2^21 / side, cube[4..6] = max, cube[7] = side,
// cube[8] = mean extent of an 8-point group (group_extent_kernel)
__global__ void bbox_final(const float* __restrict__ partial, int nparts, float* __restrict__ cube) {
    const int a = threadIdx.x;
    if (a < 6) {
        float v = partial[a];
        for (int b = 1; b < nparts; b++)
            v = a < 3 ? fminf(v, partial[b * 6 + a]) : fmaxf(v, partial[b * 6 + a]);
        cube[a < 3 ? a : a + 1] = v;
    }
    __syncthreads();
    if (a == 0) {
        float side = fmaxf(cube[4] - cube[0], fmaxf(cube[5] - cube[1], cube[6] - cube[2]));
        if (!(side > 0.f) || !(side < 3.0e38f)) side = 1.f;  // single point / degenerate / non-finite
        cube[7] = side;
        cube[3] = 2097152.0f / side;
    }
}

// cube[8] = mean over a sample of <= 4096 groups of the second largest side of the group's box:
// the scale at which the cloud is locally sampled (a surface patch of 8 points is about that
// wide).  The run-batched radius kernels size their runs of queries with it.
__global__ void __launch_bounds__(256) group_extent_kernel(const float4* __restrict__ box0, int n0,
                                                           float* __restrict__ out) {
    __shared__ float wsum[8];
    __shared__ int wcnt[8];
    const int samples = n0 < 4096 ? n0 : 4096;
    float acc = 0.f;
    int cnt = 0;
    for (int i = threadIdx.x; i < samples; i += 256) {
        const size_t g = (size_t)((long long)i * n0 / samples);
        const float4 lo = box0[2 * g], hi = box0[2 * g + 1];
        const float ex = hi.x - lo.x, ey = hi.y - lo.y, ez = hi.z - lo.z;
        const float mid = fmaxf(fminf(ex, ey), fminf(fmaxf(ex, ey), ez));  // median of the three sides
        if (mid == mid) { acc += mid; cnt++; }
    }
    for (int o = 16; o > 0; o >>= 1) {
        acc += __shfl_xor_sync(0xffffffffu, acc, o);
        cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
    }
    if ((threadIdx.x & 31) == 0) { wsum[threadIdx.x >> 5] = acc; wcnt[threadIdx.x >> 5] = cnt; }
    __syncthreads();
    if (threadIdx.x == 0) {
        float a = 0.f;
        int c = 0;
        for (int i = 0; i < 8; i++) { a += wsum[i]; c += wcnt[i]; }
        out[0] = c > 0 ? a / (float)c : 0.f;
    }
}

__global__ void morton_kernel(const float* __restrict__ xyz, int n, int stride,
                              const float* __restrict__ cube, unsigned long long* __restrict__ codes,
                              unsigned* __restrict__ perm) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float* p = xyz + (size_t)i * stride;
    codes[i] = morton63(p[0], p[1], p[2], cube, cube[3]);
    perm[i] = (unsigned)i;
}

__global__ void gather_kernel(const float* __restrict__ xyz, int n, int stride,
                              const unsigned* __restrict__ perm,
                              const unsigned long long* __restrict__ codes, float4* __restrict__ pts,
                              unsigned long long* __restrict__ gcode) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    unsigned src = perm[i];
    const float* p = xyz + (size_t)src * stride;
    pts[i] = make_float4(p[0], p[1], p[2], __int_as_float((int)src));
    if ((i & (PCNN_GROUP - 1)) == 0) gcode[i >> PCNN_GROUP_LOG] = codes[i];
}

__global__ void box_level0(const float4* __restrict__ pts, int n, int n0, float4* __restrict__ box) {
    int g = blockIdx.x * blockDim.x + threadIdx.x;
    const float inf = __int_as_float(0x7f800000);
    if (g >= n0) {
        if (g == n0 && (n0 & 1) && n0 > 1) {  // the empty box behind an odd level
            box[2 * (size_t)g] = make_float4(inf, inf, inf, 0.f);
            box[2 * (size_t)g + 1] = make_float4(-inf, -inf, -inf, 0.f);
        }
        return;
    }
    float4 lo = make_float4(inf, inf, inf, 0.f), hi = make_float4(-inf, -inf, -inf, 0.f);
    int first = g << PCNN_GROUP_LOG;
#pragma unroll
    for (int j = 0; j < PCNN_GROUP; j++) {
        if (first + j < n) {
            float4 p = pts[first + j];
            lo.x = fminf(lo.x, p.x); lo.y = fminf(lo.y, p.y); lo.z = fminf(lo.z, p.z);
            hi.x = fmaxf(hi.x, p.x); hi.y = fmaxf(hi.y, p.y); hi.z = fmaxf(hi.z, p.z);
        }
    }
    box[2 * (size_t)g] = lo;
    box[2 * (size_t)g + 1] = hi;
}

__global__ void box_up(const float4* __restrict__ child, int nchild, float4* __restrict__ parent,
                       int nparent) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nparent) {
        if (i == nparent && (nparent & 1) && nparent > 1) {  // the empty box behind an odd level
            const float inf = __int_as_float(0x7f800000);
            parent[2 * (size_t)i] = make_float4(inf, inf, inf, 0.f);
            parent[2 * (size_t)i + 1] = make_float4(-inf, -inf, -inf, 0.f);
        }
        return;
    }
    float4 lo = child[2 * (size_t)(2 * i)], hi = child[2 * (size_t)(2 * i) + 1];
    if (2 * i + 1 < nchild) {
        float4 lo2 = child[2 * (size_t)(2 * i + 1)], hi2 = child[2 * (size_t)(2 * i + 1) + 1];
        lo.x = fminf(lo.x, lo2.x); lo.y = fminf(lo.y, lo2.y); lo.z = fminf(lo.z, lo2.z);
        hi.x = fmaxf(hi.x, hi2.x); hi.y = fmaxf(hi.y, hi2.y); hi.z = fmaxf(hi.z, hi2.z);
    }
    parent[2 * (size_t)i] = lo;
    parent[2 * (size_t)i + 1] = hi;
}

static void free_index(pcnn_index* ix) {
    if (!ix) return;
    free_host_pipe(ix->pipe);
    if (ix->rperm) cudaFree(ix->rperm);
    if (ix->rperm_ev) cudaEventDestroy(ix->rperm_ev);
    cudaFree(ix->pts);
    cudaFree(ix->gcode);
    cudaFree(ix->box);
    cudaFree(ix->cube);
    if (ix->ev_begin) cudaEventDestroy(ix->ev_begin);
    if (ix->ev_end) cudaEventDestroy(ix->ev_end);
    if (ix->prof0) cudaEventDestroy(ix->prof0);
    if (ix->prof1) cudaEventDestroy(ix->prof1);
    delete ix;
}

static int build_impl(pcnn_index** out, int kind, const float* xyz_dev, int64_t n, int stride,
                      int leaf_size, double min_extent, int device, cudaStream_t s) {
    PCNN_CHECK_ARG(out != nullptr, "pcnn_index_build: out is NULL");
    *out = nullptr;
    PCNN_CHECK_ARG(kind == PCNN_KDTREE || kind == PCNN_OCTREE, "pcnn_index_build: unknown kind %d", kind);
    PCNN_CHECK_ARG(n >= 0 && n < (int64_t)0x7fffffff - 4096, "pcnn_index_build: n=%lld out of range", (long long)n);
    PCNN_CHECK_ARG(stride >= 3, "pcnn_index_build: stride must be >= 3 floats (got %d)", stride);
    PCNN_CHECK_ARG(leaf_size >= 1, "pcnn_index_build: leaf_size must be >= 1 (got %d)", leaf_size);
    PCNN_CHECK_ARG(min_extent >= 0, "pcnn_index_build: min_extent must be >= 0");
    PCNN_CHECK_ARG(n == 0 || xyz_dev != nullptr, "pcnn_index_build: xyz is NULL");

    pcnn_index* ix = new (std::nothrow) pcnn_index;
    if (!ix) { set_error("pcnn_index_build: host allocation failed"); return PCNN_ERR_NOMEM; }
    ix->kind = kind; ix->device = device; ix->n = n; ix->leaf_size = leaf_size; ix->min_extent = min_extent;

    const int ni = (int)n;
    const int n0 = (ni + PCNN_GROUP - 1) / PCNN_GROUP;
    IndexView& v = ix->view;
    v.n = ni; v.n0 = n0; v.nlev = 0;
    size_t nboxes = 0;
    for (int c = n0; c >= 1; c = (c + 1) / 2) {
        v.lev_off[v.nlev++] = (int)nboxes;
        // a level with an odd number of nodes is followed by one EMPTY box (lo = +inf, hi = -inf: every lower
        // bound against it is +inf), so that the walks load the two boxes of a pair unconditionally
        nboxes += (size_t)c + (size_t)(c > 1 ? (c & 1) : 0);
        if (c == 1) break;
    }
    if (n0 == 0) { v.nlev = 1; v.lev_off[0] = 0; nboxes = 1; }

#define PCNN_BUILD_CUDA(call)                                                                  \
    do {                                                                                       \
        cudaError_t e__ = (call);                                                              \
        if (e__ != cudaSuccess) {                                                              \
            set_error("%s:%d: %s failed: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__)); \
            free_index(ix);                                                                    \
            return e__ == cudaErrorMemoryAllocation ? PCNN_ERR_NOMEM : PCNN_ERR_CUDA;          \
        }                                                                                      \
    } while (0)

    PCNN_BUILD_CUDA(cudaMalloc(&ix->pts, sizeof(float4) * (size_t)(n0 * PCNN_GROUP + PCNN_GROUP)));
    PCNN_BUILD_CUDA(cudaMalloc(&ix->gcode, sizeof(unsigned long long) * (size_t)(n0 + 1)));
    PCNN_BUILD_CUDA(cudaMalloc(&ix->box, sizeof(float4) * 2 * nboxes));
    PCNN_BUILD_CUDA(cudaMalloc(&ix->cube, sizeof(float) * 12));
    ix->bytes = sizeof(float4) * (size_t)(n0 * PCNN_GROUP + PCNN_GROUP) +
                sizeof(unsigned long long) * (size_t)(n0 + 1) + sizeof(float4) * 2 * nboxes + 32;
    v.pts = ix->pts; v.gcode = ix->gcode; v.box = ix->box; v.cube = ix->cube;

    // owned by the handle from the moment they exist, so free_index destroys them on every error path
    PCNN_BUILD_CUDA(cudaEventCreate(&ix->ev_begin));
    PCNN_BUILD_CUDA(cudaEventCreate(&ix->ev_end));
    PCNN_BUILD_CUDA(cudaEventRecord(ix->ev_begin, s));
    // padding points of the last partial group are never read (count-guarded) but keep them defined
    PCNN_BUILD_CUDA(cudaMemsetAsync(ix->pts, 0xff, sizeof(float4) * (size_t)(n0 * PCNN_GROUP + PCNN_GROUP), s));
    if (ni > 0) {
        Scratch sc(s);
        float* partial; unsigned long long *c0, *c1; unsigned *p0, *p1, *hist;
        PCNN_BUILD_CUDA(sc.get(&partial, (size_t)BB_BLOCKS * 6));
        PCNN_BUILD_CUDA(sc.get(&c0, (size_t)ni));
        PCNN_BUILD_CUDA(sc.get(&c1, (size_t)ni));
        PCNN_BUILD_CUDA(sc.get(&p0, (size_t)ni));
        PCNN_BUILD_CUDA(sc.get(&p1, (size_t)ni));
        PCNN_BUILD_CUDA(sc.get(&hist, radix_hist_elems(ni)));
        int nb = (ni + BB_THREADS - 1) / BB_THREADS;
        if (nb > BB_BLOCKS) nb = BB_BLOCKS;
        bbox_partial<<<nb, BB_THREADS, 0, s>>>(xyz_dev, ni, stride, partial);
        bbox_final<<<1, 32, 0, s>>>(partial, nb, ix->cube);
        const int T = 256, G = (ni + T - 1) / T;
        morton_kernel<<<G, T, 0, s>>>(xyz_dev, ni, stride, ix->cube, c0, p0);
        int where = radix_sort_pairs<unsigned long long>(c0, p0, c1, p1, hist, ni, 3 * PCNN_MORTON_BITS, s);
        gather_kernel<<<G, T, 0, s>>>(xyz_dev, ni, stride, where ? p1 : p0, where ? c1 : c0, ix->pts, ix->gcode);
        box_level0<<<(n0 + 1 + T - 1) / T, T, 0, s>>>(ix->pts, ni, n0, ix->box);
        int cnt = n0;
        for (int l = 1; l < v.nlev; l++) {
            int pc = (cnt + 1) / 2;
            box_up<<<(pc + 1 + T - 1) / T, T, 0, s>>>(ix->box + 2 * (size_t)v.lev_off[l - 1], cnt,
                                                   ix->box + 2 * (size_t)v.lev_off[l], pc);
            cnt = pc;
        }
        group_extent_kernel<<<1, 256, 0, s>>>(ix->box, n0, ix->cube + 8);
        PCNN_BUILD_CUDA(cudaGetLastError());
    } else {
        PCNN_BUILD_CUDA(cudaMemsetAsync(ix->cube, 0, sizeof(float) * 12, s));
        PCNN_BUILD_CUDA(cudaMemsetAsync(ix->box, 0, sizeof(float4) * 2, s));
    }
    PCNN_BUILD_CUDA(cudaEventRecord(ix->ev_end, s));
    // The handle is usable by later work on the same stream immediately; work on any OTHER stream
    // (the host-pointer entry points run on internal streams) first waits for ev_end: wait_build().
    // build_ms needs the events to have completed, so it is resolved lazily in pcnn_index_get_info.
    ix->build_ms = -1.f;
#undef PCNN_BUILD_CUDA
    *out = ix;
    return PCNN_OK;
}

}  // namespace pcnn

// ------------------------------------------------------------------ C ABI (index part)
using namespace pcnn;

extern "C" {

const char* pcnn_last_error(void) { return g_err; }
int pcnn_version(void) { return PCNN_VERSION; }

int pcnn_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

int pcnn_index_build(pcnn_index** out, int kind, const float* xyz_dev, int64_t n, int stride,
                     int leaf_size, double min_extent, int device, pcnn_stream stream) {
    DeviceGuard g(device);
    if (!g.ok) { set_error("pcnn_index_build: cannot select CUDA device %d (no CPU fallback exists)", device); return PCNN_ERR_CUDA; }
    return build_impl(out, kind, xyz_dev, n, stride, leaf_size, min_extent, device, (cudaStream_t)stream);
}

int pcnn_index_build_host(pcnn_index** out, int kind, const float* xyz_host, int64_t n, int stride,
                          int leaf_size, double min_extent, int device) {
    DeviceGuard g(device);
    if (!g.ok) { set_error("pcnn_index_build_host: cannot select CUDA device %d (no CPU fallback exists)", device); return PCNN_ERR_CUDA; }
    PCNN_CHECK_ARG(out != nullptr, "pcnn_index_build_host: out is NULL");
    *out = nullptr;
    PCNN_CHECK_ARG(n >= 0 && stride >= 3, "pcnn_index_build_host: bad n/stride");
    PCNN_CHECK_ARG(n == 0 || xyz_host != nullptr, "pcnn_index_build_host: xyz is NULL");
    float* d = nullptr;
    size_t bytes = sizeof(float) * (size_t)n * (size_t)stride;
    cudaStream_t s = nullptr;
    PCNN_CUDA(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
    int rc = PCNN_OK;
    cudaError_t e = cudaSuccess;
    if (n > 0) {
        e = cudaMalloc(&d, bytes);
        if (e == cudaSuccess) e = cudaMemcpyAsync(d, xyz_host, bytes, cudaMemcpyHostToDevice, s);
    }
    if (e != cudaSuccess) {
        set_error("pcnn_index_build_host: staging the cloud failed: %s", cudaGetErrorString(e));
        rc = e == cudaErrorMemoryAllocation ? PCNN_ERR_NOMEM : PCNN_ERR_CUDA;
    } else {
        rc = build_impl(out, kind, d, n, stride, leaf_size, min_extent, device, s);
    }
    e = cudaStreamSynchronize(s);
    if (rc == PCNN_OK && e != cudaSuccess) {
        set_error("pcnn_index_build_host: build failed on device: %s", cudaGetErrorString(e));
        free_index(*out);
        *out = nullptr;
        rc = PCNN_ERR_CUDA;
    }
    cudaFree(d);
    cudaStreamDestroy(s);
    return rc;
}

int pcnn_index_destroy(pcnn_index* index) {
    if (!index) return PCNN_OK;
    DeviceGuard g(index->device);
    cudaDeviceSynchronize();
    free_index(index);
    return PCNN_OK;
}

int pcnn_index_get_info(const pcnn_index* index, pcnn_index_info* out) {
    PCNN_CHECK_ARG(index && out, "pcnn_index_get_info: NULL argument");
    DeviceGuard g(index->device);
    pcnn_index* ix = const_cast<pcnn_index*>(index);
    float cube[8] = {0};
    PCNN_CUDA(cudaEventSynchronize(ix->ev_end));
    if (ix->build_ms < 0) PCNN_CUDA(cudaEventElapsedTime(&ix->build_ms, ix->ev_begin, ix->ev_end));
    PCNN_CUDA(cudaMemcpy(cube, ix->cube, sizeof(cube), cudaMemcpyDeviceToHost));
    out->n_points = ix->n;
    out->kind = ix->kind;
    out->device = ix->device;
    out->n_levels = ix->view.nlev;
    out->group_size = PCNN_GROUP;
    out->device_bytes = (int64_t)ix->bytes;
    for (int a = 0; a < 3; a++) { out->bbox_min[a] = cube[a]; out->bbox_max[a] = cube[4 + a]; }
    out->build_ms = ix->build_ms;
    return PCNN_OK;
}

int pcnn_debug_warp_times(pcnn_index* index, unsigned long long* buf_dev) {
    PCNN_CHECK_ARG(index != nullptr, "pcnn_debug_warp_times: index is NULL");
    index->dbg_warp_times = buf_dev;
    return PCNN_OK;
}

int pcnn_release_cached(int device) {
    DeviceGuard g(device);
    if (!g.ok) { set_error("pcnn_release_cached: cannot select CUDA device %d", device); return PCNN_ERR_CUDA; }
    cudaMemPool_t pool = lib_pool();
    if (pool) {
        PCNN_CUDA(cudaDeviceSynchronize());
        PCNN_CUDA(cudaMemPoolTrimTo(pool, 0));
    }
    return PCNN_OK;
}

int pcnn_get_profile(const pcnn_index* index, float* main_kernel_ms) {
    PCNN_CHECK_ARG(index && main_kernel_ms, "pcnn_get_profile: NULL argument");
    PCNN_CHECK_ARG(index->opt_profile && index->prof0, "pcnn_get_profile: option \"profile\" is not set");
    DeviceGuard g(index->device);
    PCNN_CUDA(cudaEventSynchronize(index->prof1));
    PCNN_CUDA(cudaEventElapsedTime(main_kernel_ms, index->prof0, index->prof1));
    return PCNN_OK;
}

int pcnn_set_option(pcnn_index* index, const char* name, int value) {
    PCNN_CHECK_ARG(index && name, "pcnn_set_option: NULL argument");
    if (!std::strcmp(name, "sort_queries")) {
        PCNN_CHECK_ARG(value == 0 || value == 1, "pcnn_set_option: sort_queries must be 0 or 1");
        index->opt_sort_queries = value;
    } else if (!std::strcmp(name, "method")) {
        PCNN_CHECK_ARG(value >= 0 && value <= 10, "pcnn_set_option: method must be 0..10");
#ifndef PCNN_STUDY
        if (value != 0 && value != 2 && value != 3 && value != 7 && value != 9) {
            set_error("pcnn_set_option: kNN method %d is a study variant; this build ships 0 (auto), 2, 3, 7 and 9 "
                      "(build with -DPCNN_STUDY for the others)", value);
            return PCNN_ERR_UNSUPPORTED;
        }
#endif
        index->opt_method = value;
    } else if (!std::strcmp(name, "profile")) {
        PCNN_CHECK_ARG(value == 0 || value == 1, "pcnn_set_option: profile must be 0 or 1");
        DeviceGuard g(index->device);
        if (value && !index->prof0) {
            PCNN_CUDA(cudaEventCreate(&index->prof0));
            PCNN_CUDA(cudaEventCreate(&index->prof1));
        }
        index->opt_profile = value;
    } else if (!std::strcmp(name, "nvtx")) {
        PCNN_CHECK_ARG(value == 0 || value == 1, "pcnn_set_option: nvtx must be 0 or 1");
        index->opt_nvtx = value;
    } else if (!std::strcmp(name, "pdl")) {
        PCNN_CHECK_ARG(value == 0 || value == 1, "pcnn_set_option: pdl must be 0 or 1");
        index->opt_pdl = value;
    } else if (!std::strcmp(name, "icp_pdl")) {
        PCNN_CHECK_ARG(value == 0 || value == 1, "pcnn_set_option: icp_pdl must be 0 or 1");
        index->opt_icp_pdl = value;
    } else if (!std::strcmp(name, "icp_certificates")) {
        PCNN_CHECK_ARG(value == 0 || value == 1, "pcnn_set_option: icp_certificates must be 0 or 1");
        index->opt_icp_cert = value;
    } else if (!std::strcmp(name, "leaf_log")) {
        PCNN_CHECK_ARG(value >= 0 && value <= 6, "pcnn_set_option: leaf_log must be 0..6");
        index->opt_leaf_log = value;
    } else if (!std::strcmp(name, "coherent_leaf_log")) {
        PCNN_CHECK_ARG(value >= 0 && value <= 6, "pcnn_set_option: coherent_leaf_log must be 0..6");
        index->opt_coherent_leaf_log = value;
    } else if (!std::strcmp(name, "coherent_budget")) {
        PCNN_CHECK_ARG(value >= 0, "pcnn_set_option: coherent_budget must be >= 0");
        index->opt_coherent_budget = value;
    } else if (!std::strcmp(name, "radius_batch")) {
        PCNN_CHECK_ARG(value >= 0 && value <= 5, "pcnn_set_option: radius_batch must be 0 (warp per query), 1 (auto), 2 (batched runs), 3 (lock-step), 4 (stack walk) or 5 (stack walk, batched runs)");
#ifndef PCNN_STUDY
        if (value == 2 || value == 3) {
            set_error("pcnn_set_option: radius kernel %d is a study variant; this build ships 1 (auto), 0, 4 and 5 "
                      "(build with -DPCNN_STUDY for the others)", value);
            return PCNN_ERR_UNSUPPORTED;
        }
#endif
        index->opt_radius_batch = value;
    } else if (!std::strcmp(name, "radius_run_div")) {
        PCNN_CHECK_ARG(value >= 0 && value <= 1024, "pcnn_set_option: radius_run_div must be 0 (default per pass) or 1..1024");
        index->opt_radius_run_div = value;
    } else if (!std::strcmp(name, "radius_run_mult_count") || !std::strcmp(name, "radius_run_mult_fill")) {
        PCNN_CHECK_ARG(value >= 1 && value <= 100000, "pcnn_set_option: radius_run_mult_* is a percentage in 1..100000");
        (name[16] == 'c' ? index->opt_radius_run_mult_count : index->opt_radius_run_mult_fill) = value;
    } else if (!std::strcmp(name, "radius_sort_bits")) {
        PCNN_CHECK_ARG(value >= 3 && value <= 63 && value % 3 == 0, "pcnn_set_option: radius_sort_bits must be a multiple of 3 in 3..63");
        index->opt_radius_sort_bits = value;
    } else if (!std::strcmp(name, "radius_leaf_level")) {
        PCNN_CHECK_ARG(value >= 2 && value <= 6, "pcnn_set_option: radius_leaf_level must be 2..6");
        index->opt_radius_leaf_level = value;
    } else if (!std::strcmp(name, "sort_bits")) {
        PCNN_CHECK_ARG(value >= 3 && value <= 63, "pcnn_set_option: sort_bits must be in 3..63");
        index->opt_sort_bits = value;
    } else if (!std::strcmp(name, "coherent_budget2")) {
        PCNN_CHECK_ARG(value >= 0, "pcnn_set_option: coherent_budget2 must be >= 0");
        index->opt_coherent_budget2 = value;
    } else {
        set_error("pcnn_set_option: unknown option '%s'", name);
        return PCNN_ERR_INVALID;
    }
    return PCNN_OK;
}

}  // extern "C"
