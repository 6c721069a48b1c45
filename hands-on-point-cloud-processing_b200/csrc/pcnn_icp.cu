// pcnn_icp.cu -- ICP iteration as one fused kernel (sm_100a) + on-device pose update.
//
// Replaces Registration::ICPpoint2point (Homework9/hw9/src/registration.cpp:862-1011) and
// Registration::ICPpoint2plane (:710-859), whose iteration is: nanoflann 1-NN per source
// point (:772-785, :925-941) -> gate on the squared distance -> gather pairs into dense
// matrices (:788-815, :964-978) -> Kabsch SVD (:979-998) or normal equations (:818) ->
// T_total = T_delta * T_total (:849, :1002) -> transformCloudInplace (:165-178, :851, :1003).
//
// Here one kernel per iteration does, per source point: apply the pending T_delta in float32
// (same operation order as the reference, so the cloud stays bit-comparable), exact float32
// 1-NN through the Morton hierarchy, the gate, and the per-pair terms; a warp-shuffle +
// shared-memory reduction in double gives per-block partials, and the last block to finish
// folds them in a fixed order (deterministic, no floating-point atomics) and solves the 3x3
// SVD / 6x6 Cholesky on device.  The host never waits inside the loop.  The dense M x 6
// matrix A of the reference is never materialised: only the 29 scalars of A^T A, A^T b, b^T b
// and the count are carried.
#include <type_traits>
#include "pcnn_internal.cuh"
#include "pcnn_linalg.cuh"
#include "pcnn_search.cuh"
#include "pcnn_lockstep.cuh"

struct pcnn_icp_state {
    const pcnn_index* target = nullptr;
    int device = 0;
    int64_t n_src = 0;
    pcnn_icp_params prm{};
    float4* cur = nullptr;       // source cloud, current pose, Morton order, .w = global source index
    float4* nrm = nullptr;       // target normals in the index's Morton order (mode 1)
    int* prev = nullptr;         // per source point: Morton position of its last correspondence (-1 = none)
    float4* cert = nullptr;      // per source point: where it was last searched (.xyz) and the certificate radius (.w)
    int* prev2 = nullptr;        // per source point: the runner-up of that search if it matters to the certificate, else -1
    int use_cert = 1;
    int use_pdl = 1;             // programmatic dependent launch between the kernels of an iteration
    float gapk = 4.0f, move_frac = 0.03f;
    double* partials = nullptr;  // [nblocks][32]
    struct Dev* dev = nullptr;   // device-resident loop state
    int nblocks = 0;
    int acc_blocks = 0;          // grid of the sums kernel: fewer, fatter blocks = fewer partial rows to fold
    // device memory comes from the library's stream-ordered pool on the stream given to
    // pcnn_icp_create and goes back to it on the same stream: no device-wide synchronisation per
    // registration (cudaMalloc/cudaFree cost 50 ms per call next to a few GB of cached blocks)
    cudaStream_t mem_stream = nullptr;
    // peer exchange (pcnn_icp_step_peer): this rank's mailbox (cudaMalloc: IPC-exportable) and the
    // peers' mailboxes as mapped into this process
    struct PeerBox* box = nullptr;
    struct PeerBox* peer[PCNN_ICP_MAX_PEERS] = {};
    bool peer_ipc[PCNN_ICP_MAX_PEERS] = {};
    int rank = 0, world = 0;
};

namespace pcnn {

constexpr int ICP_THREADS = 128;
constexpr int ICP_NS = PCNN_ICP_NSUMS;  // 32
constexpr int ICP_FOLD_G = 16;     // partial rows folded by the last block of each group of 16 blocks ...
constexpr int ICP_MAX_GROUPS = 128; // ... and the group rows by the last group to finish (grids of up to 2048 blocks)
// sums layout: p2p  [0..2]=sum p  [3..5]=sum q  [6..14]=sum q_r*p_c  [28]=count
//              p2pl [0..20]=upper triangle of A^T A  [21..26]=A^T b  [27]=b^T b  [28]=count
//              [31] (as int64) = (global index of the last kept pair << 32) | float bits of its d2
}  // namespace pcnn

struct Dev {
    float T[16];
    float Rd[9];
    float td[3];
    double sums[pcnn::ICP_NS];
    int iters;
    int converged;
    int status;
    int done;
    int unchanged;
    float last_loss;
    float loss;
    long long last_pairs;
    unsigned block_counter;
    unsigned group_counter[pcnn::ICP_MAX_GROUPS];  // two-level fold: blocks of a group that have written their partial row
    unsigned long long peer_seq;  // iterations exchanged so far (identical on every rank)
    unsigned long long walked;    // searches run so far (the rest were certified); PCNN_ICP_STATS prints it
    unsigned long long walked_mark;
    int last_searches;            // searches of the last executed iteration
};

// One per rank, in that rank's memory; peers store into it.  Two slots: a rank can be at most
// one iteration ahead of the slowest peer (it needs that peer's flag of iteration i to finish i).
struct PeerBox {
    double sums[2][PCNN_ICP_MAX_PEERS][PCNN_ICP_NSUMS];
    unsigned long long flag[2][PCNN_ICP_MAX_PEERS];
};
struct PeerView {
    PeerBox* box[PCNN_ICP_MAX_PEERS];
    int rank, world;
};

namespace pcnn {

__device__ __forceinline__ void xform(const float* R, const float* t, float& x, float& y, float& z) {
    // ((r0*x + r1*y) + r2*z) + t, float32, no FMA: transformCloudInplace, registration.cpp:165-178
    const float nx = fadd(fadd(fadd(fmul(R[0], x), fmul(R[1], y)), fmul(R[2], z)), t[0]);
    const float ny = fadd(fadd(fadd(fmul(R[3], x), fmul(R[4], y)), fmul(R[5], z)), t[1]);
    const float nz = fadd(fadd(fadd(fmul(R[6], x), fmul(R[7], y)), fmul(R[8], z)), t[2]);
    x = nx; y = ny; z = nz;
}

__device__ bool chol6_dev(const double* N, const double* g, double* x) {
    // One thread solves the 6x6 system once per iteration, so its LATENCY is on every iteration's
    // critical path: one square root and one reciprocal per pivot, every other division a multiply
    // (6 + 6 slow operations instead of 6 + 33; 6.5 us -> see DESIGN 4.5).
    double L[36], inv[6];
#pragma unroll
    for (int i = 0; i < 36; i++) L[i] = 0;
#pragma unroll
    for (int j = 0; j < 6; j++) {
        double d = N[j * 6 + j];
#pragma unroll
        for (int k = 0; k < j; k++) d -= L[j * 6 + k] * L[j * 6 + k];
        if (!(d > 0)) return false;
        const double r = sqrt(d);
        L[j * 6 + j] = r;
        inv[j] = 1.0 / r;
#pragma unroll
        for (int i = j + 1; i < 6; i++) {
            double s = N[i * 6 + j];
#pragma unroll
            for (int k = 0; k < j; k++) s -= L[i * 6 + k] * L[j * 6 + k];
            L[i * 6 + j] = s * inv[j];
        }
    }
    double y[6];
#pragma unroll
    for (int i = 0; i < 6; i++) {
        double s = g[i];
#pragma unroll
        for (int k = 0; k < i; k++) s -= L[i * 6 + k] * y[k];
        y[i] = s * inv[i];
    }
#pragma unroll
    for (int i = 5; i >= 0; i--) {
        double s = y[i];
#pragma unroll
        for (int k = i + 1; k < 6; k++) s -= L[k * 6 + i] * x[k];
        x[i] = s * inv[i];
    }
    return true;
}

// Pose update from the (all-reduced) sums: registration.cpp:948-1003 / :820-851.
// Runs in one thread; every rank computes the identical result from identical sums.
__device__ void icp_update_dev(Dev* st, int mode, float loss_epsilon) {
    if (st->done) return;
    st->last_searches = (int)(st->walked - st->walked_mark);
    st->walked_mark = st->walked;
    const double* S = st->sums;
    const double M = S[28];
    float R[9], t[3];
    float loss = 0.f;
    bool ok = M > 0.5;
    if (mode == 0) {
        const long long packed = __double_as_longlong(S[31]);
        if (ok) { const float d2 = __int_as_float((int)(packed & 0xffffffffLL)); loss = fmul(d2, d2); }  // :939
    }
    double x[6] = {0, 0, 0, 0, 0, 0};
    if (mode == 1) {
        double N[36], g[6];
        int u = 0;
        for (int r = 0; r < 6; r++)
            for (int c = r; c < 6; c++) { N[r * 6 + c] = S[u]; N[c * 6 + r] = S[u]; u++; }
        for (int r = 0; r < 6; r++) g[r] = S[21 + r];
        ok = ok && chol6_dev(N, g, x);
        if (ok) {
            double xNx = 0, xg = 0;
            for (int r = 0; r < 6; r++) {
                for (int c = 0; c < 6; c++) xNx += x[r] * N[r * 6 + c] * x[c];
                xg += x[r] * g[r];
            }
            loss = (float)(xNx - 2.0 * xg + S[27]);  // ||A x - b||^2, :820
        }
    }
    st->last_pairs = (long long)(M + 0.5);
    st->loss = loss;
    if (fabsf(st->last_loss - loss) < loss_epsilon) st->unchanged++;
    if (st->unchanged > 15) { st->converged = 1; st->done = 1; st->last_loss = loss; return; }
    st->last_loss = loss;
    if (!ok) { st->status = 1; st->done = 1; return; }
    if (mode == 0) {
        double pc[3], qc[3], H[9];
        for (int a = 0; a < 3; a++) { pc[a] = S[a] / M; qc[a] = S[3 + a] / M; }
        for (int r = 0; r < 3; r++)
            for (int c = 0; c < 3; c++) H[r * 3 + c] = S[6 + r * 3 + c] - M * qc[r] * pc[c];  // tar_c * src_c^T
        double U[9], s[3], V[9], Rd[9];
        svd3_dev(H, U, s, V);
        for (int r = 0; r < 3; r++)
            for (int c = 0; c < 3; c++)
                Rd[r * 3 + c] = U[r * 3] * V[c * 3] + U[r * 3 + 1] * V[c * 3 + 1] + U[r * 3 + 2] * V[c * 3 + 2];
        const double det = Rd[0] * (Rd[4] * Rd[8] - Rd[5] * Rd[7]) - Rd[1] * (Rd[3] * Rd[8] - Rd[5] * Rd[6]) +
                           Rd[2] * (Rd[3] * Rd[7] - Rd[4] * Rd[6]);
        if (det < 0) {  // :990-996 as written: V * diag(1,1,det) * U^T
            const double B[3] = {1, 1, det};
            for (int r = 0; r < 3; r++)
                for (int c = 0; c < 3; c++)
                    Rd[r * 3 + c] = V[r * 3] * B[0] * U[c * 3] + V[r * 3 + 1] * B[1] * U[c * 3 + 1] +
                                    V[r * 3 + 2] * B[2] * U[c * 3 + 2];
        }
        for (int a = 0; a < 9; a++) R[a] = (float)Rd[a];
        for (int r = 0; r < 3; r++)
            t[r] = (float)(qc[r] - (Rd[r * 3] * pc[0] + Rd[r * 3 + 1] * pc[1] + Rd[r * 3 + 2] * pc[2]));
    } else {
        const float xf[6] = {(float)x[0], (float)x[1], (float)x[2], (float)x[3], (float)x[4], (float)x[5]};
        R[0] = 1; R[1] = -xf[2]; R[2] = xf[1];
        R[3] = xf[2]; R[4] = 1; R[5] = -xf[0];
        R[6] = -xf[1]; R[7] = xf[0]; R[8] = 1;  // linearised, :843
        t[0] = xf[3]; t[1] = xf[4]; t[2] = xf[5];
    }
    // T_total = T_delta * T_total in float32 (:849, :1002)
    const float D[16] = {R[0], R[1], R[2], t[0], R[3], R[4], R[5], t[1], R[6], R[7], R[8], t[2], 0, 0, 0, 1};
    float O[16];
    for (int r = 0; r < 4; r++)
        for (int c = 0; c < 4; c++)
            O[r * 4 + c] = fadd(fadd(fadd(fmul(D[r * 4], st->T[c]), fmul(D[r * 4 + 1], st->T[4 + c])),
                                     fmul(D[r * 4 + 2], st->T[8 + c])), fmul(D[r * 4 + 3], st->T[12 + c]));
    for (int i = 0; i < 16; i++) st->T[i] = O[i];
    for (int i = 0; i < 9; i++) st->Rd[i] = R[i];
    for (int i = 0; i < 3; i++) st->td[i] = t[i];
    st->iters++;
}

__global__ void icp_update_kernel(Dev* st, int mode, float loss_epsilon) {
    if (threadIdx.x == 0 && blockIdx.x == 0) icp_update_dev(st, mode, loss_epsilon);
}

__device__ __forceinline__ double warp_sum(double v) {
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ long long warp_max_ll(long long v) {
    for (int o = 16; o > 0; o >>= 1) {
        long long u = __shfl_xor_sync(0xffffffffu, v, o);
        v = u > v ? u : v;
    }
    return v;
}

// All-reduce of the folded sums through the peers' mailboxes, executed by warp 0 of the last block
// (lane = column).  Stores to every peer, system fence, release of the sequence flag; acquire of
// the flags in the own mailbox; contributions added in rank order: bit-identical on every rank.
__device__ double peer_exchange(const PeerView& pv, Dev* st, int col, double r) {
    const unsigned long long seq = st->peer_seq + 1;
    const int slot = (int)(seq & 1);
    for (int p = 0; p < pv.world; p++) {
        double* dst = &pv.box[p]->sums[slot][pv.rank][col];
        asm volatile("st.relaxed.sys.global.f64 [%0], %1;" ::"l"(dst), "d"(r) : "memory");
    }
    __threadfence_system();
    __syncwarp();
    if (col < pv.world) {
        unsigned long long* f = &pv.box[col]->flag[slot][pv.rank];
        asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(f), "l"(seq) : "memory");
    }
    PeerBox* mine = pv.box[pv.rank];
    bool ok = true;
    if (col < pv.world) {
        const unsigned long long* f = &mine->flag[slot][col];
        unsigned long long t0, t1, v;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
        for (;;) {
            asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(f) : "memory");
            if (v >= seq) break;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
            if (t1 - t0 > (unsigned long long)PCNN_ICP_PEER_TIMEOUT_MS * 1000000ull) { ok = false; break; }
            __nanosleep(200);
        }
    }
    ok = __all_sync(0xffffffffu, ok);
    // lane j acquired rank j's flag; every lane reads every rank's column below.  The warp barrier
    // orders the acquiring lanes before the others, the system fence extends that to the peers' stores.
    __syncwarp();
    __threadfence_system();
    double tot = 0.0;
    long long m = 0;
    if (ok) {
        for (int j = 0; j < pv.world; j++) {
            double x;
            asm volatile("ld.relaxed.sys.global.f64 %0, [%1];" : "=d"(x) : "l"(&mine->sums[slot][j][col]) : "memory");
            if (col == 31) { const long long xi = __double_as_longlong(x); m = (j == 0 || xi > m) ? xi : m; }
            else tot += x;
        }
        if (col == 31) tot = __longlong_as_double(m);
    } else {
        tot = r;
        if (col == 0) { st->status = 2; st->done = 1; }
    }
    if (col == 0) st->peer_seq = seq;
    return tot;
}

// MODE 0 = point-to-point (16 sums), 1 = point-to-plane (29 sums).
// KIND 0 = split phase (sums only), 1 = fused update, 2 = fused peer all-reduce + update.
// Blocks walk the source cloud with a grid stride (whole warps: tail lanes shadow the last
// point and contribute nothing), each thread keeps its sums in registers across its points.
// Register budget, measured on B200 (1M source x 1M target, us per iteration, p2plane / p2p,
// identical pose bits): search and sums in one loop with unbounded registers (112-135) 782 / 663;
// the same at 8 blocks/SM (64 registers) 536 / 523; search first, sums in a second pass over the
// block's points, 8 blocks/SM 510 / 502.  The walk is latency-bound: resident warps matter more
// than the 16-190 B of spills, and in two passes the 16 / 29 double sums are not live during it.
#ifndef PCNN_ICP_MINBLOCKS
#define PCNN_ICP_MINBLOCKS 8
#endif
constexpr int ICP_J = 4;  // slices of 128 source points a block takes together
constexpr int ICP_COOP_MAX = 4;  // a warp left with this many points to search takes them one at a time, all lanes on one (nn1_coop)

// Correspondence certificate.  A search at position c found the nearest target point x1 and a
// floor key2 under the keys of all others (Best2f).  With u = 2^-24 the float32 key of nanoflann is
// K(p, x) = D(p, x)^2 (1 + t), |t| <= 5u, D the true distance.  So D(c, x) >= sqrt(key2)(1 - 3u)
// for x != x1, and after the source point has moved to p: D(p, x) >= sqrt(key2)(1 - 3u) - |p - c|.
// If that exceeds sqrt(K(p, x1))(1 + 3u), every other point has a strictly larger KEY than x1 at p:
// x1 is still what the search would return, ties included, and the search is skipped.  The
// margins below (1e-5 relative, directed rounding) cover the u terms a hundred times over.
// The floor is chased only as far as the point is expected to travel (Best2f::gap): ICP_GAPK times
// what this iteration's T_delta moved it, and not at all while that exceeds a quarter of the
// cloud's mean 8-point-group extent -- the spacing of the target bounds d2 - d1, so a certificate
// against a larger motion would cost a wide search and fail anyway.
constexpr float ICP_GAPK = 4.0f, ICP_MOVE_MAX = 0.03f;
__device__ __forceinline__ float cert_radius(float floor_key) { return __fmul_rd(__fsqrt_rd(floor_key), 0.99999f); }
__device__ __forceinline__ bool cert_holds(const float4& p, const float4& c, float k1) {
    if (!(c.w > 0.f)) return false;
    const float dx = fabsf(p.x - c.x), dy = fabsf(p.y - c.y), dz = fabsf(p.z - c.z);
    const float m2 = __fadd_ru(__fadd_ru(__fmul_ru(dx, dx), __fmul_ru(dy, dy)), __fmul_ru(dz, dz));
    const float s = __fmul_ru(__fadd_ru(__fsqrt_ru(k1), __fsqrt_ru(m2)), 1.00001f);
    return s < c.w;
}

// Kernel 1 of an iteration: pending T_delta, certificates, searches.  64 registers, 8 blocks per SM.
__global__ void __launch_bounds__(ICP_THREADS, PCNN_ICP_MINBLOCKS)
icp_search_kernel(IndexView iv, float4* cur, int* prev, int* prev2, float4* cert, int n_src, int use_cert, float gapk,
                  float move_frac, float max_corres_dist, Dev* __restrict__ st) {
    __shared__ int s_qg[LS_QDEPTH][ICP_THREADS];
    __shared__ float s_ql[LS_QDEPTH][ICP_THREADS];
    __shared__ unsigned short s_list[ICP_J * ICP_THREADS];  // (slice of the group, thread) of a point to search
    __shared__ float s_move[ICP_J * ICP_THREADS];           // how far T_delta moved the point
    __shared__ unsigned s_co_stack[ICP_THREADS / 32][CO_STACK];  // nn1_coop: one query, the whole warp
    __shared__ unsigned s_co_gl[ICP_THREADS / 32][CO_GLIST];
    pdl_prologue();  // launched with programmatic stream serialization when the grid leaves room (pcnn_common.cuh)
    if (st->done) return;  // uniform: loop already ended (converged / failed)
    int walked_here = 0;
    const float move_max = move_frac * __ldg(iv.cube + 8);
    const float gap_min = 1e-3f * __ldg(iv.cube + 8) + 1e-6f;  // rides out the last-bit jitter of a converged pose
    const bool first_done = st->iters > 0 || st->unchanged > 0 || st->last_pairs > 0;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float Rd[9], td[3];
#pragma unroll
    for (int i = 0; i < 9; i++) Rd[i] = st->Rd[i];
#pragma unroll
    for (int i = 0; i < 3; i++) td[i] = st->td[i];
    const unsigned full = 0xffffffffu;
    constexpr int J = ICP_J;
    const int nsl = (n_src + ICP_THREADS - 1) / ICP_THREADS;  // slices of 128 consecutive (Morton-ordered) points
    const int G = (int)gridDim.x;
    // A block owns the slices blockIdx.x, + G, + 2G, ... (spread over the cloud: dense
    // regions walk longer, contiguous chunks would unbalance the blocks) and takes them J at a time:
    // apply the pending T_delta, test every point's certificate -- the loads of the J slices issued
    // together, a converged iteration is nothing but these dependent loads -- and collect the
    // points that need a search; then the block's warps search the collected points 32 at a time,
    // so a nearly converged iteration runs a few dense warps instead of every warp for its one or
    // two uncertified lanes.  The sums are not live here: the walk gets the registers.
    for (int s0 = blockIdx.x; s0 < nsl; s0 += J * G) {
        int nw = 0;  // points of THIS WARP's lanes (J slices x 32) that need a search
        unsigned short* const wlist = s_list + warp * (J * 32);
        int tj[J], ppj[J], pp2j[J];
        float4 pj[J], cj[J];
#pragma unroll
        for (int j = 0; j < J; j++) {
            const int t = (s0 + j * G) * ICP_THREADS + threadIdx.x;
            tj[j] = (s0 + j * G < nsl && t < n_src) ? t : -1;
            if (tj[j] >= 0) pj[j] = cur[tj[j]];
        }
#pragma unroll
        for (int j = 0; j < J; j++) {
            ppj[j] = -1;
            pp2j[j] = -1;
            if (tj[j] >= 0) {
                float4 p = pj[j];
                const float ox = p.x, oy = p.y, oz = p.z;
                xform(Rd, td, p.x, p.y, p.z);  // pending T_delta of the previous iteration
                cur[tj[j]] = p;
                pj[j] = p;
                s_move[j * ICP_THREADS + threadIdx.x] =
                    fmaxf(fmaxf(fabsf(p.x - ox), fabsf(p.y - oy)), fabsf(p.z - oz)) * 1.7320509f;
                if (use_cert) {
                    ppj[j] = prev[tj[j]];
                    pp2j[j] = prev2[tj[j]];
                    cj[j] = cert[tj[j]];
                }
            }
        }
        float4 qj[J];
#pragma unroll
        for (int j = 0; j < J; j++)
            if (ppj[j] >= 0) qj[j] = __ldg(iv.pts + ppj[j]);
#pragma unroll
        for (int j = 0; j < J; j++) {
            bool need = tj[j] >= 0;
            if (need && use_cert) {
                // no correspondence (pp < 0): the certificate says nothing lies within c.w of c; with
                // the gate as the key it reads "still nothing within the gate"
                float k1 = ppj[j] >= 0 ? key32(qj[j].x, qj[j].y, qj[j].z, pj[j].x, pj[j].y, pj[j].z) : max_corres_dist;
                if (pp2j[j] >= 0) {
                    // a runner-up inside the certified band (rare: near-ties): whichever of the two is
                    // nearer NOW, by (key, original index), is the correspondence
                    const float4 q2 = __ldg(iv.pts + pp2j[j]);
                    const float k2 = key32(q2.x, q2.y, q2.z, pj[j].x, pj[j].y, pj[j].z);
                    if (kless32(k2, __float_as_int(q2.w), k1, __float_as_int(qj[j].w))) {
                        prev[tj[j]] = pp2j[j];
                        prev2[tj[j]] = ppj[j];
                        k1 = k2;
                    }
                }
                need = !cert_holds(pj[j], cj[j], k1);
            }
            const unsigned m = __ballot_sync(full, need);
            if (need) wlist[nw + __popc(m & ((1u << lane) - 1u))] = (unsigned short)(j * ICP_THREADS + threadIdx.x);
            nw += __popc(m);
        }
        // The list is per WARP: nothing below waits for another warp of the block (the block-wide list of
        // round 1 cost two block barriers per 512 points, every warp waiting for the block's longest walk;
        // its purpose, packing sparse needs into dense warps, is served by nn1_coop now).
        __syncwarp();
        walked_here += nw;
        for (int b = 0; b < nw; b += 32) {  // tail lanes shadow the last entry
            if (nw - b <= ICP_COOP_MAX) {
                // a handful of points (the usual case once the pose has settled: one or two uncertified
                // points per block): one at a time with the whole warp on it, instead of one lane walking
                // for 30-60 us with 31 shadows.  Same correspondence, same kind of certificate.
                for (int r = b; r < nw; r++) {
                    const int e = wlist[r];
                    const int t = (s0 + (e / ICP_THREADS) * G) * ICP_THREADS + (e % ICP_THREADS);
                    const float4 p = cur[t];
                    const int pp = prev[t];
                    const float moved = s_move[e];
                    const float gap = (use_cert && first_done && moved < move_max) ? gapk * moved + gap_min : -1.f;
                    const int seed = pp >= 0 ? (pp >> PCNN_GROUP_LOG) : -1;
                    if (gap < 0.f) {
                        Best1f best;
                        nn1_coop(iv, p.x, p.y, p.z, seed, max_corres_dist, best, s_co_stack[warp], s_co_gl[warp]);
                        if (lane == 0) {
                            prev[t] = best.pos;
                            prev2[t] = -1;
                            cert[t] = make_float4(p.x, p.y, p.z, -1.f);
                        }
                    } else {
                        Best2f best;
                        best.gap = gap;
                        const float rr = __fadd_ru(__fmul_ru(__fsqrt_ru(max_corres_dist), 1.0001f), gap);  // as Best2f::reach
                        nn1_coop(iv, p.x, p.y, p.z, seed, __fmul_ru(rr, rr), best, s_co_stack[warp], s_co_gl[warp]);
                        if (lane == 0) {
                            prev[t] = best.pos;
                            prev2[t] = (best.pos2 >= 0 && best.key2 < best.floor_key()) ? best.pos2 : -1;
                            cert[t] = make_float4(p.x, p.y, p.z, cert_radius(best.floor_key()));
                        }
                    }
                    __syncwarp();
                }
                continue;
            }
            const int i = b + lane;
            const bool live = i < nw;
            const int e = wlist[live ? i : nw - 1];
            const int t = (s0 + (e / ICP_THREADS) * G) * ICP_THREADS + (e % ICP_THREADS);
            const float4 p = cur[t];
            const int pp = prev[t];
            const float moved = s_move[e];
            // a point searched for the first time (iteration 0 of a state) has not moved yet: no certificate
            const float gap = (use_cert && first_done && moved < move_max) ? gapk * moved + gap_min : -1.f;
            const int seed = pp >= 0 ? (pp >> PCNN_GROUP_LOG) : -1;
            if (__all_sync(full, gap < 0.f)) {
                // nobody in this warp is after a certificate (the pose still moves by more than the
                // target's spacing): the plain 1-NN walk, lighter by the runner-up and the floor.
                // The gate doubles as the search bound: a neighbour at or beyond it is dropped anyway
                Best1f best;
                nn1_lockstep<ICP_THREADS>(iv, p.x, p.y, p.z, seed, max_corres_dist, best, s_qg, s_ql, threadIdx.x);
                if (live) {
                    prev[t] = best.pos;
                    prev2[t] = -1;
                    cert[t] = make_float4(p.x, p.y, p.z, -1.f);
                }
            } else {
                Best2f best;
                best.gap = gap;
                // with a certificate in view the bound reaches `gap` beyond the gate, so that "nothing
                // within the gate" can be certified too (those searches are the widest of all: a 1 m ball)
                float bound = max_corres_dist;
                if (gap >= 0.f) {
                    const float r = __fadd_ru(__fmul_ru(__fsqrt_ru(max_corres_dist), 1.0001f), gap);  // as Best2f::reach
                    bound = __fmul_ru(r, r);
                }
                nn1_lockstep<ICP_THREADS>(iv, p.x, p.y, p.z, seed, bound, best, s_qg, s_ql, threadIdx.x);
                if (live) {
                    prev[t] = best.pos;
                    // the runner-up matters only if it lies under the floor of the rest
                    prev2[t] = (gap >= 0.f && best.pos2 >= 0 && best.key2 < best.floor_key()) ? best.pos2 : -1;
                    cert[t] = make_float4(p.x, p.y, p.z, gap >= 0.f ? cert_radius(best.floor_key()) : -1.f);
                }
            }
        }
        __syncwarp();  // the warp's list is reused
    }
    if (lane == 0 && walked_here) atomicAdd(&st->walked, (unsigned long long)walked_here);
}

// Kernel 2 of an iteration: the pairs' terms, their reduction, and (KIND) the solve / the exchange.
// Its own register budget: 16 / 29 double sums per thread next to J slices of loads in flight.
template <int MODE, int KIND>
__global__ void __launch_bounds__(ICP_THREADS, 4)
icp_accumulate_kernel(IndexView iv, const float4* cur, const float4* __restrict__ nrm, const int* prev, int n_src,
                      float max_corres_dist, float loss_epsilon, Dev* __restrict__ st, double* __restrict__ partials,
                      const PeerView pv) {
    constexpr bool FUSED = KIND != 0;
    constexpr int NT = MODE == 0 ? 16 : 29;
    constexpr int J = ICP_J;
    __shared__ double s_part[ICP_THREADS / 32][ICP_NS];
    __shared__ long long s_last[ICP_THREADS / 32];
    __shared__ bool s_is_last;
    pdl_prologue();  // launched with programmatic stream serialization when the grid leaves room (pcnn_common.cuh)
    if (st->done) return;  // uniform: loop already ended (converged / failed)
#ifdef PCNN_ICP_TIMING
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        unsigned long long t0;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
        printf("icp timing: block 0 starts at %llu\n", t0);
    }
#endif
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#ifdef PCNN_ICP_TIMING
    unsigned long long tm_a, tm_b;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tm_a));
#endif
    double v[NT];
#pragma unroll
    for (int i = 0; i < NT; i++) v[i] = 0.0;
    long long last = -1;
    const int nsl = (n_src + ICP_THREADS - 1) / ICP_THREADS;
    const int G = (int)gridDim.x;
    // points assigned to threads in a fixed pattern (the sums do not depend on which lane searched
    // what): the slices blockIdx.x, + G, ... of the search kernel, J at a time so their loads overlap
    for (int s0 = blockIdx.x; s0 < nsl; s0 += J * G) {
        float4 pj[J], qj[J], nj[J];
        int posj[J];
#pragma unroll
        for (int j = 0; j < J; j++) {
            const int t = (s0 + j * G) * ICP_THREADS + threadIdx.x;
            posj[j] = -1;
            if (s0 + j * G < nsl && t < n_src) {
                pj[j] = cur[t];
                posj[j] = prev[t];
            }
        }
#pragma unroll
        for (int j = 0; j < J; j++)
            if (posj[j] >= 0) {
                qj[j] = __ldg(iv.pts + posj[j]);
                if (MODE == 1) nj[j] = __ldg(nrm + posj[j]);
            }
#pragma unroll
        for (int j = 0; j < J; j++) {
        const bool live = posj[j] >= 0;
        const float4 p = pj[j], qpt = qj[j];
        Best1f best;
        best.init();
        best.pos = posj[j];
        if (live) best.key = key32(qpt.x, qpt.y, qpt.z, p.x, p.y, p.z);  // the key the walk found, same arithmetic
        if (live && best.pos >= 0 && best.key < max_corres_dist) {  // gate on the SQUARED distance (:783, :936)
            if (MODE == 0) {
                v[0] += p.x; v[1] += p.y; v[2] += p.z;
                v[3] += qpt.x; v[4] += qpt.y; v[5] += qpt.z;
                v[6] += (double)qpt.x * p.x; v[7] += (double)qpt.x * p.y; v[8] += (double)qpt.x * p.z;
                v[9] += (double)qpt.y * p.x; v[10] += (double)qpt.y * p.y; v[11] += (double)qpt.y * p.z;
                v[12] += (double)qpt.z * p.x; v[13] += (double)qpt.z * p.y; v[14] += (double)qpt.z * p.z;
                v[15] += 1.0;
                const long long cand =
                    ((long long)__float_as_int(p.w) << 32) | (long long)(unsigned)__float_as_int(best.key);
                last = cand > last ? cand : last;
            } else {
                const float4 nn = nj[j];
                // row of A and b in float32 exactly as written at registration.cpp:808-814
                float a[6];
                a[0] = fadd(fmul(nn.z, p.y), -fmul(nn.y, p.z));
                a[1] = fadd(fmul(nn.x, p.z), -fmul(nn.z, p.x));
                a[2] = fadd(fmul(nn.y, p.x), -fmul(nn.x, p.y));
                a[3] = nn.x; a[4] = nn.y; a[5] = nn.z;
                float b = fadd(fadd(fmul(nn.x, qpt.x), fmul(nn.y, qpt.y)), fmul(nn.z, qpt.z));
                b = fadd(b, -fmul(nn.x, p.x));
                b = fadd(b, -fmul(nn.y, p.y));
                b = fadd(b, -fmul(nn.z, p.z));
                int u = 0;
#pragma unroll
                for (int r = 0; r < 6; r++)
#pragma unroll
                    for (int c = r; c < 6; c++) v[u++] += (double)a[r] * (double)a[c];
#pragma unroll
                for (int r = 0; r < 6; r++) v[21 + r] += (double)a[r] * (double)b;
                v[27] += (double)b * (double)b;
                v[28] += 1.0;
            }
        }
        }
    }
#pragma unroll
    for (int i = 0; i < NT; i++) {
        const double s = warp_sum(v[i]);
        if (lane == 0) s_part[warp][i] = s;
    }
    if (MODE == 0) {
        const long long m = warp_max_ll(last);
        if (lane == 0) s_last[warp] = m;
    }
    __syncthreads();
    if (threadIdx.x < ICP_NS) {
        const int slot = threadIdx.x;
        int src = -1;  // which per-thread term feeds this slot (the count lives at [28] in both modes)
        if (MODE == 0) src = slot < 15 ? slot : (slot == 28 ? 15 : -1);
        else src = slot < 29 ? slot : -1;
        double s = 0.0;
        if (src >= 0) {
#pragma unroll
            for (int w2 = 0; w2 < ICP_THREADS / 32; w2++) s += s_part[w2][src];
        }
        if (MODE == 0 && slot == 31) {
            long long m = s_last[0];
            for (int w2 = 1; w2 < ICP_THREADS / 32; w2++) m = s_last[w2] > m ? s_last[w2] : m;
            s = __longlong_as_double(m);
        }
        partials[(size_t)blockIdx.x * ICP_NS + slot] = s;
    }
    // Two-level fold in a fixed order (deterministic whatever the timing; no floating atomics): the
    // last block of every group of ICP_FOLD_G blocks to arrive folds that group's partial rows into
    // one group row, and the last GROUP to finish folds the group rows.  One block folding all rows
    // (up to 296 x 32 doubles, five rounds of dependent L2 loads) took 16 us of every iteration.
#ifdef PCNN_ICP_TIMING
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tm_b));
#endif
    const int nb = (int)gridDim.x;
    const int ngroups = (nb + ICP_FOLD_G - 1) / ICP_FOLD_G;
    const int grp = blockIdx.x / ICP_FOLD_G;
    const int gfirst = grp * ICP_FOLD_G, gsize = min(ICP_FOLD_G, nb - gfirst);
    double* const grows = partials + (size_t)nb * ICP_NS;  // [ngroups][32] behind the per-block rows
    const int col = threadIdx.x & 31, part = threadIdx.x >> 5;
    constexpr int P4 = ICP_THREADS / 32;
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned ticket = atomicAdd(&st->group_counter[grp], 1u);
        s_is_last = (ticket == (unsigned)gsize - 1u);
    }
    __syncthreads();
    if (!s_is_last) return;
    __threadfence();
    {
        // level 1: <= 16 rows, four per (column, part), loads independent
        double x[ICP_FOLD_G / P4];
#pragma unroll
        for (int u = 0; u < ICP_FOLD_G / P4; u++) {
            const int r = part + u * P4;
            x[u] = r < gsize ? __ldcg(partials + (size_t)(gfirst + r) * ICP_NS + col) : (col == 31 ? __longlong_as_double(-1LL) : 0.0);
        }
        double sacc;
        if (col == 31) {
            long long m = -1;
#pragma unroll
            for (int u = 0; u < ICP_FOLD_G / P4; u++) { const long long xi = __double_as_longlong(x[u]); m = xi > m ? xi : m; }
            sacc = __longlong_as_double(m);
        } else {
            sacc = (x[0] + x[1]) + (x[2] + x[3]);
        }
        s_part[part][col] = sacc;
    }
    __syncthreads();
    if (threadIdx.x < ICP_NS) {
        double r;
        if (col == 31) {
            long long m = __double_as_longlong(s_part[0][col]);
            for (int p2 = 1; p2 < P4; p2++) { const long long xi = __double_as_longlong(s_part[p2][col]); m = xi > m ? xi : m; }
            r = __longlong_as_double(m);
        } else {
            r = ((s_part[0][col] + s_part[1][col]) + s_part[2][col]) + s_part[3][col];
        }
        grows[(size_t)grp * ICP_NS + col] = r;
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        st->group_counter[grp] = 0;  // for the next launch on this state
        const unsigned ticket = atomicAdd(&st->block_counter, 1u);
        s_is_last = (ticket == (unsigned)ngroups - 1u);
    }
    __syncthreads();
    if (!s_is_last) return;
    __threadfence();
    {
        // level 2: the group rows (<= ICP_MAX_GROUPS), eight interleaved accumulators per (column, part)
        constexpr int UN = 8;
        double acc[UN];
        long long mm[UN];
#pragma unroll
        for (int u = 0; u < UN; u++) { acc[u] = 0.0; mm[u] = -1; }
        for (int b0 = part; b0 < ngroups; b0 += P4 * UN) {
            double x[UN];
#pragma unroll
            for (int u = 0; u < UN; u++) {
                const int b = b0 + u * P4;
                x[u] = b < ngroups ? __ldcg(grows + (size_t)b * ICP_NS + col) : (col == 31 ? __longlong_as_double(-1LL) : 0.0);
            }
#pragma unroll
            for (int u = 0; u < UN; u++) {
                if (col == 31) { const long long xi = __double_as_longlong(x[u]); mm[u] = xi > mm[u] ? xi : mm[u]; }
                else acc[u] += x[u];
            }
        }
#pragma unroll
        for (int w = UN / 2; w > 0; w >>= 1)  // fixed tree
#pragma unroll
            for (int u = 0; u < w; u++) acc[u] += acc[u + w];
        long long m = -1;
#pragma unroll
        for (int u = 0; u < UN; u++) m = mm[u] > m ? mm[u] : m;
        __syncthreads();  // level 1's readers of s_part are done
        s_part[part][col] = (col == 31) ? __longlong_as_double(m) : acc[0];
    }
    __syncthreads();
    if (threadIdx.x < ICP_NS) {
        const int col = threadIdx.x;
        double r;
        if (col == 31) {
            long long m = __double_as_longlong(s_part[0][col]);
            for (int p2 = 1; p2 < ICP_THREADS / 32; p2++) {
                const long long xi = __double_as_longlong(s_part[p2][col]);
                m = xi > m ? xi : m;
            }
            r = __longlong_as_double(m);
        } else {
            r = ((s_part[0][col] + s_part[1][col]) + s_part[2][col]) + s_part[3][col];
        }
        if (KIND == 2) r = peer_exchange(pv, st, col, r);
        st->sums[col] = r;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        st->block_counter = 0;
#ifdef PCNN_ICP_TIMING
        unsigned long long tm_c, tm_d;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tm_c));
#endif
        if (FUSED) icp_update_dev(st, MODE, loss_epsilon);
#ifdef PCNN_ICP_TIMING
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tm_d));
        printf("icp timing: last block %d: its search+reduce %llu ns, fold %llu ns, solve %llu ns\n", blockIdx.x,
               tm_b - tm_a, tm_c - tm_b, tm_d - tm_c);
#endif
    }
}

}  // namespace pcnn

// ------------------------------------------------------------------ host side
namespace pcnn {

__global__ void icp_init_kernel(const float* __restrict__ src, int n, int stride, long long index_base,
                                const float* __restrict__ T0, float4* __restrict__ out) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float* p = src + (size_t)i * stride;
    float x = p[0], y = p[1], z = p[2];
    const float R[9] = {T0[0], T0[1], T0[2], T0[4], T0[5], T0[6], T0[8], T0[9], T0[10]};
    const float t[3] = {T0[3], T0[7], T0[11]};
    xform(R, t, x, y, z);  // registration.cpp:872-875
    out[i] = make_float4(x, y, z, __int_as_float((int)(index_base + i)));
}

__global__ void gather4_kernel(const float4* __restrict__ in, const unsigned* __restrict__ perm, int n,
                               float4* __restrict__ out) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = in[perm[i]];
}

__global__ void permute_normals_kernel(const float4* __restrict__ pts, int n, const float* __restrict__ normals,
                                       float4* __restrict__ out) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int o = __float_as_int(pts[i].w);
    const float* v = normals + (size_t)o * 3;
    out[i] = make_float4(v[0], v[1], v[2], 0.f);
}

__global__ void icp_state_init_kernel(Dev* st, const float* __restrict__ T0) {
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        for (int i = 0; i < 16; i++) st->T[i] = T0[i];
        st->T[12] = 0; st->T[13] = 0; st->T[14] = 0; st->T[15] = 1;  // :910-913 keeps the last row
        for (int i = 0; i < 9; i++) st->Rd[i] = (i % 4 == 0) ? 1.f : 0.f;
        for (int i = 0; i < 3; i++) st->td[i] = 0.f;
        for (int i = 0; i < ICP_NS; i++) st->sums[i] = 0.0;
        st->iters = 0; st->converged = 0; st->status = 0; st->done = 0; st->unchanged = 0;
        st->last_loss = 0.f; st->loss = 0.f; st->last_pairs = 0; st->block_counter = 0;
        for (int i = 0; i < ICP_MAX_GROUPS; i++) st->group_counter[i] = 0;
        st->peer_seq = 0;
        st->walked = 0; st->walked_mark = 0; st->last_searches = 0;
    }
}

#ifndef PCNN_NN1_STRATA
#define PCNN_NN1_STRATA 24
#endif
constexpr int NN1_STRATA = PCNN_NN1_STRATA;

__global__ void __launch_bounds__(128)
nn1_kernel(IndexView iv, const float* __restrict__ q, int nq, int stride, const unsigned* __restrict__ perm,
           int32_t* __restrict__ out_idx, float* __restrict__ out_d2) {
    __shared__ int s_qg[LS_QDEPTH][128];
    __shared__ float s_ql[LS_QDEPTH][128];
    int bid = blockIdx.x;  // launch order interleaved over the ordered batch (pcnn_knn.cu, KD_STRATA)
    if (NN1_STRATA > 1 && perm != nullptr) {
        const int per = (int)gridDim.x / NN1_STRATA;
        bid = ((int)blockIdx.x % NN1_STRATA) * per + (int)blockIdx.x / NN1_STRATA;
        if ((long long)bid * 128 >= nq) return;
    }
    const int t = bid * 128 + threadIdx.x;
    const bool live = t < nq;
    const int tt = live ? t : nq - 1;  // tail lanes shadow the last query: warps stay whole
    const int qi = perm ? (int)perm[tt] : tt;
    const float* qp = q + (size_t)qi * stride;
    Best1f best;
    nn1_lockstep<128>(iv, qp[0], qp[1], qp[2], -1, __int_as_float(0x7f800000), best, s_qg, s_ql, threadIdx.x);
    if (live) {
        out_idx[qi] = best.pos >= 0 ? best.idx : -1;
        if (out_d2) out_d2[qi] = best.key;
    }
}

// Small batches: one WARP per query (nn1_coop).  A few hundred queries cannot fill the device one per
// lane, and each would walk alone at the pace of its dependent loads; with 32 lanes per query the
// same batch finishes in a fraction of the time.  Same result, bit for bit.
constexpr int NN1_COOP_MAX_QUERIES = 2048;
__global__ void __launch_bounds__(128)
nn1_coop_kernel(IndexView iv, const float* __restrict__ q, int nq, int stride, int32_t* __restrict__ out_idx,
                float* __restrict__ out_d2) {
    __shared__ unsigned s_stack[4][CO_STACK];
    __shared__ unsigned s_gl[4][CO_GLIST];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int qi = blockIdx.x * 4 + warp;
    if (qi >= nq) return;  // warp-uniform
    const float* qp = q + (size_t)qi * stride;
    Best1f best;
    nn1_coop(iv, qp[0], qp[1], qp[2], -1, __int_as_float(0x7f800000), best, s_stack[warp], s_gl[warp]);
    if (lane == 0) {
        out_idx[qi] = best.pos >= 0 ? best.idx : -1;
        if (out_d2) out_d2[qi] = best.key;
    }
}

int launch_nn1(const pcnn_index* ix, const float* q, int64_t nq64, int stride, int32_t* out_idx,
               float* out_d2, cudaStream_t s) {
    const int nq = (int)nq64;
    if (nq == 0) return PCNN_OK;
    NvtxRange nvtx_(ix, "pcnn.nn1");
    if (nq <= NN1_COOP_MAX_QUERIES) {
        nn1_coop_kernel<<<(nq + 3) / 4, 128, 0, s>>>(ix->view, q, nq, stride, out_idx, out_d2);
        PCNN_CUDA(cudaGetLastError());
        return PCNN_OK;
    }
    Scratch sc(s);
    unsigned* perm = nullptr;
    if (ix->opt_sort_queries && nq > 1024 && ix->n > 0) {
        int rc = sort_queries(ix, q, nq, stride, &perm, sc, s);
        if (rc != PCNN_OK) return rc;
    }
    int nb = (nq + 127) / 128;
    if (NN1_STRATA > 1 && perm != nullptr) nb = ((nb + NN1_STRATA - 1) / NN1_STRATA) * NN1_STRATA;
    nn1_kernel<<<nb, 128, 0, s>>>(ix->view, q, nq, stride, perm, out_idx, out_d2);
    PCNN_CUDA(cudaGetLastError());
    return PCNN_OK;
}

static void icp_peer_detach(pcnn_icp_state* st) {
    for (int p = 0; p < PCNN_ICP_MAX_PEERS; p++) {
        if (st->peer[p] && st->peer_ipc[p]) cudaIpcCloseMemHandle(st->peer[p]);
        st->peer[p] = nullptr;
        st->peer_ipc[p] = false;
    }
    st->world = 0;
}

static void icp_free(pcnn_icp_state* st) {
    if (!st) return;
    icp_peer_detach(st);
    if (st->box) cudaFree(st->box);
    void* ptrs[7] = {st->cur, st->nrm, st->prev, st->partials, st->dev, st->cert, st->prev2};
    for (void* p : ptrs)
        if (p) cudaFreeAsync(p, st->mem_stream);
    delete st;
}

// kind: 0 = sums only, 1 = fused update, 2 = fused peer all-reduce + update (a rank without
// source points still takes part in the exchange with one block and zero sums)
static int icp_launch_accumulate(pcnn_icp_state* st, int kind, cudaStream_t s) {
    if (st->n_src == 0 && kind != 2) return PCNN_OK;
    const pcnn_index* ix = st->target;
    const int n = (int)st->n_src;
    PeerView pv{};
    if (kind == 2) {
        for (int p = 0; p < st->world; p++) pv.box[p] = st->peer[p];
        pv.rank = st->rank; pv.world = st->world;
    }
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cudaLaunchConfig_t cfg{};
    cfg.blockDim = dim3(ICP_THREADS);
    cfg.dynamicSmemBytes = 0;
    cfg.stream = s;
    cfg.attrs = attr;
    cfg.numAttrs = st->use_pdl ? 1 : 0;
    {
        NvtxRange nvtx_(ix, "pcnn.icp.search");
        cfg.gridDim = dim3(st->nblocks);
        PCNN_CUDA(cudaLaunchKernelEx(&cfg, icp_search_kernel, ix->view, st->cur, st->prev, st->prev2, st->cert, n, st->use_cert,
                                     st->gapk, st->move_frac, st->prm.max_corres_dist, st->dev));
    }
    NvtxRange nvtx_(ix, kind == 2 ? "pcnn.icp.sums+exchange+solve" : (kind == 1 ? "pcnn.icp.sums+solve" : "pcnn.icp.sums"));
    cfg.gridDim = dim3(st->acc_blocks);
#define PCNN_ICP_LAUNCH(MODE, KIND)                                                                              \
    PCNN_CUDA(cudaLaunchKernelEx(&cfg, icp_accumulate_kernel<MODE, KIND>, ix->view, (const float4*)st->cur,       \
                                 (const float4*)st->nrm, (const int*)st->prev, n, st->prm.max_corres_dist,        \
                                 st->prm.loss_epsilon, st->dev, st->partials, (const PeerView)pv))
    if (st->prm.mode == 0) {
        if (kind == 2) PCNN_ICP_LAUNCH(0, 2); else if (kind == 1) PCNN_ICP_LAUNCH(0, 1); else PCNN_ICP_LAUNCH(0, 0);
    } else {
        if (kind == 2) PCNN_ICP_LAUNCH(1, 2); else if (kind == 1) PCNN_ICP_LAUNCH(1, 1); else PCNN_ICP_LAUNCH(1, 0);
    }
#undef PCNN_ICP_LAUNCH
    PCNN_CUDA(cudaGetLastError());
    return PCNN_OK;
}

}  // namespace pcnn

using namespace pcnn;

extern "C" {

int pcnn_icp_create(pcnn_icp_state** out, const pcnn_index* target, const float* tar_normals_dev,
                    const float* src_dev, int64_t n_src, int stride, const float* init_T_host,
                    const pcnn_icp_params* params, int64_t src_index_base, pcnn_stream stream) {
    PCNN_CHECK_ARG(out && target && params && init_T_host, "pcnn_icp_create: NULL argument");
    *out = nullptr;
    PCNN_CHECK_ARG(n_src >= 0 && n_src < 0x7fffffff, "pcnn_icp_create: n_src out of range");
    PCNN_CHECK_ARG(stride >= 3, "pcnn_icp_create: stride must be >= 3");
    PCNN_CHECK_ARG(params->mode == 0 || params->mode == 1, "pcnn_icp_create: mode must be 0 or 1");
    PCNN_CHECK_ARG(params->max_iter >= 0, "pcnn_icp_create: max_iter must be >= 0");
    PCNN_CHECK_ARG(params->mode == 0 || tar_normals_dev != nullptr || target->n == 0,
                   "pcnn_icp_create: point-to-plane needs target normals");
    PCNN_CHECK_ARG(n_src == 0 || src_dev != nullptr, "pcnn_icp_create: src is NULL");
    PCNN_CHECK_ARG(src_index_base >= 0 && src_index_base + n_src < 0x7fffffff, "pcnn_icp_create: index base out of range");
    DeviceGuard g(target->device);
    if (!g.ok) { set_error("pcnn_icp_create: cannot select device %d", target->device); return PCNN_ERR_CUDA; }
    cudaStream_t s = (cudaStream_t)stream;
    pcnn_icp_state* st = new (std::nothrow) pcnn_icp_state;
    if (!st) { set_error("pcnn_icp_create: host allocation failed"); return PCNN_ERR_NOMEM; }
    st->target = target; st->device = target->device; st->n_src = n_src; st->prm = *params;
    const int n = (int)n_src;
    {
        // one block per 128-point slice up to what is resident at once (blocks per SM from the
        // occupancy calculator); beyond that the grid strides over the slices
        int per_sm = 0, sms = 148;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, (const void*)icp_search_kernel, ICP_THREADS, 0) != cudaSuccess || per_sm < 1) per_sm = 4;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, st->device);
        if (per_sm > 8) per_sm = 8;
        st->use_cert = target->opt_icp_cert;  // pcnn_set_option(target, "icp_certificates", 0|1): same result either way
        st->use_pdl = target->opt_icp_pdl;
        st->gapk = ICP_GAPK; st->move_frac = ICP_MOVE_MAX;
#ifdef PCNN_STUDY
        if (const char* e = getenv("PCNN_ICP_GAPK")) st->gapk = (float)atof(e);
        if (const char* e = getenv("PCNN_ICP_MOVE")) st->move_frac = (float)atof(e);
#endif
        st->nblocks = (n + ICP_THREADS - 1) / ICP_THREADS;
        if (st->nblocks < 1) st->nblocks = 1;
        if (st->nblocks > sms * per_sm) st->nblocks = sms * per_sm;
        // PDL pays where the grid leaves room for the successor's blocks (100k points: 34.5 -> 30.9 us per
        // iteration); on a full device the waiting blocks are in the way (1M points: 178.5 -> 184.3 us)
        if (st->nblocks >= sms * per_sm) st->use_pdl = 0;
        // the sums kernel takes ICP_J slices per block and step: one block per J slices until the
        // device is covered, so a 100k-point cloud folds ~200 partial rows instead of ~800
        const int slices = (n + ICP_THREADS - 1) / ICP_THREADS;
        st->acc_blocks = (slices + ICP_J - 1) / ICP_J;
        if (st->acc_blocks < (slices < sms ? slices : sms)) st->acc_blocks = slices < sms ? slices : sms;
        if (st->acc_blocks > 2 * sms) st->acc_blocks = 2 * sms;  // measured at 1M points: 148 / 296 / 592 / 1184 blocks -> 221 / 209 / 208 / 222 us per iteration
        if (st->acc_blocks > st->nblocks) st->acc_blocks = st->nblocks;
#ifdef PCNN_STUDY
        if (const char* e = getenv("PCNN_ICP_ACC_BLOCKS")) st->acc_blocks = atoi(e) < st->nblocks ? atoi(e) : st->nblocks;
#endif
        if (st->acc_blocks > ICP_FOLD_G * ICP_MAX_GROUPS) st->acc_blocks = ICP_FOLD_G * ICP_MAX_GROUPS;
        if (st->acc_blocks < 1) st->acc_blocks = 1;
    }
    st->mem_stream = s;
    cudaMemPool_t pool = lib_pool();
    auto palloc = [&](auto** out, size_t bytes) -> cudaError_t {
        void* p = nullptr;
        cudaError_t err = pool ? cudaMallocFromPoolAsync(&p, bytes, pool, s) : cudaMallocAsync(&p, bytes, s);
        *out = reinterpret_cast<std::remove_reference_t<decltype(*out)>>(p);
        return err;
    };
    cudaError_t e = palloc(&st->cur, sizeof(float4) * (size_t)(n + 1));
    if (e == cudaSuccess) e = palloc(&st->prev, sizeof(int) * (size_t)(n + 1));
    if (e == cudaSuccess) e = palloc(&st->cert, sizeof(float4) * (size_t)(n + 1));
    if (e == cudaSuccess) e = palloc(&st->prev2, sizeof(int) * (size_t)(n + 1));
    if (e == cudaSuccess) e = palloc(&st->partials, sizeof(double) * ICP_NS * ((size_t)st->nblocks + ICP_MAX_GROUPS));
    if (e == cudaSuccess) e = palloc(&st->dev, sizeof(Dev));
    if (e == cudaSuccess && params->mode == 1) e = palloc(&st->nrm, sizeof(float4) * (size_t)(target->n + 1));
    if (e != cudaSuccess) {
        set_error("pcnn_icp_create: device allocation failed: %s", cudaGetErrorString(e));
        icp_free(st);
        return PCNN_ERR_NOMEM;
    }
    int rc = PCNN_OK;
    {
        Scratch sc(s);
        float* T0 = nullptr;
        float4* tmp = nullptr;
        e = sc.get(&T0, 16);
        if (e == cudaSuccess) e = sc.get(&tmp, (size_t)n + 1);
        if (e == cudaSuccess) e = cudaMemcpyAsync(T0, init_T_host, sizeof(float) * 16, cudaMemcpyHostToDevice, s);
        if (e != cudaSuccess) {
            set_error("pcnn_icp_create: staging failed: %s", cudaGetErrorString(e));
            rc = PCNN_ERR_CUDA;
        } else {
            icp_state_init_kernel<<<1, 32, 0, s>>>(st->dev, T0);
            e = cudaMemsetAsync(st->prev, 0xff, sizeof(int) * (size_t)(n + 1), s);  // -1: no correspondence yet
            if (e == cudaSuccess) e = cudaMemsetAsync(st->prev2, 0xff, sizeof(int) * (size_t)(n + 1), s);
            if (e == cudaSuccess) e = cudaMemsetAsync(st->cert, 0xff, sizeof(float4) * (size_t)(n + 1), s);  // NaN: no certificate
            if (n > 0) {
                icp_init_kernel<<<(n + 255) / 256, 256, 0, s>>>(src_dev, n, stride, (long long)src_index_base, T0, tmp);
                unsigned* perm = nullptr;
                if (target->n > 0 && n > 1024) rc = sort_queries(target, (const float*)tmp, n, 4, &perm, sc, s);
                if (rc == PCNN_OK) {
                    if (perm) gather4_kernel<<<(n + 255) / 256, 256, 0, s>>>(tmp, perm, n, st->cur);
                    else e = cudaMemcpyAsync(st->cur, tmp, sizeof(float4) * (size_t)n, cudaMemcpyDeviceToDevice, s);
                }
            }
            if (rc == PCNN_OK && params->mode == 1 && target->n > 0)
                permute_normals_kernel<<<((int)target->n + 255) / 256, 256, 0, s>>>(target->pts, (int)target->n,
                                                                                     tar_normals_dev, st->nrm);
            // init_T_host may be reused by the caller as soon as we return: wait for its copy
            if (rc == PCNN_OK && (e != cudaSuccess || cudaGetLastError() != cudaSuccess)) {
                set_error("pcnn_icp_create: launch failed");
                rc = PCNN_ERR_CUDA;
            }
            cudaStreamSynchronize(s);
        }
    }
    if (rc != PCNN_OK) { icp_free(st); return rc; }
    *out = st;
    return PCNN_OK;
}

int pcnn_icp_accumulate(pcnn_icp_state* st, pcnn_stream stream) {
    PCNN_CHECK_ARG(st, "pcnn_icp_accumulate: NULL state");
    DeviceGuard g(st->device);
    return icp_launch_accumulate(st, 0, (cudaStream_t)stream);
}

double* pcnn_icp_sums_dev(pcnn_icp_state* st) { return st ? st->dev->sums : nullptr; }

int pcnn_icp_peer_export(pcnn_icp_state* st, void* handle_out, void** mailbox_dev_out) {
    PCNN_CHECK_ARG(st, "pcnn_icp_peer_export: NULL state");
    DeviceGuard g(st->device);
    if (!st->box) {
        // cudaMalloc, not the stream-ordered pool: pool memory cannot be exported over CUDA IPC
        PeerBox* b = nullptr;
        PCNN_CUDA(cudaMalloc(&b, sizeof(PeerBox)));
        cudaError_t e = cudaMemset(b, 0, sizeof(PeerBox));
        if (e == cudaSuccess) e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { cudaFree(b); set_error("pcnn_icp_peer_export: %s", cudaGetErrorString(e)); return PCNN_ERR_CUDA; }
        st->box = b;
    }
    if (handle_out) {
        static_assert(sizeof(cudaIpcMemHandle_t) == PCNN_ICP_PEER_HANDLE_BYTES, "IPC handle size");
        PCNN_CUDA(cudaIpcGetMemHandle(reinterpret_cast<cudaIpcMemHandle_t*>(handle_out), st->box));
    }
    if (mailbox_dev_out) *mailbox_dev_out = st->box;
    return PCNN_OK;
}

int pcnn_icp_peer_attach(pcnn_icp_state* st, int rank, int world, const void* handles, void* const* ptrs) {
    PCNN_CHECK_ARG(st, "pcnn_icp_peer_attach: NULL state");
    PCNN_CHECK_ARG(world >= 1 && world <= PCNN_ICP_MAX_PEERS && rank >= 0 && rank < world,
                   "pcnn_icp_peer_attach: need 0 <= rank < world <= PCNN_ICP_MAX_PEERS");
    PCNN_CHECK_ARG(st->box, "pcnn_icp_peer_attach: call pcnn_icp_peer_export first");
    PCNN_CHECK_ARG(world == 1 || handles || ptrs, "pcnn_icp_peer_attach: neither handles nor pointers given");
    PCNN_CHECK_ARG(st->target && st->target->n > 0, "pcnn_icp_peer_attach: empty target cloud");
    DeviceGuard g(st->device);
    icp_peer_detach(st);
    for (int p = 0; p < world; p++) {
        if (p == rank) { st->peer[p] = st->box; continue; }
        if (ptrs) {
            PCNN_CHECK_ARG(ptrs[p], "pcnn_icp_peer_attach: NULL peer pointer");
            cudaPointerAttributes at{};
            PCNN_CUDA(cudaPointerGetAttributes(&at, ptrs[p]));
            if (at.type == cudaMemoryTypeDevice && at.device != st->device) {
                const cudaError_t e = cudaDeviceEnablePeerAccess(at.device, 0);
                if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) {
                    icp_peer_detach(st);
                    set_error("pcnn_icp_peer_attach: no peer access %d -> %d: %s", st->device, at.device, cudaGetErrorString(e));
                    return PCNN_ERR_CUDA;
                }
                cudaGetLastError();
            }
            st->peer[p] = static_cast<PeerBox*>(ptrs[p]);
        } else {
            cudaIpcMemHandle_t h;
            std::memcpy(&h, static_cast<const char*>(handles) + (size_t)p * PCNN_ICP_PEER_HANDLE_BYTES, sizeof(h));
            void* ptr = nullptr;
            const cudaError_t e = cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess);
            if (e != cudaSuccess) {
                icp_peer_detach(st);
                set_error("pcnn_icp_peer_attach: cannot map the mailbox of rank %d: %s", p, cudaGetErrorString(e));
                return PCNN_ERR_CUDA;
            }
            st->peer[p] = static_cast<PeerBox*>(ptr);
            st->peer_ipc[p] = true;
        }
    }
    st->rank = rank; st->world = world;
    return PCNN_OK;
}

int pcnn_icp_step_peer(pcnn_icp_state* st, pcnn_stream stream) {
    PCNN_CHECK_ARG(st, "pcnn_icp_step_peer: NULL state");
    PCNN_CHECK_ARG(st->world >= 1, "pcnn_icp_step_peer: call pcnn_icp_peer_attach first");
    DeviceGuard g(st->device);
    return icp_launch_accumulate(st, 2, (cudaStream_t)stream);
}

int pcnn_icp_peer_detach(pcnn_icp_state* st) {
    PCNN_CHECK_ARG(st, "pcnn_icp_peer_detach: NULL state");
    DeviceGuard g(st->device);
    cudaDeviceSynchronize();
    icp_peer_detach(st);
    return PCNN_OK;
}

int pcnn_icp_update(pcnn_icp_state* st, pcnn_stream stream) {
    PCNN_CHECK_ARG(st, "pcnn_icp_update: NULL state");
    DeviceGuard g(st->device);
    icp_update_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(st->dev, st->prm.mode, st->prm.loss_epsilon);
    PCNN_CUDA(cudaGetLastError());
    return PCNN_OK;
}

int pcnn_icp_finish(pcnn_icp_state* st, pcnn_icp_result* out_host, pcnn_stream stream) {
    PCNN_CHECK_ARG(st && out_host, "pcnn_icp_finish: NULL argument");
    DeviceGuard g(st->device);
    Dev h;
    PCNN_CUDA(cudaMemcpyAsync(&h, st->dev, sizeof(Dev), cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    PCNN_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
#ifdef PCNN_STUDY
    if (getenv("PCNN_ICP_STATS"))
        fprintf(stderr, "pcnn_icp: %llu searches for %lld points x %d iterations (%.1f %% certified)\n", h.walked,
                (long long)st->n_src, h.iters,
                h.iters > 0 && st->n_src > 0 ? 100.0 - 100.0 * (double)h.walked / ((double)st->n_src * h.iters) : 0.0);
#endif
    std::memcpy(out_host->T, h.T, sizeof(h.T));
    out_host->iters = h.iters;
    out_host->converged = h.converged;
    out_host->status = h.status;
    out_host->last_loss = h.last_loss;
    out_host->last_pairs = h.last_pairs;
    out_host->loop_ms = 0.f;
    out_host->last_searches = h.last_searches;
    return PCNN_OK;
}

int pcnn_icp_destroy(pcnn_icp_state* st) {
    if (!st) return PCNN_OK;
    DeviceGuard g(st->device);
    cudaDeviceSynchronize();
    icp_free(st);
    return PCNN_OK;
}

int pcnn_icp(const pcnn_index* target, const float* tar_normals_dev, const float* src_dev, int64_t n_src,
             int stride, const float* init_T_host, const pcnn_icp_params* params, pcnn_icp_result* out_host,
             pcnn_stream stream) {
    PCNN_CHECK_ARG(out_host, "pcnn_icp: out is NULL");
    pcnn_icp_state* st = nullptr;
    int rc = pcnn_icp_create(&st, target, tar_normals_dev, src_dev, n_src, stride, init_T_host, params, 0, stream);
    if (rc != PCNN_OK) return rc;
    DeviceGuard g(st->device);
    if (n_src == 0 || target->n == 0) {
        // no correspondences can exist: report status 1 without iterating
        rc = pcnn_icp_finish(st, out_host, stream);
        if (rc == PCNN_OK && params->max_iter > 0) out_host->status = 1;
    } else {
        cudaEvent_t e0 = nullptr, e1 = nullptr;
        const bool timed = cudaEventCreate(&e0) == cudaSuccess && cudaEventCreate(&e1) == cudaSuccess;
        if (timed) cudaEventRecord(e0, (cudaStream_t)stream);
        for (int it = 0; it < params->max_iter && rc == PCNN_OK; it++)
            rc = icp_launch_accumulate(st, 1, (cudaStream_t)stream);  // fused update, no host sync
        if (timed) cudaEventRecord(e1, (cudaStream_t)stream);
        if (rc == PCNN_OK) rc = pcnn_icp_finish(st, out_host, stream);  // waits for the stream
        if (rc == PCNN_OK && timed) cudaEventElapsedTime(&out_host->loop_ms, e0, e1);
        if (e0) cudaEventDestroy(e0);
        if (e1) cudaEventDestroy(e1);
    }
    cudaStreamSynchronize((cudaStream_t)stream);
    icp_free(st);
    return rc;
}

int pcnn_nn1_f32(const pcnn_index* index, const float* q_dev, int64_t nq, int stride, int32_t* out_idx_dev,
                 float* out_d2_dev, pcnn_stream stream) {
    PCNN_CHECK_ARG(index && out_idx_dev, "pcnn_nn1_f32: NULL argument");
    PCNN_CHECK_ARG(nq >= 0 && nq < 0x7fffffff && stride >= 3, "pcnn_nn1_f32: bad nq/stride");
    PCNN_CHECK_ARG(nq == 0 || q_dev, "pcnn_nn1_f32: q is NULL");
    DeviceGuard g(index->device);
    return launch_nn1(index, q_dev, nq, stride, out_idx_dev, out_d2_dev, (cudaStream_t)stream);
}

}  // extern "C"
