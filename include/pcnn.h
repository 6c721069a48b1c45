/* pcnn.h -- C ABI of the B200-native nearest-neighbour + ICP hot path.
 *
 * Drop-in boundary for the nearest-neighbour path of yf26/Hands-On-Point-Cloud-Processing.
 * The reference has no FFI; its operator API for this path is a set of C++ free functions
 * and result types (citations relative to the reference tree):
 *
 *   KDTreeConstruction / KDTreeDestruction      Homework2/hw2/include/kdtree.hpp:419-435
 *   KDTreeKNNSearch / KDTreeRadiusNNSearch      kdtree.hpp:329-364 / :367-402
 *   OctreeConstruction / OctreeDestruction      Homework2/hw2/include/octree.hpp:242-277
 *   OctreeKNNSearch / OctreeRadiusNNSearch      octree.hpp:143-186 / :189-237
 *   KNNResultSet / RadiusNNResultSet            Homework2/hw2/include/resultSet.hpp:28-93 / :96-142
 *   the per-query batch loop                    Homework2/hw2/include/benchmark.hpp:30-43, :59-72
 *   Registration::ICPpoint2point / 2plane       Homework9/hw9/src/registration.cpp:862-1011 / :710-859
 *   nanoflann 1-NN as ICP uses it               registration.cpp:772-785, :925-941
 *
 * Every entry point below names the reference interface it replaces.  Plain pointers and
 * sizes only; no C++ or torch types.  include/pcnn_compat.hpp (Homework2 names) and
 * include/pcnn_registration_compat.hpp (Homework9's Registration) restore the reference's own names
 * on top of this header.
 *
 * Conventions
 *  - All functions return an int status: PCNN_OK (0) or a negative PCNN_ERR_*; the message
 *    of the last failure on the calling thread is available from pcnn_last_error().
 *  - Pointers named *_dev are device pointers on the index's device; *_host are host
 *    pointers (pageable or pinned).  `stream` is a cudaStream_t passed as void* (NULL =
 *    the legacy default stream).  Device-pointer entry points only ENQUEUE work on `stream`
 *    and return; host-pointer entry points copy in, run, copy out and return when the
 *    results are in the caller's buffers.
 *  - Points and queries are float32 with `stride` floats between consecutive points
 *    (3 = packed xyz, 4 = KITTI xyzi / float4).  The reference widens float32 input to
 *    double (test.hpp:24-29); this ABI takes the float32 originals, which is lossless.
 *  - There is no CPU fallback: without a CUDA device every compute entry point fails with
 *    PCNN_ERR_CUDA.
 *
 * Result contract
 *  - distance key  d2 = ((dx*dx) + dy*dy) + dz*dz  in IEEE double, dx = p - q widened from
 *    float32, no FMA -- the reference's own arithmetic (kdtree.hpp:341-346);
 *    reported distance = sqrt(d2), correctly rounded, as the reference stores it.
 *  - kNN rows: the k smallest under (d2, original index), ascending.  The reference breaks
 *    exact-distance ties by tree visit order (resultSet.hpp:69-90); the two differ only
 *    inside an exact tie group straddling the k-th place.  Rows with fewer than k points
 *    available are padded with index -1 and distance +inf.
 *  - radius rows: every point with sqrt(d2) <= radius (inclusive, resultSet.hpp:134), decided
 *    on the exact double key; the order inside a row is deterministic (the same in every call
 *    with the same inputs and options) but unspecified -- the reference's order is its own
 *    traversal order -- so rows compare as sets.  out_dist64 holds the
 *    exact sqrt(d2); the float32 out_dist is within 1e-6 relative of it.
 */
#ifndef PCNN_H
#define PCNN_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PCNN_VERSION 100

#define PCNN_OK 0
#define PCNN_ERR_INVALID (-1)     /* bad argument */
#define PCNN_ERR_CUDA (-2)        /* CUDA runtime error / no device */
#define PCNN_ERR_NOMEM (-3)       /* device or host allocation failed */
#define PCNN_ERR_UNSUPPORTED (-4) /* argument outside the supported range (e.g. k > PCNN_MAX_K) */
#define PCNN_ERR_CAPACITY (-5)    /* caller's output buffer too small; required size reported */

#define PCNN_KDTREE 0 /* handle created for the KDTree* entry points */
#define PCNN_OCTREE 1 /* handle created for the Octree* entry points */

#define PCNN_MAX_K 128

typedef struct pcnn_index pcnn_index;         /* opaque spatial index, per-handle ownership */
typedef struct pcnn_icp_state pcnn_icp_state; /* opaque ICP iteration state */
typedef void* pcnn_stream;                    /* cudaStream_t */

const char* pcnn_last_error(void);
int pcnn_version(void);
/* Number of visible CUDA devices (0 if none); never fails. */
int pcnn_device_count(void);

/* Scratch memory (query ordering, scans) comes from a library-owned stream-ordered pool that
 * keeps freed blocks for reuse; this returns the cached blocks of `device` to the driver. */
int pcnn_release_cached(int device);

/* ------------------------------------------------------------------ index build
 * Replaces KDTreeConstruction(db, leaf_size) (kdtree.hpp:419-429) and
 * OctreeConstruction(db, leaf_size, min_extent) (octree.hpp:242-271).
 * `kind` records which family of entry points the handle serves; both kinds return the
 * same neighbour sets.  `leaf_size` and `min_extent` are accepted for signature parity;
 * leaf_size >= 1 and min_extent >= 0 are validated, the device index chooses its own
 * bucket size.  The input cloud is copied; the caller may free it after the call returns
 * (device variant: after `stream` has passed the call).
 * Built on device: bounding box -> 63-bit Morton codes -> LSD radix sort -> float4 cloud in
 * Morton order -> implicit binary hierarchy of AABBs over 8-point groups. */
int pcnn_index_build(pcnn_index** out, int kind, const float* xyz_dev, int64_t n, int stride,
                     int leaf_size, double min_extent, int device, pcnn_stream stream);
int pcnn_index_build_host(pcnn_index** out, int kind, const float* xyz_host, int64_t n,
                          int stride, int leaf_size, double min_extent, int device);
/* Replaces KDTreeDestruction() / OctreeDestruction() (kdtree.hpp:431-435, octree.hpp:273-277),
 * but frees exactly this handle. */
int pcnn_index_destroy(pcnn_index* index);

typedef struct {
    int64_t n_points;
    int32_t kind;
    int32_t device;
    int32_t n_levels;     /* levels of the implicit AABB hierarchy */
    int32_t group_size;   /* points per base group */
    int64_t device_bytes; /* HBM held by the handle */
    float bbox_min[3];
    float bbox_max[3];
    float build_ms;       /* device time of the last build (CUDA events) */
} pcnn_index_info;
int pcnn_index_get_info(const pcnn_index* index, pcnn_index_info* out);

/* ------------------------------------------------------------------ kNN
 * Replaces, for nq queries at once, the loop body of benchmark.hpp:61-66:
 *     KNNResultSet rs(k);  K?TreeKNNSearch / OctreeKNNSearch(root, db, rs, query);
 * out_idx [nq*k] int32 original point indices; out_dist [nq*k] float32 and/or
 * out_dist64 [nq*k] double distances (either may be NULL); rows ascending by (d2, index).
 * 1 <= k <= PCNN_MAX_K. */
int pcnn_knn(const pcnn_index* index, const float* q_dev, int64_t nq, int stride, int k,
             int32_t* out_idx_dev, float* out_dist_dev, double* out_dist64_dev,
             pcnn_stream stream);
int pcnn_knn_host(const pcnn_index* index, const float* q_host, int64_t nq, int stride, int k,
                  int32_t* out_idx_host, float* out_dist_host, double* out_dist64_host);

/* Query-side knobs (all have working defaults).  `sort_queries`: 1 = order queries along
 * the Morton curve before the search (default), 0 = caller's order, e.g. already coherent.
 * (pcnn_knn_host always searches its pipeline stages in the caller's order: there the ordering costs
 * more latency than it saves search time; rows never depend on the order.)
 * `method`: 0 = auto, 2 = warp-per-query traversal (any k), 3 = brute-force tiles (k <= 16),
 * 7 = lock-step nearest-first walks (auto for k <= 16), 9 = nearest-first walks with a shared-memory
 * heap (auto for 16 < k <= 32).  A method that cannot serve the requested k falls back to 2.
 * Methods 1, 4, 5, 6, 8, 10 and radius kernels 2, 3 are measured-slower variants that exist only in
 * the study build (`make -C .../csrc STUDY=1` -> libpcnn_study.so, -DPCNN_STUDY); this library
 * answers PCNN_ERR_UNSUPPORTED for them.
 * `sort_bits` (3..63, default 32): leading Morton bits the query batch is ordered by; up to 32 the
 * keys are 32-bit (four radix passes), beyond that 64-bit (more passes than the finer order repays).
 * `radius_batch`: radius kernel, 1 = auto (default), 0 = warp per query, depth first (also the
 * fallback for clouds of 2^30 points and more), 4 = warp per query with a dense stack walk (auto
 * for sparse batches), 5 = dense stack walk per run of queries (auto for batches as dense as the
 * cloud); `radius_run_div` (0 = auto), `radius_run_mult_count` / `radius_run_mult_fill`
 * (percent of the cloud's mean 8-point-group extent) size the runs.  Every choice returns the
 * same sets.
 * `icp_certificates` (default 1): 0 makes the ICP loop search every source point in every
 * iteration; the result is the same bit for bit (tests/test_gpu_configs.py), only slower.
 * `pdl` (default 1): the query ordering of pcnn_knn and the search kernel behind it are launched with
 * programmatic stream serialization (each launch overlaps its predecessor's tail); 0 = plain launches.
 * `icp_pdl` (default 1): the two kernels of an ICP iteration are launched with programmatic
 * stream serialization (the successor's launch overlaps the predecessor) when the source cloud does
 * not fill the device; 0 = plain launches.  Same result either way.
 * `nvtx` (default 0): 1 wraps the host-side enqueue of every phase of the calls on this handle in
 * NVTX ranges (pcnn.query_order, pcnn.knn, pcnn.knn_host.stage, pcnn.radius.count / .scan / .fill,
 * pcnn.icp.search, pcnn.icp.sums[+exchange][+solve]) for a profiler timeline.
 * Returns PCNN_ERR_INVALID for unknown names or values. */
int pcnn_set_option(pcnn_index* index, const char* name, int value);
/* With option "profile" = 1 every pcnn_knn / pcnn_radius_* call on this handle brackets its
 * main search kernel (not the query ordering) with CUDA events on the call's stream;
 * this returns the device time of the most recent one in milliseconds (waits for it). */
int pcnn_get_profile(const pcnn_index* index, float* main_kernel_ms);

/* Diagnostics for load-balance studies: while buf_dev is not NULL the warp-coherent kNN kernel
 * stores, for warp w of the launch, buf_dev[2w] = its start time (globaltimer, ns) and
 * buf_dev[2w+1] = its duration in SM cycles; the buffer needs 2*ceil(nq/32) entries. */
int pcnn_debug_warp_times(pcnn_index* index, unsigned long long* buf_dev);

/* ------------------------------------------------------------------ radius search
 * Replaces, for nq queries at once, benchmark.hpp:68-71:
 *     RadiusNNResultSet rs(radius);  K?TreeRadiusNNSearch / OctreeRadiusNNSearch(...);
 * Two passes, CSR output.  pcnn_radius_count fills out_offsets_dev[nq+1] (exclusive scan of
 * the per-query hit counts, so out_offsets[nq] = total hits); if out_total_host is not NULL
 * the call waits for the stream and stores the total there.  pcnn_radius_fill writes the
 * hits of query i at [offsets[i], offsets[i+1]). */
int pcnn_radius_count(const pcnn_index* index, const float* q_dev, int64_t nq, int stride,
                      double radius, int64_t* out_offsets_dev, int64_t* out_total_host,
                      pcnn_stream stream);
int pcnn_radius_fill(const pcnn_index* index, const float* q_dev, int64_t nq, int stride,
                     double radius, const int64_t* offsets_dev, int32_t* out_idx_dev,
                     float* out_dist_dev, double* out_dist64_dev, pcnn_stream stream);
/* Host variant: offsets_host[nq+1] always filled; hits are written only if `capacity`
 * (in hits) suffices, else PCNN_ERR_CAPACITY is returned with offsets_host valid so the
 * caller can size its buffers from offsets_host[nq] and call again. */
int pcnn_radius_host(const pcnn_index* index, const float* q_host, int64_t nq, int stride,
                     double radius, int64_t* offsets_host, int32_t* out_idx_host,
                     float* out_dist_host, double* out_dist64_host, int64_t capacity);

/* Allocation-free count pass.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid it).  Same result as pcnn_radius_count -- out_offsets_dev[nq+1], the
 * exclusive scan of the per-query hit counts, bit for bit -- but the call allocates NOTHING: no
 * cudaMalloc / cudaFree, no stream-ordered pool, no library routine that allocates.  Its scratch
 * (per-query counts, scan tile sums) is a file-scope device array of PCNN_NOALLOC_SCRATCH_BYTES that
 * every GPU which loads the library holds until exit.  Limits and rules:
 *   - nq <= PCNN_NOALLOC_MAX_QUERIES, otherwise PCNN_ERR_UNSUPPORTED (split the batch);
 *   - queries are processed in the caller's order (no query ordering: that needs scratch of its own),
 *     one warp per query; meant for small and medium batches;
 *   - all calls share the one array and it keeps the last call's contents (every word a call reads it
 *     has written first, nothing needs clearing): calls that OVERLAP on the GPU -- different streams
 *     without an ordering between them, on the same device -- race on it.  Calls on one stream, or
 *     ordered by events, are safe; different devices have separate copies;
 *   - only enqueues on `stream`; the total is out_offsets_dev[nq] once the stream has passed it. */
#define PCNN_NOALLOC_MAX_QUERIES 262144
#define PCNN_NOALLOC_SCRATCH_BYTES (2 * 1024 * 1024 + 4096)
int pcnn_radius_count_noalloc(const pcnn_index* index, const float* q_dev, int64_t nq, int stride,
                              double radius, int64_t* out_offsets_dev, pcnn_stream stream);

/* ------------------------------------------------------------------ ICP
 * Replaces Registration::ICPpoint2point (registration.cpp:862-1011) and ICPpoint2plane
 * (:710-859) minus the normal-space sampling done before the loop (:728-729, :880-881):
 * the caller passes the clouds the loop should run on.  Parameter meaning follows
 * Registration::setICPparams (registration.hpp:126-137): max_corres_dist gates the SQUARED
 * 1-NN distance (registration.cpp:783, :936); the loop stops early once
 * |last_loss - loss| < loss_epsilon has held more than 15 times (:828-838, :948-958). */
typedef struct {
    int32_t max_iter;
    float max_corres_dist;
    float loss_epsilon;
    int32_t mode; /* 0 = point-to-point, 1 = point-to-plane */
} pcnn_icp_params;

typedef struct {
    float T[16];       /* row-major 4x4 T_total: R = T[0..2][0..2], t = T[0..2][3] */
    int32_t iters;     /* iterations whose update was applied */
    int32_t converged; /* 1 if the unchanged-loss rule fired */
    int32_t status;    /* 0 ok; 1 = an iteration had no correspondences or a singular system;
                          2 = a peer did not answer in time (pcnn_icp_step_peer) */
    float last_loss;
    int64_t last_pairs; /* correspondences kept in the last executed iteration */
    float loop_ms;      /* pcnn_icp / pcnn_icp_host: device time of the iteration loop alone
                           (CUDA events on the stream); 0 from pcnn_icp_finish */
    int32_t last_searches; /* source points searched in the last executed iteration; the others kept
                              their correspondence under a certificate (same result, no search) */
} pcnn_icp_result;

/* Whole loop on one GPU: no host synchronisation per iteration; one at the end to return
 * the result.  `target` indexes the target cloud; tar_normals_dev ([n_target*3] packed,
 * original target order) is required for mode 1, ignored for mode 0; init_T is a host
 * row-major 4x4 (the reference's init_R / init_t). */
int pcnn_icp(const pcnn_index* target, const float* tar_normals_dev, const float* src_dev,
             int64_t n_src, int stride, const float* init_T_host, const pcnn_icp_params* params,
             pcnn_icp_result* out_host, pcnn_stream stream);
int pcnn_icp_host(const pcnn_index* target, const float* tar_normals_host,
                  const float* src_host, int64_t n_src, int stride, const float* init_T_host,
                  const pcnn_icp_params* params, pcnn_icp_result* out_host);

/* Split-phase form for source clouds sharded over several GPUs (one process per GPU): each
 * rank holds the full target index and its slice of the source.  Per iteration:
 *     pcnn_icp_accumulate(st, stream);      fused transform + 1-NN + gate + partial sums
 *     <all-reduce(sum) PCNN_ICP_NSUMS doubles at pcnn_icp_sums_dev(st) on `stream`>
 *     pcnn_icp_update(st, stream);          solve + pose update, replicated on every rank
 * Nothing synchronises with the host until pcnn_icp_finish. */
#define PCNN_ICP_NSUMS 32
int pcnn_icp_create(pcnn_icp_state** out, const pcnn_index* target, const float* tar_normals_dev,
                    const float* src_dev, int64_t n_src, int stride, const float* init_T_host,
                    const pcnn_icp_params* params, int64_t src_index_base, pcnn_stream stream);
int pcnn_icp_accumulate(pcnn_icp_state* st, pcnn_stream stream);
double* pcnn_icp_sums_dev(pcnn_icp_state* st);
int pcnn_icp_update(pcnn_icp_state* st, pcnn_stream stream);
int pcnn_icp_finish(pcnn_icp_state* st, pcnn_icp_result* out_host, pcnn_stream stream);
int pcnn_icp_destroy(pcnn_icp_state* st);

/* Fused form of the same iteration for GPUs that can address each other's memory (NVLink /
 * NVSwitch on one box): ONE kernel per iteration does accumulate + all-reduce + update.  The
 * last block of every rank stores its 32 folded sums into a mailbox in every peer's memory
 * (plain P2P stores, then a system-scope release of a sequence flag), waits for the flags of
 * all peers in its own mailbox, adds the world's contributions in rank order (so every rank
 * holds bit-identical sums, whatever the arrival order) and runs the replicated solve.  No
 * NCCL call, no extra launch, nothing on the host inside the loop.  What replaces which
 * reference line is unchanged: the sums are those of registration.cpp:788-818 / :964-987
 * built from ALL pairs.
 *
 *     pcnn_icp_peer_export(st, handle, &ptr);      every rank; handle = 64-byte CUDA IPC handle
 *     <exchange handles (other processes) or ptrs (same process)>
 *     pcnn_icp_peer_attach(st, rank, world, handles, ptrs);
 *     max_iter x pcnn_icp_step_peer(st, stream);   every rank, same number of calls
 *     pcnn_icp_finish(...); pcnn_icp_peer_detach(st); <barrier>; pcnn_icp_destroy(st);
 *
 * A peer that does not show up within PCNN_ICP_PEER_TIMEOUT_MS ends the loop on the waiting
 * rank with status 2 (later steps return at once): the kernel never spins unbounded. */
#define PCNN_ICP_PEER_HANDLE_BYTES 64
#define PCNN_ICP_MAX_PEERS 16
#define PCNN_ICP_PEER_TIMEOUT_MS 4000
int pcnn_icp_peer_export(pcnn_icp_state* st, void* handle_out /* 64 B, may be NULL */,
                         void** mailbox_dev_out /* may be NULL */);
/* handles: world x 64 B from the ranks' pcnn_icp_peer_export (entry `rank` ignored), or NULL when
 * ptrs (world device pointers valid in THIS process, entry `rank` ignored) is given instead. */
int pcnn_icp_peer_attach(pcnn_icp_state* st, int rank, int world, const void* handles,
                         void* const* ptrs);
int pcnn_icp_step_peer(pcnn_icp_state* st, pcnn_stream stream);
/* Unmaps the peers' mailboxes; every rank detaches BEFORE any rank destroys its state. */
int pcnn_icp_peer_detach(pcnn_icp_state* st);

/* Exact float32 1-NN of every query against the index, as the ICP loop uses it
 * (nanoflann.hpp:383-408 key: ((dx*dx)+dy*dy)+dz*dz in float32, dx = q - p; lowest index
 * wins an exact tie).  out_idx [nq], out_d2 [nq].  Batches of up to 2048 queries run one WARP per
 * query (32 child boxes tested and 128 points scanned per step, per-lane partial results merged at
 * the end); larger ones one query per lane in lock step.  Same result either way. */
int pcnn_nn1_f32(const pcnn_index* index, const float* q_dev, int64_t nq, int stride,
                 int32_t* out_idx_dev, float* out_d2_dev, pcnn_stream stream);

/* ------------------------------------------------------------------ ISS keypoints
 * (the reference's batched consumer of its own radius search.)
 * Replaces ISSKeypoint::compute (Homework7/hw7/src/iss_detector.cpp:38-111) and
 * getEigenvalues (:114-152) with the setters' parameters (iss_detector.hpp:18-24):
 * neighbourhoods of every point of the indexed cloud at local_radius; scatter matrix about
 * the point itself, neighbours weighted by 1/|their own neighbourhood| when `weighted`
 * (useWeightedCovMat); eigenvalues l1 >= l2 >= l3; a point is a candidate when it has >= 3
 * neighbours, l2/l1 < gamma21, l3/l2 < gamma32 and l3 > 0; a candidate is a keypoint when it
 * has >= min_neighbors points within non_max_radius and none of them has a larger l3.
 * Neighbourhoods follow Homework7's float KD-tree: key accumulated x, y, z with every step
 * rounded to float, hit iff sqrtf(key) <= radius (kdtree.cpp:347-355, resultSet.hpp:127-136).
 * The index may be of either kind; it must have been built over the cloud to analyse.
 *   out_idx_dev     [capacity] keypoint indices (original point order), ascending, as the
 *                   reference emits them (:104-105); entries beyond capacity are dropped
 *   out_count       number of keypoints found (may exceed capacity)
 *   out_lambda3_dev optional [n]: l3 of every candidate, -1 for the others (the reference's
 *                   lambda3_vec_forall, :69)
 * pcnn_iss_detect returns after ONE host synchronisation (the count); nothing is materialised
 * per neighbour. */
typedef struct {
    float local_radius;
    float non_max_radius;
    float gamma21;
    float gamma32;
    int32_t min_neighbors;
    int32_t weighted;
} pcnn_iss_params;

int pcnn_iss_detect(const pcnn_index* index, const pcnn_iss_params* params, int32_t* out_idx_dev,
                    int64_t capacity, int64_t* out_count_host, float* out_lambda3_dev, pcnn_stream stream);
/* Same with host outputs (out_lambda3_host may be NULL). */
int pcnn_iss_detect_host(const pcnn_index* index, const pcnn_iss_params* params, int32_t* out_idx_host,
                         int64_t capacity, int64_t* out_count_host, float* out_lambda3_host);

/* ------------------------------------------------------------------ voxel-grid downsample
 *   Replaces the pcl::VoxelGrid<pcl::PointNormal> step of
 * readBinaryAndVoxelDown (Homework9/hw9/src/registration.cpp:37-42: setDownsampleAllData(true),
 * setLeafSize(v, v, v)) and Registration::VoxelGridSampling (:665-707, leaf 1.75), following
 * PCL's published algorithm (pcl/filters/impl/voxel_grid.hpp, 1.8 - 1.10): one output point per
 * occupied voxel in ascending voxel index (x fastest), position = mean of the voxel's points,
 * normal = normalised sum of their normals; a grid with more than INT_MAX cells returns the
 * input unchanged, as PCL does.  Means are formed in double and rounded once (PCL's float sums
 * depend on the order std::sort leaves a voxel's points in).
 *   xyz_dev [n*stride], normals_dev [n*normal_stride] or NULL; out_xyz_dev / out_normals_dev
 *   [capacity*3] packed; out_count = number of occupied voxels (may exceed capacity; only the
 *   first `capacity` are written).  Two host synchronisations (grid size, count). */
int pcnn_voxel_downsample(const float* xyz_dev, const float* normals_dev, int64_t n, int stride,
                          int normal_stride, float leaf_x, float leaf_y, float leaf_z,
                          float* out_xyz_dev, float* out_normals_dev, int64_t capacity,
                          int64_t* out_count_host, int device, pcnn_stream stream);

/* ------------------------------------------------------------------ normals by PCA over kNN
 *   Replaces the per-point loop of the reference's normal estimation
 * (Homework1/YuF_KIT-.../pca_normal.py:85-101 with PCA(), :17-33): for every point of the
 * indexed cloud the at most max_nn nearest neighbours within `radius` (the point itself
 * included; open3d search_hybrid_vector_3d(radius, max_nn), :93); with >= 3 of them the normal
 * is the eigenvector of the smallest eigenvalue of sum (p - mean)(p - mean)^T, unit length;
 * otherwise (0, 0, 0).  out_normals [n*3] packed, original point order.  The sign of a normal
 * is arbitrary, as it is in the reference (LAPACK's): compare up to sign.
 * 1 <= max_nn <= PCNN_MAX_K. */
int pcnn_estimate_normals(const pcnn_index* index, int max_nn, double radius, float* out_normals_dev,
                          pcnn_stream stream);
int pcnn_estimate_normals_host(const pcnn_index* index, int max_nn, double radius, float* out_normals_host);

/* ------------------------------------------------------------------ global initialisation
 *   Descriptor matching and RANSAC before ICP, Homework9/hw9/src/registration.cpp.
 *
 * pcnn_desc_nn1: exact 1-NN of every row of a [na, dim] among the rows of b [nb, dim] in float32
 * squared L2 with nanoflann's arithmetic (`result += diff * diff`, dimension by dimension), as the
 * KDTreeEigenMatrixAdaptor searches of :458-492 return it; lowest index wins an exact tie.
 * dim <= 64 (FPFHSignature33: 33).  out_idx [na] (-1 when nb == 0), out_d2 [na].
 *
 * pcnn_match_descriptors: mode 0 = findRANSACCorrespondencesInter (:437-527: pairs that are each
 * other's nearest neighbour), mode 1 = findRANSACCorrespondencesUnion (:530-611: every target's
 * and every source's pair); pairs sorted by ascending descriptor distance and the worst
 * rejection_rate of them dropped, N = floor((1 - rate) * count) in float as at :516 / :603.
 * out_pairs [capacity*2] = (source index, target index); out_count = N (may exceed capacity).
 *
 * pcnn_ransac: Registration::RANSAC (:288-434).  Hypothesis `it` (0 <= it < max_iter) samples 4
 * distinct correspondences with a counter-based generator keyed by (seed, it) -- the reference
 * seeds mt19937 from std::random_device and is not reproducible -- re-draws while the signed
 * distance of the 4th source point from the plane of the first three is <= coplanar_min (0.15 at
 * :358), solves R, t by the SVD of target * source^T with the reference's det < 0 branch, and
 * counts the correspondences with |target - (R source + t)| <= dist_threshold.  Result: the
 * hypothesis with the largest consensus, the earliest such iteration (the strict `>` of :423).
 * All hypotheses are evaluated concurrently.  src/tar_kp_dev: packed xyz of the keypoints. */
int pcnn_desc_nn1(const float* a_dev, int64_t na, const float* b_dev, int64_t nb, int dim,
                  int32_t* out_idx_dev, float* out_d2_dev, int device, pcnn_stream stream);
int pcnn_match_descriptors(const float* src_desc_dev, int64_t n_src, const float* tar_desc_dev, int64_t n_tar,
                           int dim, int mode, float rejection_rate, int32_t* out_pairs_dev, int64_t capacity,
                           int64_t* out_count_host, int device, pcnn_stream stream);
typedef struct {
    int32_t max_iter;       /* m_RANSAC_max_iter (80000 in main.cpp:86) */
    float dist_threshold;   /* m_RANSAC_dist_threshold */
    float coplanar_min;     /* 0.15 in the reference */
    int32_t reserved;
    uint64_t seed;
} pcnn_ransac_params;
typedef struct {
    float R[9];             /* row-major */
    float t[3];
    int64_t consensus;      /* max_consensus_set_size; 0 = no hypothesis had an inlier */
    int64_t best_iter;      /* the iteration that produced it, -1 if none */
} pcnn_ransac_result;
int pcnn_ransac(const float* src_kp_dev, const float* tar_kp_dev, const int32_t* pairs_dev, int64_t n_pairs,
                const pcnn_ransac_params* params, pcnn_ransac_result* out_host, int device, pcnn_stream stream);

#ifdef __cplusplus
}
#endif
#endif /* PCNN_H */
