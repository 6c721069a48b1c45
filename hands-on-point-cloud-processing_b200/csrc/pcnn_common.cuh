// pcnn_common.cuh -- shared definitions of the B200 (sm_100a) nearest-neighbour library.
//
// Data layout in HBM (one pcnn_index):
//   pts    float4[n]      the cloud in 63-bit Morton order; .w carries the ORIGINAL index
//                         (bit-cast int32), so one 16-byte load yields a candidate + its id
//   gcode  uint64[n0]     Morton code of the first point of every 8-point base group
//   box    float4[2*nb]   implicit binary hierarchy of AABBs over the base groups:
//                         level l node i covers groups [i<<l, (i+1)<<l); box[2*(off[l]+i)]
//                         = (min.xyz, -), box[2*(off[l]+i)+1] = (max.xyz, -)
// Distances follow the reference bit for bit (Homework2/hw2/include/kdtree.hpp:341-346):
// d2 = ((dx*dx)+dy*dy)+dz*dz in IEEE double from float32 inputs, each operation rounded
// (no FMA contraction -- x86-64 baseline code has none), and the AABB lower bound uses the
// SAME operation sequence so that, by monotonicity of rounding, bound <= d2 of every point
// inside the box: pruning on `bound > worst` can never lose a neighbour or a tie.
#pragma once
#include <cmath>
#include <cstdint>
#include <cuda_runtime.h>

#define PCNN_GROUP 8          // points per base group
#define PCNN_GROUP_LOG 3
#define PCNN_MAX_LEVELS 32
#define PCNN_MORTON_BITS 21   // per axis

#if defined(__CUDACC__)
#define PCNN_HD __host__ __device__ __forceinline__
#else
#define PCNN_HD inline
#endif

namespace pcnn {

struct IndexView {
    const float4* pts;
    const unsigned long long* gcode;
    const float4* box;
    int n;        // points
    int n0;       // base groups = ceil(n / 8)
    int nlev;     // levels 0..nlev-1, level nlev-1 has one node
    int lev_off[PCNN_MAX_LEVELS];
    const float* cube;  // [8]: bbox min.xyz, 2^21 / cube side, bbox max.xyz, cube side
};

#if defined(__CUDACC__)
// Programmatic dependent launch.  A kernel launched with launch_pdl(..., pdl = true) may have its blocks
// become resident while its predecessor in the stream still runs; pdl_prologue(), the FIRST thing such a
// kernel does, holds them until every prerequisite grid has completed and its memory is visible, and
// then lets the successor's launch overlap this grid.  Launched plainly, the prologue is a no-op.
// What it buys is launch latency between small dependent kernels (the 13 launches of a query ordering,
// the two kernels of an ICP iteration); it must not be used where the waiting blocks would take slots a
// full-device predecessor still needs (callers decide).
__device__ __forceinline__ void pdl_prologue() {
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, cudaStream_t s, bool pdl, Args&&... args) {
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = 0;
    cfg.stream = s;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
#endif

// ---- exact arithmetic helpers (identical results on host and device)
PCNN_HD double dmul(double a, double b) {
#if defined(__CUDA_ARCH__)
    return __dmul_rn(a, b);
#else
    volatile double r = a * b;  // volatile: keep the rounded product, no host FMA contraction
    return r;
#endif
}
PCNN_HD double dadd(double a, double b) {
#if defined(__CUDA_ARCH__)
    return __dadd_rn(a, b);
#else
    volatile double r = a + b;
    return r;
#endif
}
PCNN_HD float fmul(float a, float b) {
#if defined(__CUDA_ARCH__)
    return __fmul_rn(a, b);
#else
    volatile float r = a * b;
    return r;
#endif
}
PCNN_HD float fadd(float a, float b) {
#if defined(__CUDA_ARCH__)
    return __fadd_rn(a, b);
#else
    volatile float r = a + b;
    return r;
#endif
}

// Reference key: kdtree.hpp:341-346 / octree.hpp:152-154 without the final sqrt.
PCNN_HD double key64(float px, float py, float pz, double qx, double qy, double qz) {
    double dx = (double)px - qx, dy = (double)py - qy, dz = (double)pz - qz;
    return dadd(dadd(dmul(dx, dx), dmul(dy, dy)), dmul(dz, dz));
}
PCNN_HD double dmax0(double a, double b) {
    double m = a > b ? a : b;
    return m > 0.0 ? m : 0.0;
}
// Lower bound of key64 over every float32 point inside [lo, hi].
PCNN_HD double boxlb64(const float4& lo, const float4& hi, double qx, double qy, double qz) {
    double dx = dmax0((double)lo.x - qx, qx - (double)hi.x);
    double dy = dmax0((double)lo.y - qy, qy - (double)hi.y);
    double dz = dmax0((double)lo.z - qz, qz - (double)hi.z);
    return dadd(dadd(dmul(dx, dx), dmul(dy, dy)), dmul(dz, dz));
}
// Upper bound of key64 over the box (used by the radius `contains` fast path,
// octree.hpp:133-139): farthest corner, same monotone operation sequence.
PCNN_HD double boxub64(const float4& lo, const float4& hi, double qx, double qy, double qz) {
    double ax = qx - (double)lo.x, bx = (double)hi.x - qx;
    double ay = qy - (double)lo.y, by = (double)hi.y - qy;
    double az = qz - (double)lo.z, bz = (double)hi.z - qz;
    double dx = ax > bx ? ax : bx, dy = ay > by ? ay : by, dz = az > bz ? az : bz;
    return dadd(dadd(dmul(dx, dx), dmul(dy, dy)), dmul(dz, dz));
}

// nanoflann key as the ICP loop sees it (Homework9/hw9/include/nanoflann.hpp:383-408):
// float32, dx = q - p, accumulated x -> y -> z.
PCNN_HD float key32(float px, float py, float pz, float qx, float qy, float qz) {
    float dx = qx - px, dy = qy - py, dz = qz - pz;
    return fadd(fadd(fmul(dx, dx), fmul(dy, dy)), fmul(dz, dz));
}
PCNN_HD float fmax0(float a, float b) {
    float m = a > b ? a : b;
    return m > 0.0f ? m : 0.0f;
}
PCNN_HD float boxlb32(const float4& lo, const float4& hi, float qx, float qy, float qz) {
    float dx = fmax0(lo.x - qx, qx - hi.x);
    float dy = fmax0(lo.y - qy, qy - hi.y);
    float dz = fmax0(lo.z - qz, qz - hi.z);
    return fadd(fadd(fmul(dx, dx), fmul(dy, dy)), fmul(dz, dz));
}

#if !defined(__CUDA_ARCH__)
inline float i2f_host_prev(float f) { return nextafterf(f, -HUGE_VALF); }
inline float i2f_host_next(float f) { return nextafterf(f, HUGE_VALF); }
#endif

// ---- conservative float32 filters in front of the exact double key
// lbsq32() returns a float that is <= (1 - 2^-20) * (ax^2 + ay^2 + az^2) in real arithmetic for
// non-negative inputs that are themselves lower bounds of |dx|, |dy|, |dz|.  The exact key of
// any point is >= that real value * (1 - 2^-50), so `lbsq32 > ru32(worst)` proves the point
// (or every point of the box) cannot enter the list -- ties included.  Cheap: FP32 pipe only.
PCNN_HD float lbsq32(float ax, float ay, float az) {
#if defined(__CUDA_ARCH__)
    const float s = __fmaf_rd(az, az, __fmaf_rd(ay, ay, __fmul_rd(ax, ax)));  // every step rounds down: s <= the real sum
    return __fmul_rd(s, 0.99999905f);
#else
    const double s = (double)ax * ax + (double)ay * ay + (double)az * az;
    return (float)(s * 0.9999980926513672);  // extra margin covers the host's round-to-nearest
#endif
}
#if defined(__CUDACC__)
// The same bound without the safety factor and with fused multiply-adds, for the kNN walks: three
// instructions instead of six.  Every operation rounds DOWN, so the result s is <= the real
// ax^2 + ay^2 + az^2, and the exact double key of any point the inputs bound is >= that real value
// * (1 - 2^-50).  If s > wf in float32 then s >= wf * (1 + 2^-23), so key >= s * (1 - 2^-50) > wf
// >= the k-th key: the candidate (or the whole box) is out, ties included -- no float lies in
// [s * (1 - 2^-50), s), so a key EQUAL to the k-th never sees s > wf.  No margin is needed.
__device__ __forceinline__ float lbsq32x(float ax, float ay, float az) {
    return __fmaf_rd(az, az, __fmaf_rd(ay, ay, __fmul_rd(ax, ax)));
}
#endif
// |a - b| rounded toward zero (never above the real difference)
PCNN_HD float absdiff_rz(float a, float b) {
#if defined(__CUDA_ARCH__)
    return fabsf(__fsub_rz(a, b));
#else
    const double d = (double)a - (double)b;
    float f = (float)(d < 0 ? -d : d);
    if ((double)f > (d < 0 ? -d : d)) f = i2f_host_prev(f);
    return f;
#endif
}
// max(lo - q, q - hi, 0) with the differences rounded down
PCNN_HD float boxgap_rd(float lo, float hi, float q) {
#if defined(__CUDA_ARCH__)
    return fmaxf(fmaxf(__fsub_rd(lo, q), __fsub_rd(q, hi)), 0.0f);
#else
    const float a = absdiff_rz(lo, q), b = absdiff_rz(q, hi);
    const float m = lo > q ? a : (q > hi ? b : 0.0f);
    return m;
#endif
}
// smallest float >= w
PCNN_HD float ru32(double w) {
#if defined(__CUDA_ARCH__)
    return __double2float_ru(w);
#else
    float f = (float)w;
    if ((double)f < w) f = i2f_host_next(f);
    return f;
#endif
}

// ---- Morton codes
PCNN_HD unsigned long long spread21(unsigned long long v) {
    v &= 0x1fffffULL;
    v = (v | (v << 32)) & 0x1f00000000ffffULL;
    v = (v | (v << 16)) & 0x1f0000ff0000ffULL;
    v = (v | (v << 8)) & 0x100f00f00f00f00fULL;
    v = (v | (v << 4)) & 0x10c30c30c30c30c3ULL;
    v = (v | (v << 2)) & 0x1249249249249249ULL;
    return v;
}
PCNN_HD unsigned int quant21(float v, float mn, float inv_cell) {
    float f = (v - mn) * inv_cell;
    if (!(f > 0.0f)) f = 0.0f;               // also catches NaN
    if (f > 2097151.0f) f = 2097151.0f;
    return (unsigned int)f;
}
PCNN_HD unsigned long long morton63(float x, float y, float z, const float* bbmin, float inv_cell) {
    unsigned long long a = quant21(x, bbmin[0], inv_cell);
    unsigned long long b = quant21(y, bbmin[1], inv_cell);
    unsigned long long c = quant21(z, bbmin[2], inv_cell);
    return spread21(a) | (spread21(b) << 1) | (spread21(c) << 2);
}

PCNN_HD int f2i(float f) {
#if defined(__CUDA_ARCH__)
    return __float_as_int(f);
#else
    union { float f; int i; } u; u.f = f; return u.i;
#endif
}
PCNN_HD float i2f(int i) {
#if defined(__CUDA_ARCH__)
    return __int_as_float(i);
#else
    union { float f; int i; } u; u.i = i; return u.f;
#endif
}

// (key, index) lexicographic order: the canonical tie-break of the result contract.
PCNN_HD bool kless(double ka, int ia, double kb, int ib) {
    return ka < kb || (ka == kb && ia < ib);
}
PCNN_HD bool kless32(float ka, int ia, float kb, int ib) {
    return ka < kb || (ka == kb && ia < ib);
}

}  // namespace pcnn
