// octree_build.cu -- the octree of a scene built on the GPU, byte-identical to
// the host builder (scene.cpp) and hence to the reference's Octree::Octree (src/octree.cpp:388-450).
//
// What the reference does (octree.cpp:262-333): depth-first, per node, `grad()` over ALL dipoles
// at the node centre decides leaf / split (:267); a leaf samples `field()` over ALL dipoles at 8
// points and subtracts its residents' own terms, storing a float (:272-298); children inherit the
// dipoles within sqrt(3)*g/2 + scale*cell_r of their centre (:168-201).  O(nodes x dipoles) libm
// pow() calls on 16 std::threads: 15 s for the 2000-cell scene, 321 s for 8000 cells.
//
// Here: level-synchronous breadth-first build.  Per level four data-parallel kernels
//   criterion_kernel   one thread per node:  grad over all dipoles (sequential, reference order)
//   leaf_field_kernel  one thread per (leaf, sample point): field over all dipoles - residents
//                      (small levels: the *_group_kernel<G> variants, G lanes per point, further down)
//   spawn_kernel       one thread per child of an internal node: corner, Morton code
//   filter_kernel x2   count / fill the child's resident list from the parent's (order kept)
// with CUB exclusive scans in between; nodes live in SoA arrays per level, resident lists are
// index lists into the dipole table (double4, read uniformly by a warp -> one broadcast L1 hit).
// The depth-first order, relative child offsets and running resident cursor of the MRISCN01 /
// reference layout are produced by an O(nodes) linearisation on the host at the end.
//
// Bit-exactness against libm.  The only operation of the build the GPU cannot reproduce bit for
// bit is glibc's pow(r2, 2.5) / pow(r2, 3.5) (<= 1 ulp, not correctly rounded).  The kernels use
// r2*r2*sqrt(r2) resp. r2*r2*r2*sqrt(r2) (<= 2 ulp) and every other operation exactly as the host
// spells it (-fmad=false, IEEE sqrt and /), so each term differs from the host's by at most
// 2^-50 relative.  Both kernels carry, next to each running sum, sum|term| and sum|partial sum|,
// which bound |sum_gpu - sum_host| <= 2^-49 * sum|term| + 2^-52 * sum|partial| (derivation in
// DESIGN.md 5.3; the first part carries 2x safety, the second is the sum of both sides' half-ulp
// rounding errors and dominates: it is what sends ~n_dipoles * 2^-28 of the samples to the host).  A leaf decision is CERTIFIED when lhs -/+ bound fall on
// the same side of max_prod, a float sample when (float)(B - bound) == (float)(B + bound)
// (rounding is monotonic).  Everything else -- typically 1e-4 of the samples, no decision at all
// -- is re-evaluated by the host with libm through the seams of scene_internal.h.  The thresholds
// that are pow(., 2) in the reference are computed once by the host and passed in.
#include <cuda_runtime.h>
#include <cub/device/device_scan.cuh>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "mrisim.h"
#include "scene_internal.h"

namespace {

std::atomic<double> g_margin_scale{1.0};

constexpr double kRelTerm = 0x1p-49;            // per-term |gpu - host| / |term|, 2x the derived 2^-50
constexpr double kRelSum  = 0x1.001p-52;        // per-addition rounding of both sums: the derived 2^-52 plus second-order slack
constexpr int kThreads = 128;

template <typename T>
struct DevVec {
    T* p = nullptr; size_t n = 0;
    DevVec() = default;
    DevVec(const DevVec&) = delete; DevVec& operator=(const DevVec&) = delete;
    DevVec(DevVec&& o) noexcept : p(o.p), n(o.n) { o.p = nullptr; o.n = 0; }
    DevVec& operator=(DevVec&& o) noexcept { if (this != &o) { release(); p = o.p; n = o.n; o.p = nullptr; o.n = 0; } return *this; }
    ~DevVec() { release(); }
    void release() { if (p) cudaFree(p); p = nullptr; n = 0; }
    cudaError_t alloc(size_t count) {
        release();
        n = count;
        return cudaMalloc((void**)&p, std::max<size_t>(count, 1) * sizeof(T));
    }
};

// one level of the forest on the device
struct DevLevel {
    uint32_t n = 0;
    DevVec<double> x, y, z;             // lower corner of the node's cube
    DevVec<uint64_t> mc;                // Morton code with depth marker (no leaf bit)
    DevVec<uint32_t> off, cnt;          // resident list = pool[off .. off+cnt)
    DevVec<int32_t> pool;               // dipole indices, lists back to back in node order
};

// ... and its copy on the host, kept for the linearisation
struct HostLevel {
    uint32_t n = 0; double g = 0;
    std::vector<double> x, y, z;
    std::vector<uint64_t> mc;
    std::vector<uint8_t> leaf;
    std::vector<uint32_t> off, cnt, int_rank, Bbits, out_index;   // out_index: leaf -> node index in the linearised forest
    std::vector<int32_t> pool;
};

__device__ __forceinline__ double normsq(double x, double y, double z) { return x * x + y * y + z * z; }

// octree.cpp:228-252 + :267.  Sequential over the dipoles in table order (the reference's
// iteration order); masked terms are added as +0.0, which leaves a running sum unchanged.
__global__ void __launch_bounds__(kThreads)
criterion_kernel(const double4* __restrict__ mnp, int n_mnp, const double* __restrict__ x, const double* __restrict__ y,
                 const double* __restrict__ z, uint32_t n, double g, double grad_lim, double max_prod, double margin_scale,
                 uint8_t* __restrict__ leaf, uint32_t* __restrict__ unc_list, uint32_t* __restrict__ unc_count) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double px = x[i] + g / 2, py = y[i] + g / 2, pz = z[i] + g / 2;
    double rx = 0, ry = 0, rz = 0;       // the reference's ret_x/y/z
    double ax = 0, ay = 0, az = 0;       // sum |term|
    double sx = 0, sy = 0, sz = 0;       // sum |partial sum|
    for (int k = 0; k < n_mnp; k++) {
        const double4 m = mnp[k];
        const double dx = px - m.x, dy = py - m.y, dz = pz - m.z, M = m.w;
        const double r2 = normsq(dx, dy, dz);
        const bool on = r2 > grad_lim;
        const double divisor = r2 * r2 * r2 * sqrt(r2);              // pow(r2, 3.5) within 2 ulp
        const double xy = dx * dx + dy * dy;
        const double a = 1 * xy - 4 * dz * dz;
        const double c = 3 * xy - 2 * dz * dz;
        double tx = 3e17 * M * dx * a / divisor;
        double ty = 3e17 * M * dy * a / divisor;
        double tz = 3e17 * M * dz * c / divisor;
        tx = on ? tx : 0.0; ty = on ? ty : 0.0; tz = on ? tz : 0.0;
        rx += tx; ry += ty; rz += tz;
        ax += fabs(tx); ay += fabs(ty); az += fabs(tz);
        sx += fabs(rx); sy += fabs(ry); sz += fabs(rz);
    }
    const double grad = sqrt(normsq(rx, ry, rz));
    const double lhs = g / 2 * 1e-6 * grad;
    const double ex = kRelTerm * ax + kRelSum * sx, ey = kRelTerm * ay + kRelSum * sy, ez = kRelTerm * az + kRelSum * sz;
    const double eg = sqrt(normsq(ex, ey, ez)) + 0x1p-49 * grad;
    const double e = (g / 2 * 1e-6) * eg * margin_scale * 1.000001;
    const bool is_leaf = lhs < max_prod;
    const bool certain = isfinite(lhs) && isfinite(e) && ((lhs - e < max_prod) == (lhs + e < max_prod));
    leaf[i] = is_leaf ? 1 : 0;
    if (!certain) unc_list[atomicAdd(unc_count, 1u)] = i;
}

// octree.cpp:206-222 + :272-298: one thread per (leaf, sample point).
__global__ void __launch_bounds__(kThreads)
leaf_field_kernel(const double4* __restrict__ mnp, int n_mnp, const double* __restrict__ x, const double* __restrict__ y,
                  const double* __restrict__ z, const uint32_t* __restrict__ off, const uint32_t* __restrict__ cnt,
                  const int32_t* __restrict__ pool, const uint32_t* __restrict__ leaf_idx, uint32_t n_leaf, double g,
                  double margin_scale, uint32_t* __restrict__ Bbits, uint32_t* __restrict__ unc_list,
                  uint32_t* __restrict__ unc_count) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if ((t >> 3) >= n_leaf) return;
    const uint32_t i = leaf_idx[t >> 3], o = t & 7;
    // sic: g/4 in x but g/2 in y and z (octree.cpp:276-278)
    const double xn = x[i] + g / 4 + (double)((o >> 2) & 1) * g / 2;
    const double yn = y[i] + g / 2 + (double)((o >> 1) & 1) * g / 2;
    const double zn = z[i] + g / 2 + (double)((o >> 0) & 1) * g / 2;
    double B = 0, a = 0, s = 0;
    for (int k = 0; k < n_mnp; k++) {
        const double4 m = mnp[k];
        const double dx = xn - m.x, dy = yn - m.y, dz = zn - m.z;
        const double r2 = normsq(dx, dy, dz);
        const double divisor = r2 * r2 * sqrt(r2);                   // pow(r2, 2.5) within 1.5 ulp
        const double term = m.w * 1e11 * (2 * dz * dz - dx * dx - dy * dy) / divisor;
        B += term; a += fabs(term); s += fabs(B);
    }
    const uint32_t first = off[i], count = cnt[i];
    for (uint32_t k = 0; k < count; k++) {
        const double4 m = mnp[pool[first + k]];
        const double dx = xn - m.x, dy = yn - m.y, dz = zn - m.z;
        const double r2 = normsq(dx, dy, dz);
        const double divisor = r2 * r2 * sqrt(r2);
        const double term = m.w * 1e11 * (2 * dz * dz - dx * dx - dy * dy) / divisor;
        B -= term; a += fabs(term); s += fabs(B);
    }
    const double e = (kRelTerm * a + kRelSum * s + 0x1p-51 * fabs(B)) * margin_scale;   // last part: roundings of B -/+ e below
    const float lo = (float)(B - e), hi = (float)(B + e), Bf = (float)B;
    const bool certain = isfinite(B) && isfinite(e) && (__float_as_uint(lo) == __float_as_uint(hi));
    Bbits[(size_t)i * 8 + o] = __float_as_uint(Bf);
    if (!certain) unc_list[atomicAdd(unc_count, 1u)] = i * 8 + o;
}

// ---- warp-cooperative variants for SMALL levels --------------------------------------------------
// Near the root a level has a few hundred nodes: one thread per node then walks all dipoles alone,
// ~0.46 us per dipole, 2.5 ms per level whatever the grid size (ncu: 13 % of the warp slots busy).
// Here a WARP owns the node / sample point: the 32 lanes evaluate 32 dipoles' terms side by side,
// and the running sums take them IN TABLE ORDER through shuffles -- the order is part of the result
// (it is the host's `ret += ...`), so the values are those of the kernels above bit for bit.  The
// sum|term| bound is order-free and is reduced at the end.  ~5x the issue slots per pair of the
// thread-per-point kernels, so only used while a level is latency-bound.
// G = 32 lanes per point for the smallest levels, G = 4 for the levels in between (enough points to
// fill the machine four lanes each, too few to hide a whole dipole walk per thread).
// Measured on B200, 5190 dipoles: criterion 16 k nodes 2.2 ms (G=32) / 39 k nodes 2.6 -> 2.1 ms and 28 k nodes 2.4 -> 1.4 ms
// (thread -> G=4); leaf field 89 k points 1.83 ms (thread) vs 2.06 ms (G=4), so no G=4 band there.
constexpr uint32_t kWarpNodes = 6000, kWarpPoints = 24000;        // up to here: 32 lanes per point
constexpr uint32_t kQuadNodes = 100000, kQuadPoints = 0;          // up to here: 4 lanes per point
constexpr unsigned kFullMask = 0xffffffffu;

template <int G>
__device__ __forceinline__ double group_sum_d(double v) {
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) v += __shfl_xor_sync(kFullMask, v, o);
    return v;
}

template <int G>
__global__ void __launch_bounds__(kThreads)
criterion_group_kernel(const double4* __restrict__ mnp, int n_mnp, const double* __restrict__ x, const double* __restrict__ y,
                       const double* __restrict__ z, uint32_t n, double g, double grad_lim, double max_prod, double margin_scale,
                       uint8_t* __restrict__ leaf, uint32_t* __restrict__ unc_list, uint32_t* __restrict__ unc_count) {
    const uint32_t gt = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t lane = threadIdx.x & 31, sub = lane & (G - 1), first = lane & ~(uint32_t)(G - 1);
    if (((gt & ~31u) / G) >= n) return;                                  // whole warp past the end
    const uint32_t node = gt / G;
    const bool live = node < n;
    const uint32_t i = live ? node : n - 1;                              // idle groups shadow the last node (they still take part in the shuffles)
    const double px = x[i] + g / 2, py = y[i] + g / 2, pz = z[i] + g / 2;
    double rx = 0, ry = 0, rz = 0, ax = 0, ay = 0, az = 0, sx = 0, sy = 0, sz = 0;
    for (int k0 = 0; k0 < n_mnp; k0 += G) {
        const int cnt = min(G, n_mnp - k0);
        double tx = 0.0, ty = 0.0, tz = 0.0;
        if ((int)sub < cnt) {
            const double4 m = mnp[k0 + sub];
            const double dx = px - m.x, dy = py - m.y, dz = pz - m.z, M = m.w;
            const double r2 = normsq(dx, dy, dz);
            const bool on = r2 > grad_lim;
            const double divisor = r2 * r2 * r2 * sqrt(r2);
            const double xy = dx * dx + dy * dy;
            const double a = 1 * xy - 4 * dz * dz;
            const double c = 3 * xy - 2 * dz * dz;
            tx = 3e17 * M * dx * a / divisor;
            ty = 3e17 * M * dy * a / divisor;
            tz = 3e17 * M * dz * c / divisor;
            tx = on ? tx : 0.0; ty = on ? ty : 0.0; tz = on ? tz : 0.0;
            ax += fabs(tx); ay += fabs(ty); az += fabs(tz);
        }
        for (int j = 0; j < cnt; j++) {
            rx += __shfl_sync(kFullMask, tx, first + j); ry += __shfl_sync(kFullMask, ty, first + j); rz += __shfl_sync(kFullMask, tz, first + j);
            sx += fabs(rx); sy += fabs(ry); sz += fabs(rz);
        }
    }
    ax = group_sum_d<G>(ax) * 1.000001; ay = group_sum_d<G>(ay) * 1.000001; az = group_sum_d<G>(az) * 1.000001;
    if (sub || !live) return;
    const double grad = sqrt(normsq(rx, ry, rz));
    const double lhs = g / 2 * 1e-6 * grad;
    const double ex = kRelTerm * ax + kRelSum * sx, ey = kRelTerm * ay + kRelSum * sy, ez = kRelTerm * az + kRelSum * sz;
    const double eg = sqrt(normsq(ex, ey, ez)) + 0x1p-49 * grad;
    const double e = (g / 2 * 1e-6) * eg * margin_scale * 1.000001;
    const bool is_leaf = lhs < max_prod;
    const bool certain = isfinite(lhs) && isfinite(e) && ((lhs - e < max_prod) == (lhs + e < max_prod));
    leaf[i] = is_leaf ? 1 : 0;
    if (!certain) unc_list[atomicAdd(unc_count, 1u)] = i;
}

// G >= 4 and a multiple-of-8 point index per leaf: the 32/G points of a warp belong to ONE leaf, so the
// resident count is uniform over the warp
template <int G>
__global__ void __launch_bounds__(kThreads)
leaf_field_group_kernel(const double4* __restrict__ mnp, int n_mnp, const double* __restrict__ x, const double* __restrict__ y,
                        const double* __restrict__ z, const uint32_t* __restrict__ off, const uint32_t* __restrict__ cnt,
                        const int32_t* __restrict__ pool, const uint32_t* __restrict__ leaf_idx, uint32_t n_leaf, double g,
                        double margin_scale, uint32_t* __restrict__ Bbits, uint32_t* __restrict__ unc_list,
                        uint32_t* __restrict__ unc_count) {
    static_assert(G >= 4 && G <= 32, "a warp must stay within one leaf");
    const uint32_t gt = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t lane = threadIdx.x & 31, sub = lane & (G - 1), first_lane = lane & ~(uint32_t)(G - 1);
    const uint32_t t = gt / G;                                            // sample point
    if ((t >> 3) >= n_leaf) return;                                      // uniform over the warp
    const uint32_t i = leaf_idx[t >> 3], o = t & 7;
    const double xn = x[i] + g / 4 + (double)((o >> 2) & 1) * g / 2;
    const double yn = y[i] + g / 2 + (double)((o >> 1) & 1) * g / 2;
    const double zn = z[i] + g / 2 + (double)((o >> 0) & 1) * g / 2;
    double B = 0, a = 0, s = 0;
    const uint32_t first = off[i], count = cnt[i];
    for (int pass = 0; pass < 2; pass++) {                       // all dipoles (+), then the leaf's residents (-)
        const int total = pass == 0 ? n_mnp : (int)count;
        for (int k0 = 0; k0 < total; k0 += G) {
            const int c = min(G, total - k0);
            double term = 0.0;
            if ((int)sub < c) {
                const double4 m = mnp[pass == 0 ? k0 + (int)sub : pool[first + k0 + sub]];
                const double dx = xn - m.x, dy = yn - m.y, dz = zn - m.z;
                const double r2 = normsq(dx, dy, dz);
                const double divisor = r2 * r2 * sqrt(r2);
                term = m.w * 1e11 * (2 * dz * dz - dx * dx - dy * dy) / divisor;
                a += fabs(term);
            }
            if (pass == 0) for (int j = 0; j < c; j++) { B += __shfl_sync(kFullMask, term, first_lane + j); s += fabs(B); }
            else           for (int j = 0; j < c; j++) { B -= __shfl_sync(kFullMask, term, first_lane + j); s += fabs(B); }
        }
    }
    a = group_sum_d<G>(a) * 1.000001;
    if (sub) return;
    const double e = (kRelTerm * a + kRelSum * s + 0x1p-51 * fabs(B)) * margin_scale;
    const float lo = (float)(B - e), hi = (float)(B + e), Bf = (float)B;
    const bool certain = isfinite(B) && isfinite(e) && (__float_as_uint(lo) == __float_as_uint(hi));
    Bbits[(size_t)i * 8 + o] = __float_as_uint(Bf);
    if (!certain) unc_list[atomicAdd(unc_count, 1u)] = i * 8 + o;
}

__global__ void flag_internal_kernel(const uint8_t* __restrict__ leaf, uint32_t n, uint32_t* __restrict__ is_int) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) is_int[i] = leaf[i] ? 0u : 1u;
}

__global__ void partition_kernel(const uint8_t* __restrict__ leaf, const uint32_t* __restrict__ int_rank, uint32_t n,
                                 uint32_t* __restrict__ leaf_idx, uint32_t* __restrict__ int_idx) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (leaf[i]) leaf_idx[i - int_rank[i]] = i; else int_idx[int_rank[i]] = i;
}

// octree.cpp:303-316: child c of internal node j is node 8*rank(j)+c of the next level.
__global__ void spawn_kernel(const double* __restrict__ px, const double* __restrict__ py, const double* __restrict__ pz,
                             const uint64_t* __restrict__ pmc, const uint32_t* __restrict__ poff,
                             const uint32_t* __restrict__ pcnt, const uint32_t* __restrict__ int_idx, uint32_t n_child,
                             double g, double* __restrict__ cx, double* __restrict__ cy, double* __restrict__ cz,
                             uint64_t* __restrict__ cmc, uint32_t* __restrict__ src_off, uint32_t* __restrict__ src_cnt) {
    const uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= n_child) return;
    const uint32_t j = int_idx[c >> 3], o = c & 7;
    cx[c] = px[j] + (double)((o >> 2) & 1) * g / 2;
    cy[c] = py[j] + (double)((o >> 1) & 1) * g / 2;
    cz[c] = pz[j] + (double)((o >> 0) & 1) * g / 2;
    cmc[c] = (pmc[j] << 3) | (uint64_t)o;
    src_off[c] = poff[j]; src_cnt[c] = pcnt[j];
}

// octree.cpp:168-201 (MNP_in_space + transfer_MNPs): keep the dipoles of the source list that lie
// within the child's sphere, in source order.  FILL = false counts, FILL = true writes.
// src_pool == nullptr: the source list is the whole dipole table (subtree roots, octree.cpp:368).
template <bool FILL>
__global__ void __launch_bounds__(kThreads)
filter_kernel(const double4* __restrict__ mnp, const double* __restrict__ x, const double* __restrict__ y,
              const double* __restrict__ z, const uint32_t* __restrict__ src_off, const uint32_t* __restrict__ src_cnt,
              const int32_t* __restrict__ src_pool, uint32_t n, double g, double space_lim,
              uint32_t* __restrict__ cnt, const uint32_t* __restrict__ off, int32_t* __restrict__ pool) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double cx = x[i], cy = y[i], cz = z[i];
    const uint32_t first = src_off[i], count = src_cnt[i];
    uint32_t kept = 0;
    const uint32_t out0 = FILL ? off[i] : 0;
    for (uint32_t k = 0; k < count; k++) {
        const int32_t id = src_pool ? src_pool[first + k] : (int32_t)k;
        const double4 m = mnp[id];
        const double dx = cx + g / 2 - m.x, dy = cy + g / 2 - m.y, dz = cz + g / 2 - m.z;
        if (normsq(dx, dy, dz) < space_lim) {
            if (FILL) pool[out0 + kept] = id;
            kept++;
        }
    }
    if (!FILL) cnt[i] = kept;
}

template <typename F>
void parallel_for(size_t n, int n_threads, F&& body) {
    unsigned hw = std::thread::hardware_concurrency();
    size_t nt = n_threads > 0 ? (size_t)n_threads : (hw ? hw : 1);
    nt = std::min(nt, std::max<size_t>(n, 1));
    if (nt <= 1) { for (size_t i = 0; i < n; i++) body(i); return; }
    std::atomic<size_t> next{0};
    std::vector<std::thread> th;
    for (size_t t = 0; t < nt; t++) th.emplace_back([&] { for (size_t i; (i = next.fetch_add(1)) < n;) body(i); });
    for (auto& t : th) t.join();
}

inline double now_ms() {
    return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

inline uint32_t blocks_for(uint64_t n) { return (uint32_t)((n + kThreads - 1) / kThreads); }

}  // namespace

extern "C" void mrisim_octree_build_set_margin_scale(double scale) { g_margin_scale.store(scale > 0 ? scale : 1.0); }

namespace mrisim_host {

#define OB_CK(call)                                                                          \
    do {                                                                                     \
        cudaError_t e_ = (call);                                                             \
        if (e_ != cudaSuccess) {                                                             \
            err = std::string(#call) + ": " + cudaGetErrorString(e_) + " (no CPU fallback)"; \
            return 1;                                                                        \
        }                                                                                    \
    } while (0)

template <typename T>
static cudaError_t download(std::vector<T>& h, const DevVec<T>& d, size_t n, cudaStream_t st) {
    h.resize(n);
    if (!n) return cudaSuccess;
    return cudaMemcpyAsync(h.data(), d.p, n * sizeof(T), cudaMemcpyDeviceToHost, st);
}

static cudaError_t exclusive_sum(const uint32_t* in, uint32_t* out, uint32_t n, DevVec<uint8_t>& tmp, cudaStream_t st) {
    size_t bytes = 0;
    cudaError_t e = cub::DeviceScan::ExclusiveSum(nullptr, bytes, in, out, (int)n, st);
    if (e != cudaSuccess) return e;
    if (tmp.n < bytes) { e = tmp.alloc(bytes); if (e != cudaSuccess) return e; }
    return cub::DeviceScan::ExclusiveSum(tmp.p, bytes, in, out, (int)n, st);
}

// The float samples of the leaves are NOT read here (the libm fix-up threads may still be patching
// them): each leaf's position in the output is recorded in out_index and the samples are copied
// once the fix-ups have been joined.
static void emit_subtree(std::vector<HostLevel>& lv, const std::vector<mrisim_mnp>& mnps, size_t d, uint32_t i,
                         OctreeOut& out) {
    HostLevel& L = lv[d];
    const size_t self = out.nodes.size();
    mrisim_node nd;
    std::memset(&nd, 0, sizeof nd);
    nd.mc = L.mc[i];
    nd.res_start = (uint32_t)out.residents.size();     // running cursor for every node (MRISCN01 canonical form)
    if (L.leaf[i]) {
        nd.mc |= (1ULL << 63);
        nd.n_res = L.cnt[i];
        L.out_index[i] = (uint32_t)self;
        out.nodes.push_back(nd);
        for (uint32_t k = 0; k < L.cnt[i]; k++) out.residents.push_back(mnps[(size_t)L.pool[L.off[i] + k]]);
        return;
    }
    out.nodes.push_back(nd);
    for (uint32_t o = 0; o < 8; o++) {
        out.nodes[self].child[o] = (uint32_t)(out.nodes.size() - self);     // the reference's curr->size(), octree.cpp:320
        emit_subtree(lv, mnps, d + 1, 8 * L.int_rank[i] + o, out);
    }
}

int build_octree_gpu(const mrisim_params& p, const std::vector<mrisim_mnp>& mnps, int device, int n_threads,
                     OctreeOut& out, mrisim_octree_build_stats* stats, std::string& err) {
    const double t_begin = now_ms();
    mrisim_octree_build_stats st_{};
    int n_dev = 0;
    if (cudaGetDeviceCount(&n_dev) != cudaSuccess || device >= n_dev) {
        err = "no CUDA device " + std::to_string(device) + " (no CPU fallback)";
        return 1;
    }
    OB_CK(cudaSetDevice(device));
    cudaStream_t st;
    OB_CK(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
    struct StreamGuard { cudaStream_t s; ~StreamGuard() { cudaStreamDestroy(s); } } sg{st};
    cudaEvent_t ev0, ev1;
    OB_CK(cudaEventCreate(&ev0)); OB_CK(cudaEventCreate(&ev1));
    struct EvGuard { cudaEvent_t a, b; ~EvGuard() { cudaEventDestroy(a); cudaEventDestroy(b); } } eg{ev0, ev1};

    const double margin_scale = g_margin_scale.load();
    const int n_mnp = (int)mnps.size();
    octree_depths(p, &out.min_depth, &out.max_depth);
    const uint32_t n_roots = 1u << (3 * out.min_depth);
    const double grad_lim = octree_grad_limit(p);

    // dipole table (x, y, z, M): 32-byte records
    std::vector<double4> h_mnp((size_t)n_mnp);
    for (int k = 0; k < n_mnp; k++) h_mnp[(size_t)k] = make_double4(mnps[(size_t)k].x, mnps[(size_t)k].y, mnps[(size_t)k].z, mnps[(size_t)k].M);
    DevVec<double4> d_mnp;
    OB_CK(d_mnp.alloc((size_t)n_mnp));
    OB_CK(cudaMemcpyAsync(d_mnp.p, h_mnp.data(), (size_t)n_mnp * sizeof(double4), cudaMemcpyHostToDevice, st));

    // level min_depth: the subtree roots (octree.cpp:346-367), corners and codes from the host expressions
    std::vector<HostLevel> levels;
    levels.reserve(64);                      // the fix-up threads keep pointers into it (depth <= 21 by the Morton width)
    struct Fixers { std::vector<std::thread> th; void join() { for (auto& t : th) if (t.joinable()) t.join(); th.clear(); } ~Fixers() { join(); } } fixers;
    std::atomic<double> fix_ms{0.0};
    DevLevel cur;
    double g = 0;
    {
        HostLevel R;
        R.n = n_roots; R.x.resize(n_roots); R.y.resize(n_roots); R.z.resize(n_roots); R.mc.resize(n_roots);
        for (uint32_t i = 0; i < n_roots; i++) octree_root(p, out.min_depth, (int)i, &R.x[i], &R.y[i], &R.z[i], &g, &R.mc[i]);
        cur.n = n_roots;
        OB_CK(cur.x.alloc(n_roots)); OB_CK(cur.y.alloc(n_roots)); OB_CK(cur.z.alloc(n_roots)); OB_CK(cur.mc.alloc(n_roots));
        OB_CK(cur.off.alloc(n_roots)); OB_CK(cur.cnt.alloc(n_roots));
        OB_CK(cudaMemcpyAsync(cur.x.p, R.x.data(), n_roots * sizeof(double), cudaMemcpyHostToDevice, st));
        OB_CK(cudaMemcpyAsync(cur.y.p, R.y.data(), n_roots * sizeof(double), cudaMemcpyHostToDevice, st));
        OB_CK(cudaMemcpyAsync(cur.z.p, R.z.data(), n_roots * sizeof(double), cudaMemcpyHostToDevice, st));
        OB_CK(cudaMemcpyAsync(cur.mc.p, R.mc.data(), n_roots * sizeof(uint64_t), cudaMemcpyHostToDevice, st));
        // root resident lists: filter of the whole table
        std::vector<uint32_t> zero(n_roots, 0u), all(n_roots, (uint32_t)n_mnp);
        DevVec<uint32_t> src_off, src_cnt;
        OB_CK(src_off.alloc(n_roots)); OB_CK(src_cnt.alloc(n_roots));
        OB_CK(cudaMemcpyAsync(src_off.p, zero.data(), n_roots * sizeof(uint32_t), cudaMemcpyHostToDevice, st));
        OB_CK(cudaMemcpyAsync(src_cnt.p, all.data(), n_roots * sizeof(uint32_t), cudaMemcpyHostToDevice, st));
        const double lim = octree_space_limit(p, g);
        DevVec<uint8_t> tmp;
        OB_CK(cudaEventRecord(ev0, st));
        filter_kernel<false><<<blocks_for(n_roots), kThreads, 0, st>>>(d_mnp.p, cur.x.p, cur.y.p, cur.z.p, src_off.p, src_cnt.p,
                                                                     nullptr, n_roots, g, lim, cur.cnt.p, nullptr, nullptr);
        OB_CK(exclusive_sum(cur.cnt.p, cur.off.p, n_roots, tmp, st));
        uint32_t last_off = 0, last_cnt = 0;
        OB_CK(cudaMemcpyAsync(&last_off, cur.off.p + n_roots - 1, 4, cudaMemcpyDeviceToHost, st));
        OB_CK(cudaMemcpyAsync(&last_cnt, cur.cnt.p + n_roots - 1, 4, cudaMemcpyDeviceToHost, st));
        OB_CK(cudaStreamSynchronize(st));
        OB_CK(cur.pool.alloc((size_t)last_off + last_cnt));
        filter_kernel<true><<<blocks_for(n_roots), kThreads, 0, st>>>(d_mnp.p, cur.x.p, cur.y.p, cur.z.p, src_off.p, src_cnt.p,
                                                                    nullptr, n_roots, g, lim, nullptr, cur.off.p, cur.pool.p);
        OB_CK(cudaEventRecord(ev1, st));
        OB_CK(cudaStreamSynchronize(st));
        OB_CK(cudaGetLastError());
        float ms = 0; OB_CK(cudaEventElapsedTime(&ms, ev0, ev1)); st_.gpu_ms += ms;
        st_.kernel_launches += 3; st_.pair_evals += 2ull * n_roots * (uint64_t)n_mnp;
    }

    DevVec<uint8_t> d_leaf, scan_tmp;
    DevVec<uint32_t> d_is_int, d_int_rank, d_leaf_idx, d_int_idx, d_unc, d_unc_count, d_Bbits;
    OB_CK(d_unc_count.alloc(1));

    for (;;) {
        const uint32_t n = cur.n;
        levels.emplace_back();
        HostLevel& H = levels.back();
        H.n = n; H.g = g;
        OB_CK(d_leaf.alloc(n)); OB_CK(d_is_int.alloc(n)); OB_CK(d_int_rank.alloc(n));
        OB_CK(d_leaf_idx.alloc(n)); OB_CK(d_int_idx.alloc(n));
        OB_CK(d_unc.alloc((size_t)n * 8)); OB_CK(d_Bbits.alloc((size_t)n * 8));
        OB_CK(cudaMemsetAsync(d_unc_count.p, 0, 4, st));
        OB_CK(cudaMemsetAsync(d_Bbits.p, 0, (size_t)n * 8 * 4, st));
        OB_CK(download(H.x, cur.x, n, st)); OB_CK(download(H.y, cur.y, n, st)); OB_CK(download(H.z, cur.z, n, st));
        OB_CK(download(H.mc, cur.mc, n, st)); OB_CK(download(H.off, cur.off, n, st)); OB_CK(download(H.cnt, cur.cnt, n, st));
        OB_CK(download(H.pool, cur.pool, cur.pool.n, st));

        // ---- leaf criterion (octree.cpp:267)
        OB_CK(cudaEventRecord(ev0, st));
        if (g < kMinG) {
            OB_CK(cudaMemsetAsync(d_leaf.p, 1, n, st));
        } else {
            if (n <= kWarpNodes)
                criterion_group_kernel<32><<<blocks_for((uint64_t)n * 32), kThreads, 0, st>>>(d_mnp.p, n_mnp, cur.x.p, cur.y.p, cur.z.p, n, g,
                                                                                              grad_lim, kMaxProduct, margin_scale, d_leaf.p,
                                                                                              d_unc.p, d_unc_count.p);
            else if (n <= kQuadNodes)
                criterion_group_kernel<4><<<blocks_for((uint64_t)n * 4), kThreads, 0, st>>>(d_mnp.p, n_mnp, cur.x.p, cur.y.p, cur.z.p, n, g,
                                                                                            grad_lim, kMaxProduct, margin_scale, d_leaf.p,
                                                                                            d_unc.p, d_unc_count.p);
            else
                criterion_kernel<<<blocks_for(n), kThreads, 0, st>>>(d_mnp.p, n_mnp, cur.x.p, cur.y.p, cur.z.p, n, g, grad_lim,
                                                                   kMaxProduct, margin_scale, d_leaf.p, d_unc.p, d_unc_count.p);
            st_.kernel_launches++; st_.criterion_evals += n; st_.pair_evals += (uint64_t)n * (uint64_t)n_mnp;
        }
        OB_CK(cudaEventRecord(ev1, st));
        uint32_t n_unc = 0;
        OB_CK(cudaMemcpyAsync(&n_unc, d_unc_count.p, 4, cudaMemcpyDeviceToHost, st));
        OB_CK(download(H.leaf, d_leaf, n, st));
        OB_CK(cudaStreamSynchronize(st));
        OB_CK(cudaGetLastError());
        { float ms = 0; OB_CK(cudaEventElapsedTime(&ms, ev0, ev1)); st_.gpu_ms += ms; }
        if (n_unc) {       // decisions the bound could not certify: the host's libm decides
            const double t0 = now_ms();
            std::vector<uint32_t> unc(n_unc);
            OB_CK(cudaMemcpy(unc.data(), d_unc.p, (size_t)n_unc * 4, cudaMemcpyDeviceToHost));
            parallel_for(n_unc, n_threads, [&](size_t k) {
                const uint32_t i = unc[k];
                H.leaf[i] = octree_criterion(p, mnps, H.x[i], H.y[i], H.z[i], g) < kMaxProduct ? 1 : 0;
            });
            OB_CK(cudaMemcpyAsync(d_leaf.p, H.leaf.data(), n, cudaMemcpyHostToDevice, st));
            st_.criterion_fixups += n_unc;
            st_.fixup_ms += now_ms() - t0;
        }

        // ---- partition into leaves and internal nodes
        OB_CK(cudaEventRecord(ev0, st));
        flag_internal_kernel<<<blocks_for(n), kThreads, 0, st>>>(d_leaf.p, n, d_is_int.p);
        OB_CK(exclusive_sum(d_is_int.p, d_int_rank.p, n, scan_tmp, st));
        partition_kernel<<<blocks_for(n), kThreads, 0, st>>>(d_leaf.p, d_int_rank.p, n, d_leaf_idx.p, d_int_idx.p);
        st_.kernel_launches += 3;
        OB_CK(download(H.int_rank, d_int_rank, n, st));
        OB_CK(cudaStreamSynchronize(st));
        const uint32_t n_int = H.int_rank[n - 1] + (H.leaf[n - 1] ? 0u : 1u), n_leaf = n - n_int;

        // ---- float B samples of the leaves (octree.cpp:272-298)
        OB_CK(cudaMemsetAsync(d_unc_count.p, 0, 4, st));
        if (n_leaf) {
            const uint64_t n_pts = (uint64_t)n_leaf * 8;
            if (n_pts <= kWarpPoints)
                leaf_field_group_kernel<32><<<blocks_for(n_pts * 32), kThreads, 0, st>>>(
                    d_mnp.p, n_mnp, cur.x.p, cur.y.p, cur.z.p, cur.off.p, cur.cnt.p, cur.pool.p, d_leaf_idx.p, n_leaf, g,
                    margin_scale, d_Bbits.p, d_unc.p, d_unc_count.p);
            else if (n_pts <= kQuadPoints)
                leaf_field_group_kernel<4><<<blocks_for(n_pts * 4), kThreads, 0, st>>>(
                    d_mnp.p, n_mnp, cur.x.p, cur.y.p, cur.z.p, cur.off.p, cur.cnt.p, cur.pool.p, d_leaf_idx.p, n_leaf, g,
                    margin_scale, d_Bbits.p, d_unc.p, d_unc_count.p);
            else
                leaf_field_kernel<<<blocks_for(n_pts), kThreads, 0, st>>>(
                    d_mnp.p, n_mnp, cur.x.p, cur.y.p, cur.z.p, cur.off.p, cur.cnt.p, cur.pool.p, d_leaf_idx.p, n_leaf, g,
                    margin_scale, d_Bbits.p, d_unc.p, d_unc_count.p);
            st_.kernel_launches++; st_.b_values += (uint64_t)n_leaf * 8;
        }

        // ---- children of the internal nodes: corners, codes, resident lists
        DevLevel nxt;
        DevVec<uint32_t> src_off, src_cnt;
        const uint32_t n_child = n_int * 8;
        const double gc = g / 2;
        double lim = 0;
        if (n_child) {
            if ((uint64_t)n_int * 8 > 0x7fffffffull) { err = "octree level larger than 2^31 nodes"; return 1; }
            nxt.n = n_child;
            OB_CK(nxt.x.alloc(n_child)); OB_CK(nxt.y.alloc(n_child)); OB_CK(nxt.z.alloc(n_child)); OB_CK(nxt.mc.alloc(n_child));
            OB_CK(nxt.off.alloc(n_child)); OB_CK(nxt.cnt.alloc(n_child));
            OB_CK(src_off.alloc(n_child)); OB_CK(src_cnt.alloc(n_child));
            lim = octree_space_limit(p, gc);
            spawn_kernel<<<blocks_for(n_child), kThreads, 0, st>>>(cur.x.p, cur.y.p, cur.z.p, cur.mc.p, cur.off.p, cur.cnt.p,
                                                                 d_int_idx.p, n_child, g, nxt.x.p, nxt.y.p, nxt.z.p, nxt.mc.p,
                                                                 src_off.p, src_cnt.p);
            filter_kernel<false><<<blocks_for(n_child), kThreads, 0, st>>>(d_mnp.p, nxt.x.p, nxt.y.p, nxt.z.p, src_off.p, src_cnt.p,
                                                                         cur.pool.p, n_child, gc, lim, nxt.cnt.p, nullptr, nullptr);
            OB_CK(exclusive_sum(nxt.cnt.p, nxt.off.p, n_child, scan_tmp, st));
            st_.kernel_launches += 3;
        }
        uint32_t last_off = 0, last_cnt = 0, n_unc_b = 0;
        if (n_child) {
            OB_CK(cudaMemcpyAsync(&last_off, nxt.off.p + n_child - 1, 4, cudaMemcpyDeviceToHost, st));
            OB_CK(cudaMemcpyAsync(&last_cnt, nxt.cnt.p + n_child - 1, 4, cudaMemcpyDeviceToHost, st));
        }
        OB_CK(cudaMemcpyAsync(&n_unc_b, d_unc_count.p, 4, cudaMemcpyDeviceToHost, st));
        OB_CK(download(H.Bbits, d_Bbits, (size_t)n * 8, st));
        OB_CK(cudaStreamSynchronize(st));
        OB_CK(cudaGetLastError());
        if (n_child) {
            if ((uint64_t)last_off + last_cnt > 0xffffffffull) { err = "resident pool larger than 2^32 entries"; return 1; }
            OB_CK(nxt.pool.alloc((size_t)last_off + last_cnt));
            filter_kernel<true><<<blocks_for(n_child), kThreads, 0, st>>>(d_mnp.p, nxt.x.p, nxt.y.p, nxt.z.p, src_off.p, src_cnt.p,
                                                                        cur.pool.p, n_child, gc, lim, nullptr, nxt.off.p, nxt.pool.p);
            st_.kernel_launches++;
        }
        OB_CK(cudaEventRecord(ev1, st));
        for (uint32_t i = 0; i < n; i++) {
            if (H.leaf[i]) st_.pair_evals += 8ull * ((uint64_t)n_mnp + H.cnt[i]);
            else st_.pair_evals += 16ull * H.cnt[i];
        }

        std::vector<uint32_t> unc_b(n_unc_b);
        if (n_unc_b) OB_CK(cudaMemcpyAsync(unc_b.data(), d_unc.p, (size_t)n_unc_b * 4, cudaMemcpyDeviceToHost, st));
        OB_CK(cudaStreamSynchronize(st));
        OB_CK(cudaGetLastError());
        { float ms = 0; OB_CK(cudaEventElapsedTime(&ms, ev0, ev1)); st_.gpu_ms += ms; }
        if (n_unc_b) {     // samples whose bound straddles a float rounding boundary: the host's libm decides,
                           // on its own threads, while the GPU goes on with the next level
            HostLevel* Hp = &H;
            const double gl = g;
            st_.b_fixups += n_unc_b;
            fixers.th.emplace_back([&p, &mnps, Hp, gl, n_threads, &fix_ms, unc = std::move(unc_b)]() {
                const double t0 = now_ms();
                parallel_for(unc.size(), n_threads, [&](size_t k) {
                    const uint32_t i = unc[k] >> 3, o = unc[k] & 7;
                    const double xn = Hp->x[i] + gl / 4 + (double)((o >> 2) & 1) * gl / 2;
                    const double yn = Hp->y[i] + gl / 2 + (double)((o >> 1) & 1) * gl / 2;
                    const double zn = Hp->z[i] + gl / 2 + (double)((o >> 0) & 1) * gl / 2;
                    const float Bf = (float)octree_leaf_B(p, mnps, Hp->pool.data() + Hp->off[i], Hp->cnt[i], xn, yn, zn);
                    std::memcpy(&Hp->Bbits[(size_t)i * 8 + o], &Bf, 4);
                });
                double cur_ms = fix_ms.load();
                while (!fix_ms.compare_exchange_weak(cur_ms, cur_ms + (now_ms() - t0))) {}
            });
        }
        st_.n_leaves += n_leaf; st_.n_nodes += n; st_.n_levels++;
        if (!n_child) break;
        cur = std::move(nxt);
        g = gc;
    }

    // ---- depth-first linearisation (octree.cpp:303-322 order), subtree by subtree, beside the fix-up threads
    const double t_lin = now_ms();
    out.subtree_sizes.assign(n_roots, 0);
    out.nodes.clear(); out.residents.clear();
    out.nodes.reserve(st_.n_nodes);
    for (HostLevel& L : levels) L.out_index.assign(L.n, 0xffffffffu);
    for (uint32_t r = 0; r < n_roots; r++) {
        const size_t before = out.nodes.size();
        emit_subtree(levels, mnps, 0, r, out);
        out.subtree_sizes[r] = (uint32_t)(out.nodes.size() - before);
    }
    st_.linearise_ms = now_ms() - t_lin;
    {   // the libm re-evaluations still running: wait (only this part of them is on the critical path)
        const double t0 = now_ms();
        fixers.join();
        st_.fixup_wait_ms = now_ms() - t0;
        st_.fixup_ms = fix_ms.load();
    }
    for (const HostLevel& L : levels)                      // the (now final) float samples of every leaf
        for (uint32_t i = 0; i < L.n; i++)
            if (L.leaf[i]) std::memcpy(out.nodes[L.out_index[i]].child, &L.Bbits[(size_t)i * 8], 8 * sizeof(uint32_t));
    st_.total_ms = now_ms() - t_begin;
    if (stats) *stats = st_;
    return 0;
}

}  // namespace mrisim_host
