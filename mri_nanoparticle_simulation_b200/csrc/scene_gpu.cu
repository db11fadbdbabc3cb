// scene_gpu.cu -- the remaining O(N x cells) host loops of the scene build on the GPU
// (SURVEY.md 8f / N4): nearest-cell lookup table, water throw, Morton keys of the water sort.
// Every value is bit-identical to scene.cpp's (and the reference's) host code: the arithmetic is
// + - * / sqrt floor only, compiled with -fmad=false, and the one libm call involved
// (pow(cell_r, 2)) is evaluated by the host and passed in.
//
//   lookup_kernel     SimulationBox.cpp:262-290   one thread per lattice cube, cells scanned in order
//   candidate_kernel  BacteriaBox.cpp:80-105      attempt j owns draws 3j..3j+2 of the host stream:
//                                                 one thread per chunk of attempts, started from the
//                                                 generator state the host recorded for that chunk
//   contain_kernel    SimulationBox.cpp:161-173   one thread per attempt, all cells (as the reference)
//   compact_kernel    accepted attempts in stream order -> water_info records (CUB exclusive scan)
//   leaf_key_kernel   SimulationBox.cpp:131-147 + octree.cpp:99-128: Morton code of the leaf each
//                     water starts in (the std::sort itself stays on the host: its permutation of
//                     equal keys is libstdc++'s introsort, which is what the reference produces)
#include <cuda_runtime.h>
#include <cub/device/device_scan.cuh>

#include <algorithm>
#include <cmath>
#include <cstring>
#include <string>
#include <vector>

#include "mrisim.h"
#include "mrisim_math.h"
#include "scene_internal.h"

namespace {

constexpr int kThreads = 128;
constexpr int kChunk = 32;          // attempts per recorded generator state

template <typename T>
struct Dev {
    T* p = nullptr; size_t n = 0;
    Dev() = default;
    Dev(const Dev&) = delete; Dev& operator=(const Dev&) = delete;
    ~Dev() { if (p) cudaFree(p); }
    cudaError_t alloc(size_t count) {
        if (p) cudaFree(p);
        p = nullptr; n = count;
        return cudaMalloc((void**)&p, std::max<size_t>(count, 1) * sizeof(T));
    }
    cudaError_t upload(const T* h, size_t count) {
        cudaError_t e = alloc(count);
        if (e != cudaSuccess || !count) return e;
        return cudaMemcpy(p, h, count * sizeof(T), cudaMemcpyHostToDevice);
    }
};

inline uint32_t blocks_for(uint64_t n) { return (uint32_t)((n + kThreads - 1) / kThreads); }

__device__ __forceinline__ double normsq(double x, double y, double z) { return x * x + y * y + z * z; }

__global__ void __launch_bounds__(kThreads)
lookup_kernel(const double4* __restrict__ cells, int num_cells, int hashDim, int maxNeighbors, double cubeLength,
              double reach, int64_t h3, int32_t* __restrict__ lookup, int32_t* __restrict__ counts) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= h3) return;
    const double x = (i % hashDim) * cubeLength;
    const double y = ((i / hashDim) % (hashDim)) * cubeLength;
    const double z = i / (hashDim * hashDim) * cubeLength;
    int cnt = 0;
    for (int j = 0; j < num_cells; j++) {
        const double4 c = cells[j];
        const double dx = x - c.x, dy = y - c.y, dz = z - c.z;
        if (sqrt(normsq(dx, dy, dz)) < reach) {
            if (cnt < maxNeighbors) lookup[i * maxNeighbors + cnt] = j;
            cnt++;
        }
    }
    counts[i] = cnt;
}

__global__ void __launch_bounds__(kThreads)
candidate_kernel(const mrisim_rng* __restrict__ chunk_state, int64_t n_chunks, int64_t batch, double offset, double span,
                 double* __restrict__ cx, double* __restrict__ cy, double* __restrict__ cz) {
    const int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= n_chunks) return;
    mrisim_rng r = chunk_state[c];
    const int64_t j0 = c * kChunk, j1 = min(batch, j0 + kChunk);
    for (int64_t j = j0; j < j1; j++) {
        cx[j] = offset + mrisim_rng_u01(&r) * span;
        cy[j] = offset + mrisim_rng_u01(&r) * span;
        cz[j] = offset + mrisim_rng_u01(&r) * span;
    }
}

__global__ void __launch_bounds__(kThreads)
contain_kernel(const double4* __restrict__ cells, int num_cells, double r2, const double* __restrict__ cx,
               const double* __restrict__ cy, const double* __restrict__ cz, int64_t batch, uint32_t* __restrict__ ok) {
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= batch) return;
    const double x = cx[j], y = cy[j], z = cz[j];
    bool in = false;
    for (int i = 0; i < num_cells; i++) {
        const double4 c = cells[i];
        const double dx = x - c.x, dy = y - c.y, dz = z - c.z;
        if (normsq(dx, dy, dz) < r2) in = true;
    }
    ok[j] = in ? 0u : 1u;
}

__global__ void fill_kernel(uint32_t* __restrict__ a, int64_t n, uint32_t v) {
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j < n) a[j] = v;
}

// accepted attempt with rank r < want becomes water have + r; the attempt that fills the last
// slot is reported so that the host can rewind its generator to just after it
__global__ void compact_kernel(const double* __restrict__ cx, const double* __restrict__ cy, const double* __restrict__ cz,
                               const uint32_t* __restrict__ ok, const uint32_t* __restrict__ rank, int64_t batch,
                               int64_t want, mrisim_water* __restrict__ out, long long* __restrict__ last_attempt) {
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= batch || !ok[j]) return;
    const int64_t r = rank[j];
    if (r >= want) return;
    mrisim_water w;
    w.x = cx[j]; w.y = cy[j]; w.z = cz[j]; w.phase = 0;
    w.in_cell = 0; w._pad[0] = w._pad[1] = w._pad[2] = 0; w.nearest = -1;
    out[r] = w;
    if (r == want - 1) *last_attempt = j;
}

__device__ __forceinline__ uint64_t spread3(uint32_t v) {
    uint64_t x = v & 0x1FFFFF;
    x = (x | x << 32) & 0x1F00000000FFFFULL;
    x = (x | x << 16) & 0x1F0000FF0000FFULL;
    x = (x | x << 8)  & 0x100F00F00F00F00FULL;
    x = (x | x << 4)  & 0x10C30C30C30C30C3ULL;
    x = (x | x << 2)  & 0x1249249249249249ULL;
    return x;
}
// octree.cpp:37-72 (host flavour): floor(x / bound * 2^depth) per axis, interleaved
__device__ __forceinline__ uint64_t morton(int depth, double x, double y, double z, double bound) {
    const double size = (double)(1ull << depth);
    const uint32_t ix = (uint32_t)floor(x / bound * size);
    const uint32_t iy = (uint32_t)floor(y / bound * size);
    const uint32_t iz = (uint32_t)floor(z / bound * size);
    const uint32_t mask = depth > 16 ? 0xFFFFFFu : (depth > 8 ? 0xFFFFu : 0xFFu);
    return (spread3(ix & mask) << 2) | (spread3(iy & mask) << 1) | spread3(iz & mask);
}

__global__ void __launch_bounds__(kThreads)
leaf_key_kernel(const mrisim_water* __restrict__ waters, int64_t n, const mrisim_node* __restrict__ nodes,
                const uint64_t* __restrict__ prefix, int min_depth, double bound, uint64_t* __restrict__ keys) {
    const int64_t w = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= n) return;
    const double x = waters[w].x, y = waters[w].y, z = waters[w].z;
    uint64_t node = prefix[morton(min_depth, x, y, z, bound)];
    int d = min_depth;
    uint64_t mc;
    while (!((mc = nodes[node].mc) >> 63)) {
        const unsigned c = (unsigned)(morton(d + 1, x, y, z, bound) & 7);
        node += nodes[node].child[c];
        d++;
    }
    keys[w] = mc;
}

}  // namespace

namespace mrisim_host {

#define SG_CK(call)                                                                          \
    do {                                                                                     \
        cudaError_t e_ = (call);                                                             \
        if (e_ != cudaSuccess) {                                                             \
            err = std::string(#call) + ": " + cudaGetErrorString(e_) + " (no CPU fallback)"; \
            return 1;                                                                        \
        }                                                                                    \
    } while (0)

static int use_device(int device, std::string& err) {
    int n_dev = 0;
    if (cudaGetDeviceCount(&n_dev) != cudaSuccess || device < 0 || device >= n_dev) {
        err = "no CUDA device " + std::to_string(device) + " (no CPU fallback)";
        return 1;
    }
    SG_CK(cudaSetDevice(device));
    return 0;
}

static std::vector<double4> pad_cells(const std::vector<double>& cells) {
    std::vector<double4> c4(cells.size() / 3);
    for (size_t i = 0; i < c4.size(); i++) c4[i] = make_double4(cells[3 * i], cells[3 * i + 1], cells[3 * i + 2], 0.0);
    return c4;
}

int gpu_build_lookup(const mrisim_params& p, const std::vector<double>& cells, int device,
                     std::vector<int32_t>& lookup, int* overflow, std::string& err) {
    if (use_device(device, err)) return 1;
    const double cubeLength = p.bound / p.hashDim;
    const double diagonal = std::sqrt(3) * cubeLength;
    const int64_t h3 = (int64_t)p.hashDim * p.hashDim * p.hashDim;
    const std::vector<double4> c4 = pad_cells(cells);
    Dev<double4> d_cells; Dev<int32_t> d_lookup, d_counts;
    SG_CK(d_cells.upload(c4.data(), c4.size()));
    SG_CK(d_lookup.alloc((size_t)h3 * p.maxNeighbors)); SG_CK(d_counts.alloc((size_t)h3));
    SG_CK(cudaMemset(d_lookup.p, 0xFF, (size_t)h3 * p.maxNeighbors * sizeof(int32_t)));      // -1 padding
    lookup_kernel<<<blocks_for((uint64_t)h3), kThreads>>>(d_cells.p, p.num_cells, p.hashDim, p.maxNeighbors, cubeLength,
                                                         p.cell_r + diagonal, h3, d_lookup.p, d_counts.p);
    SG_CK(cudaGetLastError());
    lookup.resize((size_t)h3 * p.maxNeighbors);
    std::vector<int32_t> counts((size_t)h3);
    SG_CK(cudaMemcpy(lookup.data(), d_lookup.p, lookup.size() * sizeof(int32_t), cudaMemcpyDeviceToHost));
    SG_CK(cudaMemcpy(counts.data(), d_counts.p, counts.size() * sizeof(int32_t), cudaMemcpyDeviceToHost));
    *overflow = 0;
    for (int32_t c : counts) if (c >= p.maxNeighbors) *overflow = 1;     // SimulationBox.cpp:281
    return 0;
}

int gpu_throw_waters(const mrisim_params& p, const std::vector<double>& cells, mrisim_rng* gen, int64_t num_water,
                     int device, std::vector<mrisim_water>& waters, std::string& err) {
    if (use_device(device, err)) return 1;
    const double offset = (p.bound - p.water_start_bound) / 2.0;
    const double r2 = std::pow(p.cell_r, 2);                      // SimulationBox.cpp:165
    const std::vector<double4> c4 = pad_cells(cells);
    Dev<double4> d_cells;
    SG_CK(d_cells.upload(c4.data(), c4.size()));
    Dev<mrisim_water> d_out;
    SG_CK(d_out.alloc((size_t)num_water));
    Dev<mrisim_rng> d_state; Dev<double> cx, cy, cz; Dev<uint32_t> ok, rank; Dev<uint8_t> tmp; Dev<long long> d_last;
    SG_CK(d_last.alloc(1));
    int64_t have = 0;
    std::vector<mrisim_rng> state;
    while (have < num_water) {
        const int64_t want = num_water - have;
        const int64_t batch = p.avoid_intracellular_throw ? want + want / 8 + 64 : want;
        if (batch > 0x7fffffffll) { err = "water batch larger than 2^31 attempts"; return 1; }
        const int64_t n_chunks = (batch + kChunk - 1) / kChunk;
        // the host walks the stream once (3 draws per attempt) and records the state at every chunk start
        state.resize((size_t)n_chunks + 1);
        mrisim_rng r = *gen;
        for (int64_t c = 0; c < n_chunks; c++) {
            state[(size_t)c] = r;
            const int64_t draws = 3 * std::min<int64_t>(kChunk, batch - c * kChunk);
            for (int64_t k = 0; k < draws; k++) mrisim_rng_next(&r);
        }
        state[(size_t)n_chunks] = r;
        SG_CK(d_state.upload(state.data(), (size_t)n_chunks));
        SG_CK(cx.alloc((size_t)batch)); SG_CK(cy.alloc((size_t)batch)); SG_CK(cz.alloc((size_t)batch));
        SG_CK(ok.alloc((size_t)batch)); SG_CK(rank.alloc((size_t)batch));
        candidate_kernel<<<blocks_for((uint64_t)n_chunks), kThreads>>>(d_state.p, n_chunks, batch, offset, p.water_start_bound,
                                                                      cx.p, cy.p, cz.p);
        if (p.avoid_intracellular_throw)
            contain_kernel<<<blocks_for((uint64_t)batch), kThreads>>>(d_cells.p, p.num_cells, r2, cx.p, cy.p, cz.p, batch, ok.p);
        else
            fill_kernel<<<blocks_for((uint64_t)batch), kThreads>>>(ok.p, batch, 1u);
        size_t bytes = 0;
        SG_CK(cub::DeviceScan::ExclusiveSum(nullptr, bytes, ok.p, rank.p, (int)batch));
        SG_CK(tmp.alloc(bytes));
        SG_CK(cub::DeviceScan::ExclusiveSum(tmp.p, bytes, ok.p, rank.p, (int)batch));
        const long long none = -1;
        SG_CK(cudaMemcpy(d_last.p, &none, sizeof none, cudaMemcpyHostToDevice));
        compact_kernel<<<blocks_for((uint64_t)batch), kThreads>>>(cx.p, cy.p, cz.p, ok.p, rank.p, batch, want, d_out.p + have,
                                                                 d_last.p);
        SG_CK(cudaGetLastError());
        long long last = -1; uint32_t last_rank = 0, last_ok = 0;
        SG_CK(cudaMemcpy(&last, d_last.p, sizeof last, cudaMemcpyDeviceToHost));
        SG_CK(cudaMemcpy(&last_rank, rank.p + batch - 1, 4, cudaMemcpyDeviceToHost));
        SG_CK(cudaMemcpy(&last_ok, ok.p + batch - 1, 4, cudaMemcpyDeviceToHost));
        const int64_t accepted = (int64_t)last_rank + last_ok;
        if (accepted >= want) {
            // rewind to just after the attempt that filled the last slot (BacteriaBox.cpp:80-105 would stop there)
            const int64_t j = last + 1;
            mrisim_rng q = state[(size_t)(j / kChunk)];
            for (int64_t k = 0; k < 3 * (j % kChunk); k++) mrisim_rng_next(&q);
            *gen = q;
            have = num_water;
        } else {
            *gen = r;
            have += accepted;
        }
    }
    waters.resize((size_t)num_water);
    if (num_water) SG_CK(cudaMemcpy(waters.data(), d_out.p, (size_t)num_water * sizeof(mrisim_water), cudaMemcpyDeviceToHost));
    return 0;
}

int gpu_leaf_keys(const mrisim_params& p, int min_depth, const std::vector<uint64_t>& prefix,
                  const std::vector<mrisim_node>& nodes, const std::vector<mrisim_water>& waters, int device,
                  std::vector<uint64_t>& keys, std::string& err) {
    if (use_device(device, err)) return 1;
    const int64_t n = (int64_t)waters.size();
    Dev<mrisim_water> d_w; Dev<mrisim_node> d_nodes; Dev<uint64_t> d_prefix, d_keys;
    SG_CK(d_w.upload(waters.data(), waters.size()));
    SG_CK(d_nodes.upload(nodes.data(), nodes.size()));
    SG_CK(d_prefix.upload(prefix.data(), prefix.size()));
    SG_CK(d_keys.alloc((size_t)n));
    if (n) leaf_key_kernel<<<blocks_for((uint64_t)n), kThreads>>>(d_w.p, n, d_nodes.p, d_prefix.p, min_depth, p.bound, d_keys.p);
    SG_CK(cudaGetLastError());
    keys.resize((size_t)n);
    if (n) SG_CK(cudaMemcpy(keys.data(), d_keys.p, (size_t)n * sizeof(uint64_t), cudaMemcpyDeviceToHost));
    return 0;
}

}  // namespace mrisim_host
