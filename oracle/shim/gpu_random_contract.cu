// TEST INFRASTRUCTURE -- a deterministic stand-in for the reference's cuRAND wrapper ONLY
// (src/gpu_random.cu:18-72; class declared in src/gpu_random.h:23-34, which is used unchanged).
//
// Purpose: pin the walk bit for bit against the reference's OWN device code.  The reference's
// kernels (simulateDiffusion, computePhaseAccumulation, performUpdate, flipPhases, cuda_sim.cu,
// compiled where they lie) consume random numbers from two bulk buffers that gpu_rng fills per
// sprint: 5*S*N uniforms partitioned [x|y|z|coin|kick][step][walker] (cuda_sim.cu:938-944,986)
// and S*N normals [step][walker] (:946,977).  cuRAND XORWOW seeded with time(NULL) makes that
// irreproducible; here the same buffers are filled from the WALKER STREAM CONTRACT
// (include/mrisim_math.h): walker w owns XORShift(walker_seed + w) (the reference's own
// src/xorshift.h generator) and draws, per step, ux uy uz [u1 u2 at even steps] coin kick.
// Nothing else of the reference is replaced: same kernels, same launch sequence, same cuBLAS
// dgemv / thrust::reduce, same CSV.
//
// Besides filling, the shim can dump what the reference's kernels left in the buffers:
//   REF_TRACE_PREFIX=<path>   after each of the first REF_TRACE_SPRINTS (default 2) sprints:
//       <path>_s<k>.bin = int64 {magic, N, S, k}, then x, y, z [S][N] doubles (the positions
//       simulateDiffusion wrote over the consumed uniforms, cuda_sim.cu:725-727), kick [S][N]
//       doubles (computePhaseAccumulation's output over the coins, :983-988), in_cell [S][N] bytes;
//       and after the sprint loop <path>_final.bin = int64 {magic, N, 0, 0} + water_info[N].
//   REF_WALKER_SEED=<u64>     walker_seed of the contract (default 0xC0FFEE).
// The device pointers of in_cell and waters come from the utilities shim's allocation notes
// (shim/utilities/utilities.h); the end of the sprint loop is signalled by its Timer::cpuStop().
//
// Self-check: the uniforms of walkers 0 and N-1 of the first sprint are regenerated on the host
// with the reference's own XORShift<uint64_t> class and compared bit for bit (so the filler's
// xoroshiro/splitmix arithmetic is pinned to src/xorshift.h independently of mrisim_math.h).
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "mrisim_math.h"
#include "gpu_random.h"
#include "xorshift.h"
#include "rand_walk.h"
#include "parameters.h"

namespace {

#define SHIM_CK(call)                                                                          \
    do {                                                                                       \
        cudaError_t e_ = (call);                                                               \
        if (e_ != cudaSuccess) {                                                               \
            std::fprintf(stderr, "gpu_random_contract: %s: %s\n", #call, cudaGetErrorString(e_)); \
            std::exit(1);                                                                      \
        }                                                                                      \
    } while (0)

struct Note { void* p; size_t n, elem; };
std::vector<Note> g_notes;

struct State {
    int N = 0, S = 0;
    uint64_t seed = 0xC0FFEEull;
    uint64_t* s0 = nullptr; uint64_t* s1 = nullptr; double* spare = nullptr; double* normals = nullptr;
    long long steps_done = 0;
    int sprints_filled = 0;
    double* uniform_buf = nullptr;          // the caller's 5*S*N buffer of the last fill
    std::string prefix;
    int trace_sprints = 2;
    bool dumped_last = false;
} g;

__global__ void seed_kernel(uint64_t* s0, uint64_t* s1, double* spare, int n, uint64_t base) {
    const int w = blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= n) return;
    mrisim_rng r;
    mrisim_rng_seed(&r, base + (uint64_t)w);
    s0[w] = r.s0; s1[w] = r.s1; spare[w] = 0.0;
}

// one thread per walker, S steps of the contract's draw order
__global__ void fill_kernel(uint64_t* s0, uint64_t* s1, double* spare, int n, int S, int parity0,
                            double* x, double* y, double* z, double* coin, double* kick, double* normal) {
    const int w = blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= n) return;
    mrisim_rng r; r.s0 = s0[w]; r.s1 = s1[w];
    double sp = spare[w];
    for (int s = 0; s < S; s++) {
        const size_t o = (size_t)s * n + w;
        x[o] = mrisim_rng_u01(&r);
        y[o] = mrisim_rng_u01(&r);
        z[o] = mrisim_rng_u01(&r);
        double nrm;
        if (((parity0 + s) & 1) == 0) {
            const double u1 = mrisim_rng_u01(&r);
            const double u2 = mrisim_rng_u01(&r);
            mrisim_normal_pair(u1, u2, &nrm, &sp);
        } else {
            nrm = sp;
        }
        normal[o] = nrm;
        coin[o] = mrisim_rng_u01(&r);
        kick[o] = mrisim_rng_u01(&r);
    }
    s0[w] = r.s0; s1[w] = r.s1; spare[w] = sp;
}

void init_state() {
    if (g.s0) return;
    g.N = p.num_water; g.S = p.sprintSteps;
    if (const char* e = std::getenv("REF_WALKER_SEED")) g.seed = std::strtoull(e, nullptr, 0);
    if (const char* e = std::getenv("REF_TRACE_PREFIX")) g.prefix = e;
    if (const char* e = std::getenv("REF_TRACE_SPRINTS")) g.trace_sprints = std::atoi(e);
    SHIM_CK(cudaMalloc((void**)&g.s0, sizeof(uint64_t) * g.N));
    SHIM_CK(cudaMalloc((void**)&g.s1, sizeof(uint64_t) * g.N));
    SHIM_CK(cudaMalloc((void**)&g.spare, sizeof(double) * g.N));
    SHIM_CK(cudaMalloc((void**)&g.normals, sizeof(double) * (size_t)g.N * g.S));
    seed_kernel<<<(g.N + 127) / 128, 128>>>(g.s0, g.s1, g.spare, g.N, g.seed);
    SHIM_CK(cudaGetLastError());
}

const Note* find_note(size_t n, size_t elem) {
    for (const Note& t : g_notes) if (t.n == n && t.elem == elem) return &t;
    return nullptr;
}

void write_all(std::FILE* f, const void* p_, size_t bytes) {
    if (std::fwrite(p_, 1, bytes, f) != bytes) { std::fprintf(stderr, "gpu_random_contract: short write\n"); std::exit(1); }
}

// what the reference's kernels left behind for sprint k (call after they ran, before the next fill)
void dump_sprint(int k) {
    const size_t SN = (size_t)g.N * g.S;
    std::vector<double> h(SN);
    const std::string path = g.prefix + "_s" + std::to_string(k) + ".bin";
    std::FILE* f = std::fopen(path.c_str(), "wb");
    if (!f) { std::fprintf(stderr, "gpu_random_contract: cannot open %s\n", path.c_str()); std::exit(1); }
    const int64_t hdr[4] = {0x4d52495452414345ll, g.N, g.S, k};
    write_all(f, hdr, sizeof hdr);
    for (int c = 0; c < 4; c++) {                         // x, y, z, then the coins region = the kicks
        SHIM_CK(cudaMemcpy(h.data(), g.uniform_buf + (size_t)c * SN, SN * sizeof(double), cudaMemcpyDeviceToHost));
        write_all(f, h.data(), SN * sizeof(double));
    }
    const Note* ic = find_note(SN, sizeof(bool));
    if (!ic) { std::fprintf(stderr, "gpu_random_contract: in_cell array not found among the allocation notes\n"); std::exit(1); }
    std::vector<unsigned char> hb(SN);
    SHIM_CK(cudaMemcpy(hb.data(), ic->p, SN, cudaMemcpyDeviceToHost));
    write_all(f, hb.data(), SN);
    std::fclose(f);
}

void dump_final() {
    const Note* w = find_note((size_t)g.N, sizeof(water_info));
    if (!w) { std::fprintf(stderr, "gpu_random_contract: waters array not found among the allocation notes\n"); std::exit(1); }
    std::vector<water_info> h(g.N);
    SHIM_CK(cudaMemcpy(h.data(), w->p, sizeof(water_info) * g.N, cudaMemcpyDeviceToHost));
    const std::string path = g.prefix + "_final.bin";
    std::FILE* f = std::fopen(path.c_str(), "wb");
    if (!f) { std::fprintf(stderr, "gpu_random_contract: cannot open %s\n", path.c_str()); std::exit(1); }
    const int64_t hdr[4] = {0x4d52495746494e41ll, g.N, (int64_t)sizeof(water_info), g.steps_done};
    write_all(f, hdr, sizeof hdr);
    write_all(f, h.data(), sizeof(water_info) * g.N);
    std::fclose(f);
}

// first-sprint uniforms of two walkers against the reference's own generator class
void self_check(const double* dev) {
    const size_t SN = (size_t)g.N * g.S;
    const int steps = g.S < 64 ? g.S : 64;
    const int walkers[2] = {0, g.N - 1};
    for (int wi = 0; wi < 2; wi++) {
        const int w = walkers[wi];
        XORShift<uint64_t> ref(g.seed + (uint64_t)w);
        for (int s = 0; s < steps; s++) {
            double want[5];
            want[0] = ref.rand_pos_double(); want[1] = ref.rand_pos_double(); want[2] = ref.rand_pos_double();
            if ((s & 1) == 0) { ref.rand_pos_double(); ref.rand_pos_double(); }
            want[3] = ref.rand_pos_double(); want[4] = ref.rand_pos_double();
            for (int c = 0; c < 5; c++) {
                double got;
                SHIM_CK(cudaMemcpy(&got, dev + (size_t)c * SN + (size_t)s * g.N + w, sizeof got, cudaMemcpyDeviceToHost));
                if (std::memcmp(&got, &want[c], sizeof got) != 0) {
                    std::fprintf(stderr, "gpu_random_contract: self-check failed (walker %d step %d channel %d: %.17g vs xorshift.h %.17g)\n",
                                 w, s, c, got, want[c]);
                    std::exit(1);
                }
            }
        }
    }
    std::printf("gpu_random_contract: self-check against src/xorshift.h ok (walker_seed %llu)\n", (unsigned long long)g.seed);
}

}  // namespace

// hooks called by shim/utilities/utilities.h (weak there)
extern "C" void ref_shim_note_array(void* ptr, size_t n, size_t elem) { g_notes.push_back(Note{ptr, n, elem}); }
extern "C" void ref_shim_sprint_loop_done() {
    if (g.prefix.empty() || !g.s0) return;
    const int last = g.sprints_filled - 1;
    if (last >= 0 && last < g.trace_sprints && !g.dumped_last) { dump_sprint(last); g.dumped_last = true; }
    dump_final();
}

gpu_rng::gpu_rng() { gen = nullptr; }
gpu_rng::~gpu_rng() {}

// gpu_random.cu:48-51.  n = 5*S*N (cuda_sim.cu:929,976); layout [x|y|z|coin|kick][step][walker].
void gpu_rng::getUniformDoubles(size_t n, double* devData) {
    init_state();
    const size_t SN = (size_t)g.N * g.S;
    if (n != 5 * SN) { std::fprintf(stderr, "gpu_random_contract: unexpected uniform count %zu (want 5*%d*%d)\n", n, g.S, g.N); std::exit(1); }
    // the previous sprint's kernels are done with the buffers once this (legacy-stream) copy runs
    if (!g.prefix.empty() && g.sprints_filled >= 1 && g.sprints_filled - 1 < g.trace_sprints) dump_sprint(g.sprints_filled - 1);
    g.uniform_buf = devData;
    fill_kernel<<<(g.N + 127) / 128, 128>>>(g.s0, g.s1, g.spare, g.N, g.S, (int)(g.steps_done & 1),
                                             devData, devData + SN, devData + 2 * SN, devData + 3 * SN, devData + 4 * SN, g.normals);
    SHIM_CK(cudaGetLastError());
    if (g.sprints_filled == 0) self_check(devData);
    g.steps_done += g.S;
    g.sprints_filled += 1;
    g.dumped_last = false;
}

// gpu_random.cu:69-72.  n = S*N (cuda_sim.cu:930,977); the normals of the sprint just filled.
void gpu_rng::getNormalDoubles(size_t n, double* devData) {
    const size_t SN = (size_t)g.N * g.S;
    if (!g.s0 || n != SN) { std::fprintf(stderr, "gpu_random_contract: getNormalDoubles out of order or unexpected count %zu\n", n); std::exit(1); }
    SHIM_CK(cudaMemcpy(devData, g.normals, SN * sizeof(double), cudaMemcpyDeviceToDevice));
}
