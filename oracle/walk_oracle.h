/*
 * TEST INFRASTRUCTURE -- the parity oracle.  NOT product code: only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load it.
 *
 * walk_oracle: a CPU restatement, in plain C, of the device half of the reference's hot
 * path (src/cuda_sim.cu:107-286, 453-640, 688-838, 974-1024), which exists in the
 * reference only as CUDA.  Each function in walk_oracle.c cites the
 * reference lines it follows, keeps their operation order and their quirks (A.5), walks
 * the octree the reference's way (pointer forest, relative child offsets, Morton LUTs,
 * one Morton evaluation per level) and uses libm fmod/sqrt/cos.  The only thing it takes
 * from the product is the walker stream contract (include/mrisim_math.h) that replaces the
 * irreproducible cuRAND stream.
 *
 * Pinning.  The reference ships no golden vectors for this path, so the pin is the
 * reference ITSELF, run on the GPU box: oracle/_ref/ref_gpu_pin_<cfg> is the reference's unchanged
 * cuda_sim.cu (its kernels, cuBLAS dgemv, thrust::reduce, CSV line) compiled where it lies with
 * -fmad=false and linked with oracle/shim/gpu_random_contract.cu, which replaces only the 72-line
 * cuRAND wrapper (src/gpu_random.cu) by a filler that writes the walker stream contract into the
 * reference's own random buffers.  tests/test_ref_gpu_bitpin.py asserts, for three configurations,
 * that per-(step, walker) positions, membrane flags and phase kicks of this oracle, of the product
 * and of the reference's kernels are BIT-EQUAL, that the final walker state after 10^6 steps is
 * bit-equal (phases to 1e-9 rad: cuBLAS owns the dgemv summation order) and that the CSV rows print
 * identically (profiles/pytest_gpu_r02_bitpin_vs_reference_kernels.txt).  The scene arrays the walk
 * consumes come from the reference's own host code (oracle/_ref, SURVEY.md A.6 known answers), and
 * the host-side traversal below is also checked against the reference's Octree::get_voxel /
 * get_field outputs (tests/golden/probes_*.json).
 *
 * What remains a CONTRACT rather than a pin: the reference draws from cuRAND XORWOW seeded with
 * time(NULL) and has no reproducible stream; the contract (include/mrisim_math.h) substitutes the
 * reference's own src/xorshift.h generator per walker and a Box-Muller convention of its own
 * (pair drawn at even steps, v = 1 - u1, z0 = r*cos used first).  The xoroshiro/splitmix words are
 * pinned to src/xorshift.h (A.6 vectors; the shim's self-check against the XORShift class); the
 * Box-Muller convention is pinned by nothing but itself -- it is the definition of the stream.
 */
#ifndef WALK_ORACLE_H
#define WALK_ORACLE_H

#include <stdint.h>
#include "mrisim.h"

#ifdef __cplusplus
extern "C" {
#endif

/* scene exactly as SimulationBox hands it to the driver (interchange layout) */
typedef struct oracle_scene {
    const double*      cells;         /* Triple[num_cells]                        */
    int32_t            num_cells;
    const int32_t*     lookup;        /* [hashDim^3][maxNeighbors], -1 terminated */
    int32_t            min_depth, max_depth;
    const uint32_t*    subtree_sizes; /* [8^min_depth]                            */
    int32_t            n_subtrees;
    const mrisim_node* nodes;         /* all subtrees back to back                */
    uint64_t           n_nodes;
    const mrisim_mnp*  residents;
    uint64_t           n_res;
} oracle_scene;

typedef struct oracle_trace {         /* all optional, step-major [steps][n]      */
    double*   x; double* y; double* z;
    uint8_t*  in_cell;
    uint32_t* leaf;                   /* flat node index = prefix[i] + j          */
    double*   kick;                   /* per-step phase increment                 */
} oracle_trace;

/* Seed the walker streams: walker i gets XORShift(walker_seed + first_walker_id + i). */
void oracle_seed_walkers(mrisim_rng_state* rng, uint64_t n, uint64_t walker_seed, uint64_t first_walker_id);

/* One sprint (cuda_sim.cu:974-1024 without the CSV line).  waters/rng are in-out;
 * time0 = steps simulated before this sprint; signal_out gets steps/pfreq samples.
 * n_threads splits WALKERS over pthreads; results do not depend on it. */
int oracle_run_sprint(const mrisim_params* p, const oracle_scene* sc,
                      mrisim_water* waters, mrisim_rng_state* rng, uint64_t n,
                      int64_t time0, int32_t steps, double* signal_out,
                      const oracle_trace* trace, int n_threads);

/* single-point probes of the restated traversal (checked against the reference's
 * Octree::get_voxel / get_field): returns flat leaf index, writes device-semantics field */
uint32_t oracle_probe_leaf(const mrisim_params* p, const oracle_scene* sc, double x, double y, double z, uint64_t* mc_out);
double   oracle_probe_field(const mrisim_params* p, const oracle_scene* sc, double x, double y, double z);
/* updateNearest at a point: returns nearest index, writes in_cell */
int32_t  oracle_probe_nearest(const mrisim_params* p, const oracle_scene* sc, double x, double y, double z, int* in_cell);
/* the kick expression of cuda_sim.cu:632 evaluated verbatim */
double   oracle_kick_expression(const mrisim_params* p, double nD, int in_cell);
/* first `n` raw words / rand_pos_double of XORShift(seed) (xorshift.h) */
void     oracle_xorshift_words(uint64_t seed, uint64_t* out, int n);
void     oracle_xorshift_doubles(uint64_t seed, double* out, int n);
/* contract normals: n values of the Box-Muller stream fed by XORShift(seed) u1,u2,u1,u2,... */
void     oracle_normals(uint64_t seed, double* out, int n);
void     oracle_log_sincos(double v, double u, double* lg, double* sn, double* cs);

#ifdef __cplusplus
}
#endif
#endif
