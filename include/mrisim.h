/*
 * mrisim.h -- C ABI of the B200-native walker/T2-signal path.
 *
 * The reference (RabinKarp/MRI_Nanoparticle_Simulation) has no FFI of its own; its only
 * seam is C++: `void simulateWaters(std::string)` (src/cuda_sim.h:74), which consumes the
 * accessors of SimulationBox (src/SimulationBox.cpp:536-594) and compile-time knobs from
 * src/parameters.h.  This header is that seam restated as plain C: every entry point below
 * names the reference interface it replaces.  No C++/torch types cross the boundary; all
 * arrays are caller-owned and copied by the callee unless stated otherwise.
 *
 * Return convention: 0 = ok, non-zero = error (text via mrisim_last_error()).  The
 * reference aborts through HANDLE_ERROR on the one CUDA call it checks
 * (src/cuda_sim.cu:407) and ignores every other status; here every call is checked.
 *
 * There is NO CPU fallback behind any device entry point: without a CUDA device (or with
 * a failing kernel) they return an error.
 */
#ifndef MRISIM_H
#define MRISIM_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MRISIM_VERSION 100  /* 0.1.0 */

/* -------------------------------------------------------------------------------------
 * Parameters: run-time mirror of `struct ParameterStruct` (src/parameters.h:25-126) plus
 * the three file-scope constants of src/cuda_sim.cu:44-45 (pInt, pfreq).  Field names are
 * the reference's.  Derived fields (t, tcp, pfreq, in_stdev, out_stdev) are filled by
 * mrisim_params_derive() with the reference's own expressions.
 * ----------------------------------------------------------------------------------- */
enum { MRISIM_KICK_NONE = 0, MRISIM_KICK_RANDOM = 1, MRISIM_KICK_CONSTANT = 2 };

typedef struct mrisim_params {
    double  g;                  /* parameters.h:27  gyromagnetic ratio                  */
    double  scale;              /* :30  explicit-field radius in units of cell_r        */
    double  border;             /* :33  periodic-image cutoff (um)                      */
    int32_t num_threads;        /* :36  host threads for the octree build               */
    int32_t sprintSteps;        /* :42  steps per sprint (signal flush / all-reduce)    */
    int32_t num_water;          /* :46                                                  */
    int32_t num_cells;          /* :49                                                  */
    double  cell_r;             /* :50                                                  */
    double  mmoment;            /* :52                                                  */
    int32_t kick_mode;          /* :55-56  RANDOM_KICK / CONSTANT_KICK                  */
    int32_t avoid_intracellular_throw; /* :77  #define AVOID_INTRACELLULAR_THROW        */
    double  phase_stdev;        /* :61                                                  */
    double  phase_k;            /* :59 / :62                                            */
    double  bound;              /* :66                                                  */
    double  water_start_bound;  /* :71                                                  */
    int32_t hashDim;            /* :80                                                  */
    int32_t maxNeighbors;       /* :81                                                  */
    double  D_cell;             /* :84                                                  */
    double  D_extra;            /* :85                                                  */
    double  P_expr;             /* :86                                                  */
    double  tau;                /* :88                                                  */
    int32_t totaltime;          /* :91                                                  */
    int32_t t;                  /* :93  (int)(totaltime/tau)                            */
    double  taucp;              /* :94                                                  */
    int32_t tcp;                /* :95  (int)(taucp/tau)                                */
    int32_t pfreq;              /* cuda_sim.cu:44-45  (int)(1e-3/tau)                   */
    double  reflectIO;          /* :100,116                                             */
    double  reflectOI;          /* :101,117                                             */
    double  in_stdev;           /* :104,120  sqrt(pi*D_cell*tau)                        */
    double  out_stdev;          /* :105,121  sqrt(pi*D_extra*tau)                       */
} mrisim_params;

/* shipped src/parameters.h values (num_water 4032, sprintSteps 20000, 270 cells, ...) */
void mrisim_params_default(mrisim_params* p);
/* recompute t, tcp, pfreq, in_stdev, out_stdev from the primaries (parameters.h:93,95,120-121) */
void mrisim_params_derive(mrisim_params* p);
/* the commented low-permeability formulas of parameters.h:116-117 (BASELINE config 3) */
void mrisim_params_low_permeability(mrisim_params* p);

/* -------------------------------------------------------------------------------------
 * POD records.  Layout-compatible with the reference structs so that
 * SimulationBox::getWaters()/getMNPs()/getCells() pointers can be handed over directly.
 * ----------------------------------------------------------------------------------- */
typedef struct mrisim_water {   /* water_info, src/rand_walk.h:25-53 (40 bytes)          */
    double  x, y, z;
    double  phase;
    uint8_t in_cell;            /* bool */
    uint8_t _pad[3];
    int32_t nearest;
} mrisim_water;

typedef struct mrisim_mnp {     /* MNP_info, src/rand_walk.h:58-69 (40 bytes)            */
    double x, y, z, r, M;
} mrisim_mnp;

typedef struct mrisim_node {    /* oct_node, src/octree.h:28-33, resident vector flattened */
    uint64_t mc;                /* Morton code; bit 63 = leaf; bit 3*depth = depth marker */
    uint32_t child[8];          /* internal: offset RELATIVE to this node (octree.cpp:320);
                                   leaf: IEEE-754 bits of the 8 float B samples           */
    uint32_t res_start;         /* first resident in the residents array                 */
    uint32_t n_res;             /* resident count (0 for internal nodes)                 */
} mrisim_node;

/* per-walker RNG state of the walker stream contract (see mrisim_math.h) */
typedef struct mrisim_rng_state {
    uint64_t s0, s1;            /* xoroshiro128+ state (src/xorshift.h:129-141)          */
    double   spare;             /* second Box-Muller normal, consumed at odd steps       */
} mrisim_rng_state;

/* -------------------------------------------------------------------------------------
 * Host scene: replaces `BacteriaBox simBox(&gen); simBox.populateSimulation();`
 * (src/cuda_sim.cu:920-921 -> SimulationBox.cpp:77-97, BacteriaBox.cpp:49-121,
 * octree.cpp:388-450).  host_seed replaces `time(NULL) + rd()` (cuda_sim.cu:911-912).
 * n_threads <= 0 means "all hardware threads"; the result does not depend on it.
 * ----------------------------------------------------------------------------------- */
typedef struct mrisim_scene mrisim_scene;

int  mrisim_scene_build(const mrisim_params* p, uint64_t host_seed, int n_threads,
                        mrisim_scene** out);
/* Same scene, octree built on GPU `device` (replaces the 16 std::threads of
 * Octree::Octree, octree.cpp:410-414, and build_subtree/grad/field, :215-333).  The result is
 * byte-identical to mrisim_scene_build(): the leaf criterion and every float B sample are computed
 * on the GPU with a rigorous error bound against the host's libm pow(); the few values whose bound
 * straddles the decision threshold or a float rounding boundary are re-evaluated with libm on the
 * host (counted in stats).  Fails (no CPU fallback) when the device is unusable. */
typedef struct mrisim_octree_build_stats {
    double   total_ms;          /* wall time of the octree build (GPU + fix-ups + linearisation)  */
    double   gpu_ms;            /* of which: kernels, measured with CUDA events                   */
    double   fixup_ms;          /* host libm re-evaluations (thread time; they run beside the GPU work)  */
    double   linearise_ms;      /* of which: depth-first linearisation on the host                */
    uint64_t n_nodes, n_leaves, n_levels;
    uint64_t criterion_evals, criterion_fixups;   /* leaf-criterion evaluations / re-done on host */
    uint64_t b_values, b_fixups;                  /* float B samples / re-done on host            */
    uint64_t pair_evals;        /* (point, dipole) terms evaluated on the GPU                     */
    uint64_t kernel_launches;
    /* the other stages of mrisim_scene_build_gpu (csrc/scene_gpu.cu), wall ms:
     * water throw (candidates + containment + compaction), lookup table, Morton keys + std::sort */
    double   scene_throw_ms, scene_lookup_ms, scene_sort_ms;
    double   fixup_wait_ms;     /* part of fixup_ms the build had to wait for after its last kernel      */
} mrisim_octree_build_stats;
int  mrisim_scene_build_gpu(const mrisim_params* p, uint64_t host_seed, int n_threads, int device,
                            mrisim_scene** out, mrisim_octree_build_stats* stats /* may be NULL */);
/* test knob: multiplies the certified error bounds (default 1) so that the host fix-up path can be
 * exercised at will; the result must not change. */
void mrisim_octree_build_set_margin_scale(double scale);
int  mrisim_scene_load(const char* path, mrisim_scene** out);     /* MRISCN01 file */
int  mrisim_scene_save(const mrisim_scene* s, const char* path);
void mrisim_scene_free(mrisim_scene* s);
const char* mrisim_scene_last_error(void);

/* accessors -- SimulationBox::getCells/getMNPs/getMNPCount/getLookupTable/getWaters/
 * getOctree (SimulationBox.cpp:536-594).  Pointers stay valid until mrisim_scene_free. */
const double*       mrisim_scene_cells(const mrisim_scene* s, int32_t* num_cells);          /* Triple[num_cells] */
const mrisim_mnp*   mrisim_scene_mnps(const mrisim_scene* s, int32_t* num_mnps);
const int32_t*      mrisim_scene_lookup(const mrisim_scene* s, int32_t* hashDim, int32_t* maxNeighbors); /* rows contiguous, -1 padded */
const mrisim_water* mrisim_scene_waters(const mrisim_scene* s, int32_t* num_water);         /* Morton-sorted */
int  mrisim_scene_octree(const mrisim_scene* s, int32_t* min_depth, int32_t* max_depth,
                         const uint32_t** subtree_sizes, int32_t* n_subtrees,
                         const mrisim_node** nodes, uint64_t* n_nodes,
                         const mrisim_mnp** residents, uint64_t* n_res);
/* replace the waters of a scene: re-throws `num_water` waters continuing the scene's host
 * RNG stream is NOT what this does -- it throws from a fresh XORShift(seed) so that
 * benchmark-size walker sets (1M..16M) can be generated per rank without rebuilding the
 * octree.  BacteriaBox::init_waters + SimulationBox::sortWaters semantics otherwise. */
int  mrisim_scene_rethrow_waters(mrisim_scene* s, const mrisim_params* p, int64_t num_water,
                                 uint64_t seed, int sort_by_morton);
/* same waters, thrown on GPU `device`: attempts generated from recorded states of the host stream,
 * containment tested against every cell (SimulationBox.cpp:161-173), accepted in stream order,
 * leaf Morton keys from a GPU descent; only the std::sort stays on the host.  Bit-identical. */
int  mrisim_scene_rethrow_waters_gpu(mrisim_scene* s, const mrisim_params* p, int64_t num_water,
                                     uint64_t seed, int sort_by_morton, int device);

/* -------------------------------------------------------------------------------------
 * Device context: replaces the body of simulateWaters() between populateSimulation() and
 * `fout <<` (src/cuda_sim.cu:923-1024).
 * ----------------------------------------------------------------------------------- */
typedef struct mrisim_ctx mrisim_ctx;

/* field-evaluation arithmetic of the resident-dipole sum (cuda_sim.cu:107-114) */
enum {
    MRISIM_FIELD_EXACT = 0,     /* IEEE sqrt + divide, unfused: bit-identical to the oracle  */
    MRISIM_FIELD_FAST  = 1      /* rsqrt/FMA form: same positions/indices, phase within 1e-12 */
};

/* How a sprint is mapped onto the GPU.  Per-walker results (positions, flags, leaf indices, RNG
 * state, phases) are bit-identical in both; only the association order of the signal sum differs.
 *   FUSED    one thread per walker, the whole step in registers (many walkers: issue-bound).
 *   PIPELINE the reference's own split (cuda_sim.cu:976-1013: draws, diffusion, phase accumulation,
 *            block sum), for FEW walkers where one thread per walker leaves the GPU latency-bound:
 *            draws per (walker, draw-chunk) from GF(2) jump-ahead states, the position recursion
 *            with one WARP per walker (32 speculative steps per trip, membrane tests side by side),
 *            the field per (walker, step), ordered block sums.  Needs even pfreq and whole intervals.
 *   AUTO     PIPELINE up to 65 536 walkers, FUSED above (default). */
enum {
    MRISIM_PATH_AUTO = 0,
    MRISIM_PATH_FUSED = 1,
    MRISIM_PATH_PIPELINE = 2
};

/* initGPUParams (cuda_sim.cu:406-408) + `Handler handle(0)` (:916). */
int  mrisim_create(const mrisim_params* p, int device, mrisim_ctx** out);
void mrisim_destroy(mrisim_ctx* ctx);                 /* finalizeGPU (cuda_sim.cu:430-433) */
const char* mrisim_last_error(const mrisim_ctx* ctx); /* ctx may be NULL: last create error */
int  mrisim_set_field_mode(mrisim_ctx* ctx, int mode);
int  mrisim_set_path_mode(mrisim_ctx* ctx, int mode);
uint64_t mrisim_pipeline_sprints(const mrisim_ctx* ctx);   /* sprints that took the PIPELINE path */
/* Launch granularity of a sprint (no reference analogue; the reference launches one kernel per
 * sprint, cuda_sim.cu:980).  A sprint is cut into launches of `steps` timesteps; between launches
 * the walkers are re-ordered by the octree leaf they sit in so that the lanes of a warp share
 * resident lists and lookup rows.  Results per walker are unaffected (state and random stream
 * travel with the walker index); only the association order of the signal sum changes.
 * steps < 0: default (125); steps == 0: one launch per sprint, walkers in array order. */
int  mrisim_set_resort_steps(mrisim_ctx* ctx, int32_t steps);
/* The walker array is cut into `groups` contiguous ranges whose launch chains (sort, kernel, sort, ...)
 * run on separate streams, so that the tail of one range's kernel and its sort overlap the body of
 * another's.  Default 3 (B200, 1 Mi walkers: +11 % over one chain, 2..6 alike); a range never gets fewer
 * than two CTAs per SM, so few walkers run as one chain.  Per-walker results do not depend on it. */
int  mrisim_set_overlap_groups(mrisim_ctx* ctx, int32_t groups);

/* `p_array<Triple> dev_lattice(p.num_cells, simBox.getCells(), d.lattice)` (cuda_sim.cu:935,960) */
int  mrisim_set_cells(mrisim_ctx* ctx, const double* xyz, int32_t num_cells);
/* cpyLookupDevice (cuda_sim.cu:849-861): rows[hashDim^3][maxNeighbors], -1 terminated   */
int  mrisim_set_lookup(mrisim_ctx* ctx, const int32_t* rows, int32_t hashDim, int32_t maxNeighbors);
/* initOctree (cuda_sim.cu:296-369): forest of 8^min_depth DFS node vectors              */
int  mrisim_set_octree(mrisim_ctx* ctx, int32_t min_depth, int32_t max_depth,
                       const uint32_t* subtree_sizes, int32_t n_subtrees,
                       const mrisim_node* nodes, uint64_t n_nodes,
                       const mrisim_mnp* residents, uint64_t n_res);
/* the three calls above from one mrisim_scene */
int  mrisim_set_scene(mrisim_ctx* ctx, const mrisim_scene* s);

/* `p_array<water_info> dev_waters(p.num_water, simBox.getWaters(), d.waters)` (cuda_sim.cu:934,961).
 * Walker i of this call is GLOBAL walker first_walker_id + i; its random stream is
 * XORShift(walker_seed + global id) (contract in mrisim_math.h), so a run sharded over
 * several contexts/GPUs draws exactly the numbers of the unsharded run.  Resets time to 0. */
int  mrisim_set_waters(mrisim_ctx* ctx, const mrisim_water* waters, uint64_t n,
                       uint64_t walker_seed, uint64_t first_walker_id);
/* d.waters read-back (cuda_sim.cu:744) */
int  mrisim_get_waters(mrisim_ctx* ctx, mrisim_water* out, uint64_t n);
/* checkpoint/resume of the walker streams and the step counter (no reference analogue) */
int  mrisim_get_rng(mrisim_ctx* ctx, mrisim_rng_state* out, uint64_t n);
int  mrisim_set_rng(mrisim_ctx* ctx, const mrisim_rng_state* in, uint64_t n);
int64_t mrisim_get_time(const mrisim_ctx* ctx);       /* steps simulated so far (`time`, cuda_sim.cu:973,1015) */
int  mrisim_set_time(mrisim_ctx* ctx, int64_t time_steps);

/* One iteration of the sprint loop (cuda_sim.cu:974-1024) minus the CSV line: `steps`
 * timesteps (a multiple of pfreq), writing steps/pfreq signal samples
 * S_k = sum_w cos(phase_w) to signal_out (HOST memory), CPMG flips applied after sampling
 * whenever time % tcp == 0.  Synchronous. */
int  mrisim_run_sprint(mrisim_ctx* ctx, int32_t steps, double* signal_out);
/* Same, asynchronous on `cuda_stream` (a cudaStream_t; NULL is CUDA's legacy default stream,
 * as everywhere in CUDA; mrisim_ctx_stream() returns the context's own), samples written to
 * DEVICE memory signal_out_dev (steps/pfreq doubles, e.g. a buffer a collective library will
 * all-reduce).  The caller synchronises that stream before touching walker state. */
int  mrisim_run_sprint_async(mrisim_ctx* ctx, int32_t steps, double* signal_out_dev, void* cuda_stream);
void* mrisim_ctx_stream(mrisim_ctx* ctx);     /* the cudaStream_t the synchronous calls use */
/* Ordering rules of the asynchronous call.  The context records an event on `cuda_stream` at the end of
 * every sprint; the state accessors (get/set_waters, get/set_rng, get_order, checkpoints) make the
 * context's own stream wait for it, so state read after an async sprint is never stale.  Two sprints
 * launched on DIFFERENT streams are not ordered against each other -- that is the caller's job.
 * mrisim_check() waits for the last sprint launched so far and reads AND clears the device error flag
 * (a walker that left the lookup lattice, where the reference would read out of bounds, cuda_sim.cu:455-463);
 * the synchronous calls do that themselves, mrisim_set_waters() starts every walker set with a clear flag. */
int  mrisim_check(mrisim_ctx* ctx);

/* Submitted sprints -- the sprint loop without a host stall per sprint (the reference blocks on
 * thrust::reduce 20 times per sprint, cuda_sim.cu:1013).  mrisim_submit_sprint() enqueues the sprint's
 * kernels on `compute_stream` (NULL: the context's own) and, on a SIDE stream that waits for them, the
 * in-place SUM all-reduce of the steps/pfreq samples over the attached communicator (if any) and their
 * copy to pinned host memory -- so the exchange of sprint k overlaps the kernels of sprint k+1.
 * mrisim_wait_sprint() blocks until the OLDEST outstanding sprint's samples are on the host, copies them
 * to signal_out (steps/pfreq doubles: the GLOBAL signal when a communicator is attached) and reports the
 * error flag.  At most four sprints may be outstanding.  mrisim_stream_wait_sprint() makes a caller
 * stream wait for the most recently submitted sprint's exchange (device-side, for event timing). */
int  mrisim_submit_sprint(mrisim_ctx* ctx, int32_t steps, void* compute_stream);
int  mrisim_wait_sprint(mrisim_ctx* ctx, double* signal_out);
int  mrisim_stream_wait_sprint(mrisim_ctx* ctx, void* cuda_stream);
int  mrisim_pending_sprints(const mrisim_ctx* ctx);

/* Communicator of the per-sprint signal all-reduce (no reference analogue: the reference drives one
 * device, `Handler handle(0)`, cuda_sim.cu:916).  NCCL is bound at run time: an NCCL already loaded in
 * the process (e.g. PyTorch's) is used, else $MRISIM_NCCL_LIB, else the system libnccl.so.2.
 * One process per GPU: rank 0 calls mrisim_nccl_unique_id(), ships the 128 bytes to the other ranks by
 * any means, every rank calls mrisim_comm_init_rank() (collective).  One process, several GPUs: see
 * mrisim_group_* below.  mrisim_comm_attach() borrows an ncclComm_t the caller owns (NULL detaches). */
int  mrisim_nccl_unique_id(void* id_out_128_bytes);
int  mrisim_comm_init_rank(mrisim_ctx* ctx, const void* id_128_bytes, int nranks, int rank);
int  mrisim_comm_attach(mrisim_ctx* ctx, void* nccl_comm, int nranks, int rank);
int  mrisim_comm_size(const mrisim_ctx* ctx);
const char* mrisim_nccl_origin(void);         /* which NCCL was bound, and its version (or why none) */

/* walker order of the one-thread-per-walker path (a permutation sorted by octree leaf between launches;
 * it fixes the association order of the signal sum, nothing per walker).  Saved/restored by checkpoints so
 * that a resumed run prints byte-identical rows.  valid = 0: the context holds no order yet. */
int  mrisim_get_order(mrisim_ctx* ctx, uint32_t* perm, uint32_t* keys, uint64_t n, int32_t* valid);
int  mrisim_set_order(mrisim_ctx* ctx, const uint32_t* perm, const uint32_t* keys, uint64_t n);

/* what the context holds (checkpoint headers, group bookkeeping) */
uint64_t mrisim_walker_count(const mrisim_ctx* ctx);
uint64_t mrisim_first_walker_id(const mrisim_ctx* ctx);
uint64_t mrisim_walker_seed(const mrisim_ctx* ctx);
uint64_t mrisim_ctx_scene_digest(mrisim_ctx* ctx);       /* FNV-1a of the uploaded scene arrays, 0: incomplete */
const mrisim_params* mrisim_ctx_params(const mrisim_ctx* ctx);
int  mrisim_ctx_device(const mrisim_ctx* ctx);

/* -------------------------------------------------------------------------------------
 * Checkpoint / resume (no reference analogue -- its walker state does round-trip
 * through d.waters once per sprint, cuda_sim.cu:694,744, which makes a sprint boundary the natural
 * checkpoint, and the CSV row of :1022 the natural unit of output to resume after).
 * File "MRICKP01": header (parameter hash, scene digest, host/walker seeds, time, CSV byte offset,
 * walker range, shard table) + SoA walker state (x y z phase spare s0 s1 nearest in_cell) + per shard
 * the walker order.  Written to <path>.tmp and renamed.  A file written by a group of k devices can be
 * loaded by a single context or a group of any size (the walker order is restored where the shard
 * boundaries coincide, recomputed elsewhere -- it only affects the last bits of the signal sum).
 * ----------------------------------------------------------------------------------- */
typedef struct mrisim_checkpoint_info {
    uint64_t params_hash, scene_digest, host_seed, walker_seed;
    int64_t  time_steps;
    uint64_t csv_bytes;
    uint64_t n_walkers, first_walker_id;
    uint32_t n_shards, version;
} mrisim_checkpoint_info;
uint64_t mrisim_params_hash(const mrisim_params* p);      /* every knob that changes the walk (not num_threads, totaltime, t) */
int  mrisim_checkpoint_save(mrisim_ctx* ctx, const char* path, uint64_t host_seed, uint64_t csv_bytes);
int  mrisim_checkpoint_info_read(const char* path, mrisim_checkpoint_info* out);
/* load walkers [lo, lo + count) of the file into ctx (scene must be set; count = 0: all of them);
 * refuses a file whose parameter hash or scene digest differs from the context's */
int  mrisim_checkpoint_load(mrisim_ctx* ctx, const char* path, uint64_t lo, uint64_t count);
const char* mrisim_checkpoint_last_error(void);

/* -------------------------------------------------------------------------------------
 * Device group: ONE process driving several GPUs (north-star: walkers shard over the 8 B200s of a
 * box, scene replicated, one NCCL all-reduce of the signal partials per sprint).  Replaces
 * `Handler handle(0)` (cuda_sim.cu:916).  The walker array is cut into contiguous ranges (sizes differ
 * by at most one), walker i keeps the stream of its GLOBAL index, so any device count produces the
 * same per-walker results as one device.  ncclCommInitAll over the device list; n_devices == 1 needs no NCCL.
 * ----------------------------------------------------------------------------------- */
typedef struct mrisim_group mrisim_group;
int  mrisim_group_create(const mrisim_params* p, const int* devices, int n_devices, mrisim_group** out);
void mrisim_group_destroy(mrisim_group* g);
const char* mrisim_group_last_error(const mrisim_group* g);          /* g may be NULL: last create error */
int  mrisim_group_size(const mrisim_group* g);
mrisim_ctx* mrisim_group_ctx(mrisim_group* g, int i);                /* member i (field/path mode knobs, accessors) */
int  mrisim_group_set_scene(mrisim_group* g, const mrisim_scene* s);
int  mrisim_group_set_waters(mrisim_group* g, const mrisim_water* waters, uint64_t n, uint64_t walker_seed, uint64_t first_walker_id);
int  mrisim_group_get_waters(mrisim_group* g, mrisim_water* out, uint64_t n);
int  mrisim_group_get_rng(mrisim_group* g, mrisim_rng_state* out, uint64_t n);
int64_t mrisim_group_get_time(const mrisim_group* g);
/* submit on every member (kernels, then ONE grouped all-reduce on the side streams), wait for the oldest */
int  mrisim_group_submit_sprint(mrisim_group* g, int32_t steps);
int  mrisim_group_wait_sprint(mrisim_group* g, double* signal_out);
int  mrisim_group_run_sprint(mrisim_group* g, int32_t steps, double* signal_out);     /* submit + wait */
int  mrisim_group_checkpoint_save(mrisim_group* g, const char* path, uint64_t host_seed, uint64_t csv_bytes);
int  mrisim_group_checkpoint_load(mrisim_group* g, const char* path);
uint64_t mrisim_group_kernel_launches(const mrisim_group* g);

/* Test hook: a sprint that also records, for every (step, walker), position, in_cell and
 * the flat octree leaf index (= subtree_prefix[i] + j of the reference's (i, j)).  Arrays
 * are HOST memory, step-major [steps][n]; any pointer may be NULL. */
int  mrisim_trace_sprint(mrisim_ctx* ctx, int32_t steps, double* signal_out,
                         double* x, double* y, double* z, uint8_t* in_cell, uint32_t* leaf);
/* Same, plus the per-(step, walker) phase kick -- what computePhaseAccumulation writes into its
 * target buffer (cuda_sim.cu:772-804, accumulatePhase :625-640) before the dgemv sums it. */
int  mrisim_trace_sprint_kicks(mrisim_ctx* ctx, int32_t steps, double* signal_out,
                               double* x, double* y, double* z, uint8_t* in_cell, uint32_t* leaf, double* kick);

/* kernels launched by this context so far (bench.py's gpu_launches claim) */
uint64_t mrisim_kernel_launches(const mrisim_ctx* ctx);
/* device-timed duration (ms, CUDA events on the launch stream) of the last sprint's walker kernel */
float mrisim_last_sprint_kernel_ms(mrisim_ctx* ctx);

/* CSV row exactly as `fout << ((double) time * p.tau) << delim << target << endl`
 * (cuda_sim.cu:1022, parameters.h:129).  Returns the number of bytes written (incl. '\n'). */
int  mrisim_csv_row(int64_t time_steps, double tau, double target, char* buf, size_t buflen);

/* The whole of simulateWaters(fName) (cuda_sim.cu:902-1034) as one call: build the scene
 * from host_seed, upload, run max_sprints sprints (<0: all t/sprintSteps of them) on
 * `device`, write the CSV to `filename`, print the reference's banners and the elapsed
 * time of the sprint loop.  host_seed / walker_seed replace time(NULL)+rd() (:911-912) and
 * the cuRAND seed time(NULL) (gpu_random.cu:22).  The C++ drop-in
 * `void simulateWaters(std::string)` (csrc/simulate_waters.cpp) forwards here. */
int  mrisim_simulate_waters(const char* filename, const mrisim_params* p, uint64_t host_seed,
                            uint64_t walker_seed, int device, int64_t max_sprints);
/* Same on a list of devices (walkers sharded, one all-reduce per sprint overlapped with the next sprint's
 * kernels, rank-0 CSV), with checkpoints: every `checkpoint_every` sprints (0: never) the state goes to
 * `checkpoint_path` (NULL: "<filename>.ckpt"); `resume_path` (NULL: fresh run) continues a run from such
 * a file -- seeds and time come from the file, the CSV is truncated to the recorded offset and appended
 * to, and the final CSV is byte-identical to an uninterrupted run's.
 * Environment of both calls (the drop-in has no other way to take options): MRISIM_DEVICES=0,1,..
 * overrides the device (list); MRISIM_CHECKPOINT_EVERY, MRISIM_CHECKPOINT, MRISIM_RESUME as above. */
int  mrisim_simulate_waters_multi(const char* filename, const mrisim_params* p, uint64_t host_seed,
                                  uint64_t walker_seed, const int* devices, int n_devices, int64_t max_sprints,
                                  int64_t checkpoint_every, const char* checkpoint_path, const char* resume_path);

/* FP64 pipe micro-benchmark used as the roofline denominator: dependent-free DFMA chains
 * on every SM; returns achieved FP64 FMA TFLOP/s (2 flop per FMA) or <0 on error. */
double mrisim_measure_fp64_tflops(int device, int iters);

/* Device self-test of the constant-divisor division used for x/bound and x/cubeLength:
 * compares it with the IEEE divide for 3n numerators in [0, 3b); returns the number of
 * mismatches (0 expected), <0 on error. */
int64_t mrisim_selftest_div(int device, double b, uint64_t n);
/* Same for the branch-free correctly rounded sqrt / divide of the resident loop: n random x in
 * [lo, hi] (sqrt(x), t/(x*x*sqrt(x)), t/x against the compiler's IEEE sqrt and divide). */
int64_t mrisim_selftest_sqrt_div(int device, double lo, double hi, uint64_t n);

/* Same for the integer fast paths of the Morton word floor(x/bound*2^21) (cuda_sim.cu:129-135) and
 * of the lookup-lattice coordinate int(x/cubeLength) (cuda_sim.cu:455-457): 15n coordinates in
 * [0, bound), random and on both sides of every boundary, against the IEEE-divide expressions. */
int64_t mrisim_selftest_index(int device, double bound, int32_t hashDim, uint64_t n);

/* The walker order between launches (no reference analogue) is a stable radix sort written for this library
 * (csrc/mrisim.cu, sort_*_kernel).  Self-test: n pairs with `bits`-bit keys (runs of equal keys included)
 * against std::stable_sort on the host; returns the number of positions that differ, -1 on a CUDA error. */
int64_t mrisim_selftest_sort(int device, uint64_t n, int32_t bits, uint64_t seed);

/* The signal sample -- sum over n walkers of cos(phase), the reference's cosine + thrust::reduce step -- WITHOUT any
 * allocation.  This entry point has no counterpart in the reference; it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  phase_dev: n phases in DEVICE memory of `device`; signal_out_dev: one double in device memory;
 * both kernels are enqueued on `cuda_stream` (a cudaStream_t, NULL = the default stream) and the call returns
 * without synchronising.  The per-CTA partials live in a static device array of 32 KB that every GPU loading the
 * library holds until exit: n is limited to 16 777 216 walkers (larger n: nonzero return, mrisim_last_error(NULL)
 * says so; split the walkers and add the parts).  The call makes no cudaMalloc / cudaFree and calls no allocating
 * library routine; the first call per device resolves the array's address (cudaGetSymbolAddress) -- make it outside
 * a stream capture.  All calls share the array: calls that overlap on the GPU (different streams) race on it;
 * nothing in it needs clearing between calls.  Result: within 1e-6 * n of a context's signal sample (association
 * order of the sum only); deterministic for a given n.  Returns 0 on success. */
int mrisim_signal_sample_static(int device, const double* phase_dev, uint64_t n, double* signal_out_dev, void* cuda_stream);

int mrisim_version(void);

#ifdef __cplusplus
}
#endif
#endif /* MRISIM_H */
