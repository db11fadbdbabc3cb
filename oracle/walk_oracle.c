/*
 * TEST INFRASTRUCTURE -- see walk_oracle.h.  CPU restatement of the reference's device
 * functions; every function names the reference lines it follows.  Compile with
 * -ffp-contract=off (oracle/Makefile) so that no a*b+c is fused behind our back.
 */
#define _GNU_SOURCE
#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>
#include <stdio.h>

#include "walk_oracle.h"
#include "mrisim_math.h"

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

/* src/rand_walk.h:15 -- same (unparenthesised) macro as the reference */
#define NORMSQ(x, y, z) (x)*(x)+(y)*(y)+(z)*(z)

typedef struct env {
    const mrisim_params* p;
    const oracle_scene*  sc;
    uint64_t*            prefix;      /* flat index of each subtree's root */
    uint32_t             morton_x[256], morton_y[256], morton_z[256];
} env;

/* src/octree.h:83-94: "spacing out the bits in the integers from 0 to 255 and left
 * shifting them 0, 1, and 2 bits for morton_z, morton_y, and morton_x". */
static void build_morton_tables(env* e) {
    for (uint32_t i = 0; i < 256; i++) {
        uint32_t s = 0;
        for (int b = 0; b < 8; b++) s |= ((i >> b) & 1u) << (3 * b);
        e->morton_z[i] = s;
        e->morton_y[i] = s << 1;
        e->morton_x[i] = s << 2;
    }
}

static int env_init(env* e, const mrisim_params* p, const oracle_scene* sc) {
    e->p = p; e->sc = sc;
    e->prefix = (uint64_t*)malloc(sizeof(uint64_t) * ((size_t)sc->n_subtrees + 1));
    if (!e->prefix) return 1;
    e->prefix[0] = 0;
    for (int i = 0; i < sc->n_subtrees; i++) e->prefix[i + 1] = e->prefix[i] + sc->subtree_sizes[i];
    build_morton_tables(e);
    return 0;
}
static void env_free(env* e) { free(e->prefix); }

/* cuda_sim.cu:107-114 dipole_field (DEVICE version: masked inside cell_r, sq*sq*sqrt(sq)) */
static double dipole_field(const env* e, double dx, double dy, double dz, double M) {
    double sqDist = NORMSQ(dx, dy, dz);
    double divisor = sqDist * sqDist * sqrt(sqDist);
    return (sqDist > e->p->cell_r * e->p->cell_r) * M * 1e11
        * (2*dz*dz - dx*dx - dy*dy) / divisor;
}

/* cuda_sim.cu:129-161 morton_code */
static uint64_t morton_code(const env* e, int depth, double x, double y, double z) {
    uint64_t size = 1 << (depth);
    uint32_t idx_x = floor(x / e->p->bound * size);
    uint32_t idx_y = floor(y / e->p->bound * size);
    uint32_t idx_z = floor(z / e->p->bound * size);
    uint64_t answer = 0;
    if (depth > 16) {
        answer |= e->morton_z[(idx_z >> 16) & 0xFF] | e->morton_y[(idx_y >> 16) & 0xFF] | e->morton_x[(idx_x >> 16) & 0xFF];
        answer <<= 24;
    }
    if (depth > 8) {
        answer |= e->morton_z[(idx_z >> 8) & 0xFF] | e->morton_y[(idx_y >> 8) & 0xFF] | e->morton_x[(idx_x >> 8) & 0xFF];
        answer <<= 24;
    }
    answer |= e->morton_z[idx_z & 0xFF] | e->morton_y[idx_y & 0xFF] | e->morton_x[idx_x & 0xFF];
    return answer;
}

/* cuda_sim.cu:177-181 find_tree: subtree i = morton_code(min_depth) */
static uint64_t find_tree(const env* e, double wx, double wy, double wz) {
    return morton_code(e, e->sc->min_depth, wx, wy, wz);
}

/* cuda_sim.cu:198-202 find_child */
static unsigned find_child(const env* e, double wx, double wy, double wz, int d) {
    return morton_code(e, d + 1, wx, wy, wz) & 7;
}

/* cuda_sim.cu:224-234 find_node (recursive; child offsets relative to the parent) */
static const mrisim_node* find_node(const env* e, const mrisim_node* n, double wx, double wy, double wz, int d) {
    if (n->mc >> 63)
        return n;
    unsigned child_no = find_child(e, wx, wy, wz, d);
    return find_node(e, n + n->child[child_no], wx, wy, wz, d + 1);
}

/* cuda_sim.cu:249-251 get_voxel */
static const mrisim_node* get_voxel(const env* e, double wx, double wy, double wz) {
    uint64_t i = find_tree(e, wx, wy, wz);
    return find_node(e, e->sc->nodes + e->prefix[i], wx, wy, wz, e->sc->min_depth);
}

/* cuda_sim.cu:269-286 get_field */
static double get_field(const env* e, double wx, double wy, double wz, const mrisim_node* leaf) {
    uint64_t depth = 0, mc = (leaf->mc << 1) >> 1;
    while (mc >>= 3) depth++;

    unsigned child_no = find_child(e, wx, wy, wz, (int)depth);
    float Bf; memcpy(&Bf, &leaf->child[child_no], sizeof Bf);     /* B_idx union, octree.h:18-21 */
    double B = (double)Bf;

    for (uint32_t i = 0; i < leaf->n_res; i++) {
        const mrisim_mnp* np = e->sc->residents + leaf->res_start + i;
        B += dipole_field(e, wx - np->x, wy - np->y, wz - np->z, np->M);
    }
    return B;
}

/* cuda_sim.cu:453-485 updateNearest */
static void updateNearest(const env* e, mrisim_water* w) {
    const mrisim_params* p = e->p;
    double cubeLength = p->bound / p->hashDim;
    int x_idx = w->x / cubeLength;
    int y_idx = w->y / cubeLength;
    int z_idx = w->z / cubeLength;

    const int32_t* nearest = e->sc->lookup +
        (size_t)(z_idx * p->hashDim * p->hashDim + y_idx * p->hashDim + x_idx) * (size_t)p->maxNeighbors;

    double cDist = p->bound * p->bound * 3;
    int cIndex = -1;
    int guard = 0;   /* the reference relies on the -1 sentinel (assert at SimulationBox.cpp:281) */
    while (guard < p->maxNeighbors && *nearest != -1) {
        double dx = e->sc->cells[3 * *nearest + 0] - w->x;
        double dy = e->sc->cells[3 * *nearest + 1] - w->y;
        double dz = e->sc->cells[3 * *nearest + 2] - w->z;

        double dist = NORMSQ(dx, dy, dz);
        if (NORMSQ(dx, dy, dz) < cDist) {
            cDist = dist;
            cIndex = *nearest;
        }
        nearest++; guard++;
    }
    w->in_cell = (cDist < p->cell_r * p->cell_r);
    w->nearest = cIndex;
}

/* cuda_sim.cu:502-509 cell_reflect */
static int cell_reflect(const env* e, const mrisim_water* i, const mrisim_water* f, double coin) {
    int flip = (i->in_cell && (!f->in_cell) && coin < e->p->reflectIO)
            || ((!i->in_cell) && f->in_cell && coin < e->p->reflectOI);
    return flip;
}

/* cuda_sim.cu:562-586 rand_displacement */
static void rand_displacement(const env* e, const mrisim_water* w, double ux, double uy, double uz,
                              double norm, double* ox, double* oy, double* oz) {
    double dx = ux * 2 - 1.0;
    double dy = uy * 2 - 1.0;
    double dz = uz * 2 - 1.0;
    if (w->in_cell) norm *= e->p->in_stdev;
    else            norm *= e->p->out_stdev;
    double nConstant = norm / sqrt(NORMSQ(dx, dy, dz));
    dx *= nConstant; dy *= nConstant; dz *= nConstant;
    *ox = dx; *oy = dy; *oz = dz;
}

/* cuda_sim.cu:597-601 boundary_conditions */
static void boundary_conditions(const env* e, mrisim_water* w) {
    w->x = fmod(w->x + e->p->bound, e->p->bound);
    w->y = fmod(w->y + e->p->bound, e->p->bound);
    w->z = fmod(w->z + e->p->bound, e->p->bound);
}

/* cuda_sim.cu:630-636, the intracellular term, tokens kept (abs -> fabs on a double) */
static double kick_term(const mrisim_params* params, double nD, int in_cell) {
    double phase = 0;
    if (params->kick_mode == MRISIM_KICK_RANDOM) {
        phase += (in_cell) *( params->phase_stdev*sqrt(1/(fabs(nD-.5)*2)-1)*(((nD-.5)>0) - ((nD-.5))<0) * params->tau+params->phase_k*1e-3*42.58*2*M_PI*7*params->tau);
    } else if (params->kick_mode == MRISIM_KICK_CONSTANT) {
        phase += (in_cell) * params->phase_k * params->tau;
    }
    return phase;
}

/* cuda_sim.cu:625-640 accumulatePhase */
static double accumulatePhase(const env* e, double wx, double wy, double wz,
                              const mrisim_node* voxel, double nD, int in_cell) {
    const mrisim_params* params = e->p;
    double B = get_field(e, wx, wy, wz, voxel);
    double phase = kick_term(params, nD, in_cell);
    phase += B * 2 * M_PI * params->g * params->tau * 1e-3;
    return phase;
}

/* ------------------------------------------------------------------------------------- */

void oracle_seed_walkers(mrisim_rng_state* rng, uint64_t n, uint64_t walker_seed, uint64_t first_walker_id) {
    for (uint64_t i = 0; i < n; i++) {
        mrisim_rng r;
        mrisim_rng_seed(&r, walker_seed + first_walker_id + i);
        rng[i].s0 = r.s0; rng[i].s1 = r.s1; rng[i].spare = 0.0;
    }
}

typedef struct job {
    const env* e;
    mrisim_water* waters;
    mrisim_rng_state* rng;
    uint64_t n, lo, hi;
    int64_t time0;
    int32_t steps;
    double* cosbuf;               /* [steps/pfreq][n] */
    const oracle_trace* trace;
} job;

/* One walker through one sprint: simulateDiffusion (cuda_sim.cu:688-746) fused with
 * computePhaseAccumulation (:772-804), the per-pfreq dgemv block sum (:997-1006),
 * performUpdate (:814-828) and flipPhases (:833-838, trigger :1015-1020). */
static void walk_one(const job* jb, uint64_t wi) {
    const env* e = jb->e;
    const mrisim_params* p = e->p;
    mrisim_water w = jb->waters[wi];
    mrisim_rng r = { jb->rng[wi].s0, jb->rng[wi].s1 };
    double spare = jb->rng[wi].spare;
    const oracle_trace* tr = jb->trace;

    updateNearest(e, &w);                                   /* :695 */

    int64_t time = jb->time0;
    double update = 0;                                      /* dgemv with beta = 0 */
    for (int32_t i = 0; i < jb->steps; i++) {
        const int64_t k = jb->time0 + i;
        /* walker stream contract (mrisim_math.h): ux uy uz [u1 u2] coin kick */
        double ux = mrisim_rng_u01(&r), uy = mrisim_rng_u01(&r), uz = mrisim_rng_u01(&r);
        double norm;
        if ((k & 1) == 0) {
            double u1 = mrisim_rng_u01(&r), u2 = mrisim_rng_u01(&r);
            mrisim_normal_pair(u1, u2, &norm, &spare);
        } else {
            norm = spare;
        }
        double coin = mrisim_rng_u01(&r);
        double nD = mrisim_rng_u01(&r);

        mrisim_water init = w;                              /* :709 */
        double ddx, ddy, ddz;
        rand_displacement(e, &w, ux, uy, uz, norm, &ddx, &ddy, &ddz);   /* :711 */
        w.x += ddx; w.y += ddy; w.z += ddz;                 /* :712-714 */
        boundary_conditions(e, &w);                         /* :715 */
        updateNearest(e, &w);                               /* :716 */
        if (cell_reflect(e, &init, &w, coin)) w = init;     /* :720-722 */

        /* :725-728 stores (x,y,z,in_cell); :794-795 reads them back */
        const mrisim_node* voxel = get_voxel(e, w.x, w.y, w.z);
        double kick = accumulatePhase(e, w.x, w.y, w.z, voxel, nD, w.in_cell);
        update += kick;                                     /* :997-1006, ascending step order */

        if (tr) {
            size_t o = (size_t)i * jb->n + wi;
            if (tr->x) tr->x[o] = w.x;
            if (tr->y) tr->y[o] = w.y;
            if (tr->z) tr->z[o] = w.z;
            if (tr->in_cell) tr->in_cell[o] = w.in_cell;
            if (tr->leaf) tr->leaf[o] = (uint32_t)(voxel - e->sc->nodes);
            if (tr->kick) tr->kick[o] = kick;
        }

        if ((i + 1) % p->pfreq == 0) {
            double n_phase = w.phase + update;              /* :820 */
            jb->cosbuf[(size_t)(i / p->pfreq) * jb->n + wi] = cos(n_phase);   /* :823 */
            w.phase = n_phase;                              /* :826 */
            update = 0;
            time += p->pfreq;                               /* :1015 */
            if (time % p->tcp == 0) w.phase *= -1;          /* :1018-1020, :837 */
        }
    }
    jb->waters[wi] = w;                                     /* :744 */
    jb->rng[wi].s0 = r.s0; jb->rng[wi].s1 = r.s1; jb->rng[wi].spare = spare;
}

static void* worker(void* arg) {
    const job* jb = (const job*)arg;
    for (uint64_t w = jb->lo; w < jb->hi; w++) walk_one(jb, w);
    return NULL;
}

int oracle_run_sprint(const mrisim_params* p, const oracle_scene* sc,
                      mrisim_water* waters, mrisim_rng_state* rng, uint64_t n,
                      int64_t time0, int32_t steps, double* signal_out,
                      const oracle_trace* trace, int n_threads) {
    if (p->pfreq <= 0 || steps % p->pfreq != 0) return 1;
    env e;
    if (env_init(&e, p, sc)) return 2;
    const int n_int = steps / p->pfreq;
    double* cosbuf = (double*)calloc((size_t)n_int * (n ? n : 1), sizeof(double));
    if (!cosbuf) { env_free(&e); return 2; }

    if (n_threads < 1) n_threads = 1;
    if ((uint64_t)n_threads > n) n_threads = n ? (int)n : 1;
    job* jobs = (job*)calloc((size_t)n_threads, sizeof(job));
    pthread_t* th = (pthread_t*)calloc((size_t)n_threads, sizeof(pthread_t));
    for (int t = 0; t < n_threads; t++) {
        jobs[t] = (job){ &e, waters, rng, n, n * (uint64_t)t / (uint64_t)n_threads,
                         n * (uint64_t)(t + 1) / (uint64_t)n_threads, time0, steps, cosbuf, trace };
        if (n_threads == 1) worker(&jobs[t]);
        else pthread_create(&th[t], NULL, worker, &jobs[t]);
    }
    if (n_threads > 1) for (int t = 0; t < n_threads; t++) pthread_join(th[t], NULL);

    /* thrust::reduce (:1013): order unspecified in the reference; walker order here */
    for (int k = 0; k < n_int; k++) {
        double target = 0;
        for (uint64_t w = 0; w < n; w++) target += cosbuf[(size_t)k * n + w];
        signal_out[k] = target;
    }
    free(th); free(jobs); free(cosbuf); env_free(&e);
    return 0;
}

uint32_t oracle_probe_leaf(const mrisim_params* p, const oracle_scene* sc, double x, double y, double z, uint64_t* mc_out) {
    env e; env_init(&e, p, sc);
    const mrisim_node* v = get_voxel(&e, x, y, z);
    if (mc_out) *mc_out = v->mc;
    uint32_t idx = (uint32_t)(v - sc->nodes);
    env_free(&e);
    return idx;
}

double oracle_probe_field(const mrisim_params* p, const oracle_scene* sc, double x, double y, double z) {
    env e; env_init(&e, p, sc);
    double B = get_field(&e, x, y, z, get_voxel(&e, x, y, z));
    env_free(&e);
    return B;
}

int32_t oracle_probe_nearest(const mrisim_params* p, const oracle_scene* sc, double x, double y, double z, int* in_cell) {
    env e; env_init(&e, p, sc);
    mrisim_water w; memset(&w, 0, sizeof w); w.x = x; w.y = y; w.z = z;
    updateNearest(&e, &w);
    if (in_cell) *in_cell = w.in_cell;
    env_free(&e);
    return w.nearest;
}

double oracle_kick_expression(const mrisim_params* p, double nD, int in_cell) {
    return kick_term(p, nD, in_cell);
}

void oracle_xorshift_words(uint64_t seed, uint64_t* out, int n) {
    mrisim_rng r; mrisim_rng_seed(&r, seed);
    for (int i = 0; i < n; i++) out[i] = mrisim_rng_next(&r);
}
void oracle_xorshift_doubles(uint64_t seed, double* out, int n) {
    mrisim_rng r; mrisim_rng_seed(&r, seed);
    for (int i = 0; i < n; i++) out[i] = mrisim_rng_u01(&r);
}
void oracle_normals(uint64_t seed, double* out, int n) {
    mrisim_rng r; mrisim_rng_seed(&r, seed);
    for (int i = 0; i + 1 < n; i += 2) {
        double u1 = mrisim_rng_u01(&r), u2 = mrisim_rng_u01(&r);
        mrisim_normal_pair(u1, u2, &out[i], &out[i + 1]);
    }
}
void oracle_log_sincos(double v, double u, double* lg, double* sn, double* cs) {
    *lg = mrisim_log(v);
    mrisim_sincos2pi(u, sn, cs);
}
