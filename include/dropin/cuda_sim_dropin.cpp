// Glue between the reference's compile-time knobs (its own, unchanged src/parameters.h: a
// struct of static constants plus four doubles set in the constructor, and #define flags)
// and the run-time mrisim_params of the C ABI.  This file replaces src/cuda_sim.cu,
// src/gpu_random.cu, src/SimulationBox.cpp, src/BacteriaBox.cpp and src/octree.cpp in the
// reference's link line; src/T2_GPU_sim.cu stays as it is.
#include <cstdlib>
#include <ctime>
#include <random>
#include <string>

#include "cuda_sim.h"
#include "mrisim.h"

ParameterStruct p(0);                         // the global the reference defines at src/cuda_sim.cu:41

void simulateWaters(std::string fName) {
    mrisim_params q;
    mrisim_params_default(&q);
    q.g = p.g;
    q.scale = p.scale;
#ifndef FULL_BOUNDARIES
    q.border = p.border;
#endif
    q.num_threads = p.num_threads;
    q.sprintSteps = p.sprintSteps;
    q.num_water = p.num_water;
    q.num_cells = p.num_cells;
    q.cell_r = p.cell_r;
    q.mmoment = p.mmoment;
#if defined RANDOM_KICK
    q.kick_mode = MRISIM_KICK_RANDOM;
    q.phase_stdev = p.phase_stdev;
    q.phase_k = p.phase_k;
#elif defined CONSTANT_KICK
    q.kick_mode = MRISIM_KICK_CONSTANT;
    q.phase_k = p.phase_k;
#else
    q.kick_mode = MRISIM_KICK_NONE;
#endif
#ifdef AVOID_INTRACELLULAR_THROW
    q.avoid_intracellular_throw = 1;
#else
    q.avoid_intracellular_throw = 0;
#endif
    q.bound = p.bound;
    q.water_start_bound = p.water_start_bound;
    q.hashDim = p.hashDim;
    q.maxNeighbors = p.maxNeighbors;
    q.D_cell = p.D_cell;
    q.D_extra = p.D_extra;
    q.P_expr = p.P_expr;
    q.tau = p.tau;
    q.totaltime = p.totaltime;
    q.taucp = p.taucp;
    mrisim_params_derive(&q);                 // t, tcp, pfreq, in_stdev, out_stdev by the reference's formulas
    q.t = p.t;
    q.tcp = p.tcp;
    q.reflectIO = p.reflectIO;                // the four run-time members, set in ParameterStruct's constructor
    q.reflectOI = p.reflectOI;
    q.in_stdev = p.in_stdev;
    q.out_stdev = p.out_stdev;

    std::random_device rd;
    const char* hs = std::getenv("MRISIM_HOST_SEED");
    const char* ws = std::getenv("MRISIM_WALKER_SEED");
    const char* ms = std::getenv("MRISIM_MAX_SPRINTS");
    const char* dv = std::getenv("MRISIM_DEVICE");
    const uint64_t host_seed = hs ? std::strtoull(hs, nullptr, 0) : (uint64_t)std::time(nullptr) + rd();   // src/cuda_sim.cu:911-912
    const uint64_t walker_seed = ws ? std::strtoull(ws, nullptr, 0) : (uint64_t)std::time(nullptr);        // src/gpu_random.cu:22
    if (mrisim_simulate_waters(fName.c_str(), &q, host_seed, walker_seed, dv ? std::atoi(dv) : 0, ms ? std::atoll(ms) : -1))
        std::exit(1);
}
