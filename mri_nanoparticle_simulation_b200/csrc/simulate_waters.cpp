// simulate_waters.cpp -- the reference's one exported function, `void simulateWaters(std::string)`
// (src/cuda_sim.h:74, body src/cuda_sim.cu:902-1034), rebuilt on the C ABI of include/mrisim.h.
//
// Same observable behaviour: the banners on stdout, one CSV row per pfreq steps
// ("<time*tau>,<sum cos phase>", default ostream formatting, cuda_sim.cu:1022), the elapsed
// time of the sprint loop.  Differences, all deliberate: seeds are explicit, the host syncs
// once per sprint instead of once per row (40 000 -> 2 000 syncs per default run), rows are
// written per sprint, every CUDA status is checked.
#include <chrono>
#include <cstdlib>
#include <ctime>
#include <fstream>
#include <iostream>
#include <random>
#include <string>
#include <vector>

#include "mrisim.h"

namespace {
struct SceneGuard { mrisim_scene* s = nullptr; ~SceneGuard() { if (s) mrisim_scene_free(s); } };
struct CtxGuard { mrisim_ctx* c = nullptr; ~CtxGuard() { if (c) mrisim_destroy(c); } };
thread_local std::string g_sim_error;
}

extern "C" int mrisim_simulate_waters(const char* filename, const mrisim_params* pp, uint64_t host_seed,
                                      uint64_t walker_seed, int device, int64_t max_sprints) {
    if (!filename || !pp) return 1;
    const mrisim_params& p = *pp;
    std::cout << "Starting GPU Simulation..." << std::endl;                 // cuda_sim.cu:903
    std::cout << "Printing to: " << filename << std::endl;                  // :904
    std::ofstream fout(filename);                                           // :905
    if (!fout) { std::cerr << "simulateWaters: cannot open " << filename << std::endl; return 1; }
    if (p.sprintSteps <= 0 || p.pfreq <= 0 || p.sprintSteps % p.pfreq != 0) {
        std::cerr << "simulateWaters: sprintSteps must be a multiple of the printing frequency (parameters.h:41)" << std::endl;
        return 1;
    }

    SceneGuard scene;                                                       // :920-921
    const std::time_t t0 = std::time(nullptr);
    // the scene is built on the GPU the walk runs on (octree_build.cu / scene_gpu.cu: byte-identical to
    // the host build, 50x faster for dense boxes); MRISIM_HOST_SCENE=1 keeps the reference's host threads
    const char* host_scene = std::getenv("MRISIM_HOST_SCENE");
    const int rc_scene = (host_scene && host_scene[0] == '1') ? mrisim_scene_build(&p, host_seed, 0, &scene.s)
                                                              : mrisim_scene_build_gpu(&p, host_seed, 0, device, &scene.s, nullptr);
    if (rc_scene) {
        std::cerr << "simulateWaters: " << mrisim_scene_last_error() << std::endl;
        return 1;
    }
    {
        int32_t n_mnp = 0, mind = 0, maxd = 0; uint64_t n_nodes = 0;
        mrisim_scene_mnps(scene.s, &n_mnp);
        mrisim_scene_octree(scene.s, &mind, &maxd, nullptr, nullptr, nullptr, &n_nodes, nullptr, nullptr);
        const std::time_t dt = std::time(nullptr) - t0;
        std::cout << "Molecules initialized!" << std::endl;                 // BacteriaBox.cpp:104
        std::cout << "After applying boundary conditions, there are " << n_mnp << " nanoparticles." << std::endl;   // octree.cpp:404-405
        std::cout << "The tree has a minimum depth of " << mind << " and a maximum depth of " << maxd << "." << std::endl;
        std::cout << "The tree has " << (double)n_nodes << " nodes." << std::endl;
        std::cout << "Octree took " << dt / 60 << ":" << (dt % 60 < 10 ? "0" : "") << dt % 60 << " to build." << std::endl << std::endl;  // SimulationBox.cpp:90-92
    }

    CtxGuard ctx;                                                           // :924-926
    if (mrisim_create(&p, device, &ctx.c)) { std::cerr << "simulateWaters: " << mrisim_last_error(nullptr) << std::endl; return 1; }
    auto die = [&](const char* what) { std::cerr << "simulateWaters: " << what << ": " << mrisim_last_error(ctx.c) << std::endl; return 1; };
    if (mrisim_set_scene(ctx.c, scene.s)) return die("scene upload");
    std::cout << "Allocated GPU Octree!" << std::endl;                      // :368
    int32_t n_water = 0;
    const mrisim_water* waters = mrisim_scene_waters(scene.s, &n_water);    // :934,961
    if (mrisim_set_waters(ctx.c, waters, (uint64_t)n_water, walker_seed, 0)) return die("water upload");

    std::cout << "Kernel prepped!" << std::endl;                            // :968
    std::cout << "Starting GPU computation..." << std::endl;                // :969
    const auto start = std::chrono::steady_clock::now();                    // timer.cpuStart(), :970

    const int rows_per_sprint = p.sprintSteps / p.pfreq;
    std::vector<double> signal((size_t)rows_per_sprint);
    int64_t n_sprints = p.t / p.sprintSteps;                                // :974
    if (max_sprints >= 0 && max_sprints < n_sprints) n_sprints = max_sprints;
    int64_t time = 0;                                                       // :973
    char row[96];
    for (int64_t i = 0; i < n_sprints; i++) {
        if (mrisim_run_sprint(ctx.c, p.sprintSteps, signal.data())) return die("sprint");
        for (int j = 0; j < rows_per_sprint; j++) {                         // :996-1023
            time += p.pfreq;
            const int len = mrisim_csv_row(time, p.tau, signal[(size_t)j], row, sizeof row);
            if (len > 0) fout.write(row, len);
        }
        fout.flush();
    }

    const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - start).count();
    std::cout << "Kernel execution complete! Elapsed time: " << ms << " ms" << std::endl;    // :1027-1029
    fout.close();                                                           // :1033
    std::cout << "Tree freed!" << std::endl;                                // :387 (finalizeGPU)
    return 0;
}

// The reference's C++ entry point.  Parameters are the shipped parameters.h values unless
// the caller links include/dropin/cuda_sim_dropin.cpp, which fills them from the
// reference's own compile-time ParameterStruct instead.
void simulateWaters(std::string filename) {
    mrisim_params p;
    mrisim_params_default(&p);
    std::random_device rd;
    const char* hs = std::getenv("MRISIM_HOST_SEED");
    const char* ws = std::getenv("MRISIM_WALKER_SEED");
    const uint64_t host_seed = hs ? std::strtoull(hs, nullptr, 0) : (uint64_t)std::time(nullptr) + rd();   // cuda_sim.cu:911-912
    const uint64_t walker_seed = ws ? std::strtoull(ws, nullptr, 0) : (uint64_t)std::time(nullptr);        // gpu_random.cu:22
    const char* ms = std::getenv("MRISIM_MAX_SPRINTS");
    if (mrisim_simulate_waters(filename.c_str(), &p, host_seed, walker_seed, 0, ms ? std::atoll(ms) : -1)) std::exit(1);
}
