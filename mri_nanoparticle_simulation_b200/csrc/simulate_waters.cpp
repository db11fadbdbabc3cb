// simulate_waters.cpp -- the reference's one exported function, `void simulateWaters(std::string)`
// (src/cuda_sim.h:74, body src/cuda_sim.cu:902-1034), rebuilt on the C ABI of include/mrisim.h.
//
// Same observable behaviour: the banners on stdout, one CSV row per pfreq steps
// ("<time*tau>,<sum cos phase>", default ostream formatting, cuda_sim.cu:1022), the elapsed
// time of the sprint loop.  Differences, all deliberate: seeds are explicit, the host syncs
// once per sprint instead of once per row (40 000 -> 2 000 syncs per default run), rows are
// written per sprint, every CUDA status is checked.
#include <unistd.h>

#include <chrono>
#include <cstdlib>
#include <cstring>
#include <ctime>
#include <fstream>
#include <iostream>
#include <random>
#include <string>
#include <vector>

#include "mrisim.h"

namespace {
struct SceneGuard { mrisim_scene* s = nullptr; ~SceneGuard() { if (s) mrisim_scene_free(s); } };
}

namespace {
struct GroupGuard { mrisim_group* g = nullptr; ~GroupGuard() { if (g) mrisim_group_destroy(g); } };

std::vector<int> parse_devices(const char* s) {
    std::vector<int> d;
    if (!s) return d;
    const char* q = s;
    while (*q) {
        char* end = nullptr;
        const long v = std::strtol(q, &end, 10);
        if (end == q) break;
        d.push_back((int)v);
        q = (*end == ',') ? end + 1 : end;
    }
    return d;
}
}  // namespace

extern "C" int mrisim_simulate_waters_multi(const char* filename, const mrisim_params* pp, uint64_t host_seed,
                                            uint64_t walker_seed, const int* devices, int n_devices, int64_t max_sprints,
                                            int64_t checkpoint_every, const char* checkpoint_path, const char* resume_path) {
    if (!filename || !pp || !devices || n_devices < 1) return 1;
    const mrisim_params& p = *pp;
    std::cout << "Starting GPU Simulation..." << std::endl;                 // cuda_sim.cu:903
    std::cout << "Printing to: " << filename << std::endl;                  // :904
    if (p.sprintSteps <= 0 || p.pfreq <= 0 || p.sprintSteps % p.pfreq != 0) {
        std::cerr << "simulateWaters: sprintSteps must be a multiple of the printing frequency (parameters.h:41)" << std::endl;
        return 1;
    }
    // a resumed run takes its seeds, its clock and its place in the CSV from the checkpoint
    mrisim_checkpoint_info ck;
    std::memset(&ck, 0, sizeof ck);
    const bool resume = resume_path && *resume_path;
    if (resume) {
        if (mrisim_checkpoint_info_read(resume_path, &ck)) { std::cerr << "simulateWaters: " << mrisim_checkpoint_last_error() << std::endl; return 1; }
        if (ck.time_steps % p.sprintSteps != 0) { std::cerr << "simulateWaters: the checkpoint is not on a sprint boundary of this run" << std::endl; return 1; }
        host_seed = ck.host_seed; walker_seed = ck.walker_seed;
        std::cout << "Resuming from " << resume_path << " at step " << ck.time_steps << " (" << ck.n_walkers << " waters, CSV offset " << ck.csv_bytes << ")" << std::endl;
    }
    std::fstream fout;                                                      // :905
    if (resume) {
        // keep the rows written before the checkpoint, drop anything after it
        { std::ofstream touch(filename, std::ios::app); }
        if (truncate(filename, (off_t)ck.csv_bytes) != 0) { std::cerr << "simulateWaters: cannot truncate " << filename << " to the checkpoint's offset" << std::endl; return 1; }
        fout.open(filename, std::ios::in | std::ios::out | std::ios::ate);
    } else {
        fout.open(filename, std::ios::out | std::ios::trunc);
    }
    if (!fout) { std::cerr << "simulateWaters: cannot open " << filename << std::endl; return 1; }

    SceneGuard scene;                                                       // :920-921
    const std::time_t t0 = std::time(nullptr);
    // the scene is built on the first GPU of the run (octree_build.cu / scene_gpu.cu: byte-identical to
    // the host build, 50x faster for dense boxes); MRISIM_HOST_SCENE=1 keeps the reference's host threads
    const char* host_scene = std::getenv("MRISIM_HOST_SCENE");
    const int rc_scene = (host_scene && host_scene[0] == '1') ? mrisim_scene_build(&p, host_seed, 0, &scene.s)
                                                              : mrisim_scene_build_gpu(&p, host_seed, 0, devices[0], &scene.s, nullptr);
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

    GroupGuard grp;                                                         // :916,924-926 -- `Handler handle(0)` becomes a device list
    if (mrisim_group_create(&p, devices, n_devices, &grp.g)) { std::cerr << "simulateWaters: " << mrisim_group_last_error(nullptr) << std::endl; return 1; }
    auto die = [&](const char* what) { std::cerr << "simulateWaters: " << what << ": " << mrisim_group_last_error(grp.g) << std::endl; return 1; };
    if (mrisim_group_set_scene(grp.g, scene.s)) return die("scene upload");
    std::cout << "Allocated GPU Octree!" << std::endl;                      // :368
    if (n_devices > 1) std::cout << "Walkers sharded over " << n_devices << " devices (" << mrisim_nccl_origin() << ")" << std::endl;
    if (resume) {
        if (mrisim_group_checkpoint_load(grp.g, resume_path)) return die("checkpoint");
    } else {
        int32_t n_water = 0;
        const mrisim_water* waters = mrisim_scene_waters(scene.s, &n_water);    // :934,961
        if (mrisim_group_set_waters(grp.g, waters, (uint64_t)n_water, walker_seed, 0)) return die("water upload");
    }

    std::cout << "Kernel prepped!" << std::endl;                            // :968
    std::cout << "Starting GPU computation..." << std::endl;                // :969
    const auto start = std::chrono::steady_clock::now();                    // timer.cpuStart(), :970

    const int rows_per_sprint = p.sprintSteps / p.pfreq;
    std::vector<double> signal((size_t)rows_per_sprint);
    int64_t n_sprints = p.t / p.sprintSteps;                                // :974
    if (max_sprints >= 0 && max_sprints < n_sprints) n_sprints = max_sprints;
    int64_t time = mrisim_group_get_time(grp.g);                            // :973 (0 unless resumed)
    const int64_t first_sprint = time / p.sprintSteps;
    const std::string ckpt_file = (checkpoint_path && *checkpoint_path) ? std::string(checkpoint_path) : std::string(filename) + ".ckpt";
    auto boundary = [&](int64_t i) { return checkpoint_every > 0 && (i + 1) % checkpoint_every == 0; };   // checkpoint after sprint i
    char row[96];
    // one sprint is kept in flight ahead of the one whose rows are being written, so the device never
    // waits for the host; a checkpoint needs the state of a finished sprint, so nothing runs ahead of it
    if (first_sprint < n_sprints && mrisim_group_submit_sprint(grp.g, p.sprintSteps)) return die("sprint");
    for (int64_t i = first_sprint; i < n_sprints; i++) {
        const bool ahead = i + 1 < n_sprints && !boundary(i);
        if (ahead && mrisim_group_submit_sprint(grp.g, p.sprintSteps)) return die("sprint");
        if (mrisim_group_wait_sprint(grp.g, signal.data())) return die("sprint");
        for (int j = 0; j < rows_per_sprint; j++) {                         // :996-1023
            time += p.pfreq;
            const int len = mrisim_csv_row(time, p.tau, signal[(size_t)j], row, sizeof row);
            if (len > 0) fout.write(row, len);
        }
        fout.flush();
        if (boundary(i)) {
            const uint64_t csv_bytes = (uint64_t)fout.tellp();
            if (mrisim_group_checkpoint_save(grp.g, ckpt_file.c_str(), host_seed, csv_bytes)) return die("checkpoint");
            if (i + 1 < n_sprints && mrisim_group_submit_sprint(grp.g, p.sprintSteps)) return die("sprint");
        }
    }

    const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - start).count();
    std::cout << "Kernel execution complete! Elapsed time: " << ms << " ms" << std::endl;    // :1027-1029
    fout.close();                                                           // :1033
    std::cout << "Tree freed!" << std::endl;                                // :387 (finalizeGPU)
    return 0;
}

extern "C" int mrisim_simulate_waters(const char* filename, const mrisim_params* pp, uint64_t host_seed,
                                      uint64_t walker_seed, int device, int64_t max_sprints) {
    // options the reference's one-argument entry point has no room for arrive through the environment
    std::vector<int> devs = parse_devices(std::getenv("MRISIM_DEVICES"));
    if (devs.empty()) devs.push_back(device);
    const char* every = std::getenv("MRISIM_CHECKPOINT_EVERY");
    return mrisim_simulate_waters_multi(filename, pp, host_seed, walker_seed, devs.data(), (int)devs.size(), max_sprints,
                                        every ? std::atoll(every) : 0, std::getenv("MRISIM_CHECKPOINT"), std::getenv("MRISIM_RESUME"));
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
