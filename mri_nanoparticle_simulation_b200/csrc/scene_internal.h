// scene_internal.h -- seam between the host scene builder (scene.cpp, g++) and the GPU octree
// builder (octree_build.cu, nvcc).  Not part of the C ABI.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

#include "mrisim.h"
#include "mrisim_math.h"

namespace mrisim_host {

// The octree of a scene in the canonical MRISCN01 form (what build_octree() of scene.cpp fills).
struct OctreeOut {
    int32_t min_depth = 0, max_depth = 0;
    std::vector<uint32_t> subtree_sizes;   // [8^min_depth]
    std::vector<mrisim_node> nodes;        // subtrees back to back, each in depth-first order
    std::vector<mrisim_mnp> residents;     // leaf resident lists in node order
};

// Constants of `new Octree(2e-6, 5, .002, ...)` (SimulationBox.cpp:86-88).
constexpr double kMaxProduct = 2e-6, kMaxG = 5, kMinG = .002;

// ---- host (glibc) evaluations, implemented in scene.cpp with the reference's own expressions
// and libm pow(); the GPU builder uses them for thresholds and for every value it cannot certify.
void   octree_depths(const mrisim_params& p, int32_t* min_depth, int32_t* max_depth);   // octree.cpp:400-401
void   octree_root(const mrisim_params& p, int min_depth, int offset,
                   double* x, double* y, double* z, double* g, uint64_t* mc);            // octree.cpp:346-367
double octree_space_limit(const mrisim_params& p, double g);   // pow(g/2*sqrt(3) + scale*cell_r, 2), octree.cpp:173
double octree_grad_limit(const mrisim_params& p);              // pow(scale*cell_r, 2), octree.cpp:240
// g/2 * 1e-6 * grad(x+g/2, y+g/2, z+g/2)   (octree.cpp:267 with :228-252)
double octree_criterion(const mrisim_params& p, const std::vector<mrisim_mnp>& mnps,
                        double x, double y, double z, double g);
// field(xn,yn,zn) minus the residents' own terms (octree.cpp:279-296); res = indices into mnps
double octree_leaf_B(const mrisim_params& p, const std::vector<mrisim_mnp>& mnps, const int32_t* res, size_t n_res,
                     double xn, double yn, double zn);

// ---- GPU builder (octree_build.cu)
int build_octree_gpu(const mrisim_params& p, const std::vector<mrisim_mnp>& mnps, int device, int n_threads,
                     OctreeOut& out, mrisim_octree_build_stats* stats, std::string& err);

// ---- the other O(N x cells) loops of the scene build on the GPU (scene_gpu.cu)
int gpu_build_lookup(const mrisim_params& p, const std::vector<double>& cells, int device,
                     std::vector<int32_t>& lookup, int* overflow, std::string& err);
int gpu_throw_waters(const mrisim_params& p, const std::vector<double>& cells, mrisim_rng* gen, int64_t num_water,
                     int device, std::vector<mrisim_water>& waters, std::string& err);
int gpu_leaf_keys(const mrisim_params& p, int min_depth, const std::vector<uint64_t>& prefix,
                  const std::vector<mrisim_node>& nodes, const std::vector<mrisim_water>& waters, int device,
                  std::vector<uint64_t>& keys, std::string& err);

}  // namespace mrisim_host
