// TEST INFRASTRUCTURE.  Entry point for the reference's own GPU simulation (cuda_sim.cu +
// gpu_random.cu, compiled where they lie by `make -C oracle refgpu`): same call as
// T2_GPU_sim.cu:58-61, with the output file name taken from argv instead of the wall clock.
//   ref_gpu_<cfg> out.csv      prints the reference's own banner lines incl. "Elapsed time: <ms> ms"
#include <cstdio>
#include <string>
#include "utilities/utilities.h"
#include "parameters.h"
#include "cuda_sim.h"

int main(int argc, char** argv) {
    if (argc < 2) { std::fprintf(stderr, "usage: %s <out.csv>\n", argv[0]); return 2; }
    std::printf("ref config: num_water=%d num_cells=%d sprintSteps=%d t=%d tcp=%d\n", p.num_water, p.num_cells, p.sprintSteps, p.t, p.tcp);
    simulateWaters(std::string(argv[1]));
    return 0;
}
