// T2_sim_main.cpp -- stand-alone executable with the behaviour of the reference's
// src/T2_GPU_sim.cu: name the output file after the run parameters, call simulateWaters.
// (The reference's own T2_GPU_sim.cu also builds unchanged against include/dropin/; this
// file exists so that a binary can be produced without the reference tree.)
#include <ctime>
#include <string>

#include "mrisim.h"

void simulateWaters(std::string fName);

// T2_GPU_sim.cu:28-49 naming convention: tau in ps, echo time 2*taucp, border, unix time
static std::string generate_base_filename(const mrisim_params& p) {
    std::string f("tau=");
    f += std::to_string((unsigned)(p.tau * 1e9));
    f += "ps_T-e=";
    f += std::to_string((unsigned)(2 * p.taucp));
    f += "largedipole";
    f += "_border=";
    f += std::to_string((unsigned)(p.border));
    f += "um_";
    f += "impermeablemultitrial";
    f += std::to_string((unsigned)(std::time(nullptr)));
    f += ".csv";
    return f;
}

int main(int argc, char** argv) {
    mrisim_params p;
    mrisim_params_default(&p);
    simulateWaters(argc > 1 ? std::string(argv[1]) : generate_base_filename(p));
    return 0;
}
