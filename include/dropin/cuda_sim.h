/*
 * Drop-in replacement for the reference's src/cuda_sim.h: the one function its entry point
 * calls (src/T2_GPU_sim.cu:58-61).  Put this directory on the include path INSTEAD of the
 * reference's cuda_sim.h, keep the project's own parameters.h, compile
 * cuda_sim_dropin.cpp with it and link libmrisim.so.
 */
#ifndef SIM_HEADER
#define SIM_HEADER

#include <string>

#include "parameters.h"

void simulateWaters(std::string fName);      /* src/cuda_sim.h:74 */

#endif
