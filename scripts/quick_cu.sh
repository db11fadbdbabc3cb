#!/bin/bash
# compile csrc/mrisim.cu alone (20 s) and print the static SASS statistics of the hot kernel:  bash scripts/quick_cu.sh [extra -D flags]
cd "$(dirname "$0")/.."
nvcc -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -fmad=false -Xcompiler -fPIC,-ffp-contract=off,-O2 \
  -I include -I mri_nanoparticle_simulation_b200/csrc "$@" -c mri_nanoparticle_simulation_b200/csrc/mrisim.cu -o /tmp/quick_mrisim.o 2>&1 | grep -v "deprecated\|^$"
python scripts/sass_stats.py /tmp/quick_mrisim.o
