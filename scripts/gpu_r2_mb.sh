#!/bin/bash
# register-target variants of the fused kernel (CTAs per SM the allocation aims at) on the final tree
cd "$(dirname "$0")/.."
r() { echo "== $*"; env "$@" timeout 300 python scripts/profile_case.py 1048576 1000 3 2>&1 | grep -E "pre-sprint" | tail -1; }
for v in "" _mb7 _mb6; do
  L=$PWD/mri_nanoparticle_simulation_b200/libmrisim$v.so
  r MRISIM_LIB=$L
  r MRISIM_LIB=$L WORKLOAD=cfg4b_dense8000cells_1Mi_waters_per_gpu
  r MRISIM_LIB=$L WORKLOAD=default_scene_1Mi_waters_per_gpu
done
