#!/bin/bash
# walker_path_kernel: CTAs per SM the register allocation aims at (7 = 71 registers shipped; 8 = 64, 10 = 48, 12 = 40 with spills)
cd "$(dirname "$0")/.."
t() { echo "== $*"; env "$@" timeout 120 python scripts/profile_case.py $N 20000 4 0 2>/dev/null | grep -E "pre-sprint" | tail -1; }
for v in "" _pm8 _pm10 _pm12; do
  L=$PWD/mri_nanoparticle_simulation_b200/libmrisim$v.so
  for wl in cfg1_default_4032_waters:4032 cfg2_8000_waters_intracellular_kick:8000 cfg1_default_4032_waters:16384 cfg1_default_4032_waters:49152; do
    N=${wl##*:}; t MRISIM_LIB=$L WORKLOAD=${wl%%:*} NW=$N
  done
done
