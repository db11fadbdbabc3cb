#!/bin/bash
# run the short profile case over occupancy variants of the library and re-sort cadences
cd "$(dirname "$0")/.."
for v in "" _mb6 _mb5; do
  [ -f mri_nanoparticle_simulation_b200/libmrisim$v.so ] || continue
  for r in ${RESORTS:-0 1000 500 250}; do
    echo "== variant libmrisim$v.so resort=$r"
    RESORT=$r MRISIM_LIB=$PWD/mri_nanoparticle_simulation_b200/libmrisim$v.so python scripts/profile_case.py ${1:-1048576} ${2:-1000} ${3:-1} ${4:-0} 2>/dev/null | tail -3
  done
done
