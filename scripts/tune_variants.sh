#!/bin/bash
# run the short profile case with each occupancy variant of the library
cd "$(dirname "$0")/.."
for v in "" _mb4 _mb5 _mb6 _mb10; do
  echo "== variant libmrisim$v.so"
  MRISIM_LIB=$PWD/mri_nanoparticle_simulation_b200/libmrisim$v.so python scripts/profile_case.py ${1:-1048576} ${2:-1000} ${3:-1} ${4:-0} 2>/dev/null | tail -4
done
