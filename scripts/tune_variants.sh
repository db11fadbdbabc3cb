#!/bin/bash
# run the short profile case over every built variant of the library and re-sort cadences
cd "$(dirname "$0")/.."
for lib in mri_nanoparticle_simulation_b200/libmrisim*.so; do
  for r in ${RESORTS:-250}; do
    echo "== $(basename $lib) resort=$r field_mode=${4:-0}"
    RESORT=$r MRISIM_LIB=$PWD/$lib python scripts/profile_case.py ${1:-1048576} ${2:-1000} ${3:-1} ${4:-0} 2>/dev/null | tail -2
  done
done
