#!/bin/bash
# bench the short profile case over library variants:  VARIANTS="base ld" bash scripts/gpu_variants.sh <tag>
cd "$(dirname "$0")/.."
tag=${1:-var}; out=gpurun_out/$tag; mkdir -p $out
for v in ${VARIANTS:-""}; do
  lib=mri_nanoparticle_simulation_b200/libmrisim${v:+_$v}.so
  [ "$v" = "default" ] && lib=mri_nanoparticle_simulation_b200/libmrisim.so
  for g in ${WGROUPS_LIST:-1}; do for r in ${RESORTS:--1}; do
    echo "== $v groups=$g resort=$r workload=${WORKLOAD:-cfg4a}" | tee -a $out/variants.log
    WGROUPS=$g RESORT=$r MRISIM_LIB=$PWD/$lib python scripts/profile_case.py ${WALKERS:-1048576} 1000 ${PRE:-2} 0 2>/dev/null | grep pre-sprint | tee -a $out/variants.log
  done; done
done
