#!/bin/bash
# build tuning variants of the library: each argument "U:BLOCK:MINBLOCKS" -> libmrisim_u<U>b<BLOCK>m<MINBLOCKS>.so
cd "$(dirname "$0")/.."
for cfg in "$@"; do
  IFS=: read u b m <<< "$cfg"
  MRISIM_BLOCK=$b MRISIM_MIN_BLOCKS=$m MRISIM_LIB_SUFFIX=_u${u}b${b}m${m} python -m mri_nanoparticle_simulation_b200.build > /dev/null 2>&1 &
done
wait
ls mri_nanoparticle_simulation_b200/*.so
