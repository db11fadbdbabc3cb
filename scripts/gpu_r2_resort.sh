#!/bin/bash
cd "$(dirname "$0")/.."
out=gpurun_out/resort; mkdir -p $out
r() { echo "== $*" | tee -a $out/matrix.log; env "$@" timeout 300 python scripts/profile_case.py 1048576 1000 3 2>&1 | grep -E "pre-sprint" | tail -2 | tee -a $out/matrix.log; }
for rs in 125 90 62 45 31; do for g in 3 6; do r WGROUPS=$g RESORT=$rs; done; done
WORKLOAD=cfg4b_dense8000cells_1Mi_waters_per_gpu r WGROUPS=3 RESORT=62
WORKLOAD=cfg4b_dense8000cells_1Mi_waters_per_gpu r WGROUPS=6 RESORT=62
WORKLOAD=default_scene_1Mi_waters_per_gpu r WGROUPS=3 RESORT=62
WORKLOAD=default_scene_1Mi_waters_per_gpu r WGROUPS=6 RESORT=62
