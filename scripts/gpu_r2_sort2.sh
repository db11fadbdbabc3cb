#!/bin/bash
cd "$(dirname "$0")/.."
out=gpurun_out/sort2; mkdir -p $out
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "sort or launch_granularity or size_independent or multi_sprint or short_launches" > $out/pytest_sort.txt 2>&1
tail -3 $out/pytest_sort.txt
r() { echo "== $*" | tee -a $out/matrix.log; env "$@" timeout 300 python scripts/profile_case.py 1048576 1000 3 2>&1 | grep -E "pre-sprint" | tail -2 | tee -a $out/matrix.log; }
for g in 1 2 3 4 6; do r WGROUPS=$g MRISIM_MIX=0; done
for g in 1 3 4; do r WGROUPS=$g MRISIM_MIX=1; done
r WGROUPS=3 RESORT=100
r WGROUPS=4 RESORT=100
r WGROUPS=4 RESORT=80
for wl in cfg4b_dense8000cells_1Mi_waters_per_gpu default_scene_1Mi_waters_per_gpu; do
  for g in 1 3 4; do WORKLOAD=$wl r WGROUPS=$g; done
  WORKLOAD=$wl r WGROUPS=4 MRISIM_MIX=1
done
