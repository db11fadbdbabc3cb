#!/bin/bash
# round 2, session 4: own radix sort (small shared memory) + overlapping launch chains of walker groups
cd "$(dirname "$0")/.."
out=gpurun_out/sort; mkdir -p $out
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "sort or launch_granularity or size_independent or multi_sprint or short_launches" > $out/pytest_sort.txt 2>&1
tail -5 $out/pytest_sort.txt
r() { echo "== $*" | tee -a $out/matrix.log; env "$@" timeout 300 python scripts/profile_case.py 1048576 1000 3 2>&1 | grep -E "pre-sprint" | tail -2 | tee -a $out/matrix.log; }
r GROUPS=1
r GROUPS=2
r GROUPS=3
r GROUPS=4
r GROUPS=2 RESORT=100
r GROUPS=2 RESORT=62
r GROUPS=4 RESORT=62
r GROUPS=1 MRISIM_CARVEOUT=-1
r GROUPS=2 MRISIM_CARVEOUT=-1
r GROUPS=2 MRISIM_CARVEOUT=14
MRISIM_TIMELINE=$PWD/$out/timeline8_sort_g2.bin GROUPS=2 timeout 300 python scripts/profile_case.py 1048576 1000 1 2>&1 | grep -E "short" | tail -1
WORKLOAD=cfg4b_dense8000cells_1Mi_waters_per_gpu r GROUPS=1
WORKLOAD=cfg4b_dense8000cells_1Mi_waters_per_gpu r GROUPS=2
WORKLOAD=default_scene_1Mi_waters_per_gpu r GROUPS=1
WORKLOAD=default_scene_1Mi_waters_per_gpu r GROUPS=2
ls -la $out
