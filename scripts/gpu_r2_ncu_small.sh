#!/bin/bash
# ncu --set full of the library's own sort kernels (one launch each, 1 Mi walkers, 3 groups) and of walker_path_kernel (4032 waters)
cd "$(dirname "$0")/.."
out=gpurun_out/ncu_small; mkdir -p $out
ncu --set full --clock-control none --import-source on -k regex:sort_ -s 30 -c 3 -o $out/prof_sort \
    python scripts/profile_case.py 1048576 1000 0 > $out/ncu_sort.log 2>&1
WORKLOAD=cfg1_default_4032_waters ncu --set full --clock-control none --import-source on -k regex:walker_path -s 6 -c 1 -o $out/prof_path \
    python scripts/profile_case.py 4032 20000 0 > $out/ncu_path.log 2>&1
ncu -i $out/prof_sort.ncu-rep --page details > $out/sort_details.txt 2>&1
ncu -i $out/prof_path.ncu-rep --page details > $out/path_details.txt 2>&1
ls -la $out
