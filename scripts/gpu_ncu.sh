#!/bin/bash
# one ncu --set full capture of the walker kernel on the bench scene:  gpurun -- 'bash scripts/gpu_ncu.sh <tag> [field_mode]'
cd "$(dirname "$0")/.."
tag=${1:-run}; mode=${2:-0}
mkdir -p gpurun_out/$tag
python scripts/profile_case.py 1048576 1000 1 $mode > gpurun_out/$tag/plain_profile_case.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:walker_sprint -s 1 -c 1 -o gpurun_out/$tag/prof \
    python scripts/profile_case.py 1048576 1000 1 $mode > gpurun_out/$tag/ncu_full.log 2>&1
cat gpurun_out/$tag/plain_profile_case.log
