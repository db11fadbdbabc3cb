#!/bin/bash
cd "$(dirname "$0")/.."
out=gpurun_out/tl3; mkdir -p $out
r() { echo "== $*"; env "$@" TASKS=0 timeout 300 python scripts/profile_case.py 1048576 1000 2 2>&1 | grep -E "pre-sprint|short" | tail -2; }
r GROUPS=2 MRISIM_NOSORT=1 MRISIM_TIMELINE=$PWD/$out/tl_nosort_g2.bin
r GROUPS=1 MRISIM_NOSORT=1
for c in 25 50 86; do
r GROUPS=2 MRISIM_CARVEOUT=$c MRISIM_TIMELINE=$PWD/$out/tl_carve${c}_g2.bin
r GROUPS=1 MRISIM_CARVEOUT=$c
done
ls -la $out
