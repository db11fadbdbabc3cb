#!/bin/bash
cd "$(dirname "$0")/.."
out=gpurun_out/tl2; mkdir -p $out
for g in 1 2 4; do
  MRISIM_TIMELINE=$PWD/$out/timeline_groups$g.bin GROUPS=$g TASKS=0 timeout 300 python scripts/profile_case.py 1048576 1000 2 2>&1 | grep -E "pre-sprint" | tail -1
done
ls -la $out
