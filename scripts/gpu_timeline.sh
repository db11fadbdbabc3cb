#!/bin/bash
cd "$(dirname "$0")/.."
tag=${1:-tl}; out=gpurun_out/$tag; mkdir -p $out
for wl in cfg4a_dense2000cells_1Mi_waters_per_gpu cfg4b_dense8000cells_1Mi_waters_per_gpu; do
  MRISIM_TIMELINE=$PWD/$out/timeline_${wl%%_*}.bin WORKLOAD=$wl python scripts/profile_case.py 1048576 125 1 2>&1 | tail -2
done
ls -la $out
