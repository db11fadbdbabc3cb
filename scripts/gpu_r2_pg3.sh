#!/bin/bash
cd "$(dirname "$0")/.."
out=gpurun_out/pg3; mkdir -p $out
t() { echo "== $*" | tee -a $out/pg.log; env "$@" timeout 120 python scripts/profile_case.py $N 20000 4 0 2>/dev/null | grep -E "pre-sprint" | tail -2 | tee -a $out/pg.log; }
for wl in cfg1_default_4032_waters:4032 cfg2_8000_waters_intracellular_kick:8000; do
  export WORKLOAD=${wl%%:*}; N=${wl##*:}
  for c in 1 2 3; do for g in 4 8; do t MRISIM_PIPE_GROUPS=$g MRISIM_PIPE_CHUNKS=$c; done; done
  t MRISIM_PIPE_GROUPS=2 MRISIM_PIPE_CHUNKS=1
  t MRISIM_PIPE_GROUPS=12 MRISIM_PIPE_CHUNKS=2
done
