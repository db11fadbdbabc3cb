#!/bin/bash
cd "$(dirname "$0")/.."
out=gpurun_out/pg2; mkdir -p $out
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "pipelin or path_mode or full_default" > $out/pytest_pg.txt 2>&1
tail -3 $out/pytest_pg.txt
t() { echo "== $*" | tee -a $out/pg.log; env "$@" timeout 120 python scripts/profile_case.py $N 20000 4 0 2>/dev/null | grep -E "pre-sprint" | tail -2 | tee -a $out/pg.log; }
for wl in cfg1_default_4032_waters:4032 cfg2_8000_waters_intracellular_kick:8000; do
  export WORKLOAD=${wl%%:*}; N=${wl##*:}
  for g in 1 2 3 4 6 8; do t MRISIM_PIPE_GROUPS=$g; done
  t MRISIM_PIPE_GROUPS=4 MRISIM_PIPE_BUDGET_GIB=2
  t MRISIM_PIPE_GROUPS=4 MRISIM_PIPE_CARVEOUT=58
done
