#!/bin/bash
# round 2, session 4: walker_team_kernel (P2+P3+P4a fused, W walkers per warp) -- parity, then timings per variant
cd "$(dirname "$0")/.."
out=gpurun_out/team; mkdir -p $out
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "pipelined or path_mode or full_default_sprint" > $out/pytest_team.txt 2>&1
tail -5 $out/pytest_team.txt
t() { # label, env...
  echo "== $*" | tee -a $out/team.log
  env "$@" timeout 120 python scripts/profile_case.py $N 20000 3 0 2>/dev/null | grep -E "pre-sprint|short" | tail -3 | tee -a $out/team.log
}
for wl in cfg1_default_4032_waters:4032 cfg2_8000_waters_intracellular_kick:8000; do
  export WORKLOAD=${wl%%:*}; N=${wl##*:}
  t MRISIM_PIPE_TEAM=0
  for v in 0 1 2 3 4 5 6; do t MRISIM_PIPE_TEAM=1 MRISIM_TEAM_VARIANT=$v; done
  t MRISIM_PIPE_TEAM=1 MRISIM_TEAM_VARIANT=-1
done
unset WORKLOAD
for g in 1 2; do
  MRISIM_TIMELINE=$PWD/$out/timeline8_groups$g.bin WGROUPS=$g TASKS=0 timeout 300 python scripts/profile_case.py 1048576 1000 1 2>&1 | grep -E "short" | tail -1
done
ls -la $out
