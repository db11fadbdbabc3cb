#!/bin/bash
# round 2, session 4: overlapping walker groups in the pipelined path + GPU-side AoS/SoA of set/get_waters
cd "$(dirname "$0")/.."
out=gpurun_out/pg; mkdir -p $out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_group_checkpoint.py tests/test_abi.py -x -q -m gpu > $out/pytest_pg.txt 2>&1
tail -5 $out/pytest_pg.txt
t() { echo "== $*" | tee -a $out/pg.log; env "$@" timeout 120 python scripts/profile_case.py $N 20000 4 0 2>/dev/null | grep -E "pre-sprint" | tail -2 | tee -a $out/pg.log; }
for wl in cfg1_default_4032_waters:4032 cfg2_8000_waters_intracellular_kick:8000; do
  export WORKLOAD=${wl%%:*}; N=${wl##*:}
  for g in 1 2 4 8; do t MRISIM_PIPE_GROUPS=$g; done
  t MRISIM_PIPE_GROUPS=4 MRISIM_PIPE_CARVEOUT=28
  t MRISIM_PIPE_GROUPS=4 MRISIM_PIPE_CARVEOUT=72
  t MRISIM_PIPE_GROUPS=4 MRISIM_PIPE_CARVEOUT=-1
  t MRISIM_PIPE_GROUPS=6
done
export WORKLOAD=cfg1_default_4032_waters
for n in 16384 32768 49152; do N=$n; t MRISIM_PIPE_GROUPS=1; t MRISIM_PIPE_GROUPS=4; t MRISIM_PIPE_GROUPS=8; done
unset WORKLOAD
python bench.py --steps 3 --warmup 3 --no-cpu-baseline > $out/bench.json 2> $out/bench.err; cut -c1-1200 $out/bench.json
