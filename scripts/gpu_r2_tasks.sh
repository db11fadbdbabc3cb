#!/bin/bash
# round 2, session 4: ticketed tasks inside a launch -- correctness, then the timing matrix, timelines, ncu counters
cd "$(dirname "$0")/.."
out=gpurun_out/tasks; mkdir -p $out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > $out/gpu.txt
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "tasks_inside or launch_granularity or size_independent" > $out/pytest_tasks.txt 2>&1
tail -3 $out/pytest_tasks.txt
run() {  # tasks resort groups ranknear
  echo "== tasks=$1 resort=$2 groups=$3 ranknear=$4" | tee -a $out/matrix.log
  MRISIM_RANK_NEAR=$4 TASKS=$1 RESORT=$2 GROUPS=$3 timeout 300 python scripts/profile_case.py 1048576 1000 3 0 2>/dev/null | grep pre-sprint | tail -2 | tee -a $out/matrix.log
}
run 0 125 1 0
run 25 125 1 0
run 42 125 1 0
run 63 125 1 0
run 0 125 2 0
run 25 125 2 0
run 0 62 1 0
run 31 62 1 0
run 21 62 1 0
run 31 62 2 0
run 25 100 1 0
run 20 80 1 0
run 0 125 1 1
run 25 125 1 1
run 31 62 1 1
for t in 0 25; do
  MRISIM_TIMELINE=$PWD/$out/timeline_tasks$t.bin TASKS=$t timeout 300 python scripts/profile_case.py 1048576 1000 1 2>&1 | tail -1
done
for t in 0 25; do
  TASKS=$t timeout 600 ncu -k regex:walker_sprint -s 165 -c 1 --clock-control none --metrics smsp__inst_executed.sum,sm__cycles_active.avg,sm__cycles_elapsed.avg,gpu__time_duration.sum,smsp__thread_inst_executed_per_inst_executed.ratio,sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active --csv --log-file $out/ncu_tasks$t.csv python scripts/profile_case.py 1048576 1000 1 > /dev/null 2>&1
done
ls -la $out
