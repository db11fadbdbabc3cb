#!/bin/bash
# round-2 evidence visit: cfg4b parity + bench + ncu (L2 question), ncu of the default workload on HEAD, launch list
set -u
cd "$(dirname "$0")/.."
tag=${1:-r2e}; out=gpurun_out/$tag; mkdir -p $out
python -m pytest tests/test_gpu_parity.py -m gpu -x -q -rP -k "cfg4b" > $out/pytest_cfg4b.log 2>&1; echo "cfg4b exit $?"; tail -4 $out/pytest_cfg4b.log
python bench.py --workload cfg4b_dense8000cells_1Mi_waters_per_gpu --steps 3 --warmup 3 --no-cpu-baseline > $out/bench_cfg4b.json 2> $out/bench_cfg4b.err; echo "bench4b exit $?"; cut -c1-400 $out/bench_cfg4b.json
# full ncu capture, default workload (cfg4a), one 125-step launch of the second sprint
python scripts/profile_case.py 1048576 1000 1 > $out/plain_profile_case.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:walker_sprint -s 1 -c 1 -o $out/prof_cfg4a \
    python scripts/profile_case.py 1048576 1000 1 > $out/ncu_full_cfg4a.log 2>&1
cat $out/plain_profile_case.log
# same for cfg4b
WORKLOAD=cfg4b_dense8000cells_1Mi_waters_per_gpu python scripts/profile_case.py 1048576 1000 1 > $out/plain_profile_case_4b.log 2>&1 &&
WORKLOAD=cfg4b_dense8000cells_1Mi_waters_per_gpu ncu --set full --clock-control none --import-source on -k regex:walker_sprint -s 1 -c 1 -o $out/prof_cfg4b \
    python scripts/profile_case.py 1048576 1000 1 > $out/ncu_full_cfg4b.log 2>&1
cat $out/plain_profile_case_4b.log
# launch list of the default bench command
python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-small-configs > $out/plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/launches.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-small-configs > $out/ncu_launches.log 2>&1
ls -la $out
