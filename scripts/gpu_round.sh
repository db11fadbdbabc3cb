#!/bin/bash
# One GPU-box visit: parity suite, bench, launch list, full ncu capture of the hot kernel.
#   gpurun --timeout 1800 -- 'bash scripts/gpu_round.sh <tag>'
set -u
cd "$(dirname "$0")/.."
tag=${1:-run}
out=gpurun_out/$tag
mkdir -p $out
nvidia-smi --query-gpu=name,clocks.max.sm,power.limit --format=csv > $out/gpu.txt 2>&1
(time python -m pytest tests -m gpu -x -q -rP) > $out/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a $out/pytest_gpu.log
python bench.py > $out/bench.json 2> $out/bench.err; echo "bench exit $?"
cat $out/bench.json
python bench.py --field-mode fast --no-cpu-baseline --steps 3 > $out/bench_fast.json 2>> $out/bench.err
if [ "${NCU:-1}" = "1" ]; then
python bench.py --steps 2 --warmup 1 --no-cpu-baseline > $out/plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file $out/launches.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline > $out/ncu_launches.log 2>&1
# the instruction mix is a property of the walkers, not of how their launches overlap: captured with ONE group (all walkers in one launch)
WGROUPS=1 python scripts/profile_case.py 1048576 1000 1 > $out/plain_profile_case.log 2>&1 &&
WGROUPS=1 ncu --set full --clock-control none --import-source on -k regex:walker_sprint -s 165 -c 1 -o $out/prof \
    python scripts/profile_case.py 1048576 1000 1 > $out/ncu_full.log 2>&1
cat $out/plain_profile_case.log
fi
MRISIM_TIMELINE=$PWD/$out/timeline8_final.bin timeout 300 python scripts/profile_case.py 1048576 1000 1 2>&1 | grep short | tail -1
for wl in cfg4b_dense8000cells_1Mi_waters_per_gpu default_scene_1Mi_waters_per_gpu cfg3_lowperm_100k_waters; do python bench.py --workload $wl --steps 3 --warmup 3 --no-cpu-baseline --no-small-configs > $out/bench_${wl%%_*}.json 2>> $out/bench.err; cut -c1-260 $out/bench_${wl%%_*}.json; echo; done
tail -3 $out/pytest_gpu.log
