#!/bin/bash
# few-walker (pipelined) path on the final tree: per-kernel time and DRAM bytes of ONE 20 000-step sprint at 4032 waters
# (ncu serialises the launches: the shares and the bytes are the evidence, not the wall time), then the plain bench line
cd "$(dirname "$0")/.."
out=gpurun_out/pipe_launches; mkdir -p $out
export WORKLOAD=cfg1_default_4032_waters
python scripts/profile_case.py 4032 20000 2 > $out/plain.log 2>&1 && tail -2 $out/plain.log &&
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv --log-file $out/launches_cfg1.csv \
    python scripts/profile_case.py 4032 20000 0 > $out/ncu.log 2>&1
unset WORKLOAD
python bench.py > $out/bench.json 2> $out/bench.err; echo "bench exit $?"
