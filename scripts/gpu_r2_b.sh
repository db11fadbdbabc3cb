#!/bin/bash
# round-2 visit: new C-ABI features (submit/wait, checkpoint, group) + full suite + bench line
set -u
cd "$(dirname "$0")/.."
tag=${1:-r2b}
out=gpurun_out/$tag
mkdir -p $out
nvidia-smi --query-gpu=name,clocks.max.sm,power.limit --format=csv > $out/gpu.txt 2>&1
(time python -m pytest tests/test_group_checkpoint.py -m gpu -x -q -rP) > $out/pytest_group.log 2>&1; echo "group exit $?" | tee -a $out/pytest_group.log
tail -25 $out/pytest_group.log
(time python -m pytest tests -m gpu -x -q --deselect tests/test_group_checkpoint.py) > $out/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a $out/pytest_gpu.log
tail -5 $out/pytest_gpu.log
python bench.py > $out/bench.json 2> $out/bench.err; echo "bench exit $?"
tail -5 $out/bench.err
cut -c1-1500 $out/bench.json
