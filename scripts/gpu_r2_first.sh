#!/bin/bash
# round-2 first visit: the bit-level pin against the reference's kernels, the GPU suite, the bench line
set -u
cd "$(dirname "$0")/.."
tag=${1:-r2a}
out=gpurun_out/$tag
mkdir -p $out
nvidia-smi --query-gpu=name,clocks.max.sm,power.limit --format=csv > $out/gpu.txt 2>&1
(time python -m pytest tests/test_ref_gpu_bitpin.py -m gpu -x -q -rP) > $out/pytest_bitpin.log 2>&1; echo "bitpin exit $?" | tee -a $out/pytest_bitpin.log
tail -15 $out/pytest_bitpin.log
(time python -m pytest tests -m gpu -x -q -rP --deselect tests/test_ref_gpu_bitpin.py) > $out/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a $out/pytest_gpu.log
tail -3 $out/pytest_gpu.log
python bench.py > $out/bench.json 2> $out/bench.err; echo "bench exit $?"
cut -c1-600 $out/bench.json
