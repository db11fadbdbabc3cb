#!/bin/bash
# 2-GPU visit: the device group / communicator tests and the N=2 bench line
set -u
cd "$(dirname "$0")/.."
tag=${1:-r2c}
out=gpurun_out/$tag
mkdir -p $out
nvidia-smi --query-gpu=index,name --format=csv > $out/gpu.txt 2>&1
(time python -m pytest tests/test_group_checkpoint.py -m gpu -x -q -rP -k "two_device or dropin_on_two or one_process_per_gpu") > $out/pytest_group2.log 2>&1; echo "group2 exit $?" | tee -a $out/pytest_group2.log
grep -E "passed|failed|skipped|NCCL|Error|error" $out/pytest_group2.log | tail -15
NCCL_DEBUG=INFO python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 5 --warmup 3 > $out/bench_2gpu.json 2> $out/bench_2gpu.err; echo "bench2 exit $?"
grep -E "\[bench\]|NVLS|via P2P|Connected all" $out/bench_2gpu.err | head -8
cut -c1-900 $out/bench_2gpu.json
