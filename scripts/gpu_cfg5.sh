#!/bin/bash
# BASELINE configs[4] on the driver's contract: 16 Mi waters IN TOTAL over N GPUs (strong scaling), one bench line per N
#   gpurun --gpus N -- 'bash scripts/gpu_cfg5.sh <tag> "<list of N>"'
cd "$(dirname "$0")/.."
tag=${1:-cfg5}; out=gpurun_out/$tag; mkdir -p $out
for n in ${2:-1}; do
  if [ "$n" = "1" ]; then
    python bench.py --gpus 1 --workload cfg5_16Mi_waters_total --steps 3 --warmup 3 --no-cpu-baseline > $out/bench_cfg5_${n}gpu.json 2> $out/bench_cfg5_${n}gpu.err
  else
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2955$n bench.py --gpus $n \
        --workload cfg5_16Mi_waters_total --steps 3 --warmup 3 --no-cpu-baseline > $out/bench_cfg5_${n}gpu.json 2> $out/bench_cfg5_${n}gpu.err
  fi
  echo "N=$n exit $?"; cut -c1-330 $out/bench_cfg5_${n}gpu.json; grep -E "clocks" -o $out/bench_cfg5_${n}gpu.json | head -1
done
if [ -n "${WEAK:-}" ]; then
  n=$WEAK
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2966$n bench.py --gpus $n --steps 5 --warmup 3 > $out/bench_weak_${n}gpu.json 2> $out/bench_weak_${n}gpu.err
  echo "weak N=$n exit $?"; cut -c1-330 $out/bench_weak_${n}gpu.json
fi
