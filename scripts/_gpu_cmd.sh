mkdir -p gpurun_out/pipe8
for gib in 1 2 4 8; do
  export MRISIM_PIPE_BUDGET_GIB=$gib
  timeout 600 python bench.py --workload cfg1_default_4032_waters --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/pipe8/bench_cfg1_b${gib}.json 2>> gpurun_out/pipe8/bench.err
  python -c "import json,sys; d=json.load(open('gpurun_out/pipe8/bench_cfg1_b${gib}.json')); print('cfg1 budget $gib GiB',d['value'],d['ms_per_step'],d['gpu_launches'])"
  timeout 600 python bench.py --workload cfg1_default_4032_waters --walkers-per-gpu 32768 --path-mode pipeline --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/pipe8/bench_n32768_b${gib}.json 2>> gpurun_out/pipe8/bench.err
  python -c "import json,sys; d=json.load(open('gpurun_out/pipe8/bench_n32768_b${gib}.json')); print('n=32768 budget $gib GiB',d['value'],d['ms_per_step'],d['gpu_launches'])"
done
