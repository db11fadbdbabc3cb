mkdir -p gpurun_out/g8
nvidia-smi --query-gpu=index,name --format=csv,noheader | head -8 > gpurun_out/g8/gpus.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29531 scripts/sweep_config5.py --sprints 3 --warmup 1 > gpurun_out/g8/sweep_cfg5_8gpu.jsonl 2> gpurun_out/g8/sweep.err; echo "sweep exit $?"; cat gpurun_out/g8/sweep_cfg5_8gpu.jsonl | cut -c1-400; tail -3 gpurun_out/g8/sweep.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29532 bench.py --gpus 8 --steps 4 --warmup 3 > gpurun_out/g8/bench_8gpu.json 2> gpurun_out/g8/bench.err; echo "bench exit $?"; cut -c1-300 gpurun_out/g8/bench_8gpu.json
