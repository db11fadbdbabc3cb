mkdir -p gpurun_out/g4
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus 4 --steps 4 --warmup 3 > gpurun_out/g4/bench_4gpu.json 2> gpurun_out/g4/bench.err; echo "bench exit $?"; wc -l gpurun_out/g4/bench_4gpu.json; cut -c1-200 gpurun_out/g4/bench_4gpu.json
