mkdir -p gpurun_out/v12
(timeout 900 python -m pytest tests/test_octree_gpu.py -m gpu -x -q 2>&1 | tail -6) > gpurun_out/v12/pytest_octree.log; tail -3 gpurun_out/v12/pytest_octree.log
timeout 900 python scripts/octree_bench.py default dense2000 --no-host dense8000 > gpurun_out/v12/octree_bench.jsonl 2> gpurun_out/v12/octree_bench.err; cut -c1-900 gpurun_out/v12/octree_bench.jsonl; tail -3 gpurun_out/v12/octree_bench.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/v12/bench_2gpu.json 2> gpurun_out/v12/bench2.err; echo "bench2 exit $?"; wc -l gpurun_out/v12/bench_2gpu.json; head -c 150 gpurun_out/v12/bench_2gpu.json
