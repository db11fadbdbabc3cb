mkdir -p gpurun_out/oct1
(timeout 900 python -m pytest tests/test_octree_gpu.py -m gpu -x -q -rP 2>&1 | tail -60) > gpurun_out/oct1/pytest.log; tail -30 gpurun_out/oct1/pytest.log
timeout 600 python scripts/octree_bench.py default dense2000 > gpurun_out/oct1/octree_bench.jsonl 2> gpurun_out/oct1/octree_bench.err; cat gpurun_out/oct1/octree_bench.jsonl; tail -5 gpurun_out/oct1/octree_bench.err
