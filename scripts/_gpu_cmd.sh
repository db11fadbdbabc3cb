mkdir -p gpurun_out/v20
(timeout 900 python -m pytest tests/test_octree_gpu.py -m gpu -x -q -rP -k "digests" 2>&1 | grep -E "passed|failed|dense8000|Error" | cut -c1-600) > gpurun_out/v20/pytest_octree.log; cat gpurun_out/v20/pytest_octree.log
