mkdir -p gpurun_out/v19
(timeout 900 python -m pytest tests/test_octree_gpu.py -m gpu -x -q 2>&1 | tail -6) > gpurun_out/v19/pytest_octree.log; tail -2 gpurun_out/v19/pytest_octree.log
timeout 900 python scripts/octree_bench.py default dense2000 --no-host dense8000 dense8000 > gpurun_out/v19/octree_bench.jsonl 2> gpurun_out/v19/octree_bench.err
python - <<'PY'
import json
for l in open('gpurun_out/v19/octree_bench.jsonl'):
    d=json.loads(l); o=d['gpu_octree']
    print(d['scene'], d['gpu_scene_build_s'], d.get('host_scene_build_s'), d.get('byte_identical'), {k:round(o[k],1) for k in ('total_ms','gpu_ms','fixup_ms','fixup_wait_ms','linearise_ms','scene_throw_ms','scene_lookup_ms','scene_sort_ms')}, d['sha256'][:12])
PY
