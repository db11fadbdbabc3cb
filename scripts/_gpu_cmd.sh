mkdir -p gpurun_out/v16
python - > gpurun_out/v16/repeat8000.log 2>&1 <<'PY'
import sys, time
sys.path.insert(0, '.')
from mri_nanoparticle_simulation_b200 import api
p = api.ParameterStruct(num_cells=8000, hashDim=40, maxNeighbors=32, num_water=4096)
for rep in range(4):
    t0 = time.perf_counter()
    box = api.BacteriaBox(p, 11).populateSimulation(device=0)
    dt = time.perf_counter() - t0
    o = box.octree_build_stats.asdict()
    print(rep, round(dt, 3), {k: round(o[k], 1) for k in ('total_ms', 'gpu_ms', 'fixup_ms', 'fixup_wait_ms', 'linearise_ms', 'scene_throw_ms', 'scene_lookup_ms', 'scene_sort_ms')}, flush=True)
    if rep == 1:
        time.sleep(4.0)      # an idle gap, like a host build in between
PY
cat gpurun_out/v16/repeat8000.log
