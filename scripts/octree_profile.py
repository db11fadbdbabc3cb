"""One GPU build of the dense 2000-cell scene, nothing else: the target of the ncu captures of
the octree kernels (scripts/_gpu_cmd.sh).  Launch order per level: criterion, leaf_field; the
levels of this scene launch 4/29/126/308/219/5 and 4/100/698/2246/1741/37 CTAs."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mri_nanoparticle_simulation_b200 import api  # noqa: E402

cells = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
p = api.ParameterStruct(num_cells=cells, hashDim=40, maxNeighbors=16 if cells <= 2000 else 32, num_water=4096)
box = api.BacteriaBox(p, 11).populateSimulation(device=0)
print(box.octree_build_stats.asdict())
