"""Small ragged end-to-end case: GPU scene build (the octree kernel variants a small scene reaches),
GPU water throw of 333 waters (not a multiple of 32 or of any block size), one pipelined and one fused
sprint, signals printed side by side.  Written as a compute-sanitizer target (memcheck / racecheck);
the sanitizer is closed on this GPU pool, so it is run plain and the parity suite does the checking."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mri_nanoparticle_simulation_b200 import api  # noqa: E402

p = api.low_permeability(api.ParameterStruct(num_water=300, num_cells=40, bound=10, water_start_bound=6, hashDim=8, border=3,
                                              avoid_intracellular_throw=0))
box = api.BacteriaBox(p, 3).populateSimulation(device=0)
box.rethrow_waters(333, 5, device=0)                 # ragged: not a multiple of 32 or of the block sizes
w = box.getWaters().copy()
for mode in (api.PATH_PIPELINE, api.PATH_FUSED):
    ctx = api.WalkerContext(p, 0)
    ctx.set_scene(box)
    ctx.set_waters(w, 0xC0FFEE)
    ctx.set_path_mode(mode)
    sig = ctx.run_sprint(2 * p.pfreq)
    print(mode, ctx.pipeline_sprints, float(sig[0]), float(sig[-1]))
print("octree", box.octree_build_stats.asdict()["n_nodes"])
