"""Small, ncu-friendly invocation of the walker kernel on the bench scene.
   python scripts/profile_case.py [walkers] [steps] [presprints] [field_mode]
Runs `presprints` sprints of 20000 steps first (so walkers have diffused like in the timed
region of bench.py), then 3 short launches of `steps` steps -- profile those with
   ncu -k regex:walker_sprint -s <presprints> -c 1 ...
"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
from mri_nanoparticle_simulation_b200 import api

walkers = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 1000
pre = int(sys.argv[3]) if len(sys.argv) > 3 else 1
mode = int(sys.argv[4]) if len(sys.argv) > 4 else 0
wl = os.environ.get("WORKLOAD", bench.DEFAULT_WORKLOAD)
box, p, _ = bench.get_scene(api, wl, 0, 1, lambda: None, device=0)
waters = bench.throw_waters(box, p, walkers, 0, device=0)
ctx = api.WalkerContext(p, 0); ctx.set_field_mode(mode); ctx.set_resort_steps(int(os.environ.get('RESORT', '-1')))
if 'WGROUPS' in os.environ: ctx.set_overlap_groups(int(os.environ['WGROUPS']))    # default: the library's (not GROUPS: bash keeps that name to itself and ignores assignments)
ctx.set_scene(box); ctx.set_waters(waters, bench.WALKER_SEED)
for _ in range(pre):
    ctx.run_sprint(p.sprintSteps)
    print(f"pre-sprint kernel {ctx.last_sprint_kernel_ms():.1f} ms -> {walkers*p.sprintSteps/ctx.last_sprint_kernel_ms()*1e3:.3e} steps/s")
for _ in range(3):
    ctx.run_sprint(steps)
    print(f"short launch ({steps} steps) kernel {ctx.last_sprint_kernel_ms():.2f} ms -> {walkers*steps/ctx.last_sprint_kernel_ms()*1e3:.3e} steps/s")
