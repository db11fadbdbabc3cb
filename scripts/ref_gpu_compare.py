"""Informational: the reference's own GPU code (oracle/_ref/ref_gpu_*) and this repo's drop-in
simulateWaters on the same configs, same box, back to back.  Writes gpurun_out/refgpu/compare.json.
   python scripts/ref_gpu_compare.py"""
import json, os, re, subprocess, sys, tempfile, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mri_nanoparticle_simulation_b200 import api

CONFIGS = {   # name -> (binary, ParameterStruct overrides)  (GCFG_* of oracle/Makefile)
    "default_4032_waters_2ms": ("ref_gpu_short", dict(totaltime=2)),
    "65536_waters_6ms_sprint2000": ("ref_gpu_xcheck", dict(num_water=65536, sprintSteps=2000, totaltime=6)),
}
out = {"what": "reference cuda_sim.cu (cuRAND + cuBLAS + thrust, recompiled for sm_100) vs this repo's simulateWaters drop-in; "
               "rates use each program's own 'Elapsed time' banner (sprint loop only, cuda_sim.cu:970-1027)", "configs": {}}
tmp = tempfile.mkdtemp()
for name, (exe, over) in CONFIGS.items():
    p = api.ParameterStruct(**over)
    steps = p.num_water * p.t
    r = subprocess.run([os.path.join(ROOT, "oracle", "_ref", exe), os.path.join(tmp, "ref.csv")], capture_output=True, text=True, cwd=tmp)
    ref_ms = float(re.search(r"Elapsed time: ([0-9.eE+]+) ms", r.stdout).group(1))
    # ours prints the same banner from the C++ drop-in: capture it through a child process
    code = ("import sys; sys.path.insert(0, %r); from mri_nanoparticle_simulation_b200 import api; "
            "api.simulateWaters(%r, api.ParameterStruct(**%r), host_seed=12345)" % (ROOT, os.path.join(tmp, "ours.csv"), over))
    t0 = time.perf_counter()
    o = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, cwd=tmp)
    wall = time.perf_counter() - t0
    our_ms = float(re.search(r"Elapsed time: ([0-9.eE+]+) ms", o.stdout).group(1))
    out["configs"][name] = {"num_water": p.num_water, "timesteps": p.t, "sprintSteps": p.sprintSteps,
                            "reference_gpu_ms": ref_ms, "reference_gpu_walker_steps_per_s": steps / (ref_ms * 1e-3),
                            "ours_ms": our_ms, "ours_walker_steps_per_s": steps / (our_ms * 1e-3),
                            "ours_wall_s_incl_python_and_scene_build": wall, "speedup": ref_ms / our_ms}
    print(name, json.dumps(out["configs"][name]), flush=True)
os.makedirs(os.path.join(ROOT, "gpurun_out", "refgpu"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "refgpu", "compare.json"), "w"), indent=1)
