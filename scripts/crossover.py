"""Fused vs pipelined walker path over the walker count (default scene, one 20000-step sprint each).

    python scripts/crossover.py [walkers ...] > profiles/pipeline_vs_fused_crossover_rNN.json

PATH_AUTO (include/mrisim.h) switches where the two curves cross; this is the measurement behind it.
Every pair of runs must leave identical walkers (positions, phases, flags: the two paths are bit-identical
per walker by contract); the signals may differ in the last bits, the sum over walkers being grouped
differently (per-walker partials vs per-block partials)."""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from mri_nanoparticle_simulation_b200 import api  # noqa: E402

FUSED, PIPELINE = 1, 2          # MRISIM_PATH_FUSED / MRISIM_PATH_PIPELINE

counts = [int(v) for v in sys.argv[1:]] or [4032, 8192, 16384, 32768, 49152, 65536, 98304]
box, p, _ = bench.get_scene(api, "cfg1_default_4032_waters", 0, 1, lambda: None)
rows = []
for n in counts:
    waters = bench.throw_waters(box, p, n, 0)
    sigs = []
    for name, mode in (("fused", FUSED), ("pipeline", PIPELINE)):
        ctx = api.WalkerContext(p, 0)
        ctx.set_path_mode(mode)
        ctx.set_scene(box)
        ctx.set_waters(waters, bench.WALKER_SEED)
        ctx.run_sprint(p.sprintSteps)                       # warm-up sprint (buffers, jump tables)
        best = None
        for _ in range(3):
            sig = ctx.run_sprint(p.sprintSteps)
            ms = ctx.last_sprint_kernel_ms()
            best = ms if best is None else min(best, ms)
        sigs.append((ctx.get_waters().tobytes(), np.asarray(sig, dtype=np.float64).copy()))
        rows.append({"walkers": n, "path_mode": name, "ms_per_sprint": best,
                     "walker_steps_per_s": n * p.sprintSteps / (best * 1e-3)})
        del ctx
    rows[-1]["walkers_identical_to_fused"] = sigs[0][0] == sigs[1][0]
    rows[-1]["signal_max_abs_diff_over_n"] = float(np.max(np.abs(sigs[0][1] - sigs[1][1]))) / n
print(json.dumps({"scene": "default (270 cells), sprintSteps 20000, B200, best of 3 sprints after one warm-up sprint",
                  "rows": rows}, indent=1))
