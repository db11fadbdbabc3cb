"""Generate the committed golden fixtures from the REFERENCE'S OWN host code.

Run in the build container (needs /root/reference; `make -C oracle ref` first):
    python tests/golden/make_golden.py
Writes, next to this file:
  scene_<cfg>_seed<k>.scn.gz   MRISCN01 dumps written by oracle/_ref/ref_scene_<cfg>
                               (reference BacteriaBox::populateSimulation, seed k)
  scene_kats.json              sha256 of every scene array for the larger configs, plus the
                               survey-time FNV-1a-64 known answers
  probes_<cfg>_seed<k>.json    reference Octree::get_voxel()->mc and Octree::get_field() at
                               random points (octree.cpp:123-160)
Nothing here is read from /root/reference at test time.
"""
import gzip
import hashlib
import json
import os
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from mri_nanoparticle_simulation_b200 import scene_io as sio  # noqa: E402

REF = os.path.join(ROOT, "oracle", "_ref")
FULL = [("tiny", 3), ("default", 12345)]                  # committed as whole files
HASHED = [("default", 7), ("noavoid", 99), ("n8000", 5), ("n8000_noavoid", 21), ("dense2000", 11),
          ("dense8000", 11)]          # the last one is minutes of host time in the reference's octree constructor
#: how each config differs from the shipped parameters.h (mirrors oracle/Makefile CFG_*)
CONFIGS = {
    "default": {},
    "noavoid": {"avoid_intracellular_throw": 0},
    "n8000": {"num_water": 8000},
    "n8000_noavoid": {"num_water": 8000, "avoid_intracellular_throw": 0},
    "tiny": {"num_water": 256, "num_cells": 40, "bound": 10, "water_start_bound": 6, "hashDim": 8, "border": 3},
    "dense2000": {"num_water": 4096, "num_cells": 2000, "hashDim": 40, "maxNeighbors": 16},
    "dense8000": {"num_water": 4096, "num_cells": 8000, "hashDim": 40, "maxNeighbors": 32},
}


def array_digests(s: sio.Scene) -> dict:
    h = lambda a: hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()
    return {"cells": h(s.cells), "mnps": h(s.mnps), "lookup": h(s.lookup), "subtree_sizes": h(s.subtree_sizes),
            "nodes": h(s.nodes), "residents": h(s.residents), "waters": h(s.waters),
            "num_mnps": s.num_mnps, "n_nodes": s.n_nodes, "n_res": int(len(s.residents)),
            "min_depth": s.min_depth, "max_depth": s.max_depth}


def run_ref(cfg, seed, out, probes=None, probe_out=None):
    cmd = [os.path.join(REF, f"ref_scene_{cfg}"), str(seed), out]
    if probes:
        cmd += [probes, probe_out]
    subprocess.check_call(cmd, stdout=subprocess.DEVNULL)


def main():
    kats = {"configs": CONFIGS, "scenes": {}}
    with tempfile.TemporaryDirectory() as td:
        for cfg, seed in FULL + HASHED:
            path = os.path.join(td, f"{cfg}_{seed}.scn")
            probes = probe_out = None
            if (cfg, seed) in FULL:
                rng = np.random.default_rng(seed)
                bound = CONFIGS[cfg].get("bound", 30)
                pts = rng.random((1500, 3)) * bound
                probes, probe_out = os.path.join(td, "p.bin"), os.path.join(td, "po.bin")
                pts.tofile(probes)
            run_ref(cfg, seed, path, probes, probe_out)
            raw = open(path, "rb").read()
            s = sio.loads(raw)
            kats["scenes"][f"{cfg}_seed{seed}"] = array_digests(s)
            if (cfg, seed) in FULL:
                with gzip.GzipFile(os.path.join(HERE, f"scene_{cfg}_seed{seed}.scn.gz"), "wb", mtime=0) as g:
                    g.write(raw)
                po = np.fromfile(probe_out, dtype=[("mc", "<u8"), ("field", "<f8")])
                json.dump({"points": [[float.hex(float(v)) for v in p] for p in pts],
                           "mc": [int(v) for v in po["mc"]],
                           "field": [float.hex(float(v)) for v in po["field"]]},
                          open(os.path.join(HERE, f"probes_{cfg}_seed{seed}.json"), "w"))
            if (cfg, seed) == ("default", 12345):
                # the survey-time known answers, recomputed from this dump
                parts = []
                for n in s.nodes:
                    parts.append(n["mc"].tobytes() + n["child"].tobytes() + np.uint32(n["n_res"]).tobytes()
                                 + s.residents[n["res_start"]:n["res_start"] + n["n_res"]].tobytes())
                kats["survey_A6"] = {
                    "cells_fnv": "%016x" % sio.fnv1a64(s.cells.tobytes()),
                    "mnps_fnv": "%016x" % sio.fnv1a64(s.mnps.tobytes()),
                    "lookup_fnv": "%016x" % sio.fnv1a64(s.lookup.tobytes()),
                    "octree_fnv": "%016x" % sio.fnv1a64(b"".join(parts)),
                    "waters_xyz_fnv": "%016x" % sio.fnv1a64(np.ascontiguousarray(s.waters[:, :3]).tobytes()),
                    "expected": {"cells_fnv": "2ca9e12af715eefb", "mnps_fnv": "f755475a761c2316",
                                 "lookup_fnv": "ff3f5d661bb2b2be", "octree_fnv": "0cf04bd4fcde3e42",
                                 "waters_xyz_fnv": "0a3dd7b776238bef"},
                }
                assert all(kats["survey_A6"][k] == v for k, v in kats["survey_A6"]["expected"].items()), kats["survey_A6"]
    json.dump(kats, open(os.path.join(HERE, "scene_kats.json"), "w"), indent=1, sort_keys=True)
    print("golden fixtures written to", HERE)


if __name__ == "__main__":
    main()
