"""Octree build: host threads (the reference's algorithm, csrc/scene.cpp) vs GPU (csrc/octree_build.cu).

    python scripts/octree_bench.py [default dense2000 dense8000] [--no-host dense8000]

Prints one JSON line per scene: wall time of the whole scene build either way, the GPU builder's
counters, and whether the two MRISCN01 files are byte-identical.  The octree dominates the scene
build (15 s at 2000 cells, 321 s at 8000 cells on 8 host cores)."""
import hashlib
import json
import os
import sys
import tempfile
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mri_nanoparticle_simulation_b200 import api  # noqa: E402

SCENES = {
    "default": (dict(), 12345),
    "dense2000": (dict(num_cells=2000, hashDim=40, maxNeighbors=16, num_water=4096), 11),
    "dense8000": (dict(num_cells=8000, hashDim=40, maxNeighbors=32, num_water=4096), 11),
}


def build(cfg, seed, device):
    p = api.ParameterStruct(**cfg)
    t0 = time.perf_counter()
    box = api.BacteriaBox(p, seed).populateSimulation(device=device)
    dt = time.perf_counter() - t0
    with tempfile.TemporaryDirectory() as d:
        f = os.path.join(d, "s.scn")
        box.save(f)
        digest = hashlib.sha256(open(f, "rb").read()).hexdigest()
    oc = box.getOctree()
    return dt, digest, len(oc.nodes), len(oc.residents), getattr(box, "octree_build_stats", None)


def waters_bench(n):
    """mrisim_scene_rethrow_waters (host threads) vs _gpu on the dense2000 scene, n waters."""
    cfg, seed = SCENES["dense2000"]
    p = api.ParameterStruct(**cfg)
    box = api.BacteriaBox(p, seed).populateSimulation(device=0)
    box.rethrow_waters(4096, 5, device=0)                      # warm-up
    row = {"waters": n, "scene": "dense2000"}
    digests = []
    for name, dev in (("gpu", 0), ("host", None)):
        for sort in (False, True):
            t0 = time.perf_counter()
            box.rethrow_waters(n, 5, sort_by_morton=sort, device=dev)
            row[f"{name}_{'sorted' if sort else 'throw_only'}_s"] = round(time.perf_counter() - t0, 4)
        digests.append(hashlib.sha256(box.getWaters().tobytes()).hexdigest())
    row["byte_identical"] = digests[0] == digests[1]
    row["host_threads"] = os.cpu_count()
    print(json.dumps(row), flush=True)


def main():
    args = sys.argv[1:]
    if args and args[0] == "--waters":
        for n in args[1:]:
            waters_bench(int(n))
        return
    no_host = set()
    if "--no-host" in args:
        k = args.index("--no-host")
        no_host = set(args[k + 1:])
        args = args[:k]
    names = (args or ["default", "dense2000"]) + [n for n in no_host if n not in args]
    build({}, 1, 0)                                   # context creation / module load outside the timings
    for name in names:
        cfg, seed = SCENES[name]
        g_s, g_digest, n_nodes, n_res, st = build(cfg, seed, 0)
        row = {"scene": name, "seed": seed, "n_nodes": n_nodes, "n_residents": n_res,
               "gpu_scene_build_s": round(g_s, 4), "gpu_octree": st.asdict()}
        st_d = st.asdict()
        row["gpu_pair_evals_per_s"] = st_d["pair_evals"] / (st_d["gpu_ms"] * 1e-3) if st_d["gpu_ms"] else None
        if name not in no_host:
            h_s, h_digest, _, _, _ = build(cfg, seed, None)
            row.update(host_scene_build_s=round(h_s, 4), host_threads=os.cpu_count(),
                       byte_identical=(h_digest == g_digest), speedup_scene_build=round(h_s / g_s, 2))
        row["sha256"] = g_digest
        print(json.dumps(row), flush=True)


if __name__ == "__main__":
    main()
