"""BASELINE.json configs[4]: 16 Mi waters, totaltime 200 ms, mmoment x phase_stdev sweep on N B200s.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
        --master-port 29531 scripts/sweep_config5.py [--waters 16777216] [--sprints 4] [--cells 270]

A full 200 ms point is 16 Mi x 2e8 = 3.4e15 walker-steps (about 4.5 h on 8 GPUs at the measured
rate), so every sweep point is timed over a fixed window of sprints and
reported as walker-steps/s; `--full-point` runs ONE point to the end if the budget allows.

Per mmoment the octree is REBUILT (the leaf criterion uses |grad B|), on the GPU
(mrisim_scene_build_gpu, byte-identical to the host build); phase_stdev is a no-op under the
reference's operator-precedence quirk D4, which the sweep shows instead of assuming:
the three phase_stdev points of one mmoment must print identical signals.

One process per GPU, walkers sharded by global index, scene replicated, one 20-double NCCL
all-reduce per sprint.  One JSON line per sweep point on rank 0.
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--waters", type=int, default=1 << 24)
    ap.add_argument("--sprints", type=int, default=4)
    ap.add_argument("--warmup", type=int, default=1)
    ap.add_argument("--cells", type=int, default=270)
    ap.add_argument("--mmoments", default="1.75e-17,3.5e-17,7e-17")
    ap.add_argument("--stdev-factors", default="0.5,1,2")
    args = ap.parse_args()

    import torch
    import torch.distributed as dist
    from mri_nanoparticle_simulation_b200 import api
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=dev)
    barrier = (lambda: dist.barrier()) if world > 1 else (lambda: None)
    n_rank = args.waters // world
    over = dict(totaltime=200, num_water=64)
    if args.cells != 270:
        over.update(num_cells=args.cells, hashDim=40, maxNeighbors=16 if args.cells <= 2000 else 32)
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)

    for mm in [float(v) for v in args.mmoments.split(",")]:
        p0 = api.ParameterStruct(mmoment=mm, **over)
        t0 = time.perf_counter()
        box = api.BacteriaBox(p0, 12345).populateSimulation(device=local)       # every rank builds the same bytes on its own GPU
        build_s = time.perf_counter() - t0
        st = box.octree_build_stats.asdict()
        t0 = time.perf_counter()
        box.rethrow_waters(n_rank, 7000 + rank, True, p0, device=local)
        throw_s = time.perf_counter() - t0
        waters = box.getWaters().copy()
        first_signals = None
        for f in [float(v) for v in args.stdev_factors.split(",")]:
            p = p0.copy()
            p.phase_stdev = p0.phase_stdev * f
            steps, n_int = p.sprintSteps, p.sprintSteps // p.pfreq
            ctx = api.WalkerContext(p, local)
            ctx.set_scene(box)
            ctx.set_waters(waters, 0xC0FFEE, rank * n_rank)
            sig = torch.zeros(n_int, dtype=torch.float64, device=dev)
            rows = []

            def sprint():
                ctx.run_sprint_async(steps, sig.data_ptr(), stream.cuda_stream)
                if world > 1:
                    dist.all_reduce(sig)
            for _ in range(args.warmup):
                sprint()
                rows.append(sig.cpu().numpy().copy())
            torch.cuda.synchronize(); barrier()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(stream)
            for _ in range(args.sprints):
                sprint()
                rows.append(sig.cpu().numpy().copy())          # the CSV rows of this sprint (D2H of 20 doubles, as the driver does)
            b.record(stream); b.synchronize()
            ms = torch.tensor([a.elapsed_time(b)], dtype=torch.float64, device=dev)
            if world > 1:
                dist.all_reduce(ms, op=dist.ReduceOp.MAX)
            torch.cuda.synchronize(); barrier()
            import numpy as np
            allrows = np.concatenate(rows)
            if first_signals is None:
                first_signals = allrows
            if rank == 0:
                n_total = n_rank * world
                print(json.dumps({
                    "config": "cfg5_sweep", "n_gpus": world, "waters_total": n_total, "num_cells": p.num_cells,
                    "totaltime_ms": p.totaltime, "t_steps": p.t, "mmoment": mm, "phase_stdev": p.phase_stdev,
                    "phase_stdev_factor": f, "octree_nodes": st["n_nodes"], "scene_build_gpu_s": round(build_s, 4),
                    "octree_gpu_ms": round(st["total_ms"], 2), "octree_b_fixups": st["b_fixups"],
                    "water_throw_gpu_s": round(throw_s, 4),
                    "sprints_timed": args.sprints, "steps_per_sprint": steps,
                    "walker_steps_per_s": n_total * steps * args.sprints / (float(ms.item()) * 1e-3),
                    "ms_per_sprint": float(ms.item()) / args.sprints,
                    "signal_over_N_first_last": [float(allrows[0]) / n_total, float(allrows[-1]) / n_total],
                    "signal_identical_to_first_stdev_point": bool(allrows.tobytes() == first_signals.tobytes()),
                    "full_point_estimate_h": p.t / steps * (float(ms.item()) / args.sprints) * 1e-3 / 3600.0,
                }), flush=True)
            del ctx
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
