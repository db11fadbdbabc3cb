"""Walker sharding across ranks (one process per GPU) and the per-sprint signal all-reduce.

The reference is single-GPU (src/cuda_sim.cu:916, ``Handler handle(0)``).  Walkers never
interact and the scene is read-only, so the path shards by walker: the Morton-sorted walker
array (SimulationBox.cpp:131-147) is cut into ``world`` contiguous ranges, the scene is
replicated, and each walker keeps the random stream of its GLOBAL index
(``first_walker_id`` of ``mrisim_set_waters``), so a sharded run draws exactly the numbers
of the unsharded run.  The only exchange is one SUM all-reduce of ``sprintSteps/pfreq``
doubles per sprint (NCCL over NVLink on the GPU box, gloo in the CPU tests).
"""
from __future__ import annotations

from typing import Callable, Tuple

import numpy as np


def shard_range(n: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous [lo, hi) of the sorted walker array owned by ``rank``; sizes differ by <= 1."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError("bad rank/world")
    return n * rank // world, n * (rank + 1) // world


def allreduce_signal(partial, group=None):
    """In-place SUM all-reduce of a rank's signal partial (torch tensor, any device)."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(partial, op=dist.ReduceOp.SUM, group=group)
    return partial


def run_sharded_sprints(sprint_fn: Callable[[int], "np.ndarray"], n_sprints: int, steps: int, device="cpu", group=None):
    """Drive ``n_sprints`` sprints on this rank and return the GLOBAL signal rows.

    ``sprint_fn(steps)`` runs one sprint over this rank's walkers and returns its partial
    signal (numpy array or torch tensor of steps/pfreq doubles).  Used with the CUDA context
    on the GPU box and with the oracle in the gloo CPU tests.
    """
    import torch
    rows = []
    for _ in range(n_sprints):
        part = sprint_fn(steps)
        t = part if isinstance(part, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(part, dtype=np.float64)).to(device)
        allreduce_signal(t, group)
        rows.append(t.detach().cpu().numpy().copy())
    return np.concatenate(rows) if rows else np.zeros(0)
