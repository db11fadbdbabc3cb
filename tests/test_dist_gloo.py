"""The N>1 path on CPU: world_size-2 gloo, one process per rank, walkers sharded by
contiguous range with global stream indices, one SUM all-reduce of the signal partials per
sprint -- equal to the unsharded run.  The oracle stands in for the GPU context here."""
import os
import socket

import numpy as np
import pytest
import torch.multiprocessing as mp

import oracle_lib as ol
from mri_nanoparticle_simulation_b200 import api, dist

TINY = dict(num_water=160, num_cells=40, bound=10, water_start_bound=6, hashDim=8, border=3)
STEPS, SPRINTS, SEED = 1000, 3, 0xC0FFEE


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, out_path):
    import torch.distributed as td
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    td.init_process_group("gloo", rank=rank, world_size=world)
    p = api.ParameterStruct(**TINY)
    box = api.BacteriaBox(p, 3).populateSimulation()          # scene replicated: same seed on every rank
    orc = ol.oracle_for_box(box, p)
    w_all = box.getWaters()
    lo, hi = dist.shard_range(len(w_all), rank, world)
    w = w_all[lo:hi].copy(); rng = orc.seed(hi - lo, SEED, lo)
    state = {"t": 0}

    def sprint(steps):
        sig, _ = orc.run_sprint(w, rng, state["t"], steps, n_threads=1)
        state["t"] += steps
        return sig

    rows = dist.run_sharded_sprints(sprint, SPRINTS, STEPS)
    if rank == 0:
        np.save(out_path, rows)
    td.destroy_process_group()


def test_world2_gloo_matches_unsharded(tmp_path):
    out = str(tmp_path / "rows.npy")
    mp.spawn(_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    rows = np.load(out)
    p = api.ParameterStruct(**TINY)
    box = api.BacteriaBox(p, 3).populateSimulation()
    orc = ol.oracle_for_box(box, p)
    w = box.getWaters().copy(); rng = orc.seed(len(w), SEED)
    ref = []
    for k in range(SPRINTS):
        s, _ = orc.run_sprint(w, rng, k * STEPS, STEPS, n_threads=1)
        ref.append(s)
    ref = np.concatenate(ref)
    assert rows.shape == ref.shape
    assert np.max(np.abs(rows - ref)) <= 1e-12 * len(w)        # reassociation of the walker sum only


def test_shard_ranges_partition():
    for n in (0, 1, 7, 4032, 1 << 20):
        for world in (1, 2, 3, 8):
            r = [dist.shard_range(n, k, world) for k in range(world)]
            assert r[0][0] == 0 and r[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(r, r[1:]))
            assert max(hi - lo for lo, hi in r) - min(hi - lo for lo, hi in r) <= 1
    with pytest.raises(ValueError):
        dist.shard_range(10, 2, 2)
