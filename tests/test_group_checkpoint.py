"""Submitted sprints, on-disk checkpoint/resume and the multi-device group -- all through the C ABI.

CPU part (-m "not gpu"): the MRICKP01 reader refuses garbage, truncated and inconsistent files
without touching a device; the parameter hash ignores exactly the knobs that do not change a walk.
GPU part: see each test.  Multi-device tests need >= 2 GPUs and skip on a 1-GPU box
(run them with `gpurun --gpus 2`).
"""
import os
import struct
import subprocess
import sys

import numpy as np
import pytest

from mri_nanoparticle_simulation_b200 import api

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SEED = 0xC0FFEE


def n_gpus():
    try:
        import torch
        return torch.cuda.device_count() if torch.cuda.is_available() else 0
    except Exception:
        return 0


# ---------------------------------------------------------------------------------------- CPU
def _header(n=4, n_shards=1, magic=b"MRICKP01", version=1, time_steps=2000):
    return struct.pack("<8sII4Qq3Q6Q", magic, version, n_shards, 11, 22, 33, 44, time_steps, 55, n, 0, *([0] * 6))


def test_checkpoint_reader_refuses_bad_files(tmp_path):
    p = str(tmp_path / "x.ckpt")
    with pytest.raises(api.MrisimError, match="cannot open"):
        api.checkpoint_info(p)
    open(p, "wb").write(b"not a checkpoint at all")
    with pytest.raises(api.MrisimError, match="not an MRICKP01 file"):
        api.checkpoint_info(p)
    n = 4
    shard = struct.pack("<QQII", 0, n, 0, 0)
    body = bytes(5 * n * 8 + 2 * n * 8 + n * 4 + 8)
    open(p, "wb").write(_header(n) + shard + body)
    info = api.checkpoint_info(p)
    assert (info.n_walkers, info.time_steps, info.csv_bytes, info.host_seed, info.walker_seed, info.n_shards) == (n, 2000, 55, 33, 44, 1)
    open(p, "wb").write(_header(n) + shard + body[:-3])
    with pytest.raises(api.MrisimError, match="file size"):
        api.checkpoint_info(p)
    open(p, "wb").write(_header(n) + struct.pack("<QQII", 1, n, 0, 0) + body)
    with pytest.raises(api.MrisimError, match="shard table"):
        api.checkpoint_info(p)
    open(p, "wb").write(_header(n, n_shards=0) + body)
    with pytest.raises(api.MrisimError, match="implausible"):
        api.checkpoint_info(p)
    open(p, "wb").write(_header(n, version=7) + shard + body)
    with pytest.raises(api.MrisimError, match="not an MRICKP01 file"):
        api.checkpoint_info(p)


def test_params_hash_tracks_the_walk_only():
    base = api.params_hash(api.ParameterStruct())
    assert api.params_hash(api.ParameterStruct(num_threads=3)) == base
    assert api.params_hash(api.ParameterStruct(totaltime=7)) == base       # a resumed run may be extended
    for kw in (dict(mmoment=7e-17), dict(sprintSteps=10000), dict(taucp=4.0), dict(reflectIO=0.5), dict(num_cells=271),
               dict(kick_mode=api.KICK_CONSTANT), dict(D_extra=1.7)):
        assert api.params_hash(api.ParameterStruct(**kw)) != base, kw


# ---------------------------------------------------------------------------------------- GPU
def _scene(p, seed, n=None):
    box = api.BacteriaBox(p, seed).populateSimulation()
    if n is not None:
        box.rethrow_waters(n, 99, True, p)
    return box, box.getWaters().copy()


def _ctx(p, box, w, mode=api.PATH_AUTO, first=0):
    ctx = api.WalkerContext(p, 0)
    ctx.set_path_mode(mode)
    ctx.set_scene(box)
    ctx.set_waters(w, SEED, first)
    return ctx


@pytest.mark.gpu
@pytest.mark.parametrize("n,mode", [(4032, api.PATH_PIPELINE), (4032, api.PATH_FUSED), (70000, api.PATH_AUTO)])
def test_submitted_sprints_equal_synchronous_ones(n, mode):
    """mrisim_submit_sprint x3 (ring, side-stream D2H) then wait x3 == three mrisim_run_sprint calls, bit for bit."""
    p = api.low_permeability(api.ParameterStruct(avoid_intracellular_throw=0, sprintSteps=2000))
    box, w = _scene(p, 12345, n)
    a = _ctx(p, box, w, mode)
    want = [a.run_sprint(2000) for _ in range(3)]
    b = _ctx(p, box, w, mode)
    for _ in range(3):
        b.submit_sprint(2000)
    assert b.pending_sprints == 3
    got = [b.wait_sprint(2000) for _ in range(3)]
    assert b.pending_sprints == 0
    for x, y in zip(want, got):
        assert x.tobytes() == y.tobytes()
    assert a.get_waters().tobytes() == b.get_waters().tobytes() and a.get_rng().tobytes() == b.get_rng().tobytes()
    b.check()
    with pytest.raises(api.MrisimError, match="no sprint outstanding"):
        b.wait_sprint(2000)
    for _ in range(4):
        b.submit_sprint(1000)
    with pytest.raises(api.MrisimError, match="outstanding"):
        b.submit_sprint(1000)
    with pytest.raises(api.MrisimError, match="outstanding"):
        b.set_waters(w, SEED)
    for _ in range(4):
        b.wait_sprint(1000)


@pytest.mark.gpu
@pytest.mark.parametrize("n,mode", [(4032, api.PATH_PIPELINE), (70000, api.PATH_FUSED)])
def test_checkpoint_roundtrip_bit_exact(n, mode, tmp_path):
    """save after 2 sprints -> a fresh context loads -> the next 2 sprints print the same BITS as the
    uninterrupted context's (walker state, streams, time and -- fused path -- the walker order)."""
    p = api.low_permeability(api.ParameterStruct(avoid_intracellular_throw=0, sprintSteps=2000, taucp=0.003))
    box, w = _scene(p, 12345, n)
    a = _ctx(p, box, w, mode)
    for _ in range(2):
        a.run_sprint(2000)
    path = str(tmp_path / "a.ckpt")
    a.checkpoint_save(path, host_seed=12345, csv_bytes=77)
    info = api.checkpoint_info(path)
    assert (info.n_walkers, info.time_steps, info.csv_bytes, info.host_seed, info.walker_seed) == (n, 4000, 77, 12345, SEED)
    assert info.params_hash == api.params_hash(p)
    want = [a.run_sprint(2000) for _ in range(2)]
    b = api.WalkerContext(p, 0)
    b.set_path_mode(mode)
    with pytest.raises(api.MrisimError, match="set the scene"):
        b.checkpoint_load(path)
    b.set_scene(box)
    b.checkpoint_load(path)
    assert b.time == 4000 and b.n == n
    got = [b.run_sprint(2000) for _ in range(2)]
    for x, y in zip(want, got):
        assert x.tobytes() == y.tobytes(), "resumed sprint differs from the uninterrupted one"
    assert a.get_waters().tobytes() == b.get_waters().tobytes() and a.get_rng().tobytes() == b.get_rng().tobytes()
    # a shard of the file: walkers [lo, lo+count) with their global ids
    lo, cnt = n // 3, n // 2
    c = api.WalkerContext(p, 0)
    c.set_path_mode(mode)
    c.set_scene(box)
    c.checkpoint_load(path, lo, cnt)
    c.run_sprint(2000); c.run_sprint(2000)
    assert c.get_waters().tobytes() == a.get_waters()[lo:lo + cnt].tobytes()
    # refused: other parameters, other scene
    q = api.low_permeability(api.ParameterStruct(avoid_intracellular_throw=0, sprintSteps=2000, taucp=0.003, mmoment=7e-17))
    d = api.WalkerContext(q, 0)
    d.set_scene(box)
    with pytest.raises(api.MrisimError, match="different parameters"):
        d.checkpoint_load(path)
    box2, _ = _scene(p, 4321, 64)
    e = api.WalkerContext(p, 0)
    e.set_scene(box2)
    with pytest.raises(api.MrisimError, match="different scene"):
        e.checkpoint_load(path)


@pytest.mark.gpu
@pytest.mark.parametrize("n_water", [4032, 60000])
def test_simulate_waters_killed_and_resumed_is_byte_identical(n_water, tmp_path):
    """The drop-in with MRISIM_CHECKPOINT_EVERY / MRISIM_RESUME semantics: a run stopped after 4 of 7
    sprints (checkpoint every 2; rows of a sprint after the last checkpoint are already in the CSV, as
    after a kill) and resumed produces the very bytes of an uninterrupted run -- pipelined path (4032
    waters) and fused path (60 000)."""
    p = api.ParameterStruct(num_water=n_water, sprintSteps=4000, totaltime=1, taucp=0.006)
    full, part = str(tmp_path / "full.csv"), str(tmp_path / "part.csv")
    api.simulateWaters(full, p, host_seed=12345, walker_seed=SEED, max_sprints=7)
    ck = str(tmp_path / "run.ckpt")
    api.simulateWaters(part, p, host_seed=12345, walker_seed=SEED, max_sprints=5, checkpoint_every=2, checkpoint_path=ck)
    info = api.checkpoint_info(ck)
    assert info.time_steps == 4 * 4000 and info.csv_bytes < os.path.getsize(part)      # sprint 5's rows are past the checkpoint
    # resume ignores the seeds it is given: they come from the file
    api.simulateWaters(part, p, host_seed=1, walker_seed=2, max_sprints=7, checkpoint_every=2, checkpoint_path=ck, resume=ck)
    assert open(full, "rb").read() == open(part, "rb").read()
    assert len(open(full).read().splitlines()) == 7 * 4
    assert api.checkpoint_info(ck).time_steps == 6 * 4000


@pytest.mark.gpu
def test_single_device_group_equals_context(tmp_path):
    """The group code path on one device (no NCCL needed): same bits as a plain context."""
    p = api.low_permeability(api.ParameterStruct(avoid_intracellular_throw=0, sprintSteps=2000))
    box, w = _scene(p, 12345, 5000)
    a = _ctx(p, box, w)
    want = [a.run_sprint(2000) for _ in range(2)]
    g = api.WalkerGroup(p, [0])
    g.set_scene(box)
    g.set_waters(w, SEED)
    got = [g.run_sprint(2000) for _ in range(2)]
    for x, y in zip(want, got):
        assert x.tobytes() == y.tobytes()
    assert g.get_waters().tobytes() == a.get_waters().tobytes() and g.get_rng().tobytes() == a.get_rng().tobytes()
    assert g.time == 4000 and len(g) == 1 and g.kernel_launches > 0
    path = str(tmp_path / "g.ckpt")
    g.checkpoint_save(path, 1, 2)
    b = api.WalkerContext(p, 0)
    b.set_scene(box)
    b.checkpoint_load(path)
    assert b.run_sprint(2000).tobytes() == g.run_sprint(2000).tobytes()


@pytest.mark.gpu
@pytest.mark.parametrize("n,mode", [(8000, api.PATH_AUTO), (120001, api.PATH_AUTO)])
def test_two_device_group_matches_one_device(n, mode, tmp_path):
    """2 GPUs, ONE process, NCCL all-reduce on the side streams: the shard states concatenate to the
    one-device state bit for bit; the signal agrees to 1e-12 relative (association order of the sum only);
    a checkpoint written by the 2-device group resumes on ONE device with the same per-walker bits."""
    if n_gpus() < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    p = api.low_permeability(api.ParameterStruct(avoid_intracellular_throw=0, sprintSteps=2000, taucp=0.003))
    box, w = _scene(p, 12345, n)
    one = _ctx(p, box, w, mode)
    want = [one.run_sprint(2000) for _ in range(3)]
    g = api.WalkerGroup(p, [0, 1])
    for i in range(2):
        g.member(i).set_path_mode(mode)
    g.set_scene(box)
    g.set_waters(w, SEED)
    g.submit_sprint(2000); g.submit_sprint(2000)
    got = [g.wait_sprint(2000), g.wait_sprint(2000)]
    path = str(tmp_path / "g2.ckpt")
    g.checkpoint_save(path, 12345, 0)
    assert api.checkpoint_info(path).n_shards == 2
    got.append(g.run_sprint(2000))
    for x, y in zip(want, got):
        assert np.max(np.abs(x - y)) <= 1e-12 * n
    assert g.get_waters().tobytes() == one.get_waters().tobytes() and g.get_rng().tobytes() == one.get_rng().tobytes()
    r = api.WalkerContext(p, 1)
    r.set_path_mode(mode)
    r.set_scene(box)
    r.checkpoint_load(path)
    assert r.time == 4000
    r.run_sprint(2000)
    assert r.get_waters().tobytes() == one.get_waters().tobytes()
    # and a 2-device group resumes its own file with its walker orders restored: same signal BITS
    g2 = api.WalkerGroup(p, [0, 1])
    for i in range(2):
        g2.member(i).set_path_mode(mode)
    g2.set_scene(box)
    g2.checkpoint_load(path)
    assert g2.run_sprint(2000).tobytes() == got[2].tobytes()
    print("\nNCCL:", api.nccl_origin())


@pytest.mark.gpu
def test_dropin_on_two_devices_prints_the_one_device_csv(tmp_path):
    """`simulateWaters` sharded over MRISIM_DEVICES=0,1 (the drop-in's way to take a device list): rows and
    time stamps equal the one-device CSV, values to the 6 printed digits."""
    if n_gpus() < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    p = api.ParameterStruct(num_water=100000, sprintSteps=4000, totaltime=1, taucp=0.006)
    one, two = str(tmp_path / "one.csv"), str(tmp_path / "two.csv")
    api.simulateWaters(one, p, host_seed=12345, walker_seed=SEED, max_sprints=5)
    os.environ["MRISIM_DEVICES"] = "0,1"
    try:
        api.simulateWaters(two, p, host_seed=12345, walker_seed=SEED, max_sprints=5)
    finally:
        del os.environ["MRISIM_DEVICES"]
    a = [l.split(",") for l in open(one).read().splitlines()]
    b = [l.split(",") for l in open(two).read().splitlines()]
    assert [r[0] for r in a] == [r[0] for r in b] and len(a) == 20
    va, vb = np.array([float(r[1]) for r in a]), np.array([float(r[1]) for r in b])
    assert np.all(np.abs(va - vb) <= 1.01 * 10.0 ** (np.floor(np.log10(np.abs(va))) - 5))
    assert sum(x[1] == y[1] for x, y in zip(a, b)) >= 19


_RANK_SCRIPT = r"""
import os, sys, numpy as np
sys.path.insert(0, {root!r}); sys.path.insert(0, os.path.join({root!r}, "tests"))
import torch, torch.distributed as dist
from mri_nanoparticle_simulation_b200 import api, dist as shard
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank)
dist.init_process_group("gloo")
p = api.low_permeability(api.ParameterStruct(avoid_intracellular_throw=0, sprintSteps=2000))
box = api.BacteriaBox(p, 12345).populateSimulation()
box.rethrow_waters(9001, 99, True, p)
w = box.getWaters().copy()
lo, hi = shard.shard_range(len(w), rank, world)
ids = [api.nccl_unique_id() if rank == 0 else None]
dist.broadcast_object_list(ids, src=0)
ctx = api.WalkerContext(p, rank)
ctx.set_scene(box); ctx.set_waters(w[lo:hi], 0xC0FFEE, lo)
ctx.comm_init_rank(ids[0], world, rank)
assert ctx.comm_size == world
for _ in range(2): ctx.submit_sprint(2000)
sig = np.concatenate([ctx.wait_sprint(2000), ctx.wait_sprint(2000)])
np.save(os.path.join({out!r}, f"sig{{rank}}.npy"), sig)
np.save(os.path.join({out!r}, f"w{{rank}}.npy"), ctx.get_waters())
dist.barrier(); dist.destroy_process_group()
"""


@pytest.mark.gpu
def test_one_process_per_gpu_comm_init_rank(tmp_path):
    """The multi-PROCESS form (what bench.py drives under torch.distributed.run): rank 0 makes the NCCL
    unique id, it travels through the launcher's store, every rank attaches its own communicator with
    mrisim_comm_init_rank; the all-reduced signal is identical on both ranks and equals one device's."""
    if n_gpus() < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    script = tmp_path / "rank.py"
    script.write_text(_RANK_SCRIPT.format(root=ROOT, out=str(tmp_path)))
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                          "--master-port", "29517", str(script)], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-3000:]
    s0, s1 = np.load(tmp_path / "sig0.npy"), np.load(tmp_path / "sig1.npy")
    assert s0.tobytes() == s1.tobytes()
    p = api.low_permeability(api.ParameterStruct(avoid_intracellular_throw=0, sprintSteps=2000))
    box, w = _scene(p, 12345, 9001)
    one = _ctx(p, box, w)
    want = np.concatenate([one.run_sprint(2000), one.run_sprint(2000)])
    assert np.max(np.abs(want - s0)) <= 1e-12 * len(w)
    w0, w1, ref_w = np.load(tmp_path / "w0.npy"), np.load(tmp_path / "w1.npy"), one.get_waters()
    for f in ("x", "y", "z", "phase", "in_cell", "nearest"):       # field by field: .npy does not keep the struct's padding bytes
        assert np.concatenate([w0[f], w1[f]]).tobytes() == ref_w[f].tobytes(), f
