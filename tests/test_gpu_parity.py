"""Parity of the CUDA path (through the C ABI) against the oracle on a B200.

Bar (BASELINE.json north_star): RNG words, cell placement, octree node/leaf indices
bit-exact; positions and T2 signal within tolerance (max |dS| / N <= 1e-6).  In the default
EXACT field mode positions, phases, membrane flags and per-step leaf indices are bit-exact
too; in the FAST field mode everything discrete stays bit-exact and the phases agree to
1e-12 rad.
"""
import ctypes as C
import json
import os

import numpy as np
import pytest

import oracle_lib as ol
from mri_nanoparticle_simulation_b200 import api, dist

pytestmark = pytest.mark.gpu

SIGNAL_TOL = 1e-6          # max_k |S_gpu(T_k) - S_oracle(T_k)| / N   (BASELINE.json)
FAST_PHASE_TOL = 1e-12     # rad, FAST field mode vs oracle after one sprint
SEED = 0xC0FFEE
TINY = dict(num_water=256, num_cells=40, bound=10, water_start_bound=6, hashDim=8, border=3)
KATS = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "scene_kats.json")))


def make(params, seed):
    box = api.BacteriaBox(params, seed).populateSimulation()
    return box, ol.oracle_for_box(box, params)


def gpu_ctx(params, box, waters, first=0, mode=api.FIELD_EXACT):
    ctx = api.WalkerContext(params, 0)
    ctx.set_field_mode(mode)
    ctx.set_scene(box)
    ctx.set_waters(waters, SEED, first)
    return ctx


def assert_state_equal(got, want, fields=("x", "y", "z", "phase", "in_cell", "nearest")):
    for f in fields:
        assert got[f].tobytes() == want[f].tobytes(), f"walker field {f} differs from the oracle"


CASES = {
    "tiny": (api.ParameterStruct(**TINY), 3, 3000),
    "tiny_noavoid_lowperm": (api.low_permeability(api.ParameterStruct(avoid_intracellular_throw=0, **TINY)), 5, 3000),
    "tiny_impermeable": (api.ParameterStruct(avoid_intracellular_throw=0, reflectIO=1.0, reflectOI=1.0, **TINY), 6, 2000),
    "tiny_constant_kick": (api.ParameterStruct(avoid_intracellular_throw=0, kick_mode=api.KICK_CONSTANT,
                                               phase_k=2 * 3.14 * 42 * 12 * 5e-3, **TINY), 7, 2000),
    "default_cfg1": (api.ParameterStruct(), 12345, 2000),
    "cfg2_noavoid_8000": (api.ParameterStruct(num_water=8000, avoid_intracellular_throw=0), 21, 1000),
    "ragged_n": (api.ParameterStruct(**{**TINY, "num_water": 77}), 8, 2000),
    "single_walker": (api.ParameterStruct(**{**TINY, "num_water": 1}), 9, 2000),
}


@pytest.mark.parametrize("name", sorted(CASES))
def test_traced_sprint_bit_exact(name):
    """Every (step, walker): position, in_cell, flat leaf index; final state; RNG state."""
    p, seed, steps = CASES[name]
    box, orc = make(p, seed)
    w0 = box.getWaters().copy()
    ctx = gpu_ctx(p, box, w0)
    sig, tr = ctx.trace_sprint(steps)
    want, wrng = w0.copy(), orc.seed(len(w0), SEED)
    osig, otr = orc.run_sprint(want, wrng, 0, steps, trace=True)
    for k in ("x", "y", "z", "in_cell", "leaf"):
        assert np.array_equal(tr[k], otr[k]), f"{name}: per-step {k} differs"
    assert_state_equal(ctx.get_waters(), want)
    assert ctx.get_rng().tobytes() == wrng.tobytes()
    assert np.max(np.abs(sig - osig)) / len(w0) <= SIGNAL_TOL
    assert ctx.time == steps
    if "noavoid" in name or "impermeable" in name or "kick" in name:
        assert tr["in_cell"].any(), "case is meant to exercise intracellular walkers"


@pytest.mark.parametrize("name", sorted(CASES))
@pytest.mark.parametrize("field_mode", [api.FIELD_EXACT, api.FIELD_FAST])
def test_pipelined_path_bit_exact(name, field_mode):
    """The few-walker PIPELINE path (jump-ahead draws, serial position kernel, per-step field, ordered
    block sums) against the oracle over two chained sprints: final walkers, RNG state incl. the spare
    normal, signal; and against the FUSED path, whose walkers it must equal bit for bit."""
    p, seed, _ = CASES[name]
    box, orc = make(p, seed)
    w0 = box.getWaters().copy()
    ctx = gpu_ctx(p, box, w0, mode=field_mode)
    ctx.set_path_mode(api.PATH_PIPELINE)
    sig = np.concatenate([ctx.run_sprint(2000), ctx.run_sprint(1000)])
    assert ctx.pipeline_sprints == 2 and ctx.time == 3000
    fused = gpu_ctx(p, box, w0, mode=field_mode)
    fused.set_path_mode(api.PATH_FUSED)
    fsig = np.concatenate([fused.run_sprint(2000), fused.run_sprint(1000)])
    assert fused.pipeline_sprints == 0
    got, grng = ctx.get_waters(), ctx.get_rng()
    assert got.tobytes() == fused.get_waters().tobytes() and grng.tobytes() == fused.get_rng().tobytes()
    assert np.max(np.abs(sig - fsig)) <= 1e-9 * len(w0)
    want, wrng = w0.copy(), orc.seed(len(w0), SEED)
    osig = np.concatenate([orc.run_sprint(want, wrng, 0, 2000)[0], orc.run_sprint(want, wrng, 2000, 1000)[0]])
    assert grng.tobytes() == wrng.tobytes()
    if field_mode == api.FIELD_EXACT:
        assert_state_equal(got, want)
    else:
        assert_state_equal(got, want, fields=("x", "y", "z", "in_cell", "nearest"))
        assert np.max(np.abs(got["phase"] - want["phase"])) <= FAST_PHASE_TOL
    assert np.max(np.abs(sig - osig)) / len(w0) <= SIGNAL_TOL


@pytest.mark.parametrize("n", [16384, 30000])
def test_pipelined_path_mid_sizes_equal_fused(n):
    """Walker counts where the pipelined path picks draw-chunks that are NOT a multiple of its 8-step
    store tile (50 steps at 16 Ki walkers) and cuts a sprint into several launch groups: every walker
    and RNG state must equal the fused path's (which the other tests pin to the oracle), low-permeability
    membranes with walkers starting inside cells so that crossings end speculative trips."""
    p = api.low_permeability(api.ParameterStruct(avoid_intracellular_throw=0))
    box, _ = make(p, 12345)
    box.rethrow_waters(n, 99, True, p)
    w0 = box.getWaters().copy()
    out = []
    for mode in (api.PATH_PIPELINE, api.PATH_FUSED):
        ctx = gpu_ctx(p, box, w0)
        ctx.set_path_mode(mode)
        sig = np.concatenate([ctx.run_sprint(3000), ctx.run_sprint(1000)])
        out.append((ctx.get_waters(), ctx.get_rng(), sig, ctx.pipeline_sprints))
    assert out[0][3] == 2 and out[1][3] == 0
    assert out[0][0].tobytes() == out[1][0].tobytes() and out[0][1].tobytes() == out[1][1].tobytes()
    assert np.max(np.abs(out[0][2] - out[1][2])) <= 1e-9 * n
    assert int(out[0][0]["in_cell"].sum()) > 0


@pytest.mark.parametrize("variant", ["split", 0, 1, 2, 3, 4, 5, 6])
def test_pipelined_path_variants_equal_fused(variant, monkeypatch):
    """Every variant of the pipelined path's walk stage -- the split kernels (the default: positions and kicks through
    HBM) and the experimental walker_team_kernel with 1..4 walkers per warp, roomy and tight register targets -- leaves the walkers, RNG
    states and block sums of the fused kernel, bit for bit: low-permeability membranes with walkers inside cells
    (crossings end speculative trips), a ragged walker count, two chained sprints, intracellular kicks."""
    if variant != "split":
        monkeypatch.setenv("MRISIM_PIPE_TEAM", "1")
        monkeypatch.setenv("MRISIM_TEAM_VARIANT", str(variant))
    p = api.low_permeability(api.ParameterStruct(avoid_intracellular_throw=0))
    box, _ = make(p, 12345)
    box.rethrow_waters(5003, 77, True, p)
    w0 = box.getWaters().copy()
    out = []
    for mode in (api.PATH_PIPELINE, api.PATH_FUSED):
        ctx = gpu_ctx(p, box, w0)
        ctx.set_path_mode(mode)
        sig = np.concatenate([ctx.run_sprint(3000), ctx.run_sprint(2000)])
        out.append((ctx.get_waters(), ctx.get_rng(), sig, ctx.pipeline_sprints))
    assert out[0][3] == 2 and out[1][3] == 0
    assert out[0][0].tobytes() == out[1][0].tobytes() and out[0][1].tobytes() == out[1][1].tobytes()
    assert np.max(np.abs(out[0][2] - out[1][2])) <= 1e-9 * len(w0)
    assert int(out[0][0]["in_cell"].sum()) > 0


def test_path_mode_auto_and_fallbacks():
    """AUTO takes the pipelined path for few walkers and the fused one for many, for traces, and for
    sprints that do not start on an even step; mixing the paths sprint by sprint changes no walker."""
    p, seed, _ = CASES["tiny_noavoid_lowperm"]
    box, orc = make(p, seed)
    w0 = box.getWaters().copy()
    ctx = gpu_ctx(p, box, w0)
    ctx.run_sprint(1000)
    assert ctx.pipeline_sprints == 1
    ctx.trace_sprint(1000)                                   # traces use the fused kernel
    assert ctx.pipeline_sprints == 1
    ctx.set_path_mode(api.PATH_FUSED); ctx.run_sprint(1000)
    ctx.set_path_mode(api.PATH_AUTO); ctx.run_sprint(2000)
    assert ctx.pipeline_sprints == 2
    want, wrng = w0.copy(), orc.seed(len(w0), SEED)
    for t0, n in ((0, 1000), (1000, 1000), (2000, 1000), (3000, 2000)):
        orc.run_sprint(want, wrng, t0, n)
    assert_state_equal(ctx.get_waters(), want)
    assert ctx.get_rng().tobytes() == wrng.tobytes()
    with pytest.raises(api.MrisimError, match="unknown mode"):
        ctx.set_path_mode(7)


@pytest.mark.parametrize("resort,groups", [(0, 1), (300, 1), (1000, 2), (70, 3)])
def test_launch_granularity_and_overlap_groups_do_not_change_walkers(resort, groups):
    """A sprint cut into launches of any length (block sums carried across launches), walkers
    re-ordered by leaf in between, the walker array split into overlapping launch chains: every
    walker's state is the oracle's bit for bit; the signal differs by reassociation only."""
    p = api.low_permeability(api.ParameterStruct(num_cells=60, bound=12, water_start_bound=8, hashDim=10, border=3,
                                                  avoid_intracellular_throw=0, num_water=64))
    box, orc = make(p, 17)
    box.rethrow_waters(80000, 31)
    w0 = box.getWaters().copy()
    ctx = gpu_ctx(p, box, w0)
    ctx.set_resort_steps(resort); ctx.set_overlap_groups(groups)
    sig = np.concatenate([ctx.run_sprint(2000), ctx.run_sprint(1000)])
    got, grng = ctx.get_waters(), ctx.get_rng()
    sel = np.arange(0, len(w0), 41)
    sub = w0[sel].copy()
    rng = np.concatenate([orc.seed(1, SEED, int(i)) for i in sel])
    orc.run_sprint(sub, rng, 0, 2000); orc.run_sprint(sub, rng, 2000, 1000)
    assert_state_equal(got[sel], sub)
    assert grng[sel].tobytes() == rng.tobytes()
    ref = gpu_ctx(p, box, w0)
    ref.set_resort_steps(0); ref.set_overlap_groups(1)
    rsig = np.concatenate([ref.run_sprint(2000), ref.run_sprint(1000)])
    assert ref.get_waters().tobytes() == got.tobytes()
    assert np.max(np.abs(sig - rsig)) <= 1e-9 * len(w0)


def test_traced_sprint_with_short_launches():
    """Per-step trace (positions, flags, leaf indices) with 130-step launches and re-ordering."""
    p, seed, _ = CASES["tiny_noavoid_lowperm"]
    box, orc = make(p, seed)
    w0 = box.getWaters().copy()
    ctx = gpu_ctx(p, box, w0)
    ctx.set_resort_steps(130)
    sig, tr = ctx.trace_sprint(2000)
    want, wrng = w0.copy(), orc.seed(len(w0), SEED)
    osig, otr = orc.run_sprint(want, wrng, 0, 2000, trace=True)
    for k in ("x", "y", "z", "in_cell", "leaf"):
        assert np.array_equal(tr[k], otr[k]), f"per-step {k} differs"
    assert_state_equal(ctx.get_waters(), want)
    assert np.max(np.abs(sig - osig)) / len(w0) <= SIGNAL_TOL


def test_leaf_indices_map_to_reference_subtree_offsets():
    """flat leaf index == subtree_prefix[i] + j of the reference's (subtree i, vector offset j)."""
    p, seed, _ = CASES["default_cfg1"]
    box, orc = make(p, seed)
    ctx = gpu_ctx(p, box, box.getWaters().copy())
    _, tr = ctx.trace_sprint(1000)
    oct_ = box.getOctree()
    prefix = oct_.subtree_prefix
    leaf = tr["leaf"][-1]
    assert ((oct_.nodes["mc"][leaf] >> np.uint64(63)) == 1).all()
    i = np.searchsorted(prefix, leaf, side="right") - 1
    assert ((leaf - prefix[i]) < oct_.subtree_sizes[i]).all()
    for w in range(0, len(leaf), 97):
        idx, mc = orc.probe_leaf(tr["x"][-1, w], tr["y"][-1, w], tr["z"][-1, w])
        assert idx == leaf[w] and mc == int(oct_.nodes["mc"][leaf[w]])


def test_full_default_sprint_cfg1():
    """BASELINE configs[0]: shipped parameters.h, one full sprint (20000 steps, 20 rows)."""
    p, seed, _ = CASES["default_cfg1"]
    box, orc = make(p, seed)
    w0 = box.getWaters().copy()
    ctx = gpu_ctx(p, box, w0)
    sig = ctx.run_sprint(p.sprintSteps)
    want, wrng = w0.copy(), orc.seed(len(w0), SEED)
    osig, _ = orc.run_sprint(want, wrng, 0, p.sprintSteps)
    assert_state_equal(ctx.get_waters(), want)
    assert ctx.get_rng().tobytes() == wrng.tobytes()
    assert len(sig) == 20 and np.max(np.abs(sig - osig)) / len(w0) <= SIGNAL_TOL


def test_cfg3_low_permeability_100k_subsample():
    """BASELINE configs[2] (AVOID_INTRACELLULAR_THROW + low-permeability reflectIO/OI): 100k
    waters on the GPU, the first 4096 of them against the oracle (streams are per walker)."""
    p = api.low_permeability(api.ParameterStruct())
    box, orc = make(p, 12345)
    box.rethrow_waters(100000, 4242)
    w0 = box.getWaters().copy()
    ctx = gpu_ctx(p, box, w0)
    sig = ctx.run_sprint(4000)
    got = ctx.get_waters()
    sub, rng = w0[:4096].copy(), orc.seed(4096, SEED)
    orc.run_sprint(sub, rng, 0, 4000)
    assert_state_equal(got[:4096], sub)
    assert 0.99 * len(w0) < sig[0] <= len(w0)


def test_dense_scene_cfg4a_subsample():
    """BASELINE configs[3] scene (2000 cells, hashDim 40, maxNeighbors 16; reference digest in
    tests/golden/scene_kats.json): 64k walkers on the GPU, 2048 of them against the oracle."""
    p = api.ParameterStruct(**KATS["configs"]["dense2000"])
    box, orc = make(p, 11)
    box.rethrow_waters(65536, 7000)
    w0 = box.getWaters().copy()
    ctx = gpu_ctx(p, box, w0)
    ctx.run_sprint(2000)
    got = ctx.get_waters()
    sel = np.arange(0, 65536, 32)
    sub = w0[sel].copy()
    rng = np.concatenate([orc.seed(1, SEED, int(i)) for i in sel])
    orc.run_sprint(sub, rng, 0, 2000)
    assert_state_equal(got[sel], sub)


def test_dense_scene_cfg4b_subsample():
    """SURVEY.md 8d config 4b (BASELINE configs[3], octree-heavy): 8000 cells, 21 385 dipoles, 455 008 octree
    nodes, leaves with more than 100 residents, lookup rows with up to 15 candidates -- the reference loops stressed
    hardest are get_field's resident loop (cuda_sim.cu:280-283) and updateNearest's scan (:470-481).  The scene is
    built on the GPU (its digests equal the reference's own host code's, tests/test_octree_gpu.py); 64k walkers
    thrown anywhere (so that some sit inside cells) run on the GPU, 1024 of them against the oracle -- final
    state bit-exact -- and a traced sprint of 512 walkers per (step, walker) incl. leaf indices and kicks."""
    p = api.ParameterStruct(avoid_intracellular_throw=0, **KATS["configs"]["dense8000"])
    box = api.BacteriaBox(p, 11).populateSimulation(device=0)
    orc = ol.oracle_for_box(box, p)
    nodes = box.getOctree().nodes
    leaf = (nodes["mc"] >> np.uint64(63)) != 0
    assert len(nodes) == 455008 and int(nodes["n_res"][leaf].max()) >= 100
    assert int((box.getLookupTable() != -1).sum(axis=1).max()) >= 15
    box.rethrow_waters(65536, 7000, True, p, device=0)
    w0 = box.getWaters().copy()
    ctx = gpu_ctx(p, box, w0)
    ctx.run_sprint(2000)
    got = ctx.get_waters()
    sel = np.arange(0, 65536, 64)
    sub = w0[sel].copy()
    rng = np.concatenate([orc.seed(1, SEED, int(i)) for i in sel])
    orc.run_sprint(sub, rng, 0, 2000)
    assert_state_equal(got[sel], sub)
    assert int(got["in_cell"].sum()) > 1000, "walkers thrown anywhere in a box that is 6 % cells"
    # per-step trace of a contiguous block (walkers 4096..4607 keep their global stream ids)
    first, cnt = 4096, 512
    tctx = gpu_ctx(p, box, w0[first:first + cnt], first=first)
    _, tr = tctx.trace_sprint(1000, kicks=True)
    want, wrng = w0[first:first + cnt].copy(), orc.seed(cnt, SEED, first)
    _, otr = orc.run_sprint(want, wrng, 0, 1000, trace=True, trace_kick=True)
    for k in ("x", "y", "z", "in_cell", "leaf"):
        assert np.array_equal(tr[k], otr[k]), f"cfg4b: per-step {k} differs"
    assert tr["kick"].tobytes() == otr["kick"].tobytes(), "cfg4b: per-step phase kick differs"
    n_res_at = nodes["n_res"][otr["leaf"]]
    print(f"\ncfg4b: residents per step along the traced walk: mean {n_res_at.mean():.2f}, max {int(n_res_at.max())}; "
          f"in-cell share {tr['in_cell'].mean():.3f}")


def test_multi_sprint_with_cpmg_flip_and_resume():
    """Sprints chain through device-resident state; flips land after sampling; a checkpoint
    (waters + rng + time) restored into a fresh context continues bit-identically."""
    p = api.ParameterStruct(tcp=3000, **TINY)
    box, orc = make(p, 3)
    w0 = box.getWaters().copy()
    ctx = gpu_ctx(p, box, w0)
    want, wrng = w0.copy(), orc.seed(len(w0), SEED)
    for k in range(4):                                   # flips at t = 3000 and 6000
        sig = ctx.run_sprint(2000)
        osig, _ = orc.run_sprint(want, wrng, 2000 * k, 2000)
        assert np.max(np.abs(sig - osig)) / len(w0) <= SIGNAL_TOL
        assert_state_equal(ctx.get_waters(), want)
        if k == 1:
            ck_w, ck_r, ck_t = ctx.get_waters(), ctx.get_rng(), ctx.time
    ctx2 = gpu_ctx(p, box, ck_w)
    ctx2.set_rng(ck_r); ctx2.time = ck_t
    ctx2.run_sprint(2000); ctx2.run_sprint(2000)
    assert_state_equal(ctx2.get_waters(), want)


def test_sharded_contexts_sum_to_unsharded():
    """Walker sharding (dist.shard_range + first_walker_id): per-shard states concatenate to
    the unsharded run bit-exactly; signal partials add up to it (reassociation only)."""
    p, seed, _ = CASES["tiny_noavoid_lowperm"]
    box, _ = make(p, seed)
    w0 = box.getWaters().copy()
    whole = gpu_ctx(p, box, w0)
    s = whole.run_sprint(2000)
    total, parts = np.zeros_like(s), []
    for rank in range(3):
        lo, hi = dist.shard_range(len(w0), rank, 3)
        c = gpu_ctx(p, box, w0[lo:hi], first=lo)
        total += c.run_sprint(2000)
        parts.append(c.get_waters())
    assert_state_equal(np.concatenate(parts), whole.get_waters())
    assert np.max(np.abs(total - s)) <= 1e-12 * len(w0)


def test_fast_field_mode_within_tolerance():
    p, seed, _ = CASES["default_cfg1"]
    box, orc = make(p, seed)
    w0 = box.getWaters().copy()
    ctx = gpu_ctx(p, box, w0, mode=api.FIELD_FAST)
    sig = ctx.run_sprint(p.sprintSteps)
    want, wrng = w0.copy(), orc.seed(len(w0), SEED)
    osig, _ = orc.run_sprint(want, wrng, 0, p.sprintSteps)
    got = ctx.get_waters()
    assert_state_equal(got, want, fields=("x", "y", "z", "in_cell", "nearest"))
    assert ctx.get_rng().tobytes() == wrng.tobytes()
    assert np.max(np.abs(got["phase"] - want["phase"])) <= FAST_PHASE_TOL
    assert np.max(np.abs(sig - osig)) / len(w0) <= SIGNAL_TOL


def test_async_sprint_into_device_buffer_on_torch_stream():
    import torch
    p, seed, _ = CASES["tiny"]
    box, _ = make(p, seed)
    w0 = box.getWaters().copy()
    a = gpu_ctx(p, box, w0)
    ref = a.run_sprint(2000)
    b = gpu_ctx(p, box, w0)
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        out = torch.zeros(2, dtype=torch.float64, device="cuda")
        b.run_sprint_async(2000, out.data_ptr(), st.cuda_stream)
    st.synchronize()
    assert np.array_equal(out.cpu().numpy(), ref)
    assert_state_equal(b.get_waters(), a.get_waters())


def test_size_independent_properties_1M_walkers():
    """At bench size (1 Mi walkers, dense scene is covered by the subsample test): positions
    in the box, membrane flag consistent with a fresh nearest-cell test, phases finite,
    S(0+) ~ N, determinism (same inputs -> identical bits), linearity over shards."""
    p = api.ParameterStruct(avoid_intracellular_throw=0)
    box, orc = make(p, 12345)
    box.rethrow_waters(1 << 20, 99)
    w0 = box.getWaters().copy()
    ctx = gpu_ctx(p, box, w0)
    s1 = ctx.run_sprint(1000)
    g1 = ctx.get_waters()
    for k in "xyz":
        assert (g1[k] >= 0).all() and (g1[k] < p.bound).all()
    assert np.isfinite(g1["phase"]).all()
    assert 0.999 * len(w0) < s1[0] <= len(w0)
    for i in range(0, len(w0), 4099):
        near, ic = orc.probe_nearest(g1["x"][i], g1["y"][i], g1["z"][i])
        assert ic == g1["in_cell"][i] and near == g1["nearest"][i]
    ctx2 = gpu_ctx(p, box, w0)
    s2 = ctx2.run_sprint(1000)
    assert np.array_equal(s1, s2) and ctx2.get_waters().tobytes() == g1.tobytes()
    half = len(w0) // 2
    a = gpu_ctx(p, box, w0[:half]); b = gpu_ctx(p, box, w0[half:], first=half)
    assert abs((a.run_sprint(1000) + b.run_sprint(1000) - s1)[0]) <= 1e-9 * len(w0)


def test_device_selftests_of_exact_arithmetic():
    """The two value-exact rewrites the kernel relies on, against the IEEE operations on the
    device: constant-divisor division and the branch-free sqrt/divide of the resident loop."""
    L = api.lib()
    for b in (30.0, 1.5, 0.75, 10.0, 1.25, 0.7300001):
        assert L.mrisim_selftest_div(0, b, 50_000_000) == 0
    assert L.mrisim_selftest_sqrt_div(0, 0.3, 2700.0, 100_000_000) == 0
    assert L.mrisim_selftest_sqrt_div(0, 1e-3, 1e6, 100_000_000) == 0
    assert L.mrisim_selftest_sqrt_div(0, 1e-31, 4.0, 100_000_000) == 0      # |d|^2 of the displacement
    # integer fast paths of the Morton word and the lookup-lattice coordinate
    for bound, h in ((30.0, 20), (30.0, 40), (10.0, 8), (7.3, 1290), (100.0, 3)):
        assert L.mrisim_selftest_index(0, bound, h, 20_000_000) == 0


@pytest.mark.parametrize("n,bits", [(1, 8), (31, 3), (2048, 8), (2049, 17), (100_003, 17), (1 << 20, 17), (3_000_001, 19), (1 << 20, 32)])
def test_walker_order_sort_is_a_stable_sort(n, bits):
    """The library's own radix sort (it replaced CUB's so that it can share an SM with walker CTAs) against
    std::stable_sort: every (key, value) pair in place, ragged sizes, tiles larger than 2048, keys of 3..32 bits."""
    assert api.lib().mrisim_selftest_sort(0, n, bits, 1234 + n) == 0


def test_error_paths():
    p, seed, _ = CASES["tiny"]
    box, _ = make(p, seed)
    ctx = api.WalkerContext(p, 0)
    with pytest.raises(api.MrisimError, match="scene incomplete|no waters"):
        ctx.run_sprint(1000)
    ctx.set_scene(box)
    with pytest.raises(api.MrisimError, match="no waters"):
        ctx.run_sprint(1000)
    ctx.set_waters(box.getWaters().copy(), SEED)
    with pytest.raises(api.MrisimError, match="multiple of pfreq"):
        ctx.run_sprint(1500)
    with pytest.raises(api.MrisimError):
        api.WalkerContext(p, 99)
    bad = box.getWaters().copy()
    bad["x"][0] = -1e-9                       # the reference would index its lookup table out of bounds
    with pytest.raises(api.MrisimError, match="outside the periodic box"):
        ctx.set_waters(bad, SEED)
    assert ctx.kernel_launches >= 1


def test_simulate_waters_dropin_csv(tmp_path):
    """mrisim_simulate_waters == simulateWaters(fName): CSV rows `time*tau,signal` with the
    reference's ostream formatting, equal to the oracle's rows at 6 significant digits."""
    p = api.ParameterStruct(sprintSteps=2000, **TINY)
    out = str(tmp_path / "signal.csv")
    api.simulateWaters(out, p, host_seed=3, walker_seed=SEED, device=0, max_sprints=2)
    rows = open(out).read().splitlines()
    assert len(rows) == 4
    box, orc = make(p, 3)
    w, rng = box.getWaters().copy(), orc.seed(p.num_water, SEED)
    want = np.concatenate([orc.run_sprint(w, rng, 2000 * k, 2000)[0] for k in range(2)])
    for k, line in enumerate(rows):
        assert line == api.csv_row(1000 * (k + 1), p.tau, want[k]).rstrip("\n")
    assert rows[0].startswith("0.001,")
