"""BIT-LEVEL pin of the walk against the reference's OWN device code, run on this box.

oracle/_ref/ref_gpu_pin_<cfg> (`make -C oracle refgpupin`) is the reference's cuda_sim.cu --
simulateDiffusion, computePhaseAccumulation, performUpdate, flipPhases, its cuBLAS dgemv and
thrust::reduce, its CSV line -- compiled where it lies with -fmad=false, linked with the
reference's host scene code and ONE replaced translation unit: oracle/shim/gpu_random_contract.cu
stands in for the 72-line cuRAND wrapper (src/gpu_random.cu) and fills the reference's random
buffers from the walker stream contract (include/mrisim_math.h) instead of XORWOW(time(NULL)).
The shim dumps what the reference's kernels left in their buffers (per-(step, walker) positions,
membrane flags, phase kicks) and the final water_info array.

Asserted, for three configurations:
  * per-(step, walker) x, y, z, in_cell and phase kick of the first two sprints: BIT-EQUAL between
    the reference binary, the product (mrisim_trace_sprint_kicks through the C ABI) and
    oracle/walk_oracle.c;
  * final positions, in_cell, nearest after the whole run (500 sprints, 10^6 steps, CPMG flips):
    BIT-EQUAL reference vs product; phases within 1e-9 rad (the reference sums each 1000-kick block
    with cuBLAS dgemv, whose association order is cuBLAS's own);
  * the CSV: same rows and time stamps; values equal to the printed 6 significant digits (a unit
    in the last printed digit is allowed where the two summation orders straddle a rounding
    boundary; counted and bounded).
Nothing is skipped by default: ~10 s of GPU time per configuration.
"""
import os
import subprocess

import numpy as np
import pytest

import oracle_lib as ol
from mri_nanoparticle_simulation_b200 import _cabi as cabi
from mri_nanoparticle_simulation_b200 import api

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "oracle", "_ref")
WALKER_SEED = 0xC0FFEE
TRACE_SPRINTS = 2

# name -> (host seed, ParameterStruct overrides) mirroring PGCFG_<name> of oracle/Makefile
CONFIGS = {
    "pin_default": (12345, dict(sprintSteps=2000, totaltime=1, taucp=0.004)),
    "pin_lowperm": (21, dict(num_water=8000, sprintSteps=2000, totaltime=1, taucp=0.006, avoid_intracellular_throw=0,
                             kick_mode=api.KICK_CONSTANT, phase_k=2 * 3.14 * 42 * 12 * 5e-3, _low_permeability=True)),
    "pin_dense": (11, dict(num_water=4096, num_cells=2000, hashDim=40, maxNeighbors=16, sprintSteps=2000, totaltime=1,
                           taucp=0.004, avoid_intracellular_throw=0)),
}


def params_for(name):
    seed, kw = CONFIGS[name]
    kw = dict(kw)
    low = kw.pop("_low_permeability", False)
    p = api.ParameterStruct(**kw)
    if low:
        api.low_permeability(p)
    return seed, p


def read_sprint_dump(path):
    with open(path, "rb") as f:
        magic, n, s, k = np.frombuffer(f.read(32), dtype=np.int64)
        assert magic == 0x4D52495452414345, "not a sprint dump of gpu_random_contract.cu"
        cnt = int(n) * int(s)
        out = {c: np.frombuffer(f.read(8 * cnt), dtype=np.float64).reshape(int(s), int(n)) for c in ("x", "y", "z", "kick")}
        out["in_cell"] = np.frombuffer(f.read(cnt), dtype=np.uint8).reshape(int(s), int(n))
    return out


def read_final_dump(path):
    with open(path, "rb") as f:
        magic, n, item, steps = np.frombuffer(f.read(32), dtype=np.int64)
        assert magic == 0x4D52495746494E41 and item == 40
        return np.frombuffer(f.read(40 * int(n)), dtype=cabi.WATER_DTYPE), int(steps)


def read_csv(path):
    rows = [l.split(",") for l in open(path).read().splitlines() if l]
    return [r[0] for r in rows], [r[1] for r in rows]


def same_bits(a, b):
    return np.array_equal(np.ascontiguousarray(a).view(np.uint64), np.ascontiguousarray(b).view(np.uint64))


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(CONFIGS))
def test_walk_bit_equal_to_reference_device_code(name, tmp_path):
    exe = os.path.join(REF, f"ref_gpu_{name}")
    if not os.path.exists(exe):
        pytest.skip(f"oracle/_ref/ref_gpu_{name} not built (make -C oracle refgpupin needs /root/reference)")
    seed, p = params_for(name)
    N, S = p.num_water, p.sprintSteps
    assert p.t == 1_000_000 and p.pfreq == 1000 and p.tcp in (4000, 6000)

    # ---- the reference's device code on the contract stream
    prefix = str(tmp_path / "ref")
    env = dict(os.environ, REF_TRACE_PREFIX=prefix, REF_TRACE_SPRINTS=str(TRACE_SPRINTS), REF_WALKER_SEED=str(WALKER_SEED))
    ref_csv = str(tmp_path / "ref.csv")
    out = subprocess.run([exe, ref_csv], capture_output=True, text=True, timeout=900, cwd=str(tmp_path), env=env)
    assert out.returncode == 0, (out.stdout[-2000:], out.stderr[-2000:])
    assert "self-check against src/xorshift.h ok" in out.stdout

    # ---- the product: same scene (byte-identical builder), same stream
    box = api.BacteriaBox(p, seed).populateSimulation()
    w0 = box.getWaters().copy()
    assert len(w0) == N
    ctx = api.WalkerContext(p, 0)
    ctx.set_scene(box)
    ctx.set_waters(w0, WALKER_SEED)
    orc = ol.oracle_for_box(box, p)
    owat, orng = w0.copy(), orc.seed(N, WALKER_SEED)
    n_in = 0
    for k in range(TRACE_SPRINTS):
        ref = read_sprint_dump(f"{prefix}_s{k}.bin")
        _, tr = ctx.trace_sprint(S, kicks=True)
        _, otr = orc.run_sprint(owat, orng, k * S, S, trace=True, trace_kick=True)
        for c in ("x", "y", "z"):
            assert same_bits(ref[c], tr[c]), f"{name} sprint {k}: per-step {c} differs between the reference's kernels and the product"
            assert same_bits(ref[c], otr[c]), f"{name} sprint {k}: per-step {c} differs between the reference's kernels and the oracle"
        assert np.array_equal(ref["in_cell"], tr["in_cell"]) and np.array_equal(ref["in_cell"], otr["in_cell"])
        # phase kicks: bit-equal, NaN payloads included (the 0.5-uniform corner of cuda_sim.cu:632)
        nan_ref, nan_our = np.isnan(ref["kick"]), np.isnan(tr["kick"])
        assert np.array_equal(nan_ref, nan_our) and np.array_equal(nan_ref, np.isnan(otr["kick"]))
        assert same_bits(np.where(nan_ref, 0.0, ref["kick"]), np.where(nan_our, 0.0, tr["kick"])), \
            f"{name} sprint {k}: per-step phase kick differs between the reference's kernels and the product"
        assert same_bits(np.where(nan_ref, 0.0, ref["kick"]), np.where(nan_ref, 0.0, otr["kick"])), \
            f"{name} sprint {k}: per-step phase kick differs between the reference's kernels and the oracle"
        n_in += int(ref["in_cell"].sum())
        del ref, tr, otr
    if name != "pin_default":
        assert n_in > 0, "configuration is meant to exercise intracellular walkers"

    # ---- the rest of the run on the product's fast paths, then final state and CSV
    sig = []
    while ctx.time < p.t:
        sig.append(ctx.run_sprint(20000 if (p.t - ctx.time) >= 20000 else S))
    got = ctx.get_waters()
    ref_final, ref_steps = read_final_dump(f"{prefix}_final.bin")
    assert ref_steps == p.t == ctx.time
    for c in ("x", "y", "z", "in_cell", "nearest"):
        assert got[c].tobytes() == ref_final[c].tobytes(), f"{name}: final {c} differs from the reference's kernels after {p.t} steps"
    finite = np.isfinite(ref_final["phase"])
    assert np.array_equal(finite, np.isfinite(got["phase"]))
    dphi = np.max(np.abs(got["phase"][finite] - ref_final["phase"][finite])) if finite.any() else 0.0
    assert dphi <= 1e-9, f"{name}: final phases differ by {dphi} rad"

    our_csv = str(tmp_path / "ours.csv")
    api.simulateWaters(our_csv, p, host_seed=seed, walker_seed=WALKER_SEED, device=0)
    t_ref, s_ref = read_csv(ref_csv)
    t_our, s_our = read_csv(our_csv)
    assert t_ref == t_our and len(t_ref) == p.t // p.pfreq
    a, b = np.array([float(v) for v in s_ref]), np.array([float(v) for v in s_our])
    assert np.array_equal(np.isnan(a), np.isnan(b))
    ok = ~np.isnan(a)
    # 6 significant digits printed: one unit of the last printed digit at most, and only rarely
    ulp6 = 10.0 ** (np.floor(np.log10(np.maximum(np.abs(a[ok]), 1e-300))) - 5)
    assert np.all(np.abs(a[ok] - b[ok]) <= 1.01 * ulp6)
    same = sum(x == y for x, y in zip(s_ref, s_our))
    assert same >= 0.99 * len(s_ref), f"only {same} of {len(s_ref)} CSV rows print identically"
    print(f"\n{name}: {TRACE_SPRINTS}x{S} steps x {N} walkers bit-equal (reference kernels == product == oracle), "
          f"{n_in} intracellular (step, walker) entries; after {p.t} steps final positions/flags bit-equal, "
          f"max |dphase| = {dphi:.2e} rad; CSV rows identical as printed: {same}/{len(s_ref)}")
