"""Invariants of the oracle's walk (CPU): independence from threading, sprint splitting,
sharding by walker with global stream indices, CPMG flips, impermeable membranes."""
import numpy as np
import pytest

import oracle_lib as ol
from mri_nanoparticle_simulation_b200 import api, dist

TINY = dict(num_water=192, num_cells=40, bound=10, water_start_bound=6, hashDim=8, border=3)


@pytest.fixture(scope="module")
def tiny():
    p = api.low_permeability(api.ParameterStruct(avoid_intracellular_throw=0, **TINY))
    box = api.BacteriaBox(p, 3).populateSimulation()
    return p, box, ol.oracle_for_box(box, p)


def run(orc, waters, seed, steps, time0=0, first=0, threads=1, rng=None):
    w = waters.copy()
    r = orc.seed(len(w), seed, first) if rng is None else rng.copy()
    sig, _ = orc.run_sprint(w, r, time0, steps, n_threads=threads)
    return sig, w, r


def same(a, b):
    return all(a[f].tobytes() == b[f].tobytes() for f in a.dtype.names)


def test_thread_count_independent(tiny):
    p, box, orc = tiny
    w0 = box.getWaters()
    s1, a1, r1 = run(orc, w0, 7, 2000, threads=1)
    s4, a4, r4 = run(orc, w0, 7, 2000, threads=5)
    assert np.array_equal(s1, s4) and same(a1, a4) and r1.tobytes() == r4.tobytes()


def test_two_half_sprints_equal_one(tiny):
    p, box, orc = tiny
    w0 = box.getWaters()
    s, a, r = run(orc, w0, 7, 4000)
    s1, a1, r1 = run(orc, w0, 7, 1000)
    s2, a2, r2 = run(orc, a1, 7, 3000, time0=1000, rng=r1)
    assert np.array_equal(np.concatenate([s1, s2]), s) and same(a2, a) and r2.tobytes() == r.tobytes()


def test_sharding_by_walker_reproduces_unsharded_run(tiny):
    p, box, orc = tiny
    w0 = box.getWaters()
    s, a, _ = run(orc, w0, 11, 2000)
    parts, total = [], np.zeros_like(s)
    for rank in range(3):
        lo, hi = dist.shard_range(len(w0), rank, 3)
        sp, ap, _ = run(orc, w0[lo:hi], 11, 2000, first=lo)
        parts.append(ap); total += sp
    assert same(np.concatenate(parts), a)
    assert np.max(np.abs(total - s)) <= 1e-12 * len(w0)


def test_cpmg_flip_after_sampling(tiny):
    """phase *= -1 when time % tcp == 0, AFTER the sample of that row (cuda_sim.cu:1013-1020)."""
    p, box, _ = tiny
    q = p.copy(); q.tcp = 2000
    orc = ol.oracle_for_box(box, q)
    w0 = box.getWaters()
    s_a, a, _ = run(orc, w0, 5, 2000)
    orc_nf = ol.oracle_for_box(box, p)                 # tcp = 5.5e6: no flip in 2000 steps
    s_b, b, _ = run(orc_nf, w0, 5, 2000)
    assert np.array_equal(s_a, s_b)                    # the row at the flip uses the pre-flip phases
    assert np.array_equal(a["phase"], -b["phase"]) and np.array_equal(a["x"], b["x"])


def test_impermeable_membrane_keeps_walkers_on_their_side(tiny):
    p, box, _ = tiny
    q = p.copy(); q.reflectIO = 1.0; q.reflectOI = 1.0
    orc = ol.oracle_for_box(box, q)
    w = box.getWaters().copy(); r = orc.seed(len(w), 3)
    _, tr = orc.run_sprint(w, r, 0, 3000, trace=True)
    ic = tr["in_cell"]
    assert ic[0].sum() > 0 and (ic[0] == 0).sum() > 0          # both populations present
    assert (ic == ic[0][None, :]).all()
    # and with fully permeable membranes some walkers do cross
    q.reflectIO = q.reflectOI = 0.0
    orc2 = ol.oracle_for_box(box, q)
    w = box.getWaters().copy(); r = orc2.seed(len(w), 3)
    _, tr2 = orc2.run_sprint(w, r, 0, 3000, trace=True)
    assert (tr2["in_cell"] != tr2["in_cell"][0][None, :]).any()


def test_positions_stay_in_box_and_steps_have_contract_length(tiny):
    p, box, orc = tiny
    w = box.getWaters().copy(); r = orc.seed(len(w), 9)
    _, tr = orc.run_sprint(w, r, 0, 1000, trace=True)
    for k in "xyz":
        assert (tr[k] >= 0).all() and (tr[k] < p.bound).all()
    d = np.sqrt(sum(np.diff(tr[k], axis=0) ** 2 for k in "xyz"))
    d = d[d < p.bound / 2]                                       # ignore periodic wraps
    assert d.max() < 6.5 * p.out_stdev and 0.5 * p.in_stdev < d.mean() < p.out_stdev


def test_signal_starts_at_n_and_decays(tiny):
    p, box, orc = tiny
    s, _, _ = run(orc, box.getWaters(), 1, 5000)
    n = len(box.getWaters())
    assert 0.999 * n < s[0] <= n and s[-1] < s[0]


def test_steps_must_be_multiple_of_pfreq(tiny):
    p, box, orc = tiny
    with pytest.raises(RuntimeError):
        run(orc, box.getWaters(), 1, 1500)
