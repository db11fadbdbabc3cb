"""mrisim_signal_sample_static: the signal sample (sum of cos(phase) over the walkers) from an entry point that
allocates nothing -- its scratch is a static device array of the library (include/mrisim.h).

Checked through the C ABI against the existing entry point's output: the last signal sample of mrisim_run_sprint is
the sum of cos(phase) over the walkers mrisim_get_waters then returns (the CPMG flip after the sample only changes
the sign of a phase, cos is even), so the two must agree within the suite's signal tolerance (1e-6 * N: the sums
are associated differently).  Two calls in a row, on different inputs, because all calls share the scratch array.

Collected last on purpose: a failure here hides no other test.
"""
import ctypes as C

import numpy as np
import pytest

from mri_nanoparticle_simulation_b200 import api

pytestmark = pytest.mark.gpu

SIGNAL_TOL = 1e-6          # |S_static - S_sprint| / N, the tolerance of tests/test_gpu_parity.py (BASELINE.json)
MAX_WALKERS = 1 << 24      # the limit include/mrisim.h states


def _static_sample(torch, phases_host, n=None):
    """Device copies made by torch (the test's plumbing); the entry point itself gets plain device pointers."""
    ph = torch.from_numpy(np.ascontiguousarray(phases_host, dtype=np.float64)).cuda()
    out = torch.full((1,), float("nan"), dtype=torch.float64, device="cuda")
    n = len(phases_host) if n is None else n
    rc = api.lib().mrisim_signal_sample_static(0, C.c_void_p(ph.data_ptr()), n, C.c_void_p(out.data_ptr()),
                                               C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0, api.lib().mrisim_last_error(None).decode()
    torch.cuda.synchronize()
    return float(out.item())


def test_static_signal_sample_matches_the_sprint_signal_twice():
    import torch
    p = api.low_permeability(api.ParameterStruct(num_water=512, num_cells=40, bound=10, water_start_bound=6,
                                                  hashDim=8, border=3, avoid_intracellular_throw=0))
    box = api.BacteriaBox(p, 3).populateSimulation()
    box.rethrow_waters(9001, 5)                           # three CTAs of the static kernel, the last one ragged
    waters = box.getWaters().copy()
    n = len(waters)
    ctx = api.WalkerContext(p, 0)
    ctx.set_scene(box)
    ctx.set_waters(waters, 0xC0FFEE)
    samples = []
    for call in range(2):                                 # two calls in a row: the scratch array keeps the first call's partials
        sig = ctx.run_sprint(2 * p.pfreq)                 # existing entry point: mrisim_run_sprint
        phases = ctx.get_waters()["phase"]
        got = _static_sample(torch, phases)
        assert abs(got - sig[-1]) / n <= SIGNAL_TOL, f"call {call}: static {got!r} vs sprint {sig[-1]!r}"
        assert abs(got - float(np.cos(phases).sum())) / n <= SIGNAL_TOL
        samples.append(got)
    assert samples[0] != samples[1]                       # the walkers moved on in between: the inputs did differ
    # a shorter input right after a longer one: slots the longer call wrote must not leak into the sum
    short = ctx.get_waters()["phase"][:1000]
    assert abs(_static_sample(torch, short) - float(np.cos(short).sum())) / len(short) <= SIGNAL_TOL
    # the same input twice: bit-identical (fixed association order)
    assert _static_sample(torch, phases) == samples[1]
    ctx.close()


def test_static_signal_sample_rejects_what_it_cannot_hold():
    import torch
    L = api.lib()
    buf = torch.zeros(8, dtype=torch.float64, device="cuda")
    ptr = C.c_void_p(buf.data_ptr())
    assert L.mrisim_signal_sample_static(0, ptr, MAX_WALKERS + 1, ptr, None) != 0        # refused before any launch
    msg = L.mrisim_last_error(None).decode()
    assert "exceeds the static scratch array" in msg and str(MAX_WALKERS) in msg
    assert L.mrisim_signal_sample_static(0, None, 8, ptr, None) != 0
    assert "null pointer" in L.mrisim_last_error(None).decode()
    assert L.mrisim_signal_sample_static(0, ptr, 0, ptr, None) != 0
    assert L.mrisim_signal_sample_static(99, ptr, 8, ptr, None) != 0
    assert "device index" in L.mrisim_last_error(None).decode()
