"""The walker stream contract (include/mrisim_math.h) and the kick expression."""
import ctypes as C
import math

import numpy as np

import oracle_lib as ol
from mri_nanoparticle_simulation_b200 import api


def test_kick_expression_verbatim_equals_closed_form():
    """cuda_sim.cu:632 evaluated token for token (oracle) == in_cell*K for nD != 0.5, NaN for
    nD == 0.5 whatever in_cell is -- the form the kernel uses."""
    p = api.ParameterStruct()
    K = 9.3638310632897375e-06
    L = ol.lib()
    grid = [0.0, 2.0 ** -52, 1e-7, 0.1, 0.25, 0.49999, 0.5 - 2.0 ** -53, 0.5 + 2.0 ** -52, 0.50001, 0.75, 0.9, 1 - 2.0 ** -52]
    rng = np.random.default_rng(0)
    grid += list(rng.integers(0, 2 ** 52, 20000) / 2.0 ** 52)
    for nD in grid:
        if nD == 0.5:
            continue
        assert L.oracle_kick_expression(C.byref(p), nD, 1) == K
        assert L.oracle_kick_expression(C.byref(p), nD, 0) == 0.0
    assert math.isnan(L.oracle_kick_expression(C.byref(p), 0.5, 1))
    assert math.isnan(L.oracle_kick_expression(C.byref(p), 0.5, 0))
    # phase_stdev is a no-op under the reference's operator precedence
    q = api.ParameterStruct(phase_stdev=123.0)
    assert L.oracle_kick_expression(C.byref(q), 0.3, 1) == K
    # CONSTANT_KICK branch (cuda_sim.cu:635)
    c = api.ParameterStruct(kick_mode=api.KICK_CONSTANT, phase_k=2 * 3.14 * 42 * 12 * 5e-3)
    assert L.oracle_kick_expression(C.byref(c), 0.3, 1) == c.phase_k * c.tau
    assert L.oracle_kick_expression(C.byref(c), 0.3, 0) == 0.0


def test_log_and_sincos_accuracy():
    L = ol.lib()
    rng = np.random.default_rng(1)
    lg, sn, cs = C.c_double(), C.c_double(), C.c_double()
    worst_l = worst_t = 0.0
    vs = list(1.0 - rng.integers(0, 2 ** 52, 20000) / 2.0 ** 52) + [1.0, 2.0 ** -52, 0.5, 0.70710678, 0.70710679]
    us = list(rng.integers(0, 2 ** 52, 20000) / 2.0 ** 52) + [0.0, 0.125, 0.25, 0.375, 0.5, 0.75, 1 - 2.0 ** -52]
    for v, u in zip(vs, us + us[:len(vs) - len(us)] if len(us) < len(vs) else us):
        L.oracle_log_sincos(v, u, C.byref(lg), C.byref(sn), C.byref(cs))
        ref = math.log(v)
        if ref != 0.0:
            worst_l = max(worst_l, abs(lg.value - ref) / abs(ref))
        else:
            assert lg.value == 0.0
        worst_t = max(worst_t, abs(sn.value - math.sin(2 * math.pi * u)), abs(cs.value - math.cos(2 * math.pi * u)))
        assert abs(sn.value ** 2 + cs.value ** 2 - 1.0) < 1e-15
    assert worst_l < 4e-16 and worst_t < 1.5e-15


def test_normal_stream_moments():
    z = np.zeros(1 << 21)
    ol.lib().oracle_normals(7, z.ctypes.data, len(z))
    assert abs(z.mean()) < 3e-3 and abs(z.std() - 1) < 3e-3
    assert abs(((z - z.mean()) ** 4).mean() / z.var() ** 2 - 3) < 2e-2
    assert abs(np.corrcoef(z[0::2], z[1::2])[0, 1]) < 3e-3
    assert np.isfinite(z).all()


def test_uniform_grid_and_range():
    d = np.zeros(100000)
    ol.lib().oracle_xorshift_doubles(99, d.ctypes.data, len(d))
    assert (d >= 0).all() and (d < 1).all()
    assert ((d * 2.0 ** 52) % 1 == 0).all()          # rand_pos_double lives on the 2^-52 grid
