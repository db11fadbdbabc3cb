"""The C-ABI library loads, exports every symbol include/mrisim.h declares, and refuses to
compute without a CUDA device (no fallback).  No compute calls here."""
import ctypes as C
import os
import re

import pytest

from mri_nanoparticle_simulation_b200 import api, _cabi as cabi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, "include", "mrisim.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(mrisim_[a-z0-9_]+)\s*\(", text)))


def test_every_declared_symbol_is_exported():
    L = C.CDLL(api.LIB_PATH)
    declared = _declared()
    assert len(declared) >= 35
    missing = [s for s in declared if not hasattr(L, s)]
    assert not missing, f"declared in mrisim.h but not exported: {missing}"
    assert sorted(api.ABI_SYMBOLS) == declared, "api.ABI_SYMBOLS and include/mrisim.h disagree"


def test_cpp_dropin_symbol_exported():
    # void simulateWaters(std::string) -- the reference's only exported function (cuda_sim.h:74)
    import subprocess
    out = subprocess.check_output(["nm", "-D", "--defined-only", "-C", api.LIB_PATH], text=True)
    assert re.search(r"\bT simulateWaters\(std::(__cxx11::)?basic_string", out)


def test_struct_layouts_match_header():
    assert C.sizeof(cabi.Params) == 192      # static_assert in csrc/scene.cpp
    assert cabi.WATER_DTYPE.itemsize == 40      # water_info, rand_walk.h:25-53
    assert cabi.MNP_DTYPE.itemsize == 40        # MNP_info
    assert cabi.NODE_DTYPE.itemsize == 48       # oct_node with flattened residents
    assert cabi.RNG_DTYPE.itemsize == 24


def test_shipped_parameters():
    """mrisim_params_default == the shipped src/parameters.h."""
    p = api.ParameterStruct()
    assert (p.num_water, p.sprintSteps, p.num_cells, p.hashDim, p.maxNeighbors) == (4032, 20000, 270, 20, 13)
    assert (p.t, p.tcp, p.pfreq) == (40000000, 5500000, 1000)
    assert (p.reflectIO, p.reflectOI, p.avoid_intracellular_throw, p.kick_mode) == (0.0, 0.0, 1, cabi.KICK_RANDOM)
    assert p.in_stdev == 0.00132009145325097 and p.out_stdev == 0.0022865341663977239
    assert abs(p.phase_stdev - 0.0474768) < 1e-15
    q = api.low_permeability(api.ParameterStruct())
    assert q.reflectIO == 0.999561484166706 and q.reflectOI == 0.9997468303722928


def test_csv_row_format():
    """`fout << time*tau << "," << target << endl` with default ostream precision (cuda_sim.cu:1022)."""
    assert api.csv_row(1000, 1e-6, 4031.98123) == "0.001,4031.98\n"
    assert api.csv_row(40000000, 1e-6, -3.2e-5) == "40,-3.2e-05\n"
    assert api.csv_row(5500000, 1e-6, 1234567.0) == "5.5,1.23457e+06\n"
    assert api.csv_row(1000, 1e-6, float("nan")) in ("0.001,nan\n", "0.001,-nan\n")


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present; the refusal path is for CPU-only boxes")
    with pytest.raises(api.MrisimError, match="no CUDA device|no CPU fallback"):
        api.WalkerContext(api.ParameterStruct())
    with pytest.raises(api.MrisimError):
        api.measure_fp64_tflops(0)


def test_product_does_not_import_oracle():
    """Nothing under the package may reference oracle/ (the oracle is test infrastructure)."""
    pkg = os.path.join(ROOT, "mri_nanoparticle_simulation_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cpp", ".cu", ".h")):
                src = open(os.path.join(dp, f), errors="ignore").read()
                assert "oracle_lib" not in src and "walk_oracle" not in src and "liboracle" not in src, f
