"""The reference's UNCHANGED entry point (src/T2_GPU_sim.cu) and its own parameters.h compile
and link against include/dropin + libmrisim.so.  Needs the reference tree, so it runs in the
build container only (skipped on the GPU box); the run itself is covered by
tests/test_gpu_parity.py::test_simulate_waters_dropin_csv through the same C entry."""
import os
import subprocess

import pytest

from mri_nanoparticle_simulation_b200 import api

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_MAIN = "/root/reference/src/T2_GPU_sim.cu"
PATCHED = os.path.join(ROOT, "oracle", "_ref", "src")        # parameters.h with the one-token STCONST fix


@pytest.mark.skipif(not os.path.exists(REF_MAIN), reason="reference tree not present")
def test_reference_main_links_against_dropin(tmp_path):
    if not os.path.exists(os.path.join(PATCHED, "parameters.h")):
        subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "_ref/src/.stamp"])
    # only parameters.h may come from the reference tree: give the compiler a directory
    # that holds nothing else (the patched tree also has the reference's other headers)
    inc = tmp_path / "params"; inc.mkdir()
    os.symlink(os.path.join(PATCHED, "parameters.h"), inc / "parameters.h")
    # quote-includes resolve next to the including file first, so the entry point is reached
    # through a symlink that sits beside the (one-token-fixed) parameters.h -- as in a
    # maintainer's tree after applying the STCONST fix
    os.symlink(REF_MAIN, inc / "T2_GPU_sim.cu")
    exe = str(tmp_path / "gpu_sim.out")
    pkg = os.path.dirname(api.LIB_PATH)
    cmd = ["g++", "-std=c++17", "-O2", "-x", "c++", str(inc / "T2_GPU_sim.cu"), os.path.join(ROOT, "include", "dropin", "cuda_sim_dropin.cpp"),
           "-I", os.path.join(ROOT, "include", "dropin"), "-I", str(inc), "-I", os.path.join(ROOT, "include"),
           "-I", "/usr/local/cuda/include", "-L", pkg, "-lmrisim", f"-Wl,-rpath,{pkg}", "-o", exe]
    subprocess.check_call(cmd)
    out = subprocess.check_output(["nm", "-C", exe], text=True)
    assert " T simulateWaters(" in out and " T generate_base_filename" in out and " B p" in out
