"""In-tree build of libmrisim.so (host scene builder + sm_100a kernels + C ABI).

``python -m mri_nanoparticle_simulation_b200.build`` or ``__graft_entry__.build()``.
nvcc cross-compiles for sm_100a without a GPU; the .so is git-ignored but travels with the
working tree to the GPU box.
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys

PKG = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG)
CSRC = os.path.join(PKG, "csrc")
INC = os.path.join(ROOT, "include")
OBJ = os.path.join(PKG, "build" + os.environ.get("MRISIM_LIB_SUFFIX", ""))
# tuning builds: MRISIM_MIN_BLOCKS=<n> MRISIM_LIB_SUFFIX=_mb<n> python -m ...build  (api: MRISIM_LIB=<path>)
_SUFFIX = os.environ.get("MRISIM_LIB_SUFFIX", "")
LIB = os.path.join(PKG, f"libmrisim{_SUFFIX}.so")
EXE = os.path.join(PKG, "gpu_sim.out")

# -fmad=false / -ffp-contract=off: the arithmetic discipline that makes positions and
# octree indices bit-identical to the oracle (see include/mrisim_math.h)
NVCC_FLAGS = ["-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3",
              "-fmad=false", "-Xcompiler", "-fPIC,-ffp-contract=off,-O2", "-I", INC] + \
             [f"-D{k}={os.environ[k]}" for k in ("MRISIM_MIN_BLOCKS", "MRISIM_BLOCK") if os.environ.get(k)] + \
             os.environ.get("MRISIM_DEFS", "").split()            # tuning builds: extra -D switches of csrc/mrisim.cu
CUDA_INC = os.path.join(os.environ.get("CUDA_HOME", "/usr/local/cuda"), "include")      # group.cpp: cuda_runtime.h for <nccl.h>
CXX_FLAGS = ["-std=c++17", "-O2", "-ffp-contract=off", "-fPIC", "-pthread", "-Wall", "-Wextra", "-I", INC, "-I", CUDA_INC]

CU_SOURCES = ["mrisim.cu", "octree_build.cu", "scene_gpu.cu"]
CXX_SOURCES = ["scene.cpp", "simulate_waters.cpp", "checkpoint.cpp", "group.cpp"]


def _nvcc() -> str:
    for c in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if c and os.path.exists(c):
            return c
    raise RuntimeError("nvcc not found")


def _digest(paths, extra="") -> str:
    h = hashlib.sha256(extra.encode())
    for p in sorted(paths):
        with open(p, "rb") as f:
            h.update(f.read())
    return h.hexdigest()


def _run(cmd, verbose):
    if verbose:
        print("+", " ".join(cmd), flush=True)
    subprocess.check_call(cmd)


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile every source for sm_100a and link libmrisim.so; returns its path."""
    os.makedirs(OBJ, exist_ok=True)
    headers = [os.path.join(INC, f) for f in os.listdir(INC) if os.path.isfile(os.path.join(INC, f))] + \
              [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".h")]
    sources = [os.path.join(CSRC, f) for f in CU_SOURCES + CXX_SOURCES if os.path.exists(os.path.join(CSRC, f))]
    stamp = os.path.join(OBJ, "stamp")
    # flags enter the digest with the checkout's path taken out: the tree is built here and RUN from another
    # directory on the GPU box, where a path-dependent digest would trigger a pointless one-minute rebuild
    dig = _digest(headers + sources, " ".join(NVCC_FLAGS + CXX_FLAGS).replace(ROOT, "."))
    if not force and os.path.exists(LIB) and os.path.exists(stamp) and open(stamp).read() == dig:
        return LIB
    objs = []
    for f in CU_SOURCES:
        o = os.path.join(OBJ, f + ".o")
        _run([_nvcc()] + NVCC_FLAGS + ["-c", os.path.join(CSRC, f), "-o", o], verbose)
        objs.append(o)
    for f in CXX_SOURCES:
        src = os.path.join(CSRC, f)
        if not os.path.exists(src):
            continue
        o = os.path.join(OBJ, f + ".o")
        _run(["g++"] + CXX_FLAGS + ["-c", src, "-o", o], verbose)
        objs.append(o)
    _run([_nvcc(), "-shared", "-o", LIB] + objs + ["-Xlinker", "--no-undefined", "-lpthread", "-ldl"], verbose)
    main_src = os.path.join(CSRC, "T2_sim_main.cpp")
    if os.path.exists(main_src):
        _run(["g++"] + CXX_FLAGS + ["-I", CSRC, main_src, "-o", EXE, "-L", PKG, "-lmrisim", "-Wl,-rpath,$ORIGIN"], verbose)
    with open(stamp, "w") as f:
        f.write(dig)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True))
