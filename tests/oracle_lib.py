"""ctypes binding of the parity oracle (oracle/_build/liboracle_walk.so).

TEST INFRASTRUCTURE: imported only by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs -- never by the product package.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

from mri_nanoparticle_simulation_b200 import _cabi as cabi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB_PATH = os.path.join(ROOT, "oracle", "_build", "liboracle_walk.so")


class OracleScene(C.Structure):
    _fields_ = [
        ("cells", C.c_void_p), ("num_cells", C.c_int32),
        ("lookup", C.c_void_p),
        ("min_depth", C.c_int32), ("max_depth", C.c_int32),
        ("subtree_sizes", C.c_void_p), ("n_subtrees", C.c_int32),
        ("nodes", C.c_void_p), ("n_nodes", C.c_uint64),
        ("residents", C.c_void_p), ("n_res", C.c_uint64),
    ]


class OracleTrace(C.Structure):
    _fields_ = [(k, C.c_void_p) for k in ("x", "y", "z", "in_cell", "leaf", "kick")]


def build(force: bool = False) -> str:
    if force or not os.path.exists(LIB_PATH):
        subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "port"])
    return LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(LIB_PATH)
        L.oracle_run_sprint.restype = C.c_int
        L.oracle_run_sprint.argtypes = [C.POINTER(cabi.Params), C.POINTER(OracleScene), C.c_void_p, C.c_void_p,
                                        C.c_uint64, C.c_int64, C.c_int32, C.c_void_p, C.POINTER(OracleTrace), C.c_int]
        L.oracle_seed_walkers.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint64]
        L.oracle_probe_leaf.restype = C.c_uint32
        L.oracle_probe_leaf.argtypes = [C.POINTER(cabi.Params), C.POINTER(OracleScene), C.c_double, C.c_double, C.c_double, C.POINTER(C.c_uint64)]
        L.oracle_probe_field.restype = C.c_double
        L.oracle_probe_field.argtypes = [C.POINTER(cabi.Params), C.POINTER(OracleScene), C.c_double, C.c_double, C.c_double]
        L.oracle_probe_nearest.restype = C.c_int32
        L.oracle_probe_nearest.argtypes = [C.POINTER(cabi.Params), C.POINTER(OracleScene), C.c_double, C.c_double, C.c_double, C.POINTER(C.c_int)]
        L.oracle_kick_expression.restype = C.c_double
        L.oracle_kick_expression.argtypes = [C.POINTER(cabi.Params), C.c_double, C.c_int]
        L.oracle_xorshift_words.argtypes = [C.c_uint64, C.c_void_p, C.c_int]
        L.oracle_xorshift_doubles.argtypes = [C.c_uint64, C.c_void_p, C.c_int]
        L.oracle_normals.argtypes = [C.c_uint64, C.c_void_p, C.c_int]
        L.oracle_log_sincos.argtypes = [C.c_double, C.c_double] + [C.POINTER(C.c_double)] * 3
        _lib = L
    return _lib


class Oracle:
    """Holds numpy views of a scene (scene_io.Scene) in the layout the oracle walks."""

    def __init__(self, scene, params: "cabi.Params"):
        self.params = params
        self.scene = scene
        self._keep = dict(
            cells=np.ascontiguousarray(scene.cells, dtype=np.float64),
            lookup=np.ascontiguousarray(scene.lookup, dtype=np.int32),
            sizes=np.ascontiguousarray(scene.subtree_sizes, dtype=np.uint32),
            nodes=np.ascontiguousarray(scene.nodes),
            residents=np.ascontiguousarray(scene.residents),
        )
        k = self._keep
        self.sc = OracleScene(
            k["cells"].ctypes.data, scene.num_cells, k["lookup"].ctypes.data,
            scene.min_depth, scene.max_depth, k["sizes"].ctypes.data, scene.n_subtrees,
            k["nodes"].ctypes.data, scene.n_nodes, k["residents"].ctypes.data, len(k["residents"]))

    def seed(self, n, walker_seed, first_id=0):
        rng = np.zeros(n, dtype=cabi.RNG_DTYPE)
        lib().oracle_seed_walkers(rng.ctypes.data, n, walker_seed, first_id)
        return rng

    def run_sprint(self, waters, rng, time0, steps, trace=False, n_threads=None, trace_kick=False):
        """waters: (n,) WATER_DTYPE in/out; rng: (n,) RNG_DTYPE in/out."""
        n = len(waters)
        n_int = steps // self.params.pfreq
        sig = np.zeros(n_int, dtype=np.float64)
        tr = None
        out = {}
        if trace:
            out = dict(x=np.zeros((steps, n)), y=np.zeros((steps, n)), z=np.zeros((steps, n)),
                       in_cell=np.zeros((steps, n), np.uint8), leaf=np.zeros((steps, n), np.uint32))
            if trace_kick:
                out["kick"] = np.zeros((steps, n))
            tr = OracleTrace(*[out[k].ctypes.data if k in out else None for k in ("x", "y", "z", "in_cell", "leaf", "kick")])
        if n_threads is None:
            n_threads = os.cpu_count() or 1
        rc = lib().oracle_run_sprint(C.byref(self.params), C.byref(self.sc), waters.ctypes.data, rng.ctypes.data,
                                     n, time0, steps, sig.ctypes.data, C.byref(tr) if tr else None, n_threads)
        if rc:
            raise RuntimeError(f"oracle_run_sprint failed rc={rc}")
        return sig, out

    def probe_leaf(self, x, y, z):
        mc = C.c_uint64()
        idx = lib().oracle_probe_leaf(C.byref(self.params), C.byref(self.sc), x, y, z, C.byref(mc))
        return idx, mc.value

    def probe_field(self, x, y, z):
        return lib().oracle_probe_field(C.byref(self.params), C.byref(self.sc), x, y, z)

    def probe_nearest(self, x, y, z):
        ic = C.c_int()
        n = lib().oracle_probe_nearest(C.byref(self.params), C.byref(self.sc), x, y, z, C.byref(ic))
        return n, ic.value


class _SceneView:
    """Adapter: the arrays of a product BacteriaBox in the attribute layout Oracle expects."""

    def __init__(self, box):
        oct_ = box.getOctree()
        self.cells = box.getCells()
        self.lookup = box.getLookupTable()
        self.subtree_sizes = oct_.subtree_sizes
        self.nodes = oct_.nodes
        self.residents = oct_.residents
        self.num_cells = len(self.cells)
        self.min_depth, self.max_depth = oct_.min_depth, oct_.max_depth
        self.n_subtrees = len(oct_.subtree_sizes)
        self.n_nodes = len(oct_.nodes)


def oracle_for_box(box, params) -> Oracle:
    return Oracle(_SceneView(box), params)
