"""Host-side mirror of the reference interface for the walker path, over the C ABI.

Names follow the reference: ``ParameterStruct`` (src/parameters.h), ``BacteriaBox`` with
``populateSimulation`` / ``getCells`` / ``getWaters`` / ``getMNPs`` / ``getLookupTable`` /
``getOctree`` / ``getMNPCount`` (src/SimulationBox.h:50-93, src/BacteriaBox.h:26-38) and
``simulateWaters(filename)`` (src/cuda_sim.h:74).  Everything here is a thin ctypes veneer on
``libmrisim.so``; there is no Python or CPU implementation of the walk behind it -- if the
library is missing or no CUDA device is present the device calls raise.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import numpy as np

from . import _cabi as cabi
from ._cabi import FIELD_EXACT, FIELD_FAST, PATH_AUTO, PATH_FUSED, PATH_PIPELINE, KICK_CONSTANT, KICK_NONE, KICK_RANDOM, OctreeBuildStats, Params  # noqa: F401

PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("MRISIM_LIB") or os.path.join(PKG, "libmrisim.so")   # MRISIM_LIB: tuning builds only

_lib = None


class MrisimError(RuntimeError):
    pass


def lib():
    """Load libmrisim.so (building it is ``__graft_entry__.build()``'s job); fail loudly."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise MrisimError(f"{LIB_PATH} not built: run `python -m mri_nanoparticle_simulation_b200.build` "
                          "(there is no fallback implementation)")
    L = C.CDLL(LIB_PATH)
    vp, i32, i64, u64, dbl = C.c_void_p, C.c_int32, C.c_int64, C.c_uint64, C.c_double
    PP = C.POINTER(Params)
    sig = {
        "mrisim_version": (C.c_int, []),
        "mrisim_params_default": (None, [PP]),
        "mrisim_params_derive": (None, [PP]),
        "mrisim_params_low_permeability": (None, [PP]),
        "mrisim_scene_build": (C.c_int, [PP, u64, C.c_int, C.POINTER(vp)]),
        "mrisim_scene_build_gpu": (C.c_int, [PP, u64, C.c_int, C.c_int, C.POINTER(vp), C.POINTER(OctreeBuildStats)]),
        "mrisim_octree_build_set_margin_scale": (None, [C.c_double]),
        "mrisim_scene_load": (C.c_int, [C.c_char_p, C.POINTER(vp)]),
        "mrisim_scene_save": (C.c_int, [vp, C.c_char_p]),
        "mrisim_scene_free": (None, [vp]),
        "mrisim_scene_last_error": (C.c_char_p, []),
        "mrisim_scene_cells": (vp, [vp, C.POINTER(i32)]),
        "mrisim_scene_mnps": (vp, [vp, C.POINTER(i32)]),
        "mrisim_scene_lookup": (vp, [vp, C.POINTER(i32), C.POINTER(i32)]),
        "mrisim_scene_waters": (vp, [vp, C.POINTER(i32)]),
        "mrisim_scene_octree": (C.c_int, [vp, C.POINTER(i32), C.POINTER(i32), C.POINTER(vp), C.POINTER(i32),
                                          C.POINTER(vp), C.POINTER(u64), C.POINTER(vp), C.POINTER(u64)]),
        "mrisim_scene_rethrow_waters": (C.c_int, [vp, PP, i64, u64, C.c_int]),
        "mrisim_scene_rethrow_waters_gpu": (C.c_int, [vp, PP, i64, u64, C.c_int, C.c_int]),
        "mrisim_create": (C.c_int, [PP, C.c_int, C.POINTER(vp)]),
        "mrisim_destroy": (None, [vp]),
        "mrisim_last_error": (C.c_char_p, [vp]),
        "mrisim_set_field_mode": (C.c_int, [vp, C.c_int]),
        "mrisim_set_path_mode": (C.c_int, [vp, C.c_int]),
        "mrisim_pipeline_sprints": (C.c_uint64, [vp]),
        "mrisim_set_resort_steps": (C.c_int, [vp, i32]),
        "mrisim_set_overlap_groups": (C.c_int, [vp, i32]),
        "mrisim_set_cells": (C.c_int, [vp, vp, i32]),
        "mrisim_set_lookup": (C.c_int, [vp, vp, i32, i32]),
        "mrisim_set_octree": (C.c_int, [vp, i32, i32, vp, i32, vp, u64, vp, u64]),
        "mrisim_set_scene": (C.c_int, [vp, vp]),
        "mrisim_set_waters": (C.c_int, [vp, vp, u64, u64, u64]),
        "mrisim_get_waters": (C.c_int, [vp, vp, u64]),
        "mrisim_get_rng": (C.c_int, [vp, vp, u64]),
        "mrisim_set_rng": (C.c_int, [vp, vp, u64]),
        "mrisim_get_time": (i64, [vp]),
        "mrisim_set_time": (C.c_int, [vp, i64]),
        "mrisim_run_sprint": (C.c_int, [vp, i32, vp]),
        "mrisim_run_sprint_async": (C.c_int, [vp, i32, vp, vp]),
        "mrisim_ctx_stream": (vp, [vp]),
        "mrisim_trace_sprint": (C.c_int, [vp, i32, vp, vp, vp, vp, vp, vp]),
        "mrisim_trace_sprint_kicks": (C.c_int, [vp, i32, vp, vp, vp, vp, vp, vp, vp]),
        "mrisim_kernel_launches": (u64, [vp]),
        "mrisim_last_sprint_kernel_ms": (C.c_float, [vp]),
        "mrisim_csv_row": (C.c_int, [i64, dbl, dbl, C.c_char_p, C.c_size_t]),
        "mrisim_measure_fp64_tflops": (dbl, [C.c_int, C.c_int]),
        "mrisim_selftest_div": (i64, [C.c_int, dbl, u64]),
        "mrisim_selftest_sqrt_div": (i64, [C.c_int, dbl, dbl, u64]),
        "mrisim_selftest_index": (i64, [C.c_int, dbl, i32, u64]),
        "mrisim_selftest_sort": (i64, [C.c_int, u64, i32, u64]),
        "mrisim_signal_sample_static": (C.c_int, [C.c_int, vp, u64, vp, vp]),
        "mrisim_simulate_waters": (C.c_int, [C.c_char_p, PP, u64, u64, C.c_int, i64]),
        "mrisim_simulate_waters_multi": (C.c_int, [C.c_char_p, PP, u64, u64, C.POINTER(C.c_int), C.c_int, i64, i64, C.c_char_p, C.c_char_p]),
        "mrisim_check": (C.c_int, [vp]),
        "mrisim_submit_sprint": (C.c_int, [vp, i32, vp]),
        "mrisim_wait_sprint": (C.c_int, [vp, vp]),
        "mrisim_stream_wait_sprint": (C.c_int, [vp, vp]),
        "mrisim_pending_sprints": (C.c_int, [vp]),
        "mrisim_nccl_unique_id": (C.c_int, [vp]),
        "mrisim_comm_init_rank": (C.c_int, [vp, vp, C.c_int, C.c_int]),
        "mrisim_comm_attach": (C.c_int, [vp, vp, C.c_int, C.c_int]),
        "mrisim_comm_size": (C.c_int, [vp]),
        "mrisim_nccl_origin": (C.c_char_p, []),
        "mrisim_get_order": (C.c_int, [vp, vp, vp, u64, C.POINTER(i32)]),
        "mrisim_set_order": (C.c_int, [vp, vp, vp, u64]),
        "mrisim_walker_count": (u64, [vp]),
        "mrisim_first_walker_id": (u64, [vp]),
        "mrisim_walker_seed": (u64, [vp]),
        "mrisim_ctx_scene_digest": (u64, [vp]),
        "mrisim_ctx_params": (PP, [vp]),
        "mrisim_ctx_device": (C.c_int, [vp]),
        "mrisim_params_hash": (u64, [PP]),
        "mrisim_checkpoint_save": (C.c_int, [vp, C.c_char_p, u64, u64]),
        "mrisim_checkpoint_info_read": (C.c_int, [C.c_char_p, C.POINTER(cabi.CheckpointInfo)]),
        "mrisim_checkpoint_load": (C.c_int, [vp, C.c_char_p, u64, u64]),
        "mrisim_checkpoint_last_error": (C.c_char_p, []),
        "mrisim_group_create": (C.c_int, [PP, C.POINTER(C.c_int), C.c_int, C.POINTER(vp)]),
        "mrisim_group_destroy": (None, [vp]),
        "mrisim_group_last_error": (C.c_char_p, [vp]),
        "mrisim_group_size": (C.c_int, [vp]),
        "mrisim_group_ctx": (vp, [vp, C.c_int]),
        "mrisim_group_set_scene": (C.c_int, [vp, vp]),
        "mrisim_group_set_waters": (C.c_int, [vp, vp, u64, u64, u64]),
        "mrisim_group_get_waters": (C.c_int, [vp, vp, u64]),
        "mrisim_group_get_rng": (C.c_int, [vp, vp, u64]),
        "mrisim_group_get_time": (i64, [vp]),
        "mrisim_group_submit_sprint": (C.c_int, [vp, i32]),
        "mrisim_group_wait_sprint": (C.c_int, [vp, vp]),
        "mrisim_group_run_sprint": (C.c_int, [vp, i32, vp]),
        "mrisim_group_checkpoint_save": (C.c_int, [vp, C.c_char_p, u64, u64]),
        "mrisim_group_checkpoint_load": (C.c_int, [vp, C.c_char_p]),
        "mrisim_group_kernel_launches": (u64, [vp]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)          # AttributeError here = the ABI and the header disagree
        fn.restype, fn.argtypes = res, args
    _lib = L
    return L


#: every symbol include/mrisim.h declares (tests check the library exports each one)
ABI_SYMBOLS = [
    "mrisim_version", "mrisim_params_default", "mrisim_params_derive", "mrisim_params_low_permeability",
    "mrisim_scene_build", "mrisim_scene_build_gpu", "mrisim_octree_build_set_margin_scale", "mrisim_scene_load", "mrisim_scene_save", "mrisim_scene_free",
    "mrisim_scene_last_error", "mrisim_scene_cells", "mrisim_scene_mnps", "mrisim_scene_lookup",
    "mrisim_scene_waters", "mrisim_scene_octree", "mrisim_scene_rethrow_waters", "mrisim_scene_rethrow_waters_gpu",
    "mrisim_create", "mrisim_destroy", "mrisim_last_error", "mrisim_set_field_mode", "mrisim_set_path_mode", "mrisim_pipeline_sprints", "mrisim_set_resort_steps", "mrisim_set_overlap_groups",
    "mrisim_set_cells", "mrisim_set_lookup", "mrisim_set_octree", "mrisim_set_scene",
    "mrisim_set_waters", "mrisim_get_waters", "mrisim_get_rng", "mrisim_set_rng",
    "mrisim_get_time", "mrisim_set_time", "mrisim_run_sprint", "mrisim_run_sprint_async", "mrisim_ctx_stream",
    "mrisim_trace_sprint", "mrisim_trace_sprint_kicks", "mrisim_kernel_launches", "mrisim_last_sprint_kernel_ms",
    "mrisim_csv_row", "mrisim_simulate_waters", "mrisim_measure_fp64_tflops", "mrisim_selftest_div", "mrisim_selftest_sqrt_div", "mrisim_selftest_index", "mrisim_selftest_sort", "mrisim_signal_sample_static",
    "mrisim_simulate_waters_multi", "mrisim_check", "mrisim_submit_sprint", "mrisim_wait_sprint", "mrisim_stream_wait_sprint", "mrisim_pending_sprints",
    "mrisim_nccl_unique_id", "mrisim_comm_init_rank", "mrisim_comm_attach", "mrisim_comm_size", "mrisim_nccl_origin",
    "mrisim_get_order", "mrisim_set_order", "mrisim_walker_count", "mrisim_first_walker_id", "mrisim_walker_seed",
    "mrisim_ctx_scene_digest", "mrisim_ctx_params", "mrisim_ctx_device", "mrisim_params_hash",
    "mrisim_checkpoint_save", "mrisim_checkpoint_info_read", "mrisim_checkpoint_load", "mrisim_checkpoint_last_error",
    "mrisim_group_create", "mrisim_group_destroy", "mrisim_group_last_error", "mrisim_group_size", "mrisim_group_ctx",
    "mrisim_group_set_scene", "mrisim_group_set_waters", "mrisim_group_get_waters", "mrisim_group_get_rng", "mrisim_group_get_time",
    "mrisim_group_submit_sprint", "mrisim_group_wait_sprint", "mrisim_group_run_sprint",
    "mrisim_group_checkpoint_save", "mrisim_group_checkpoint_load", "mrisim_group_kernel_launches",
]


def ParameterStruct(**overrides) -> Params:
    """The shipped ``parameters.h`` values; keyword overrides use the reference's knob names.
    Derived knobs (t, tcp, pfreq, in_stdev, out_stdev) are recomputed unless overridden."""
    p = Params()
    lib().mrisim_params_default(C.byref(p))
    derived = {"t", "tcp", "pfreq", "in_stdev", "out_stdev"}
    for k, v in overrides.items():
        if k not in derived:
            if not hasattr(p, k):
                raise AttributeError(f"ParameterStruct has no knob {k!r}")
            setattr(p, k, v)
    lib().mrisim_params_derive(C.byref(p))
    for k, v in overrides.items():
        if k in derived:
            setattr(p, k, v)
    return p


def low_permeability(p: Params) -> Params:
    """reflectIO / reflectOI from the commented formulas of parameters.h:116-117."""
    lib().mrisim_params_low_permeability(C.byref(p))
    return p


class Octree:
    """View of the hashed octree forest (reference ``Octree``: min_depth, max_depth, space)."""

    def __init__(self, min_depth, max_depth, subtree_sizes, nodes, residents):
        self.min_depth, self.max_depth = min_depth, max_depth
        self.subtree_sizes, self.nodes, self.residents = subtree_sizes, nodes, residents

    @property
    def subtree_prefix(self) -> np.ndarray:
        out = np.zeros(len(self.subtree_sizes) + 1, dtype=np.uint64)
        np.cumsum(self.subtree_sizes, out=out[1:])
        return out


def _view(ptr, dtype, count):
    if not count:
        return np.zeros(0, dtype=dtype)
    buf = (C.c_char * (np.dtype(dtype).itemsize * count)).from_address(ptr)
    return np.frombuffer(buf, dtype=dtype, count=count)


class BacteriaBox:
    """``BacteriaBox simBox(&gen); simBox.populateSimulation();`` (src/cuda_sim.cu:920-921).

    ``seed`` replaces the reference's ``time(NULL) + rd()`` (src/cuda_sim.cu:911-912).
    Accessors return numpy views into library-owned memory, valid while this object lives.
    """

    def __init__(self, params: Params, seed: int, n_threads: int = 0):
        self.params, self.seed, self.n_threads = params.copy(), int(seed), n_threads
        self._h = C.c_void_p()

    @classmethod
    def load(cls, path: str, params: Optional[Params] = None) -> "BacteriaBox":
        self = cls(params if params is not None else ParameterStruct(), 0)
        if lib().mrisim_scene_load(path.encode(), C.byref(self._h)):
            raise MrisimError(lib().mrisim_scene_last_error().decode())
        return self

    def populateSimulation(self, device: Optional[int] = None) -> "BacteriaBox":
        """``device=None``: the octree is built by the host threads (as the reference does,
        octree.cpp:410-414); ``device=k``: on GPU k (csrc/octree_build.cu), byte-identical result,
        ``self.octree_build_stats`` then holds the builder's counters."""
        if self._h:
            return self
        if device is None:
            rc = lib().mrisim_scene_build(C.byref(self.params), self.seed, self.n_threads, C.byref(self._h))
        else:
            self.octree_build_stats = OctreeBuildStats()
            rc = lib().mrisim_scene_build_gpu(C.byref(self.params), self.seed, self.n_threads, int(device),
                                              C.byref(self._h), C.byref(self.octree_build_stats))
        if rc:
            raise MrisimError(lib().mrisim_scene_last_error().decode())
        return self

    def save(self, path: str) -> None:
        if lib().mrisim_scene_save(self._handle(), path.encode()):
            raise MrisimError(lib().mrisim_scene_last_error().decode())

    def rethrow_waters(self, num_water: int, seed: int, sort_by_morton: bool = True, params: Optional[Params] = None,
                       device: Optional[int] = None):
        pp = C.byref(params) if params is not None else None
        if device is None:
            rc = lib().mrisim_scene_rethrow_waters(self._handle(), pp, num_water, seed, int(sort_by_morton))
        else:
            rc = lib().mrisim_scene_rethrow_waters_gpu(self._handle(), pp, num_water, seed, int(sort_by_morton), int(device))
        if rc:
            raise MrisimError(lib().mrisim_scene_last_error().decode())

    def _handle(self):
        if not self._h:
            raise AssertionError("populateSimulation() has not been called")   # reference: assert(populated)
        return self._h

    def getCells(self) -> np.ndarray:
        n = C.c_int32()
        p = lib().mrisim_scene_cells(self._handle(), C.byref(n))
        return _view(p, np.float64, n.value * 3).reshape(-1, 3)

    def getMNPs(self) -> np.ndarray:
        n = C.c_int32()
        p = lib().mrisim_scene_mnps(self._handle(), C.byref(n))
        return _view(p, cabi.MNP_DTYPE, n.value)

    def getMNPCount(self) -> int:
        return len(self.getMNPs())

    def getLookupTable(self) -> np.ndarray:
        h, m = C.c_int32(), C.c_int32()
        p = lib().mrisim_scene_lookup(self._handle(), C.byref(h), C.byref(m))
        return _view(p, np.int32, h.value ** 3 * m.value).reshape(-1, m.value)

    def getWaters(self) -> np.ndarray:
        n = C.c_int32()
        p = lib().mrisim_scene_waters(self._handle(), C.byref(n))
        return _view(p, cabi.WATER_DTYPE, n.value)

    def getOctree(self) -> Octree:
        mind, maxd, nsub = C.c_int32(), C.c_int32(), C.c_int32()
        sizes, nodes, res = C.c_void_p(), C.c_void_p(), C.c_void_p()
        nn, nr = C.c_uint64(), C.c_uint64()
        lib().mrisim_scene_octree(self._handle(), C.byref(mind), C.byref(maxd), C.byref(sizes), C.byref(nsub),
                                  C.byref(nodes), C.byref(nn), C.byref(res), C.byref(nr))
        return Octree(mind.value, maxd.value, _view(sizes.value, np.uint32, nsub.value),
                      _view(nodes.value, cabi.NODE_DTYPE, nn.value), _view(res.value, cabi.MNP_DTYPE, nr.value))

    def __del__(self):
        try:
            if self._h and _lib is not None:
                _lib.mrisim_scene_free(self._h)
                self._h = C.c_void_p()
        except Exception:
            pass


class WalkerContext:
    """Device context of the sprint loop (src/cuda_sim.cu:923-1024 minus the CSV line)."""

    def __init__(self, params: Params, device: int = 0, _borrowed=None):
        self.params = params.copy()
        self._h = C.c_void_p()
        self._owned = _borrowed is None
        if _borrowed is not None:                      # a member of a WalkerGroup: the group owns the handle
            self._h = C.c_void_p(_borrowed)
            self.n = int(lib().mrisim_walker_count(self._h))
            return
        if lib().mrisim_create(C.byref(self.params), device, C.byref(self._h)):
            raise MrisimError(lib().mrisim_last_error(None).decode())
        self.n = 0

    def _ck(self, rc):
        if rc:
            raise MrisimError(lib().mrisim_last_error(self._h).decode())

    def set_field_mode(self, mode: int):
        self._ck(lib().mrisim_set_field_mode(self._h, mode))

    def set_path_mode(self, mode: int):
        """PATH_AUTO (default) / PATH_FUSED / PATH_PIPELINE, see include/mrisim.h."""
        self._ck(lib().mrisim_set_path_mode(self._h, mode))

    @property
    def pipeline_sprints(self) -> int:
        return int(lib().mrisim_pipeline_sprints(self._h))

    def set_resort_steps(self, steps: int):
        """Steps per kernel launch; walkers are re-ordered by octree leaf between launches
        (<0: default = 125, 0: one launch per sprint in array order)."""
        self._ck(lib().mrisim_set_resort_steps(self._h, steps))

    def set_overlap_groups(self, groups: int):
        """Walker ranges whose launch chains overlap on separate streams (default 1)."""
        self._ck(lib().mrisim_set_overlap_groups(self._h, groups))

    def set_scene(self, box: BacteriaBox):
        self._ck(lib().mrisim_set_scene(self._h, box._handle()))

    def set_cells(self, cells: np.ndarray):
        c = np.ascontiguousarray(cells, dtype=np.float64)
        self._ck(lib().mrisim_set_cells(self._h, c.ctypes.data, len(c)))

    def set_lookup(self, rows: np.ndarray, hashDim: int):
        r = np.ascontiguousarray(rows, dtype=np.int32)
        self._ck(lib().mrisim_set_lookup(self._h, r.ctypes.data, hashDim, r.shape[1]))

    def set_octree(self, min_depth, max_depth, subtree_sizes, nodes, residents):
        s = np.ascontiguousarray(subtree_sizes, dtype=np.uint32)
        n = np.ascontiguousarray(nodes)
        r = np.ascontiguousarray(residents)
        assert n.dtype == cabi.NODE_DTYPE and r.dtype.itemsize == 40
        self._ck(lib().mrisim_set_octree(self._h, min_depth, max_depth, s.ctypes.data, len(s), n.ctypes.data, len(n),
                                         r.ctypes.data if len(r) else None, len(r)))

    def set_waters(self, waters: np.ndarray, walker_seed: int, first_walker_id: int = 0):
        w = np.ascontiguousarray(waters)
        assert w.dtype == cabi.WATER_DTYPE
        self._ck(lib().mrisim_set_waters(self._h, w.ctypes.data, len(w), walker_seed, first_walker_id))
        self.n = len(w)

    def get_waters(self) -> np.ndarray:
        out = np.zeros(self.n, dtype=cabi.WATER_DTYPE)
        self._ck(lib().mrisim_get_waters(self._h, out.ctypes.data, self.n))
        return out

    def get_rng(self) -> np.ndarray:
        out = np.zeros(self.n, dtype=cabi.RNG_DTYPE)
        self._ck(lib().mrisim_get_rng(self._h, out.ctypes.data, self.n))
        return out

    def set_rng(self, rng: np.ndarray):
        r = np.ascontiguousarray(rng)
        assert r.dtype == cabi.RNG_DTYPE
        self._ck(lib().mrisim_set_rng(self._h, r.ctypes.data, len(r)))

    @property
    def time(self) -> int:
        return lib().mrisim_get_time(self._h)

    @time.setter
    def time(self, t: int):
        self._ck(lib().mrisim_set_time(self._h, t))

    def run_sprint(self, steps: Optional[int] = None, out: Optional[np.ndarray] = None) -> np.ndarray:
        steps = self.params.sprintSteps if steps is None else steps
        if out is None:
            out = np.zeros(max(steps // max(self.params.pfreq, 1), 1), dtype=np.float64)
        self._ck(lib().mrisim_run_sprint(self._h, steps, out.ctypes.data))
        return out

    def run_sprint_async(self, steps: int, signal_dev_ptr: int, stream: int = 0):
        """signal_dev_ptr: device pointer to steps/pfreq doubles (e.g. tensor.data_ptr());
        stream: a cudaStream_t handle (e.g. torch.cuda.current_stream().cuda_stream; 0 is
        CUDA's legacy default stream)."""
        self._ck(lib().mrisim_run_sprint_async(self._h, steps, signal_dev_ptr, stream))

    def trace_sprint(self, steps: int, kicks: bool = False):
        n = self.n
        sig = np.zeros(steps // self.params.pfreq, dtype=np.float64)
        tr = dict(x=np.zeros((steps, n)), y=np.zeros((steps, n)), z=np.zeros((steps, n)),
                  in_cell=np.zeros((steps, n), np.uint8), leaf=np.zeros((steps, n), np.uint32))
        if kicks:
            tr["kick"] = np.zeros((steps, n))
        self._ck(lib().mrisim_trace_sprint_kicks(self._h, steps, sig.ctypes.data, tr["x"].ctypes.data, tr["y"].ctypes.data,
                                                 tr["z"].ctypes.data, tr["in_cell"].ctypes.data, tr["leaf"].ctypes.data,
                                                 tr["kick"].ctypes.data if kicks else None))
        return sig, tr

    # ---- submitted sprints, communicator, checkpoints (include/mrisim.h) ---------------------------
    def check(self):
        """Wait for the last sprint launched so far; read and clear the device error flag."""
        self._ck(lib().mrisim_check(self._h))

    def submit_sprint(self, steps: Optional[int] = None, stream: int = 0):
        """Enqueue a sprint (kernels on ``stream`` or the context's own; all-reduce over the attached
        communicator and D2H of the samples on the side stream) and return at once."""
        steps = self.params.sprintSteps if steps is None else steps
        self._ck(lib().mrisim_submit_sprint(self._h, steps, stream or None))

    def wait_sprint(self, steps: Optional[int] = None) -> np.ndarray:
        """Samples of the oldest outstanding sprint (global signal when a communicator is attached)."""
        steps = self.params.sprintSteps if steps is None else steps
        out = np.zeros(max(steps // max(self.params.pfreq, 1), 1), dtype=np.float64)
        self._ck(lib().mrisim_wait_sprint(self._h, out.ctypes.data))
        return out

    def stream_wait_sprint(self, stream: int):
        self._ck(lib().mrisim_stream_wait_sprint(self._h, stream))

    @property
    def pending_sprints(self) -> int:
        return int(lib().mrisim_pending_sprints(self._h))

    def comm_init_rank(self, unique_id: bytes, nranks: int, rank: int):
        """Collective: every rank passes the 128 bytes rank 0 got from ``nccl_unique_id()``."""
        assert len(unique_id) == 128
        buf = C.create_string_buffer(unique_id, 128)
        self._ck(lib().mrisim_comm_init_rank(self._h, buf, nranks, rank))

    @property
    def comm_size(self) -> int:
        return int(lib().mrisim_comm_size(self._h))

    def get_order(self):
        perm = np.zeros(self.n, np.uint32); keys = np.zeros(self.n, np.uint32); valid = C.c_int32()
        self._ck(lib().mrisim_get_order(self._h, perm.ctypes.data, keys.ctypes.data, self.n, C.byref(valid)))
        return (perm, keys) if valid.value else None

    def set_order(self, perm: np.ndarray, keys: np.ndarray):
        pm = np.ascontiguousarray(perm, np.uint32); ky = np.ascontiguousarray(keys, np.uint32)
        self._ck(lib().mrisim_set_order(self._h, pm.ctypes.data, ky.ctypes.data, len(pm)))

    def checkpoint_save(self, path: str, host_seed: int = 0, csv_bytes: int = 0):
        if lib().mrisim_checkpoint_save(self._h, path.encode(), host_seed, csv_bytes):
            raise MrisimError(lib().mrisim_checkpoint_last_error().decode())

    def checkpoint_load(self, path: str, lo: int = 0, count: int = 0):
        if lib().mrisim_checkpoint_load(self._h, path.encode(), lo, count):
            raise MrisimError(lib().mrisim_checkpoint_last_error().decode())
        self.n = int(lib().mrisim_walker_count(self._h))

    @property
    def kernel_launches(self) -> int:
        return lib().mrisim_kernel_launches(self._h)

    def last_sprint_kernel_ms(self) -> float:
        return lib().mrisim_last_sprint_kernel_ms(self._h)

    def close(self):
        if self._h and _lib is not None and self._owned:
            _lib.mrisim_destroy(self._h)
        self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class WalkerGroup:
    """One process, several GPUs: walkers sharded over ``devices``, scene replicated, one NCCL
    all-reduce of the signal partials per sprint on a side stream (``mrisim_group_*``; replaces the
    reference's ``Handler handle(0)``, src/cuda_sim.cu:916)."""

    def __init__(self, params: Params, devices):
        self.params = params.copy()
        self.devices = [int(d) for d in devices]
        self._h = C.c_void_p()
        arr = (C.c_int * len(self.devices))(*self.devices)
        if lib().mrisim_group_create(C.byref(self.params), arr, len(self.devices), C.byref(self._h)):
            raise MrisimError(lib().mrisim_group_last_error(None).decode())
        self.n = 0

    def _ck(self, rc):
        if rc:
            raise MrisimError(lib().mrisim_group_last_error(self._h).decode())

    def member(self, i: int) -> WalkerContext:
        h = lib().mrisim_group_ctx(self._h, i)
        if not h:
            raise IndexError(i)
        return WalkerContext(self.params, _borrowed=h)

    def __len__(self):
        return int(lib().mrisim_group_size(self._h))

    def set_scene(self, box: BacteriaBox):
        self._ck(lib().mrisim_group_set_scene(self._h, box._handle()))

    def set_waters(self, waters: np.ndarray, walker_seed: int, first_walker_id: int = 0):
        w = np.ascontiguousarray(waters)
        assert w.dtype == cabi.WATER_DTYPE
        self._ck(lib().mrisim_group_set_waters(self._h, w.ctypes.data, len(w), walker_seed, first_walker_id))
        self.n = len(w)

    def get_waters(self) -> np.ndarray:
        out = np.zeros(self.n, dtype=cabi.WATER_DTYPE)
        self._ck(lib().mrisim_group_get_waters(self._h, out.ctypes.data, self.n))
        return out

    def get_rng(self) -> np.ndarray:
        out = np.zeros(self.n, dtype=cabi.RNG_DTYPE)
        self._ck(lib().mrisim_group_get_rng(self._h, out.ctypes.data, self.n))
        return out

    @property
    def time(self) -> int:
        return lib().mrisim_group_get_time(self._h)

    def submit_sprint(self, steps: Optional[int] = None):
        self._ck(lib().mrisim_group_submit_sprint(self._h, self.params.sprintSteps if steps is None else steps))

    def wait_sprint(self, steps: Optional[int] = None) -> np.ndarray:
        steps = self.params.sprintSteps if steps is None else steps
        out = np.zeros(max(steps // max(self.params.pfreq, 1), 1), dtype=np.float64)
        self._ck(lib().mrisim_group_wait_sprint(self._h, out.ctypes.data))
        return out

    def run_sprint(self, steps: Optional[int] = None) -> np.ndarray:
        self.submit_sprint(steps)
        return self.wait_sprint(steps)

    def checkpoint_save(self, path: str, host_seed: int = 0, csv_bytes: int = 0):
        self._ck(lib().mrisim_group_checkpoint_save(self._h, path.encode(), host_seed, csv_bytes))

    def checkpoint_load(self, path: str):
        self._ck(lib().mrisim_group_checkpoint_load(self._h, path.encode()))
        info = checkpoint_info(path)
        self.n = int(info.n_walkers)

    @property
    def kernel_launches(self) -> int:
        return int(lib().mrisim_group_kernel_launches(self._h))

    def close(self):
        if self._h and _lib is not None:
            _lib.mrisim_group_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def nccl_unique_id() -> bytes:
    buf = C.create_string_buffer(128)
    if lib().mrisim_nccl_unique_id(buf):
        raise MrisimError(lib().mrisim_last_error(None).decode())
    return buf.raw


def nccl_origin() -> str:
    return lib().mrisim_nccl_origin().decode()


def checkpoint_info(path: str) -> "cabi.CheckpointInfo":
    info = cabi.CheckpointInfo()
    if lib().mrisim_checkpoint_info_read(path.encode(), C.byref(info)):
        raise MrisimError(lib().mrisim_checkpoint_last_error().decode())
    return info


def params_hash(p: Params) -> int:
    return int(lib().mrisim_params_hash(C.byref(p)))


def csv_row(time_steps: int, tau: float, target: float) -> str:
    """One line of the signal CSV, formatted as src/cuda_sim.cu:1022 does."""
    buf = C.create_string_buffer(128)
    n = lib().mrisim_csv_row(time_steps, tau, target, buf, len(buf))
    if n < 0:
        raise MrisimError("csv_row: buffer too small")
    return buf.value.decode()


def measure_fp64_tflops(device: int = 0, iters: int = 1 << 16) -> float:
    v = lib().mrisim_measure_fp64_tflops(device, iters)
    if v < 0:
        raise MrisimError("mrisim_measure_fp64_tflops failed (no CUDA device?)")
    return v


def simulateWaters(filename: str, params: Optional[Params] = None, host_seed: int = 12345,
                   walker_seed: int = 0xC0FFEE, device: int = 0, max_sprints: int = -1, devices=None,
                   checkpoint_every: int = 0, checkpoint_path: Optional[str] = None, resume: Optional[str] = None) -> None:
    """``simulateWaters(fName)`` (src/cuda_sim.cu:902-1034) through the C++ drop-in: builds the
    scene, runs t/sprintSteps sprints on ``device`` (or sharded over ``devices``) and writes the signal
    CSV to ``filename``; optional on-disk checkpoints every ``checkpoint_every`` sprints and ``resume``."""
    p = params if params is not None else ParameterStruct()
    if devices is None and not checkpoint_every and not resume:
        rc = lib().mrisim_simulate_waters(filename.encode(), C.byref(p), host_seed, walker_seed, device, max_sprints)
    else:
        devs = [int(d) for d in (devices if devices is not None else [device])]
        arr = (C.c_int * len(devs))(*devs)
        rc = lib().mrisim_simulate_waters_multi(filename.encode(), C.byref(p), host_seed, walker_seed, arr, len(devs), max_sprints,
                                                checkpoint_every, checkpoint_path.encode() if checkpoint_path else None,
                                                resume.encode() if resume else None)
    if rc:
        raise MrisimError("simulateWaters failed (see stderr)")
