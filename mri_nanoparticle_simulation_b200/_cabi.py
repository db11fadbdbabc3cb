"""ctypes mirror of include/mrisim.h (struct layouts and numpy dtypes)."""
from __future__ import annotations

import ctypes as C

import numpy as np

KICK_NONE, KICK_RANDOM, KICK_CONSTANT = 0, 1, 2
FIELD_EXACT, FIELD_FAST = 0, 1
PATH_AUTO, PATH_FUSED, PATH_PIPELINE = 0, 1, 2     # include/mrisim.h


class Params(C.Structure):
    """``mrisim_params`` -- run-time mirror of the reference's ``ParameterStruct``
    (src/parameters.h:25-126); same field names."""
    _fields_ = [
        ("g", C.c_double), ("scale", C.c_double), ("border", C.c_double),
        ("num_threads", C.c_int32), ("sprintSteps", C.c_int32),
        ("num_water", C.c_int32), ("num_cells", C.c_int32),
        ("cell_r", C.c_double), ("mmoment", C.c_double),
        ("kick_mode", C.c_int32), ("avoid_intracellular_throw", C.c_int32),
        ("phase_stdev", C.c_double), ("phase_k", C.c_double),
        ("bound", C.c_double), ("water_start_bound", C.c_double),
        ("hashDim", C.c_int32), ("maxNeighbors", C.c_int32),
        ("D_cell", C.c_double), ("D_extra", C.c_double), ("P_expr", C.c_double),
        ("tau", C.c_double),
        ("totaltime", C.c_int32), ("t", C.c_int32),
        ("taucp", C.c_double),
        ("tcp", C.c_int32), ("pfreq", C.c_int32),
        ("reflectIO", C.c_double), ("reflectOI", C.c_double),
        ("in_stdev", C.c_double), ("out_stdev", C.c_double),
    ]

    def copy(self) -> "Params":
        q = Params()
        C.memmove(C.byref(q), C.byref(self), C.sizeof(Params))
        return q


class OctreeBuildStats(C.Structure):
    """``mrisim_octree_build_stats`` (include/mrisim.h)."""
    _fields_ = [("total_ms", C.c_double), ("gpu_ms", C.c_double), ("fixup_ms", C.c_double), ("linearise_ms", C.c_double),
                ("n_nodes", C.c_uint64), ("n_leaves", C.c_uint64), ("n_levels", C.c_uint64),
                ("criterion_evals", C.c_uint64), ("criterion_fixups", C.c_uint64),
                ("b_values", C.c_uint64), ("b_fixups", C.c_uint64), ("pair_evals", C.c_uint64),
                ("kernel_launches", C.c_uint64),
                ("scene_throw_ms", C.c_double), ("scene_lookup_ms", C.c_double), ("scene_sort_ms", C.c_double),
                ("fixup_wait_ms", C.c_double)]

    def asdict(self) -> dict:
        return {k: getattr(self, k) for k, _ in self._fields_}


class CheckpointInfo(C.Structure):
    """``mrisim_checkpoint_info`` (include/mrisim.h)."""
    _fields_ = [("params_hash", C.c_uint64), ("scene_digest", C.c_uint64), ("host_seed", C.c_uint64), ("walker_seed", C.c_uint64),
                ("time_steps", C.c_int64), ("csv_bytes", C.c_uint64), ("n_walkers", C.c_uint64), ("first_walker_id", C.c_uint64),
                ("n_shards", C.c_uint32), ("version", C.c_uint32)]


#: ``mrisim_water`` == reference ``water_info`` (src/rand_walk.h:25-53), 40 bytes
WATER_DTYPE = np.dtype({
    "names": ["x", "y", "z", "phase", "in_cell", "nearest"],
    "formats": ["<f8", "<f8", "<f8", "<f8", "u1", "<i4"],
    "offsets": [0, 8, 16, 24, 32, 36],
    "itemsize": 40,
})
#: ``mrisim_rng_state``
RNG_DTYPE = np.dtype([("s0", "<u8"), ("s1", "<u8"), ("spare", "<f8")])
#: ``mrisim_mnp`` == reference ``MNP_info``
MNP_DTYPE = np.dtype([("x", "<f8"), ("y", "<f8"), ("z", "<f8"), ("r", "<f8"), ("M", "<f8")])
#: ``mrisim_node``
NODE_DTYPE = np.dtype([("mc", "<u8"), ("child", "<u4", (8,)), ("res_start", "<u4"), ("n_res", "<u4")])

assert WATER_DTYPE.itemsize == 40 and NODE_DTYPE.itemsize == 48 and MNP_DTYPE.itemsize == 40
