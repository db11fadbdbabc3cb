"""mri_nanoparticle_simulation_b200 -- B200-native walker / T2-signal path.

Only what the hot path needs: ``csrc/`` (sm_100a kernels, host scene builder, C ABI), the
ctypes mirror of the reference interface (``api``), walker sharding (``dist``) and the scene
interchange file (``scene_io``).
"""
from .api import (BacteriaBox, MrisimError, Octree, ParameterStruct, WalkerContext, csv_row,  # noqa: F401
                  low_permeability, measure_fp64_tflops, simulateWaters)
from ._cabi import FIELD_EXACT, FIELD_FAST, KICK_CONSTANT, KICK_NONE, KICK_RANDOM, Params  # noqa: F401

__all__ = ["BacteriaBox", "MrisimError", "Octree", "ParameterStruct", "WalkerContext", "csv_row",
           "low_permeability", "measure_fp64_tflops", "simulateWaters", "Params",
           "FIELD_EXACT", "FIELD_FAST", "KICK_NONE", "KICK_RANDOM", "KICK_CONSTANT"]
