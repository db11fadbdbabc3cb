"""Reader/writer for the ``MRISCN01`` scene interchange file.

The file carries exactly what the reference's GPU driver pulls out of ``SimulationBox``
before the sprint loop (reference ``src/cuda_sim.cu:926-964``): cell centres
(``getCells``), MNPs incl. periodic images (``getMNPs``), the nearest-cell lookup table
(``getLookupTable``), the octree forest (``getOctree``: per-subtree DFS node vectors with
their resident lists) and the Morton-sorted waters (``getWaters``).  Both the product scene
builder (``csrc/scene.cpp``) and the oracle-side dump of the reference's own host code
(``oracle/ref_scene_dump.cpp``) write it, so scene parity is a byte comparison.
"""
from __future__ import annotations

import dataclasses
import struct

import numpy as np

MAGIC = b"MRISCN01"

#: one octree node as exchanged over the C ABI (``mrisim_node`` in include/mrisim.h);
#: mirrors reference ``oct_node`` (octree.h:28-33) with the resident vector flattened.
NODE_DTYPE = np.dtype(
    [("mc", "<u8"), ("child", "<u4", (8,)), ("res_start", "<u4"), ("n_res", "<u4")]
)
#: reference ``MNP_info`` (rand_walk.h:58-69): x, y, z, r, M
MNP_DTYPE = np.dtype([("x", "<f8"), ("y", "<f8"), ("z", "<f8"), ("r", "<f8"), ("M", "<f8")])

_HDR = struct.Struct("<8sQ8iQQ4dq")


@dataclasses.dataclass
class Scene:
    seed: int
    num_cells: int
    num_mnps: int
    hash_dim: int
    max_neighbors: int
    min_depth: int
    max_depth: int
    num_water: int
    n_subtrees: int
    bound: float
    cell_r: float
    mmoment: float
    water_start_bound: float
    avoid_intracellular_throw: int
    cells: np.ndarray          # (num_cells, 3) f8
    mnps: np.ndarray           # (num_mnps,) MNP_DTYPE
    lookup: np.ndarray         # (hash_dim**3, max_neighbors) i4
    subtree_sizes: np.ndarray  # (n_subtrees,) u4
    nodes: np.ndarray          # (n_nodes,) NODE_DTYPE
    residents: np.ndarray      # (n_res,) MNP_DTYPE
    waters: np.ndarray         # (num_water, 4) f8: x, y, z, phase

    @property
    def n_nodes(self) -> int:
        return int(self.nodes.shape[0])

    @property
    def subtree_prefix(self) -> np.ndarray:
        """flat node index of subtree i's root: reference (i, j) -> prefix[i] + j."""
        out = np.zeros(self.n_subtrees + 1, dtype=np.uint64)
        np.cumsum(self.subtree_sizes, out=out[1:])
        return out


def loads(buf: bytes) -> Scene:
    (magic, seed, ncell, nmnp, hdim, maxn, mind, maxd, nwat, nsub, nnodes, nres,
     bound, cell_r, mmoment, wsb, avoid) = _HDR.unpack_from(buf, 0)
    if magic != MAGIC:
        raise ValueError("not an MRISCN01 scene file")
    off = _HDR.size

    def take(dtype, count, shape=None):
        nonlocal off
        a = np.frombuffer(buf, dtype=dtype, count=count, offset=off).copy()
        off += a.nbytes
        return a.reshape(shape) if shape is not None else a

    cells = take("<f8", ncell * 3, (ncell, 3))
    mnps = take(MNP_DTYPE, nmnp)
    lookup = take("<i4", hdim ** 3 * maxn, (hdim ** 3, maxn))
    sizes = take("<u4", nsub)
    nodes = take(NODE_DTYPE, nnodes)
    residents = take(MNP_DTYPE, nres)
    waters = take("<f8", nwat * 4, (nwat, 4))
    if off != len(buf):
        raise ValueError(f"trailing bytes in scene file: {len(buf) - off}")
    return Scene(seed, ncell, nmnp, hdim, maxn, mind, maxd, nwat, nsub, bound, cell_r,
                 mmoment, wsb, avoid, cells, mnps, lookup, sizes, nodes, residents, waters)


def load(path: str) -> Scene:
    with open(path, "rb") as f:
        return loads(f.read())


def dumps(s: Scene) -> bytes:
    hdr = _HDR.pack(MAGIC, s.seed, s.num_cells, s.num_mnps, s.hash_dim, s.max_neighbors,
                    s.min_depth, s.max_depth, s.num_water, s.n_subtrees, s.n_nodes,
                    int(s.residents.shape[0]), s.bound, s.cell_r, s.mmoment,
                    s.water_start_bound, s.avoid_intracellular_throw)
    parts = [hdr, np.ascontiguousarray(s.cells, "<f8").tobytes(), s.mnps.tobytes(),
             np.ascontiguousarray(s.lookup, "<i4").tobytes(), s.subtree_sizes.astype("<u4").tobytes(),
             s.nodes.tobytes(), s.residents.tobytes(), np.ascontiguousarray(s.waters, "<f8").tobytes()]
    return b"".join(parts)


def fnv1a64(data: bytes) -> int:
    """FNV-1a-64 as used for the survey-time known answers."""
    h = 1469598103934665603
    # vectorising FNV is not possible (serial dependency); keep inputs small (<~2 MB)
    for b in data:
        h = ((h ^ b) * 1099511628211) & 0xFFFFFFFFFFFFFFFF
    return h
