"""GPU octree build (csrc/octree_build.cu) against the reference's host build.

The GPU-built scene must be byte-identical, as a whole MRISCN01 file, to (i) the committed dumps
of the reference's own host code (tests/golden/), (ii) the sha256 digests of the larger configs,
(iii) the product host builder on fresh seeds -- including when the certified error bounds are
inflated so that a large share of the values goes through the host libm fix-up path."""
import gzip
import hashlib
import json
import os

import numpy as np
import pytest

from mri_nanoparticle_simulation_b200 import api, scene_io as sio

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = os.path.join(HERE, "golden")
KATS = json.load(open(os.path.join(GOLD, "scene_kats.json")))


def dump(cfg, seed, tmp_path, device=None, tag=""):
    p = api.ParameterStruct(**KATS["configs"][cfg])
    box = api.BacteriaBox(p, seed).populateSimulation(device=device)
    out = str(tmp_path / f"{cfg}_{seed}_{tag}.scn")
    box.save(out)
    return open(out, "rb").read(), getattr(box, "octree_build_stats", None)


def test_refuses_without_a_device():
    """No CPU fallback behind the GPU entry point: a bad device index is an error."""
    p = api.ParameterStruct(**KATS["configs"]["tiny"])
    with pytest.raises(api.MrisimError, match="no CUDA device|CUDA|cuda"):
        api.BacteriaBox(p, 3).populateSimulation(device=63)


def test_gpu_rethrow_refuses_without_a_device():
    p = api.ParameterStruct(**KATS["configs"]["tiny"])
    box = api.BacteriaBox(p, 3).populateSimulation()
    before = box.getWaters().tobytes()
    with pytest.raises(api.MrisimError, match="no CUDA device|CUDA|cuda"):
        box.rethrow_waters(100, 5, device=63)
    with pytest.raises(api.MrisimError, match="device must be >= 0"):
        box.rethrow_waters(100, 5, device=-1)
    assert box.getWaters().tobytes() == before          # a refused call leaves the scene alone


@pytest.mark.gpu
@pytest.mark.parametrize("cfg,seed", [("tiny", 3), ("default", 12345)])
def test_gpu_build_byte_identical_to_reference_dump(cfg, seed, tmp_path):
    want = gzip.open(os.path.join(GOLD, f"scene_{cfg}_seed{seed}.scn.gz")).read()
    got, st = dump(cfg, seed, tmp_path, device=0)
    assert got == want
    assert st.n_nodes == sio.loads(got).n_nodes and st.kernel_launches > 0 and st.pair_evals > 0
    print(f"\n{cfg}: {st.asdict()}")


@pytest.mark.gpu
@pytest.mark.parametrize("key", ["default_seed7", "noavoid_seed99", "n8000_seed5", "dense2000_seed11", "dense8000_seed11"])
def test_gpu_build_digests_of_larger_configs(key, tmp_path):
    """Digests written by the REFERENCE'S own host code (tests/golden/make_golden.py).  dense8000 is
    SURVEY.md 8d config 4b -- 21 322 dipoles, 455 008 nodes: six minutes of the reference's octree
    constructor on 8 cores, under a second here."""
    cfg, seed = key.rsplit("_seed", 1)
    raw, st = dump(cfg, int(seed), tmp_path, device=0)
    s = sio.loads(raw)
    want = KATS["scenes"][key]
    h = lambda a: hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()
    for f in ("cells", "mnps", "lookup", "subtree_sizes", "nodes", "residents", "waters"):
        assert h(getattr(s, f)) == want[f], f"{key}: {f} differs from the reference"
    print(f"\n{key}: {st.asdict()}")


@pytest.mark.gpu
@pytest.mark.parametrize("cfg,seed", [("tiny", 8), ("tiny", 9), ("tiny", 17), ("default", 2024), ("noavoid", 31337)])
def test_gpu_build_equals_host_build_on_fresh_seeds(cfg, seed, tmp_path):
    host, _ = dump(cfg, seed, tmp_path, tag="host")
    gpu, st = dump(cfg, seed, tmp_path, device=0, tag="gpu")
    assert gpu == host
    # the certification is doing the work, not the fix-up path
    assert st.b_fixups <= max(8, st.b_values // 500) and st.criterion_fixups <= max(2, st.criterion_evals // 500)


@pytest.mark.gpu
@pytest.mark.parametrize("scale", [1e4, 1e9])
def test_fixup_path_gives_the_same_bytes(scale, tmp_path):
    """Inflated bounds push many samples (1e4) or nearly everything (1e9) through the host libm
    re-evaluation: the file must not change, only the counters."""
    host, _ = dump("default", 12345, tmp_path, tag="host")
    L = api.lib()
    L.mrisim_octree_build_set_margin_scale(scale)
    try:
        gpu, st = dump("default", 12345, tmp_path, device=0, tag=f"gpu{scale:g}")
    finally:
        L.mrisim_octree_build_set_margin_scale(1.0)
    assert gpu == host
    assert st.b_fixups > 0
    if scale >= 1e9:
        assert st.b_fixups > st.b_values // 2 and st.criterion_fixups > 0


@pytest.mark.gpu
def test_gpu_built_scene_drives_the_walk(tmp_path):
    """A GPU-built scene goes straight into the walker context and gives the signal of the
    host-built one (same bytes in, same bytes out)."""
    p = api.ParameterStruct(**KATS["configs"]["tiny"])
    sig = []
    for dev in (None, 0):
        box = api.BacteriaBox(p, 3).populateSimulation(device=dev)
        ctx = api.WalkerContext(p, 0)
        ctx.set_scene(box)
        ctx.set_waters(box.getWaters().copy(), 0xC0FFEE)
        sig.append(np.asarray(ctx.run_sprint(2 * p.pfreq)).copy())
    assert sig[0].tobytes() == sig[1].tobytes()


# ---- the other scene-build stages on the GPU (csrc/scene_gpu.cu) ----------------

@pytest.mark.gpu
@pytest.mark.parametrize("n,avoid", [(5000, 1), (5000, 0), (1, 1), (0, 1), (200_000, 1)])
def test_gpu_water_throw_equals_host(n, avoid):
    """Same seed -> same waters, byte for byte, sorted or not (BacteriaBox.cpp:80-105,
    SimulationBox.cpp:131-147,161-173)."""
    p = api.ParameterStruct(**KATS["configs"]["tiny"])
    p.avoid_intracellular_throw = avoid
    box = api.BacteriaBox(p, 3).populateSimulation()
    for sort in (False, True):
        # tobytes() of the library-owned view: whole 40-byte records, padding included
        # (a numpy .copy() of a padded structured array leaves the padding undefined)
        box.rethrow_waters(n, 99, sort_by_morton=sort, params=p)
        host = box.getWaters().tobytes()
        box.rethrow_waters(n, 99, sort_by_morton=sort, params=p, device=0)
        gpu = box.getWaters().tobytes()
        assert len(gpu) == 40 * n and gpu == host


@pytest.mark.gpu
def test_gpu_scene_build_leaves_the_host_stream_where_the_host_build_does():
    """The water throw rewinds the host generator to just after the last accepted attempt; a
    second scene stage drawing from it (none today) would see the same stream.  Checked through
    what depends on it: the whole scene file, for a config whose throw needs several batches
    (1 water: batch 65, and 300 000 waters in the 40-cell box)."""
    import tempfile
    for nw in (1, 300_000):
        p = api.ParameterStruct(**dict(KATS["configs"]["tiny"], num_water=nw))
        blobs = []
        for dev in (None, 0):
            box = api.BacteriaBox(p, 21).populateSimulation(device=dev)
            with tempfile.NamedTemporaryFile(suffix=".scn") as f:
                box.save(f.name)
                blobs.append(open(f.name, "rb").read())
        assert blobs[0] == blobs[1]


@pytest.mark.gpu
def test_gpu_lookup_overflow_is_reported():
    p = api.ParameterStruct(num_cells=2000, num_water=16)       # default hashDim=20 / maxNeighbors=13 overflow
    with pytest.raises(api.MrisimError, match="maxNeighbors"):
        api.BacteriaBox(p, 1).populateSimulation(device=0)
