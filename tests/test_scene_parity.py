"""The product host scene builder (csrc/scene.cpp) reproduces the reference's
BacteriaBox::populateSimulation() byte for byte: committed dumps of the reference's own host
code (tests/golden/), sha256 digests of the larger configs, and -- when oracle/_ref is
present (build container and GPU box) -- fresh seeds against the reference binary itself."""
import gzip
import hashlib
import json
import os
import subprocess

import numpy as np
import pytest

from mri_nanoparticle_simulation_b200 import api, scene_io as sio

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = os.path.join(HERE, "golden")
REF = os.path.join(os.path.dirname(HERE), "oracle", "_ref")
KATS = json.load(open(os.path.join(GOLD, "scene_kats.json")))


def build_and_dump(cfg, seed, tmp_path, n_threads=0):
    p = api.ParameterStruct(**KATS["configs"][cfg])
    box = api.BacteriaBox(p, seed, n_threads).populateSimulation()
    out = str(tmp_path / f"{cfg}_{seed}.scn")
    box.save(out)
    return open(out, "rb").read()


@pytest.mark.parametrize("cfg,seed", [("tiny", 3), ("default", 12345)])
def test_byte_identical_to_reference_dump(cfg, seed, tmp_path):
    want = gzip.open(os.path.join(GOLD, f"scene_{cfg}_seed{seed}.scn.gz")).read()
    got = build_and_dump(cfg, seed, tmp_path)
    assert got == want


@pytest.mark.parametrize("key", ["default_seed7", "noavoid_seed99", "n8000_seed5", "n8000_noavoid_seed21", "dense2000_seed11"])
def test_digests_of_larger_configs(key, tmp_path):
    cfg, seed = key.rsplit("_seed", 1)
    s = sio.loads(build_and_dump(cfg, int(seed), tmp_path))
    want = KATS["scenes"][key]
    h = lambda a: hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()
    for f in ("cells", "mnps", "lookup", "subtree_sizes", "nodes", "residents", "waters"):
        assert h(getattr(s, f)) == want[f], f"{key}: {f} differs from the reference"
    assert (s.num_mnps, s.n_nodes, s.min_depth, s.max_depth) == (want["num_mnps"], want["n_nodes"], want["min_depth"], want["max_depth"])


def test_thread_count_does_not_matter(tmp_path):
    assert build_and_dump("tiny", 17, tmp_path, n_threads=1) == build_and_dump("tiny", 17, tmp_path, n_threads=5)


@pytest.mark.parametrize("cfg,seed", [("default", 2024), ("noavoid", 31337), ("tiny", 8), ("tiny", 9), ("n8000", 77)])
def test_fresh_seeds_against_reference_binary(cfg, seed, tmp_path):
    exe = os.path.join(REF, f"ref_scene_{cfg}")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref not built (needs the reference tree: make -C oracle ref)")
    ref_out = str(tmp_path / "ref.scn")
    subprocess.check_call([exe, str(seed), ref_out], stdout=subprocess.DEVNULL)
    assert build_and_dump(cfg, seed, tmp_path) == open(ref_out, "rb").read()


def test_save_load_roundtrip(tmp_path):
    raw = build_and_dump("tiny", 3, tmp_path)
    box = api.BacteriaBox.load(str(tmp_path / "tiny_3.scn"))
    out = str(tmp_path / "again.scn")
    box.save(out)
    assert open(out, "rb").read() == raw
    assert sio.dumps(sio.loads(raw)) == raw


def test_lookup_overflow_is_an_error_not_an_assert():
    """The reference asserts ncells.size() < maxNeighbors (SimulationBox.cpp:281); the C ABI
    returns an error instead."""
    p = api.ParameterStruct(num_cells=2000, num_water=16)       # default hashDim=20 / maxNeighbors=13 overflow
    with pytest.raises(api.MrisimError, match="maxNeighbors"):
        api.BacteriaBox(p, 1).populateSimulation()


def test_accessors_before_populate_raise():
    box = api.BacteriaBox(api.ParameterStruct(), 1)
    with pytest.raises(AssertionError):          # reference: assert(populated), SimulationBox.cpp:537
        box.getWaters()


def test_rethrow_waters_semantics():
    """rethrow = init_waters + sortWaters from a fresh XORShift(seed): inside the start cube,
    outside every cell when AVOID_INTRACELLULAR_THROW, sorted by leaf Morton code."""
    p = api.ParameterStruct(**KATS["configs"]["tiny"])
    box = api.BacteriaBox(p, 3).populateSimulation()
    box.rethrow_waters(5000, 99)
    w = box.getWaters()
    assert len(w) == 5000
    lo = (p.bound - p.water_start_bound) / 2
    for f in "xyz":
        assert (w[f] >= lo).all() and (w[f] < lo + p.water_start_bound).all()
    cells = box.getCells()
    d2 = ((np.stack([w["x"], w["y"], w["z"]], 1)[:, None, :] - cells[None]) ** 2).sum(2).min(1)
    assert (d2 >= p.cell_r ** 2).all()
    import oracle_lib as ol
    orc = ol.oracle_for_box(box, p)
    mcs = [orc.probe_leaf(*q)[1] for q in zip(w["x"], w["y"], w["z"])]
    assert mcs == sorted(mcs)


def test_scene_load_refuses_malformed_files(tmp_path):
    """mrisim_scene_load trusts nothing in the header: counts are checked for sign and bounds and the implied
    payload size must equal the file's before anything is allocated; node links and lookup entries are checked."""
    import struct
    p = api.ParameterStruct(num_water=256, num_cells=40, bound=10, water_start_bound=6, hashDim=8, border=3)
    good = str(tmp_path / "good.scn")
    api.BacteriaBox(p, 3).populateSimulation().save(good)
    raw = bytearray(open(good, "rb").read())
    api.BacteriaBox.load(good, p)                                  # sanity: the untouched file loads
    hdr = struct.Struct("<8sQ8iQQ4dq")
    fields = list(hdr.unpack_from(raw, 0))

    def variant(name, **kw):
        f = list(fields)
        idx = dict(num_cells=2, num_mnps=3, hashDim=4, maxNeighbors=5, min_depth=6, max_depth=7, num_water=8, n_subtrees=9, n_nodes=10, n_res=11)
        for k, v in kw.items():
            f[idx[k]] = v
        out = bytearray(raw)
        hdr.pack_into(out, 0, *f)
        path = str(tmp_path / (name + ".scn"))
        open(path, "wb").write(out)
        return path

    for name, kw in [("neg_cells", dict(num_cells=-5)), ("huge_hash", dict(hashDim=2000)), ("huge_nodes", dict(n_nodes=1 << 40)),
                     ("bad_depth", dict(min_depth=9)), ("bad_subtrees", dict(n_subtrees=7)), ("more_water", dict(num_water=fields[8] + 1)),
                     ("neg_water", dict(num_water=-1))]:
        with pytest.raises(api.MrisimError, match="malformed scene file"):
            api.BacteriaBox.load(variant(name, **kw), p)
    trunc = str(tmp_path / "trunc.scn")
    open(trunc, "wb").write(raw[:-8])
    with pytest.raises(api.MrisimError, match="file size"):
        api.BacteriaBox.load(trunc, p)
    # a lookup entry naming a cell that does not exist
    off = hdr.size + fields[2] * 24 + fields[3] * 40
    broken = bytearray(raw)
    struct.pack_into("<i", broken, off, fields[2] + 3)
    path = str(tmp_path / "lookup.scn")
    open(path, "wb").write(broken)
    with pytest.raises(api.MrisimError, match="lookup table"):
        api.BacteriaBox.load(path, p)
