"""bench.py pieces that can be checked without a GPU: the reference arm's Python mirror of the shipped
parameters equals mrisim_params_default byte for byte; the reference arm prints one JSON line with the
contract's keys, takes its scene from the reference's own host code when oracle/_ref is there, and never
loads the product library."""
import ctypes as C
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def test_shipped_params_mirror_equals_the_library_defaults():
    import bench
    from mri_nanoparticle_simulation_b200 import api, _cabi as cabi
    for over in (dict(), dict(num_cells=2000, hashDim=40, maxNeighbors=16), dict(num_water=8000, avoid_intracellular_throw=0)):
        a, b = bench.shipped_params(cabi, **over), api.ParameterStruct(**over)
        assert bytes(a) == bytes(b), over
    assert set(bench.STRONG) <= set(bench.WORKLOADS) and set(bench.SMALL_CONFIGS) <= set(bench.WORKLOADS)


def test_kernel_source_digest_guards_the_instruction_mix(tmp_path, monkeypatch):
    import bench
    d = bench.kernel_source_digest()
    assert len(d) == 64
    mix = bench.load_instr_mix()
    assert mix is None or mix.get("stale") or mix["source_digest"] == d


@pytest.mark.slow
def test_reference_arm_line(tmp_path):
    env = dict(os.environ, MRISIM_REFARM_BUDGET_S="1.0")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0"],
                         capture_output=True, text=True, timeout=900, env=env)
    assert out.returncode == 0, out.stderr[-2000:]
    line = json.loads(out.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["metric"] == "walker-steps/sec" and line["higher_is_better"] is True
    assert line["cpu_baseline"]["kind"] == "port" and line["cpu_baseline"]["cores"] == line["config"]["host_cores"] == (os.cpu_count() or 1)
    assert line["e2e"]["value"] == line["value"] and line["e2e"]["h2d_bytes_per_step"] == 0
    if os.path.exists(os.path.join(ROOT, "oracle", "_ref", "ref_scene_dense2000_32k")):
        assert line["config"]["scene_by"].startswith("oracle/_ref")
        assert "libmrisim" not in out.stderr
