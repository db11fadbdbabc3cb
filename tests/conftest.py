import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    config.addinivalue_line("markers", "slow: minutes of GPU time (deselect with -m 'gpu and not slow' while iterating)")


@pytest.fixture(scope="session", autouse=True)
def _built():
    """libmrisim.so and the oracle port must exist; build them if the tree is fresh."""
    from mri_nanoparticle_simulation_b200 import build as b
    b.build()
    import oracle_lib
    oracle_lib.build()
