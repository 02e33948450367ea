import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200)")
    config.addinivalue_line("markers", "slow: long-running, opt-in with PAC_SLOW=1")


def pytest_collection_modifyitems(config, items):
    if os.environ.get("PAC_SLOW") == "1":
        return
    skip = pytest.mark.skip(reason="set PAC_SLOW=1 to run")
    for it in items:
        if "slow" in it.keywords:
            it.add_marker(skip)


@pytest.fixture(scope="session")
def kats():
    import numpy as np
    return np.load(os.path.join(GOLDEN, "kats.npz"))


@pytest.fixture(scope="session")
def ctx48():
    from audio_codec_b200 import _lib
    return _lib.default_context(48000)
