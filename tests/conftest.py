import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN


@pytest.fixture(scope="session")
def sp_weights():
    from glue_factory_b200 import synth

    return synth.superpoint_open_state_dict()


@pytest.fixture(scope="session")
def lg_weights():
    from glue_factory_b200 import synth

    return synth.lightglue_state_dict()
