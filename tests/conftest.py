import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def lib():
    """The in-tree C-ABI library; built on demand (nvcc cross-compiles without a GPU)."""
    import __graft_entry__ as ge
    ge.build()
    import kubeflow_controller_b200 as pkg
    return pkg.load_library()


@pytest.fixture(scope="session")
def pkg(lib):
    import kubeflow_controller_b200 as pkg
    return pkg


@pytest.fixture(scope="session")
def kat():
    import json
    with open(os.path.join(ROOT, "tests", "golden", "kat_survey.json")) as f:
        return json.load(f)
