import importlib.util
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def load_package():
    """The package directory is `mg-gcn_b200` (hyphen): register it as `mg_gcn_b200`."""
    if "mg_gcn_b200" in sys.modules:
        return sys.modules["mg_gcn_b200"]
    spec = importlib.util.spec_from_file_location(
        "mg_gcn_b200", os.path.join(ROOT, "mg-gcn_b200", "__init__.py"),
        submodule_search_locations=[os.path.join(ROOT, "mg-gcn_b200")])
    mod = importlib.util.module_from_spec(spec)
    sys.modules["mg_gcn_b200"] = mod
    spec.loader.exec_module(mod)
    return mod


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def mg():
    mod = load_package()
    if not os.path.exists(mod.LIB_PATH):
        mod.build()
    mod.lib()
    return mod


@pytest.fixture(scope="session")
def orc():
    from oracle import oracle
    oracle.lib()
    return oracle


@pytest.fixture(scope="session")
def cuda():
    import torch
    if not torch.cuda.is_available():
        pytest.fail("this test is marked gpu and needs a CUDA device; there is no CPU fallback")
    torch.cuda.set_device(0)
    return torch
