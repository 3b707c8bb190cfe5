import importlib.util
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))


def load_package():
    """Imports cartographer-1_b200/ (the directory name is not a Python identifier)."""
    name = "cartographer_1_b200"
    if name in sys.modules:
        return sys.modules[name]
    spec = importlib.util.spec_from_file_location(
        name, os.path.join(ROOT, "cartographer-1_b200", "__init__.py"))
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def csm():
    return load_package()


@pytest.fixture(scope="session")
def gpu_ctx(csm):
    ctx = csm.Context(0)
    yield ctx
    ctx.close()
