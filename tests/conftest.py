import importlib.util
import os
import sys

import pytest

os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")  # as the library does when it is loaded first
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))


def load_package():
    """Imports the product's ctypes binding (directory `dla-future_b200`) as module `dlaf_b200`."""
    if "dlaf_b200" in sys.modules:
        return sys.modules["dlaf_b200"]
    path = os.path.join(ROOT, "dla-future_b200", "__init__.py")
    spec = importlib.util.spec_from_file_location("dlaf_b200", path,
                                                  submodule_search_locations=[os.path.dirname(path)])
    mod = importlib.util.module_from_spec(spec)
    sys.modules["dlaf_b200"] = mod
    spec.loader.exec_module(mod)
    return mod


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def pkg():
    return load_package()


@pytest.fixture(scope="session")
def L(pkg):
    return pkg.lib()
