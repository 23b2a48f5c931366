import importlib.util
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def load_product():
    """import dect-nr-plus-sdr_b200/__init__.py (directory name is not an identifier)"""
    if "dectnrp_b200" in sys.modules:
        return sys.modules["dectnrp_b200"]
    path = os.path.join(ROOT, "dect-nr-plus-sdr_b200", "__init__.py")
    spec = importlib.util.spec_from_file_location("dectnrp_b200", path)
    mod = importlib.util.module_from_spec(spec)
    sys.modules["dectnrp_b200"] = mod
    spec.loader.exec_module(mod)
    return mod


@pytest.fixture(scope="session")
def product():
    mod = load_product()
    if not os.path.exists(mod.LIB_PATH):
        mod.build()
    return mod


# packet geometries used across the tests: (u, b, type, len, tm_mode, mcs, Z)
CONFIGS = {
    "C1_loopback": (1, 1, 1, 1, 0, 2, 2048),
    "C2_b16_siso_mcs7": (1, 16, 1, 1, 0, 7, 6144),
    "C3_b16_txdiv4_mcs7": (1, 16, 1, 2, 5, 7, 6144),
    "small_b2_mcs4": (1, 2, 0, 3, 0, 4, 2048),
    "txdiv2_b4_mcs3": (2, 4, 1, 1, 1, 3, 2048),
}
