import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    config.addinivalue_line("markers", "slow: larger CPU cases")


@pytest.fixture(scope="session")
def golden_dir():
    return os.path.join(ROOT, "tests", "golden")


@pytest.fixture(autouse=True)
def _library_options_back_to_default():
    """Tests pin implementations with cg_set_option (per thread); nothing pinned leaks into the next test."""
    yield
    from fv3net_b200 import _lib

    if _lib._LIB is not None:
        for name, value in (("plevel_impl", "auto"), ("plevel_nansafe", "0"), ("mappm_arrays", "auto"), ("mappm_impl", "auto"),
                            ("exposed_area_impl", "auto")):
            _lib.set_option(name, value)
