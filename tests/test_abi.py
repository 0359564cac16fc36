"""The drop-in boundary: libcoarsen_b200.so loads (no GPU needed) and exports every entry
point include/coarsen_b200.h declares; the ctypes binding lists exactly those.  No compute
call is made here."""
import ctypes
import os
import re

import pytest

from fv3net_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "coarsen_b200.h")


def declared_functions():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)           # drop comments
    return sorted(set(re.findall(r"\b(cg_[a-z0-9_]+)\s*\(", text)))


@pytest.fixture(scope="module")
def handle():
    if not os.path.exists(_lib.LIB_PATH):
        _lib.build()
    return _lib.lib()


def test_header_declares_the_path():
    names = declared_functions()
    for required in ("cg_weighted_block_average", "cg_block_reduce", "cg_block_median", "cg_block_mode",
                     "cg_block_upsample", "cg_pressure_at_interface", "cg_mask_weights", "cg_mappm",
                     "cg_pressure_level_average", "cg_area_above_fine_surface", "cg_last_error"):
        assert required in names


def test_library_exports_every_declared_symbol(handle):
    raw = ctypes.CDLL(_lib.LIB_PATH)
    for name in declared_functions():
        assert hasattr(raw, name), f"{name} declared in the header but not exported"


def test_binding_matches_header(handle):
    assert sorted(_lib.SIGNATURES) == declared_functions()


def test_host_only_calls(handle):
    assert handle.cg_version() >= 100
    assert isinstance(handle.cg_last_error(), bytes)
    n = handle.cg_pressure_level_workspace_bytes(26, 6, 79, 384, 384, 8)
    assert n >= 0


def test_header_cites_the_reference_interfaces():
    text = open(HEADER).read()
    for cite in ("coarsen.py", "regridz.py", "mappm.f90"):
        assert cite in text
