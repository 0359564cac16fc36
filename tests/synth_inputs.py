"""Regenerates the reference regression test's synthetic inputs without synth/xarray.

external/synth/synth/core.py:63-67: every variable is
``np.random.seed(0); np.random.uniform(lo, hi, shape).astype(dtype)`` with the
ranges of external/vcm/tests/test_coarsen_restarts.py:24-29 (default ±1000).
Shapes/dtypes come from tests/golden/restart_schemas.json (extracted from the
reference's schema files by tests/golden/make_golden.py).
"""
import json
import os

import numpy as np

RANGES = {"delp": (3, 5), "area": (0.5, 1), "dx": (0.5, 1), "dy": (0.5, 1)}
DEFAULT_RANGE = (-1000, 1000)
_HERE = os.path.dirname(os.path.abspath(__file__))


def load_schemas():
    with open(os.path.join(_HERE, "golden", "restart_schemas.json")) as f:
        return json.load(f)


def generate(category, var, squeeze_time=True):
    """Returns (array, dims)."""
    v = load_schemas()[category]["variables"][var]
    lo, hi = RANGES.get(var, DEFAULT_RANGE)
    np.random.seed(0)
    a = np.random.uniform(low=lo, high=hi, size=v["shape"]).astype(np.dtype(v["dtype"]))
    dims = list(v["dims"])
    if squeeze_time and "Time" in dims:
        ax = dims.index("Time")
        a = np.squeeze(a, axis=ax)
        dims.pop(ax)
    return a, dims


def golden(tag, category, var, squeeze_time=True):
    z = np.load(os.path.join(_HERE, "golden", "coarsen_restarts_reference.npz"))
    a = z[f"{tag}/{category}/{var}"]
    dims = [str(d) for d in z[f"{tag}/{category}/{var}@dims"]]
    if squeeze_time and "Time" in dims:
        ax = dims.index("Time")
        a = np.squeeze(a, axis=ax)
        dims.pop(ax)
    return a, dims


def dataset(category):
    """The synthetic Dataset of one restart category (or "grid") as the reference's test
    builds it (synth.generate: variables with their attrs, coordinates from the schema's
    values), as an fv3net_b200.labeled.Dataset on the host."""
    from fv3net_b200.labeled import DataArray, Dataset

    schema = load_schemas()[category]
    coords = {}
    for k, c in schema["coords"].items():
        v = np.asarray(c["value"])
        coords[k] = DataArray(v, tuple(c["dims"]), None, k, c.get("attrs", {}))
    out = Dataset(coords=coords)
    for var, v in schema["variables"].items():
        a, dims = generate(category, var, squeeze_time=False)
        own = {k: c for k, c in coords.items() if all(d in dims for d in c.dims)}
        out[var] = DataArray(a, tuple(dims), own, var, v.get("attrs", {}))
    return out


def golden_meta(tag, category):
    with open(os.path.join(_HERE, "golden", "coarsen_restarts_meta.json")) as f:
        return json.load(f)[f"{tag}/{category}"]
