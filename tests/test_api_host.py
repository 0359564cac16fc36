"""Host logic of the reference-facing API (no GPU): coordinate arithmetic, metadata rules,
argument checking and error messages - mirrors the non-numerical half of
external/vcm/tests/test_cubedsphere.py and tests/test_regridz.py.  Nothing here computes on
data: the last test pins that asking for it without a CUDA device fails loudly.
"""
from functools import partial

import numpy as np
import pytest

from fv3net_b200 import labeled as lab
from fv3net_b200.labeled import DataArray, Dataset
from fv3net_b200.cubedsphere import coarsen as C
from fv3net_b200.cubedsphere import regridz as R


# ---- test_cubedsphere.py:148-174
@pytest.mark.parametrize(
    ("coarsening_factor", "input_coordinate", "expected_coordinate"),
    [
        (2, [1.0, 2.0, 3.0, 4.0], [1.0, 2.0]),
        (2, [5.0, 6.0, 7.0, 8.0], [3.0, 4.0]),
        (2, [1.0, 2.0, 3.0], [1.0, 2.0]),
        (2, [3.0, 4.0, 5.0], [2.0, 3.0]),
    ],
    ids=["cell centers; first subtile", "cell centers; second subtile",
         "cell interfaces; first subtile", "cell interfaces; second subtile"],
)
def test_coarsen_coords(coarsening_factor, input_coordinate, expected_coordinate):
    da = DataArray(input_coordinate, ["x"], [input_coordinate], name="x")
    result = C.coarsen_coords(coarsening_factor, da, ["x"])
    assert result["x"].dtype == np.float32
    np.testing.assert_array_equal(result["x"].values, np.array(expected_coordinate, np.float32))
    np.testing.assert_array_equal(result["x"].coords["x"].values, result["x"].values)


# ---- test_cubedsphere.py:177-203
@pytest.mark.parametrize("coarsened_object_type", ["DataArray", "Dataset"])
def test_add_coordinates(coarsened_object_type):
    x = np.array([1.0, 2.0]).astype(np.float32)
    y = np.array([3.0, 4.0]).astype(np.float32)
    reference_obj = DataArray([[1, 1], [1, 1]], ["x", "y"], [x, y], name="foo")
    coarsened_obj = DataArray([[1]], ["x", "y"], None, name="foo")
    if coarsened_object_type == "Dataset":
        coarsened_obj = coarsened_obj.to_dataset()
    result = C.add_coordinates(reference_obj, coarsened_obj, 2, ["x", "y"])
    assert type(result) is type(coarsened_obj)
    np.testing.assert_array_equal(result.coords["x"].values, np.array([1.0], np.float32))
    np.testing.assert_array_equal(result.coords["y"].values, np.array([2.0], np.float32))
    assert result.coords["x"].dtype == np.float32
    assert "x" not in coarsened_obj.coords                      # input untouched


@pytest.mark.parametrize("subtile", [0, 1, 2])
def test_coarsen_coords_coord_func_subtiles(subtile):
    """test_cubedsphere.py:437-456: coarse subtile coordinates continue the global index."""
    c = (np.array([1.0, 2.0, 3.0, 4.0]) + 4.0 * subtile).astype(np.float32)
    got = C._reduce_coordinate(c, 2, C.coarsen_coords_coord_func, pad=False)
    np.testing.assert_array_equal(got, (np.array([1.0, 2.0]) + 2.0 * subtile).astype(np.float32))
    assert got.dtype == np.float32
    stag = (np.array([1.0, 2.0, 3.0, 4.0, 5.0]) + 4.0 * subtile).astype(np.float32)
    got = C._reduce_coordinate(stag, 2, C.coarsen_coords_coord_func, pad=True)
    np.testing.assert_array_equal(got, (np.array([1.0, 2.0, 3.0]) + 2.0 * subtile).astype(np.float32))


@pytest.mark.parametrize(
    "coord_func, name, expected",
    [
        ("mean", "x", [0.5, 2.5]),
        (np.mean, "x", [0.5, 2.5]),
        ({"x": np.max, "y": "median"}, "x", [1.0, 3.0]),
        ({"x": np.max, "y": "median"}, "y", [0.5, 2.5]),
        ({"x": np.min}, "y", [0.5, 2.5]),                       # unnamed coordinate -> "mean"
    ],
)
def test_coord_func_forms(coord_func, name, expected):
    """test_cubedsphere.py:300-328: str, callable, dict, dict with a dim missing."""
    fn = C._coord_function_for(coord_func, name)
    got = C._reduce_coordinate(np.arange(4, dtype=np.float64), 2, fn, pad=False)
    np.testing.assert_array_equal(got, expected)
    assert got.dtype == np.float64


def test_coordinate_padding_matches_boundary_pad():
    got = C._reduce_coordinate(np.array([1.0, 2.0, 3.0]), 2, "mean", pad=True)   # nan-skipping
    np.testing.assert_array_equal(got, [1.5, 3.0])
    got = C._reduce_coordinate(np.array([1.0, 2.0, 3.0]), 2, np.max, pad=True)   # cval = nan
    assert got[0] == 2.0 and np.isnan(got[1])
    with pytest.raises(ValueError, match="Could not coarsen a dimension of size 3"):
        C._reduce_coordinate(np.array([1.0, 2.0, 3.0]), 2, "mean", pad=False)


def test_carry_coords_keeps_untouched_and_drops_foreign():
    da = DataArray(np.zeros((4, 4, 2), np.float32), ["x", "y", "z"],
                   {"x": np.arange(1, 5, dtype=np.float32), "z": [10.0, 20.0],
                    "t": ((), np.float64(3.0))}, name="foo")
    coords = C._carry_coords(da, ("x", "y", "z"), {"x": 2, "y": 2}, C.coarsen_coords_coord_func)
    assert set(coords) == {"x", "z", "t"}
    np.testing.assert_array_equal(coords["x"].values, np.array([1, 2], np.float32))
    np.testing.assert_array_equal(coords["z"].values, [10.0, 20.0])


def test_propagate_attrs():
    ref = Dataset({"a": DataArray([1.0], ["x"], attrs={"units": "m"})}, attrs={"test": "a"})
    out = Dataset({"a": DataArray([2.0], ["x"])})
    out = C._propagate_attrs(ref, out)
    assert out.attrs == {"test": "a"} and out["a"].attrs == {"units": "m"}


# ---- argument checking happens before anything touches a device
@pytest.mark.parametrize("fn", [C.edge_weighted_block_average, None])
def test_bad_edge(fn):
    da = DataArray(np.zeros((2, 3)), ["x_dim", "y_dim"])
    with pytest.raises(ValueError, match="'edge' most be either 'x' or 'y'; got z."):
        if fn is None:
            C.block_edge_coarsen(da, 2, x_dim="x_dim", y_dim="y_dim", edge="z")
        else:
            fn(da, da, 2, x_dim="x_dim", y_dim="y_dim", edge="z")


def test_not_divisible_is_refused_like_boundary_exact():
    da = DataArray(np.zeros((3, 4), np.float32), ["x", "y"])
    with pytest.raises(ValueError, match="Could not coarsen a dimension of size 3 with window 2"):
        C.weighted_block_average(da, da, 2, x_dim="x", y_dim="y")


def test_arbitrary_reduction_callable_is_refused():
    with pytest.raises(NotImplementedError, match="no CPU fallback"):
        C._classify_reduction(np.ptp)
    assert C._classify_reduction(np.median)[0] == "median"
    assert C._classify_reduction(np.nansum)[:2] == ("sum", True)
    kernel, _, kwargs = C._classify_reduction(partial(C._mode_reduce, nan_policy="omit"))
    assert kernel == "mode" and kwargs == {"nan_policy": "omit"}


@pytest.mark.parametrize("kernel,dtype,expected", [
    ("sum", np.int64, np.int64), ("sum", np.int32, np.int64), ("min", np.int32, np.int32),
    ("mean", np.int64, np.float64), ("median", np.float32, np.float32), ("sum", np.float32, np.float32),
])
def test_result_dtype_follows_numpy(kernel, dtype, expected):
    assert C._result_dtype(kernel, dtype) == np.dtype(expected)


def test_block_reduce_without_the_dims_returns_input_unchanged():
    """coarsen.py:432-436, test_cubedsphere.py:341-357 (`bar` has no horizontal dims)."""
    bar = DataArray([1, 2, 3], ["t"], name="bar", attrs={"units": "m"})
    assert C._xarray_block_reduce_dataarray(bar, {"x": 2, "y": 2}, np.median) is bar


def _input_dataarray(shape, z_dim_name="z"):
    data = np.arange(np.prod(shape)).reshape(shape).astype(np.float32)
    dims = ["t", "x", "y", z_dim_name]
    return DataArray(data, dims[len(dims) - len(shape):], None, name="foo")


# ---- test_regridz.py:73-98
def test_regrid_vertical_invalid_dimension_names():
    p_in, f_in, p_out = _input_dataarray((4, 6)), _input_dataarray((4, 5)), _input_dataarray((4, 3))
    with pytest.raises(ValueError, match="must not be equal"):
        R.regrid_vertical(p_in, f_in, p_out, z_dim_center="z", z_dim_outer="z")


def test_regrid_vertical_invalid_columns():
    p_in = _input_dataarray((4, 6), "z_outer")
    f_in = _input_dataarray((3, 5), "z_center")
    p_out = _input_dataarray((4, 3), "z_outer")
    with pytest.raises(ValueError, match="must be same size"):
        R.regrid_vertical(p_in, f_in, p_out, z_dim_center="z_center", z_dim_outer="z_outer")


def test_regrid_vertical_invalid_vertical_dimension_size():
    p_in = _input_dataarray((4, 6), "z_outer")
    f_in = _input_dataarray((4, 3), "z_center")
    p_out = _input_dataarray((4, 3), "z_outer")
    with pytest.raises(ValueError, match="one shorter than p_in"):
        R.regrid_vertical(p_in, f_in, p_out, z_dim_center="z_center", z_dim_outer="z_outer")


def test_signatures_match_the_reference():
    """keyword names and defaults of the public functions."""
    import inspect

    def defaults(fn):
        return {k: v.default for k, v in inspect.signature(fn).parameters.items()
                if v.default is not inspect.Parameter.empty}

    assert defaults(C.weighted_block_average) == {
        "x_dim": "xaxis_1", "y_dim": "yaxis_2", "coord_func": C.coarsen_coords_coord_func}
    assert defaults(C.edge_weighted_block_average) == {
        "x_dim": "xaxis_1", "y_dim": "yaxis_1", "edge": "x", "coord_func": C.coarsen_coords_coord_func}
    assert defaults(C.block_coarsen) == {
        "x_dim": "xaxis_1", "y_dim": "yaxis_1", "method": "sum",
        "coord_func": C.coarsen_coords_coord_func, "func_kwargs": None}
    assert defaults(C.block_edge_coarsen)["method"] == "sum"
    assert list(inspect.signature(C.block_upsample).parameters) == ["obj", "upsampling_factor", "dims"]
    assert defaults(R.regrid_vertical) == {
        "iv": 1, "kord": 1, "z_dim_center": "zaxis_1", "z_dim_outer": "zaxis_2"}
    assert defaults(R.regrid_to_area_weighted_pressure) == {
        "x_dim": "xaxis_1", "y_dim": "yaxis_2", "z_dim": "zaxis_1"}
    assert defaults(R.regrid_to_edge_weighted_pressure) == {
        "x_dim": "xaxis_1", "y_dim": "yaxis_1", "z_dim": "zaxis_1", "edge": "x"}
    import fv3net_b200.cubedsphere as pkg
    for name in ["block_coarsen", "block_edge_sum", "block_median", "coarsen_coords",
                 "edge_weighted_block_average", "horizontal_block_reduce", "weighted_block_average",
                 "xarray_block_reduce", "block_upsample", "shift_edge_var_to_center", "regrid_vertical"]:
        assert callable(getattr(pkg, name)), name


def test_labelled_identical_semantics():
    a = DataArray(np.array([1.0, np.nan]), ["x"], {"x": [1, 2]}, "a", {"u": 1})
    assert lab.identical(a, a.copy())
    assert not lab.identical(a, a.rename("b"))
    assert not lab.identical(a, a.astype(np.float32))
    assert lab.identical(DataArray([1, 2], ["x"]), DataArray([1.0, 2.0], ["x"]), check_dtype=False)


def test_compute_without_cuda_fails_loudly():
    """There is no CPU path for data: without a CUDA device the call raises."""
    torch = pytest.importorskip("torch")
    if torch.cuda.is_available():
        pytest.skip("CUDA device present")
    from fv3net_b200._lib import CoarsenLibraryError

    da = DataArray(np.ones((4, 4), np.float32), ["x", "y"], name="foo")
    with pytest.raises((CoarsenLibraryError, RuntimeError, AssertionError)):
        C.weighted_block_average(da, da, 2, x_dim="x", y_dim="y")


def test_staging_host_helpers_without_a_gpu():
    """The host-side pieces of the staging layer: binding is a no-op (None, affinity untouched) where NVML or a
    CUDA device is missing, and the input tree flattens / unflattens losslessly."""
    import os

    from fv3net_b200 import staging

    before = os.sched_getaffinity(0) if hasattr(os, "sched_getaffinity") else None
    import torch
    if not torch.cuda.is_available():
        assert staging.bind_host_thread(0) is None
        if before is not None:
            assert os.sched_getaffinity(0) == before
    tree = {"middle3d": {"T": np.zeros((1, 2, 8, 8), "f4")}, "middle2d": {"LHTFLsfc": np.ones((1, 8, 8), "f4")},
            "begin": {"T": np.zeros((1, 2, 8, 8), "f4")}, "end": {"T": np.ones((1, 2, 8, 8), "f4")},
            "delp": np.full((1, 2, 8, 8), 5.0, "f4"), "area": np.ones((1, 8, 8), "f4"),
            "exposed_area_fine": np.ones((1, 8, 8), "f4")}
    flat = staging._flatten(tree)
    assert list(flat)[-3:] == [("delp", None), ("area", None), ("exposed_area_fine", None)]
    back = staging._unflatten(flat)
    assert back["delp"] is tree["delp"] and back["end"]["T"] is tree["end"]["T"]
    with pytest.raises(Exception):
        staging.BudgetPipeline().run(tree)              # no CUDA device, no CPU fallback
