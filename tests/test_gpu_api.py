"""The reference's own unit tests for the coarse-graining API, re-stated against the drop-in
(`fv3net_b200.cubedsphere`): same inputs, same expected values, same identical-including-
dtype bar (external/vcm/tests/test_cubedsphere.py, test_regridz.py, test_mappm.py).  Where
the reference compares with xarray (absent here) the expected object is built from the CPU
oracle instead.  Every data value below is computed by a CUDA kernel through the C ABI.
"""
import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

import oracle  # noqa: E402
from oracle import recipes  # noqa: E402
from fv3net_b200.labeled import DataArray, Dataset, assert_identical, merge  # noqa: E402


@pytest.fixture(scope="module", autouse=True)
def _needs_cuda():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")


from fv3net_b200.cubedsphere import (  # noqa: E402
    block_coarsen, block_edge_sum, block_median, block_upsample, edge_weighted_block_average,
    horizontal_block_reduce, pressure_level_average, regrid_to_area_weighted_pressure,
    regrid_to_edge_weighted_pressure, regrid_vertical, shift_edge_var_to_center,
    weighted_block_average,
)
from fv3net_b200.cubedsphere.coarsen import (  # noqa: E402
    _block_mode, _xarray_block_reduce_dataarray, block_edge_coarsen, block_upsample_like,
)
from fv3net_b200.cubedsphere.regridz import _mask_weights  # noqa: E402
from fv3net_b200 import thermo  # noqa: E402


# ---- test_cubedsphere.py:206-236
@pytest.mark.parametrize("object_type", ["DataArray", "Dataset"])
def test_weighted_block_average(object_type):
    dims = ["x", "y"]
    attrs = {"units": "m"}
    ds_attrs = {"test": "a"}
    data = DataArray(np.array([[2.0, 6.0], [6.0, 2.0]]), dims, None, "foo", attrs)
    if object_type == "Dataset":
        data = data.to_dataset()
        data.attrs = ds_attrs
    weights = DataArray(np.array([[6.0, 2.0], [2.0, 6.0]]), dims)
    expected = DataArray(np.array([[3.0]]), dims, None, "foo", attrs)
    if object_type == "Dataset":
        expected = expected.to_dataset()
        expected.attrs = ds_attrs
    result = weighted_block_average(data, weights, 2, x_dim="x", y_dim="y")
    assert_identical(result, expected)


# ---- test_cubedsphere.py:239-258
@pytest.mark.parametrize(
    ("data", "spacing", "factor", "edge", "expected_data"),
    [
        ([[2, 6, 2], [6, 2, 6]], [[6, 2, 6], [2, 6, 2]], 2, "x", [[3.0, 3.0]]),
        ([[2, 6], [6, 2], [2, 6]], [[6, 2], [2, 6], [6, 2]], 2, "y", [[3.0], [3.0]]),
    ],
    ids=["edge='x'", "edge='y'"],
)
def test_edge_weighted_block_average(data, spacing, factor, edge, expected_data):
    dims = ["x_dim", "y_dim"]
    attrs = {"units": "m"}
    da = DataArray(data, dims, None, None, attrs)
    weights = DataArray(spacing, dims)
    expected = DataArray(expected_data, dims, None, None, attrs)
    result = edge_weighted_block_average(da, weights, factor, x_dim="x_dim", y_dim="y_dim", edge=edge)
    assert_identical(result, expected)


@pytest.fixture()
def input_dataarray():
    shape = (4, 4, 2)
    data = np.arange(np.prod(shape)).reshape(shape).astype(np.float32)
    return DataArray(data, ["x", "y", "z"], None, "foo", {"units": "m"})


@pytest.fixture()
def input_dataset(input_dataarray):
    bar = DataArray([1, 2, 3], ["t"], name="bar", attrs={"units": "m"})
    ds = merge([input_dataarray, bar])
    ds.attrs = {"test": "a"}
    return ds


def _blocked(a, f=2):
    """(x, y, z) -> windows over x and y, numpy."""
    nx, ny, nz = a.shape
    return a.reshape(nx // f, f, ny // f, f, nz)


# ---- test_cubedsphere.py:275-290 (regtest checksum there; values here)
@pytest.mark.parametrize("reduction_function", [np.mean, np.median])
def test_xarray_block_reduce_dataarray(reduction_function, input_dataarray):
    result = _xarray_block_reduce_dataarray(input_dataarray, {"x": 2, "y": 2}, reduction_function)
    expected = reduction_function(_blocked(input_dataarray.values), axis=(1, 3))
    assert_identical(result, DataArray(expected, ["x", "y", "z"], None, "foo", {"units": "m"}))


# ---- test_cubedsphere.py:300-328
@pytest.mark.parametrize(
    "coord_func, expected_x, expected_y",
    [
        ("mean", [0.5, 2.5], [0.5, 2.5]),
        (np.mean, [0.5, 2.5], [0.5, 2.5]),
        ({"x": np.max, "y": "median"}, [1.0, 3.0], [0.5, 2.5]),
        ({"x": np.min}, [0.0, 2.0], [0.5, 2.5]),
    ],
    ids=["single str", "single function", "dict mapping coord name to str or function",
         "dict with a dimension missing"],
)
def test_block_reduce_dataarray_coordinates(input_dataarray, coord_func, expected_x, expected_y):
    for dim, size in input_dataarray.sizes.items():
        input_dataarray[dim] = np.arange(size, dtype=np.float64)
    result = _xarray_block_reduce_dataarray(input_dataarray, {"x": 2, "y": 2}, np.median,
                                            coord_func=coord_func)
    expected = DataArray(np.median(_blocked(input_dataarray.values), axis=(1, 3)), ["x", "y", "z"],
                         {"x": np.array(expected_x), "y": np.array(expected_y),
                          "z": np.arange(2, dtype=np.float64)}, "foo", {"units": "m"})
    assert_identical(result, expected)


# ---- test_cubedsphere.py:331-383
def test_horizontal_block_reduce_dataarray(input_dataarray):
    expected = _xarray_block_reduce_dataarray(input_dataarray, {"x": 2, "y": 2, "z": 1}, np.median)
    result = horizontal_block_reduce(input_dataarray, 2, np.median, "x", "y")
    assert_identical(result, expected)


def test_horizontal_block_reduce_dataset(input_dataset):
    expected_foo = _xarray_block_reduce_dataarray(input_dataset.foo, {"x": 2, "y": 2, "z": 1}, np.median)
    expected = merge([expected_foo, input_dataset.bar])      # bar: no horizontal dims, unchanged
    expected.attrs = {"test": "a"}
    result = horizontal_block_reduce(input_dataset, 2, np.median, "x", "y")
    assert_identical(result, expected)


def test_block_median(input_dataarray):
    expected = DataArray(np.median(_blocked(input_dataarray.values), axis=(1, 3)), ["x", "y", "z"],
                         None, "foo", {"units": "m"})
    assert_identical(block_median(input_dataarray, 2, "x", "y"), expected)
    assert_identical(block_coarsen(input_dataarray, 2, "x", "y", method="median"), expected)


def test_block_coarsen(input_dataarray):
    expected = DataArray(np.min(_blocked(input_dataarray.values), axis=(1, 3)), ["x", "y", "z"],
                         None, "foo", {"units": "m"})
    assert_identical(block_coarsen(input_dataarray, 2, "x", "y", "min"), expected)


# ---- test_cubedsphere.py:386-419
@pytest.mark.parametrize(
    ("data", "factor", "edge", "expected_data"),
    [([[2, 6, 2], [6, 2, 6]], 2, "x", [[8, 8]]), ([[2, 6], [6, 2], [2, 6]], 2, "y", [[8], [8]])],
    ids=["edge='x'", "edge='y'"],
)
def test_block_edge_sum(data, factor, edge, expected_data):
    dims = ["x_dim", "y_dim"]
    da = DataArray(data, dims, None, None, {"units": "m"})
    expected = DataArray(expected_data, dims, None, None, {"units": "m"})
    assert_identical(block_edge_sum(da, factor, x_dim="x_dim", y_dim="y_dim", edge=edge), expected)


@pytest.mark.parametrize(
    ("data", "factor", "edge", "expected_data"),
    [([[2, 6, 2], [6, 2, 6]], 2, "x", [[2, 2]]), ([[2, 6], [6, 2], [2, 6]], 2, "y", [[2], [2]])],
    ids=["edge='x'", "edge='y'"],
)
def test_block_edge_coarsen(data, factor, edge, expected_data):
    dims = ["x_dim", "y_dim"]
    da = DataArray(data, dims, None, None, {"units": "m"})
    expected = DataArray(expected_data, dims, None, None, {"units": "m"})
    result = block_edge_coarsen(da, factor, x_dim="x_dim", y_dim="y_dim", edge=edge, method="min")
    assert_identical(result, expected)


# ---- test_cubedsphere.py:422-614: subtile coordinates
@pytest.fixture(params=[np.float32, np.float64])
def dtype(request):
    return request.param


@pytest.fixture(params=[0, 1, 2])
def subtile_x(request):
    return request.param


@pytest.fixture(params=[0, 2])
def subtile_y(request):
    return request.param


def _subtile(n, subtile):
    return (np.arange(1.0, n + 1) + 4.0 * subtile).astype(np.float32)


def _coarse_subtile(n, subtile):
    return (np.arange(1.0, n + 1) + 2.0 * subtile).astype(np.float32)


@pytest.fixture()
def with_subtile_coordinates(subtile_x, subtile_y, dtype):
    data = np.arange(32).reshape(4, 4, 2).astype(dtype)
    return DataArray(data, ["x", "y", "z"], {"x": _subtile(4, subtile_x), "y": _subtile(4, subtile_y)}, "foo")


def _check_coords(result, nx, ny, subtile_x, subtile_y):
    for dim, n, s in (("x", nx, subtile_x), ("y", ny, subtile_y)):
        assert result[dim].dtype == np.float32
        np.testing.assert_array_equal(result[dim].values, _coarse_subtile(n, s))
    assert "z" not in result.coords


def test_coarsening_functions_with_coordinates(with_subtile_coordinates, subtile_x, subtile_y):
    da = with_subtile_coordinates
    for result in (
        weighted_block_average(da, da, 2, "x", "y"),
        horizontal_block_reduce(da, 2, np.mean, "x", "y"),
        block_median(da, 2, "x", "y"),
        block_coarsen(da, 2, "x", "y", "sum"),
    ):
        _check_coords(result, 2, 2, subtile_x, subtile_y)
        assert result.dtype == da.dtype


@pytest.fixture()
def with_staggered_subtile_coordinates(subtile_x, subtile_y, dtype):
    data = np.arange(40).reshape(5, 4, 2).astype(dtype)
    return DataArray(data, ["x", "y", "z"], {"x": _subtile(5, subtile_x), "y": _subtile(4, subtile_y)}, "foo")


def test_edge_functions_with_coordinates(with_staggered_subtile_coordinates, subtile_x, subtile_y):
    da = with_staggered_subtile_coordinates
    avg = edge_weighted_block_average(da, da, 2, "x", "y", edge="y")
    tot = block_edge_sum(da, 2, "x", "y", edge="y")
    for result in (avg, tot):
        _check_coords(result, 3, 2, subtile_x, subtile_y)
        assert result.shape == (3, 2, 2)
    a = da.values
    np.testing.assert_array_equal(tot.values, (a[::2, 0::2] + a[::2, 1::2]))
    w = a[::2]
    with np.errstate(invalid="ignore"):
        exp = (w[:, 0::2] ** 2 + w[:, 1::2] ** 2) / (w[:, 0::2] + w[:, 1::2])
    np.testing.assert_allclose(avg.values, exp, rtol=1e-6)


# ---- test_cubedsphere.py:731-768
@pytest.mark.parametrize("via", ["_block_mode", "block_coarsen"])
def test_block_mode(via):
    data = np.array([[0.0, 0.0, 1.0, 1.0], [0.0, 0.0, 1.0, 1.0], [1.0, 1.0, 0.0, 0.0],
                     [1.0, 1.0, 0.0, np.nan]])
    da = DataArray(data, ["x", "y"], attrs={"units": "m"})
    expected = DataArray(np.array([[0.0, 1.0], [1.0, 0.0]]), ["x", "y"], attrs={"units": "m"})
    if via == "_block_mode":
        result = _block_mode(da, 2, x_dim="x", y_dim="y", nan_policy="omit")
    else:
        result = block_coarsen(da, 2, x_dim="x", y_dim="y", method="mode",
                               func_kwargs={"nan_policy": "omit"})
    assert_identical(result, expected)


# ---- test_cubedsphere.py:771-816
def test_block_upsample_dataset():
    foo = DataArray([[1, 2], [3, 4]], ["xt", "yt"], name="foo")
    u = DataArray([[1, 2, 3], [4, 5, 6]], ["xt", "y"], name="u")
    expected = merge([
        DataArray([[1, 1, 2, 2], [1, 1, 2, 2], [3, 3, 4, 4], [3, 3, 4, 4]], ["xt", "yt"], name="foo"),
        DataArray([[1, 1, 2, 2, 3], [1, 1, 2, 2, 3], [4, 4, 5, 5, 6], [4, 4, 5, 5, 6]], ["xt", "y"], name="u"),
    ])
    assert_identical(block_upsample(merge([foo, u]), 2, dims=["xt", "y", "yt"]), expected)


@pytest.mark.parametrize(
    ("data", "expected_data"),
    [
        ([[1, 2], [3, 4]], [[1, 1, 2, 2], [1, 1, 2, 2], [3, 3, 4, 4], [3, 3, 4, 4]]),
        ([[1, 2, 3], [4, 5, 6]], [[1, 1, 2, 2, 3], [1, 1, 2, 2, 3], [4, 4, 5, 5, 6], [4, 4, 5, 5, 6]]),
    ],
)
def test_block_upsample_dataarray(data, expected_data):
    foo = DataArray(data, ["x", "y"], name="foo")
    expected = DataArray(expected_data, ["x", "y"], name="foo")
    assert_identical(block_upsample(foo, 2, dims=["x", "y"]), expected)


def test_block_upsample_like_takes_reference_coordinates():
    ref = DataArray(np.zeros((4, 4), np.float32), ["x", "y"], {"x": [1.0, 2, 3, 4], "y": [5.0, 6, 7, 8]})
    coarse = DataArray(np.array([[1.0, 2.0], [3.0, 4.0]], np.float32), ["x", "y"], {"x": [1.0, 2.0]}, "c")
    up = block_upsample_like(coarse, ref, x_dim="x", y_dim="y")
    np.testing.assert_array_equal(up.values, np.kron(coarse.values, np.ones((2, 2), np.float32)))
    np.testing.assert_array_equal(up["x"].values, [1.0, 2, 3, 4])
    np.testing.assert_array_equal(up["y"].values, [5.0, 6, 7, 8])


# ---- test_cubedsphere.py:57-106
def test_shift_edge_var_to_center():
    y_edge = DataArray(np.array([[30.0, 40, 50]]).T.reshape(1, 3, 1) * np.ones((1, 3, 2)),
                       ["tile", "grid_y", "grid_xt"])
    got = shift_edge_var_to_center(y_edge)
    assert got.dims == ("tile", "grid_yt", "grid_xt")
    np.testing.assert_array_equal(got.values, np.array([[[35.0, 35.0], [45.0, 45.0]]]))
    with pytest.raises(ValueError, match="centered on one horizontal axis"):
        shift_edge_var_to_center(DataArray(np.zeros((2, 2)), ["a", "b"]))


# ================================================================= regridz (test_regridz.py)
def _input_dataarray(shape, z_dim_name="z"):
    data = np.arange(np.prod(shape)).reshape(shape).astype(np.float32)
    dims = ["t", "x", "y", z_dim_name]
    return DataArray(data, dims[len(dims) - len(shape):], None, "foo")


# ---- test_regridz.py:22-71 (shapes of the dask cases, eager here), :101-110
@pytest.mark.parametrize(
    "p_in_shape, f_in_shape, p_out_shape, expected_shape",
    [((4, 6), (4, 5), (4, 6), (4, 5)), ((4, 4, 6), (4, 4, 5), (4, 4, 6), (4, 4, 5)),
     ((4, 4, 6), (4, 4, 5), (4, 4, 3), (4, 4, 2))],
    ids=["2d", "3d", "3d-new-nlevels"],
)
def test_regrid_vertical_shapes_and_values(p_in_shape, f_in_shape, p_out_shape, expected_shape):
    p_in = _input_dataarray(p_in_shape, "z_outer")
    f_in = _input_dataarray(f_in_shape, "z_center").assign_attrs({"units": "m"})
    p_out = _input_dataarray(p_out_shape, "z_outer")
    f_out = regrid_vertical(p_in, f_in, p_out, z_dim_center="z_center", z_dim_outer="z_outer")
    assert f_out.shape == expected_shape and f_out.dims == f_in.dims
    assert f_out.attrs == {"units": "m"} and f_out.dtype == np.float32
    ncol = int(np.prod(p_in_shape[:-1]))
    exp = oracle.mappm(p_in.values.reshape(ncol, -1), f_in.values.reshape(ncol, -1),
                       p_out.values.reshape(ncol, -1))
    np.testing.assert_array_equal(f_out.values.reshape(ncol, -1), exp)


def test_regrid_vertical_z_anywhere_matches_z_last():
    """The kernel strides through columns: (z, y, x) and (y, x, z) layouts agree bit for bit."""
    rng = np.random.RandomState(3)
    delp = rng.uniform(3, 5, (6, 4, 5)).astype(np.float32)
    p_in = np.concatenate([np.full((1, 4, 5), 300.0, np.float32), 300 + np.cumsum(delp, 0)]).astype(np.float32)
    p_out = (p_in * rng.uniform(0.9, 1.1, (1, 4, 5))).astype(np.float32)
    f = rng.uniform(-1000, 1000, (6, 4, 5)).astype(np.float32)
    first = regrid_vertical(DataArray(p_in, ["zo", "y", "x"]), DataArray(f, ["zc", "y", "x"]),
                            DataArray(p_out, ["zo", "y", "x"]), z_dim_center="zc", z_dim_outer="zo")
    last = regrid_vertical(DataArray(p_in.transpose(1, 2, 0), ["y", "x", "zo"]),
                           DataArray(f.transpose(1, 2, 0), ["y", "x", "zc"]),
                           DataArray(p_out.transpose(1, 2, 0), ["y", "x", "zo"]),
                           z_dim_center="zc", z_dim_outer="zo")
    assert first.dims == ("zc", "y", "x") and last.dims == ("y", "x", "zc")
    np.testing.assert_array_equal(first.values, last.values.transpose(2, 0, 1))


# ---- test_mappm.py:5-44 through the public entry point
@pytest.mark.parametrize(
    "p_in, f_in, p_out, expected, decimal",
    [
        ([0.0, 1.0, 2.0, 3.0, 4.0, 5.0], [0.0, 1.0, 2.0, 3.0, 4.0], [0.5, 1.2, 2.4, 2.8, 3.2, 4.5],
         [0.35, 1.3, 2.1, 2.5, 3.35], 5),
        ([1.0, 2.0, 3.0, 4.0, 5.0], [1.5, 2.5, 3.5, 4.5], [0.0, 2.5, 3.5, 4.5, 50.0],
         [1.5, 3.0, 4.0, 4.502747], 5),
        ([1.0, 2.0, 3.0, 2.0, 5.0], [np.nan] * 4, [0.0, 2.5, 3.5, 4.5, 50.0], [np.nan] * 4, 7),
    ],
    ids=["identity", "out_of_bounds", "nans"],
)
def test_mappm_known_answers(p_in, f_in, p_out, expected, decimal):
    f_out = regrid_vertical(DataArray(np.asarray(p_in)[None, :], ["c", "zo"]),
                            DataArray(np.asarray(f_in)[None, :], ["c", "zc"]),
                            DataArray(np.asarray(p_out)[None, :], ["c", "zo"]),
                            z_dim_center="zc", z_dim_outer="zo")
    assert f_out.dtype == np.float32
    np.testing.assert_almost_equal(f_out.values, np.asarray([expected], np.float32), decimal=decimal)


# ---- test_regridz.py:113-128
def test__mask_weights():
    weights = _input_dataarray((2, 2, 1)).isel(z=0)
    phalf_coarse_on_fine = _input_dataarray((2, 2, 3))
    phalf_fine = _input_dataarray((2, 2, 3))
    pc, pf = phalf_coarse_on_fine.values.copy(), phalf_fine.values.copy()
    pc[:, :, 0], pc[:, :, 1], pc[:, :, 2] = 1.0, 2.0, 3.0
    pf[:, :, 0], pf[:, :, 1] = 1.0, 2.0
    pf[0, :, 2], pf[1, :, 2] = 2.5, 3.5
    expected = np.broadcast_to(weights.values[:, :, None], (2, 2, 2)).copy()
    expected[0, :, 1] = 0.0
    masked = _mask_weights(weights, phalf_coarse_on_fine.copy(pc), phalf_fine.copy(pf),
                           dim_center="z", dim_outer="z")
    assert masked.dims == ("x", "y", "z")
    np.testing.assert_array_equal(masked.values, expected)


def test_pressure_at_interface_and_friends():
    rng = np.random.RandomState(0)
    delp = rng.uniform(300, 500, (2, 7, 4, 4)).astype(np.float32)
    da = DataArray(delp, ["tile", "pfull", "y", "x"], {"pfull": np.arange(7.0), "y": np.arange(4.0)})
    pi = thermo.pressure_at_interface(da)
    assert pi.dims == ("tile", "phalf", "y", "x") and "pfull" not in pi.coords and "y" in pi.coords
    np.testing.assert_array_equal(pi.values, oracle.pressure_at_interface(delp, axis=1))
    zlast = thermo.pressure_at_interface(DataArray(delp.transpose(0, 2, 3, 1), ["tile", "y", "x", "z"]),
                                         dim_center="z", dim_outer="zi")
    np.testing.assert_array_equal(zlast.values, pi.values.transpose(0, 2, 3, 1))
    ps = thermo.surface_pressure_from_delp(da, vertical_dim="pfull")
    assert ps.attrs == {"long_name": "surface pressure", "units": "Pa"}
    np.testing.assert_allclose(ps.values, delp.sum(1) + 300.0, rtol=1e-6)
    pm = thermo.pressure_at_midpoint_log(da, dim="pfull")
    np.testing.assert_allclose(pm.values, oracle.pressure_at_midpoint_log(delp, axis=1), rtol=2e-5)


def test_thermo_helpers_run_on_the_library_bit_level():
    """surface_pressure_from_delp, pressure_at_midpoint_log and shift_edge_var_to_center call CUDA kernels (not
    torch expressions): launch counter moves, results equal the numpy restatements of the reference
    (vertically_dependent.py:153-208, coarsen.py:54-80) bit for bit, except the logarithm, whose last bit is the
    math library's: |d log p| <= 2 ulp of log(1e5) on each side of the difference."""
    from fv3net_b200 import _lib, synthetic as syn
    rng = np.random.RandomState(5)
    delp = syn.delp_from_ps(syn.surface_pressure((2, 8, 8), seed=3))          # (2, 79, 8, 8) float32
    delp[0, 5, 2, 3] = np.nan                                                  # xarray's sum skips it
    da = DataArray(delp, ["tile", "pfull", "y", "x"])
    n0 = _lib.lib().cg_launch_count()
    ps = thermo.surface_pressure_from_delp(da, vertical_dim="pfull")
    np.testing.assert_array_equal(ps.values, oracle.surface_pressure_from_delp(delp, axis=1))
    zlast = thermo.surface_pressure_from_delp(DataArray(delp.transpose(0, 2, 3, 1), ["tile", "y", "x", "z"]))
    np.testing.assert_array_equal(zlast.values, ps.values)                     # layout does not change the order of additions
    delp = np.nan_to_num(delp, nan=400.0).astype(np.float32)
    pm = thermo.pressure_at_midpoint_log(DataArray(delp, ["tile", "pfull", "y", "x"]), dim="pfull")
    exp = oracle.pressure_at_midpoint_log(delp, axis=1)
    pi = oracle.pressure_at_interface(delp, axis=1)
    dlogp = np.diff(np.log(pi.astype(np.float64)), axis=1)
    bound = 4 * np.spacing(np.float32(np.log(1.1e5))) / dlogp                  # relative: 2 ulp per logarithm
    assert (np.abs(pm.values - exp) <= bound * np.abs(exp) + 2 * np.spacing(np.abs(exp))).all()
    pm64 = thermo.pressure_at_midpoint_log(DataArray(delp.astype(np.float64), ["tile", "pfull", "y", "x"]), dim="pfull")
    np.testing.assert_allclose(pm64.values, oracle.pressure_at_midpoint_log(delp.astype(np.float64), axis=1), rtol=1e-12)
    u = rng.uniform(-30, 30, (2, 5, 9, 8)).astype(np.float32)
    got = shift_edge_var_to_center(DataArray(u, ["tile", "z", "grid_y", "grid_xt"]))
    np.testing.assert_array_equal(got.values, (np.float32(0.5) * (u[:, :, 1:] + u[:, :, :-1])))
    v = rng.uniform(-30, 30, (2, 5, 8, 9))
    got = shift_edge_var_to_center(DataArray(v, ["tile", "z", "grid_yt", "grid_x"]))
    assert got.dims == ("tile", "z", "grid_yt", "grid_xt")
    np.testing.assert_array_equal(got.values, 0.5 * (v[..., 1:] + v[..., :-1]))
    assert _lib.lib().cg_launch_count() - n0 >= 6


# ================================================================= compositions vs the oracle
def _column_inputs(nt, n, seed, nvar=2):
    from fv3net_b200 import synthetic as syn

    delp = syn.delp_from_ps(syn.surface_pressure((nt, n, n), seed=seed))
    area = syn.area((nt, n, n), seed=seed + 1)
    fields = [syn.field(delp.shape, seed + 10 + i, scale=10.0 ** i) for i in range(nvar)]
    return delp, area, fields


def test_regrid_to_area_weighted_pressure_then_average_equals_fused():
    """The reference's two-step recipe (fine-grid remap, then masked-area block average,
    coarsen_restarts.py:152-225) against the oracle, and the fused single-kernel form against
    both, masked and unmasked."""
    from fv3net_b200 import synthetic as syn

    nt, n, f = 2, 16, 8
    delp, area, fields = _column_inputs(nt, n, seed=5)
    dims = ["tile", "zaxis_1", "yaxis_2", "xaxis_1"]
    ds = Dataset({f"v{i}": DataArray(a, dims, attrs={"i": i}) for i, a in enumerate(fields)}, attrs={"k": 1})
    d_delp, d_area = DataArray(delp, dims, name="delp"), DataArray(area, dims[:1] + dims[2:], name="area")

    # The remap is ill-conditioned in the coarse pressures (one fp32 ulp of p ~ 1e5 Pa against
    # ~50 Pa layers near the top moves a remapped value by ~1e-4 of the layer-to-layer jump),
    # so the oracle is given OUR coarse delp, itself checked against the oracle's first.
    dc = weighted_block_average(d_delp, d_area, f, x_dim="xaxis_1", y_dim="yaxis_2")
    np.testing.assert_allclose(dc.values, oracle.weighted_block_average(delp, area[:, None], f), rtol=2e-6)
    regridded, masked = regrid_to_area_weighted_pressure(ds, d_delp, d_area, f)
    exp_fields, exp_masked = recipes.regrid_given_delp(fields, delp, dc.values, area)
    np.testing.assert_array_equal(masked.values, exp_masked)
    for i, e in enumerate(exp_fields):
        tol = 3e-5 * np.abs(fields[i]).max()          # north_star PPM tolerance
        assert np.abs(regridded[f"v{i}"].values - e).max() <= tol
        assert regridded[f"v{i}"].attrs == {"i": i}
    assert regridded.attrs == {"k": 1}

    two_step = weighted_block_average(regridded, masked, f, x_dim="xaxis_1", y_dim="yaxis_2")
    for mask in (True, False):
        fused, dcoarse = pressure_level_average(ds, d_delp, d_area, f, mask_below_surface=mask)
        np.testing.assert_allclose(dcoarse.values, dc.values, rtol=2e-6)
        exp_fields, exp_masked = recipes.regrid_given_delp(fields, delp, dcoarse.values, area)
        for i, e in enumerate(exp_fields):
            w = exp_masked if mask else np.broadcast_to(area[:, None], e.shape)
            with np.errstate(invalid="ignore"):
                exp = oracle.weighted_block_average(e, w, f)
            tol = 3e-5 * np.abs(fields[i]).max()
            got = fused[f"v{i}"].values
            np.testing.assert_array_equal(np.isnan(got), np.isnan(exp))
            assert np.nanmax(np.abs(got - exp)) <= tol
            if mask:
                assert np.nanmax(np.abs(two_step[f"v{i}"].values - exp)) <= 2 * tol
            assert fused[f"v{i}"].attrs == {"i": i} and fused[f"v{i}"].dims == tuple(dims)


@pytest.mark.parametrize("edge", ["x", "y"])
def test_regrid_to_edge_weighted_pressure(edge):
    """u with dx (edge='x', staggered in y) and v with dy (edge='y') across the six faces
    (regridz.py:58-112); the face-halo rule is the one pinned by the golden restarts."""
    rng = np.random.RandomState(11)
    n, nz, f = 8, 6, 2
    delp = rng.uniform(3, 5, (6, nz, n, n)).astype(np.float32)
    shape = (6, nz, n + 1, n) if edge == "x" else (6, nz, n, n + 1)
    wind = rng.uniform(-1000, 1000, shape).astype(np.float32)
    length = rng.uniform(0.5, 1, (6,) + shape[2:]).astype(np.float32)
    x_dim, y_dim = ("xaxis_1", "yaxis_1") if edge == "x" else ("xaxis_2", "yaxis_2")
    d_delp = DataArray(delp, ["tile", "zaxis_1", "yaxis_2", "xaxis_1"], {"tile": np.arange(6)})
    ds = Dataset({"w": DataArray(wind, ["tile", "zaxis_1", y_dim, x_dim])})
    d_len = DataArray(length, ["tile", y_dim, x_dim])
    regridded, masked = regrid_to_edge_weighted_pressure(ds, d_delp, d_len, f, x_dim=x_dim, y_dim=y_dim,
                                                        edge=edge)
    exp_fields, exp_masked = recipes.regrid_to_edge_weighted_pressure([wind], delp, length, f, edge=edge)
    np.testing.assert_array_equal(masked.values, exp_masked)
    assert np.abs(regridded["w"].values - exp_fields[0]).max() <= 3e-5 * 1000


# ================================================================= the f2py ABI (test_mappm.py)
def test_mappm_module_is_a_drop_in_for_the_f2py_one():
    """external/vcm/tests/test_mappm.py:5-44 verbatim, against `fv3net_b200.mappm`."""
    from fv3net_b200 import mappm

    p_in = np.asarray([0.0, 1.0, 2.0, 3.0, 4.0, 5.0])[None, :]
    f_in = np.asarray([0.0, 1.0, 2.0, 3.0, 4.0])[None, :]
    p_out = np.asarray([0.5, 1.2, 2.4, 2.8, 3.2, 4.5])[None, :]
    n_columns, iv, kord, dummy_ptop = 1.0, 1.0, 1.0, 0.0
    result = mappm.mappm(p_in, f_in, p_out, 1, n_columns, iv, kord, dummy_ptop)
    assert result.dtype == np.float32
    np.testing.assert_almost_equal(result, np.asarray([[0.35, 1.3, 2.1, 2.5, 3.35]], np.float32), decimal=5)

    p_in = np.asarray([1.0, 2.0, 3.0, 4.0, 5.0])[None, :]
    f_in = np.asarray([1.5, 2.5, 3.5, 4.5])[None, :]
    p_out = np.asarray([0.0, 2.5, 3.5, 4.5, 50.0])[None, :]
    result = mappm.mappm(p_in, f_in, p_out, 1, n_columns, iv, kord, dummy_ptop)
    np.testing.assert_almost_equal(result, np.asarray([[1.5, 3.0, 4.0, 4.502747]], np.float32), decimal=5)

    p_in = np.asarray([1.0, 2.0, 3.0, 2.0, 5.0])[None, :]
    f_in = np.asarray([np.nan, np.nan, np.nan, np.nan])[None, :]
    result = mappm.mappm(p_in, f_in, p_out, 1, n_columns, iv, kord, dummy_ptop)
    np.testing.assert_almost_equal(result, np.asarray([[np.nan] * 4], np.float32))


def test_mappm_module_many_columns_bit_exact():
    from fv3net_b200 import mappm

    rng = np.random.RandomState(2)
    n, km, kn = 1000, 79, 40
    pe1 = 300 + np.concatenate([np.zeros((n, 1)), np.cumsum(rng.uniform(50, 2000, (n, km)), 1)], 1)
    pe2 = 300 + np.concatenate([np.zeros((n, 1)), np.cumsum(rng.uniform(100, 4000, (n, kn)), 1)], 1)
    q1 = rng.uniform(-1000, 1000, (n, km))
    got = mappm.mappm(pe1, q1, pe2, 1, n, 1, 1, 0.0)
    np.testing.assert_array_equal(got, oracle.mappm(pe1, q1, pe2))
