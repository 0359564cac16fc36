"""The xarray / dask boundary of the drop-in API (fv3net_b200.labeled.adapt / restore): xarray in -> xarray out
with the same dims, coords, attrs and name; lazy (dask) in -> lazy out (external/vcm/tests/test_regridz.py:55-71).
xarray and dask are not installed in this image, so the boundary is driven with duck-typed stand-ins
(tests/standins) that carry exactly the attributes the boundary touches."""
import numpy as np
import pytest

from fv3net_b200 import labeled as lab
from tests.standins import fake_dask_array as fda, fake_xarray as fxr


@pytest.fixture
def xr(monkeypatch):
    monkeypatch.setattr(lab, "_xr", fxr)
    return fxr


def _cube(xr, lazy=False, name="T"):
    rng = np.random.RandomState(0)
    data = rng.uniform(-5, 5, (2, 3, 8, 8)).astype(np.float32)
    coords = {"tile": np.arange(2), "z": xr.Variable(("z",), np.arange(3.0), {"units": "level"}),
              "y": np.arange(1.0, 9.0), "x": np.arange(1.0, 9.0)}
    return xr.DataArray(fda.from_array(data, chunks=(1, 3, 8, 8)) if lazy else data,
                        dims=("tile", "z", "y", "x"), coords=coords, name=name, attrs={"units": "K"}), data


def test_adapt_and_restore_round_trip(xr):
    """Host only: every piece of metadata survives the boundary in both directions."""
    src, data = _cube(xr)
    inner = lab.adapt(src)
    assert isinstance(inner, lab.DataArray) and inner.dims == src.dims and inner.name == "T" and inner.attrs == {"units": "K"}
    assert inner.coords["z"].attrs == {"units": "level"}
    np.testing.assert_array_equal(inner.values, data)
    out = lab.restore(inner, src)
    assert isinstance(out, xr.DataArray) and out.dims == src.dims and out.name == "T" and out.attrs == {"units": "K"}
    assert list(out.coords) == list(src.coords) and out.coords["z"].attrs == {"units": "level"}
    np.testing.assert_array_equal(out.values, data)
    ds = xr.Dataset({"a": src, "b": src}, attrs={"title": "t"})
    inner = lab.adapt(ds)
    assert isinstance(inner, lab.Dataset) and list(inner.data_vars) == ["a", "b"] and inner.attrs == {"title": "t"}
    back = lab.restore(inner, ds)
    assert isinstance(back, xr.Dataset) and list(back.data_vars) == ["a", "b"] and back.attrs == {"title": "t"}
    assert lab.restore((inner["a"], inner["b"]), src)[1].dims == src.dims          # tuples of results
    a = inner["a"]
    assert lab.restore(a, a) is a                                                  # labelled in -> labelled out


def test_lazy_in_lazy_out_on_the_host(xr):
    src, data = _cube(xr, lazy=True)
    before = fda.Array.computed
    inner = lab.adapt(src)
    assert fda.Array.computed == before + 1                                        # the boundary computes, once
    out = lab.restore(inner, src)
    assert isinstance(out.data, fda.Array) and out.data.chunks == data.shape
    np.testing.assert_array_equal(out.compute().values, data)


@pytest.mark.gpu
def test_api_functions_through_the_xarray_boundary(xr):
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from fv3net_b200.cubedsphere import block_median, regrid_vertical, weighted_block_average

    src, data = _cube(xr)
    area = xr.DataArray(np.random.RandomState(1).uniform(0.5, 1, (2, 8, 8)).astype(np.float32),
                        dims=("tile", "y", "x"), coords={"y": np.arange(1.0, 9.0), "x": np.arange(1.0, 9.0)})
    got = weighted_block_average(src, area, 4, x_dim="x", y_dim="y")
    ref = weighted_block_average(lab.adapt(src), lab.adapt(area), 4, x_dim="x", y_dim="y")
    assert isinstance(got, xr.DataArray) and got.dims == ref.dims and got.name == "T" and got.attrs == {"units": "K"}
    assert set(got.coords) == set(ref.coords)
    for k in ref.coords:
        np.testing.assert_array_equal(got.coords[k].values, ref.coords[k].values)
    np.testing.assert_array_equal(got.values, ref.values)
    ds = xr.Dataset({"a": src, "b": src}, attrs={"title": "t"})
    got = weighted_block_average(ds, area, 4, x_dim="x", y_dim="y")
    assert isinstance(got, xr.Dataset) and got.attrs == {"title": "t"}
    np.testing.assert_array_equal(got["b"].values, ref.values)
    med = block_median(src, 4, x_dim="x", y_dim="y")
    assert isinstance(med, xr.DataArray) and med.dims == src.dims
    np.testing.assert_array_equal(med.values, block_median(lab.adapt(src), 4, x_dim="x", y_dim="y").values)


@pytest.mark.gpu
@pytest.mark.parametrize("lazy", [False, True])
def test_regrid_vertical_lazy_in_lazy_out(xr, lazy):
    """external/vcm/tests/test_regridz.py:55-71."""
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from fv3net_b200.cubedsphere import regrid_vertical

    rng = np.random.RandomState(2)
    nz = 6
    dp = rng.uniform(300, 500, (4, 5, nz)).astype(np.float32)
    p_in = np.concatenate([np.full((4, 5, 1), 300.0, np.float32), 300 + np.cumsum(dp, -1)], -1)
    p_out = np.concatenate([np.full((4, 5, 1), 300.0, np.float32), 300 + np.cumsum(dp[..., ::-1], -1)], -1)
    f_in = rng.uniform(-1, 1, (4, 5, nz)).astype(np.float32)

    def wrap(a, zname):
        d = fda.from_array(a, chunks=a.shape) if lazy else a
        return xr.DataArray(d, dims=("x", "y", zname), coords={"x": np.arange(4.0)}, name="f", attrs={"long_name": "field"})

    from fv3net_b200 import _lib
    computed, launched = fda.Array.computed, _lib.lib().cg_launch_count()
    out = regrid_vertical(wrap(p_in, "z_outer"), wrap(f_in, "z_center"), wrap(p_out, "z_outer"),
                          z_dim_center="z_center", z_dim_outer="z_outer")
    assert isinstance(out, xr.DataArray) and out.dims == ("x", "y", "z_center")
    assert isinstance(out.data, fda.Array) == lazy
    assert out.shape == (4, 5, nz) and out.name == "f" and out.attrs == {"long_name": "field"}
    if lazy:                # nothing was materialised and no kernel ran when the call was made ...
        assert fda.Array.computed == computed and _lib.lib().cg_launch_count() == launched
        with pytest.raises(ValueError):          # ... but the shape checks did (regridz.py:334-350)
            regrid_vertical(wrap(p_in[:, :, :-1], "z_outer"), wrap(f_in, "z_center"), wrap(p_out, "z_outer"),
                            z_dim_center="z_center", z_dim_outer="z_outer")
    eager = regrid_vertical(lab.DataArray(p_in, ["x", "y", "z_outer"]), lab.DataArray(f_in, ["x", "y", "z_center"], name="f"),
                            lab.DataArray(p_out, ["x", "y", "z_outer"]), z_dim_center="z_center", z_dim_outer="z_outer")
    np.testing.assert_array_equal(out.compute().values if lazy else out.values, eager.values)
    if lazy:                # ... the three inputs and the result were materialised by compute(), on the GPU
        assert fda.Array.computed == computed + 4 and _lib.lib().cg_launch_count() > launched
