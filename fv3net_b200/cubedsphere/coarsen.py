"""Cubed-sphere block reductions: the reference's `vcm.cubedsphere.coarsen` API
(external/vcm/vcm/cubedsphere/coarsen.py), same names, signatures, defaults, dims / coords
/ attrs / name behaviour and error messages, with every reduction over data running as a
hand-written CUDA kernel behind the C ABI (include/coarsen_b200.h).

Inputs are `fv3net_b200.labeled.DataArray` / `Dataset` (numpy or device-resident torch data)
or real xarray objects where xarray is installed (adapted at the boundary, xarray out).
Inputs are never modified.  Only 1-D coordinate arithmetic (a few hundred integers per
call) runs on the host; there is no CPU path for the data.
"""
from collections import OrderedDict
from functools import partial

import numpy as np
import torch

from .. import labeled as lab
from .. import ops
from ..labeled import DataArray, Dataset
from . import _layout as lay
from .constants import COORD_X_OUTER, COORD_Y_OUTER

NUM_TILES = 6
SUBTILE_FILE_PATTERN = "{prefix}.tile{tile:d}.nc.{subtile:04d}"
STAGGERED_DIMS = [COORD_X_OUTER, COORD_Y_OUTER]


# ------------------------------------------------------------------ coordinates (host, integer)
def coarsen_coords_coord_func(coordinate, axis=-1):
    """coarsen.py:109-132: subtile index of the first point of every window; `coordinate` is
    the (n_windows, window) reshaping xarray hands to a coord_func; `axis` is not used."""
    coordinate = np.asarray(coordinate)
    return ((coordinate[:, 0] - 1) // coordinate.shape[1] + 1).astype(int).astype(np.float32)


def coarsen_coords(coarsening_factor, reference_subtile, dims):
    """coarsen.py:83-106: {dim: coarse coordinate} for restart / diagnostic subtiles."""
    reference_subtile = lab.adapt(reference_subtile)
    result = {}
    for dim in dims:
        c = np.asarray(reference_subtile[dim].values)[::coarsening_factor]
        c = ((c - 1) // coarsening_factor + 1).astype(int).astype(np.float32)
        result[dim] = DataArray(c, (dim,), {dim: c}, name=dim)
    return result


def add_coordinates(reference_obj, coarsened_obj, coarsening_factor, dims):
    """coarsen.py:135-161."""
    ref, coarse = lab.adapt(reference_obj), lab.adapt(coarsened_obj)
    coords = {d: c.values for d, c in coarsen_coords(coarsening_factor, ref, dims).items()}
    result = coarse.assign_coords(coords)
    if isinstance(coarse, DataArray):
        result = result.rename(coarse.name)
    return lab.restore(result, coarsened_obj)


def _propagate_attrs(reference_obj, obj):
    """coarsen.py:164-180."""
    if isinstance(reference_obj, Dataset):
        for variable in reference_obj:
            if variable in obj.data_vars:
                obj.data_vars[variable].attrs = dict(reference_obj.data_vars[variable].attrs)
    obj.attrs = dict(reference_obj.attrs)
    return obj


def _is_dict_like(value):
    return hasattr(value, "keys") and hasattr(value, "__getitem__")


_NAN_SKIPPING = {"sum": np.nansum, "mean": np.nanmean, "median": np.nanmedian, "min": np.nanmin,
                 "max": np.nanmax, "std": np.nanstd, "var": np.nanvar, "prod": np.nanprod}


def _windows_of(values, window, pad):
    """(n_windows, window) view of a 1-D coordinate; `pad` = xarray's boundary='pad' / the
    cval=nan padding of _skimage.block_reduce (coarsen.py:304-318)."""
    values = np.asarray(values)
    extra = (-len(values)) % window
    if extra:
        if not pad:
            lay.check_exact(len(values), window, "coordinate")
        values = np.concatenate([values.astype(np.result_type(values.dtype, np.float32)
                                               if values.dtype.kind != "f" else values.dtype),
                                 np.full(extra, np.nan)])
    return values.reshape(-1, window)


def _reduce_coordinate(values, window, function, pad):
    """One 1-D coordinate through a coord_func given as a string or a callable
    (coarsen.py:304-318; xarray Variable.coarsen for the string form)."""
    blocked = _windows_of(values, window, pad)
    if isinstance(function, str):
        if function not in _NAN_SKIPPING:
            raise ValueError(f"unsupported coord_func {function!r}")
        fn = _NAN_SKIPPING[function] if blocked.dtype.kind == "f" else getattr(np, function)
        return fn(blocked, axis=1)
    return np.asarray(function(blocked, axis=(1,)))


def _coord_function_for(coord_func, name):
    if _is_dict_like(coord_func):
        return coord_func.get(name, "mean")
    return coord_func


def _carry_coords(reference, result_dims, windows, coord_func, pad=False, drop=()):
    """Coordinates of `reference` for a result whose dims `windows` were block-reduced:
    1-D coordinates along a reduced dim go through coord_func (xarray's default for a
    coordinate that coord_func does not name is "mean"); the others are kept."""
    out = OrderedDict()
    for name, c in reference.coords.items():
        if name in drop or any(d not in result_dims for d in c.dims):
            continue
        hit = [d for d in c.dims if d in windows]
        if not hit:
            out[name] = c
        elif c.ndim == 1:
            fn = _coord_function_for(coord_func, name)
            out[name] = DataArray(_reduce_coordinate(c.values, windows[hit[0]], fn, pad), c.dims,
                                  None, name, c.attrs)
        else:
            raise NotImplementedError(f"multi-dimensional coordinate {name!r} along a coarsened dim")
    return out


# ------------------------------------------------------------------ weighted averages (a1, a2)
def _as_list(obj):
    """(list of data variables, rebuild function) for a DataArray or a Dataset."""
    if isinstance(obj, Dataset):
        names = list(obj.data_vars)
        return [obj[k] for k in names], names
    if not isinstance(obj, DataArray):
        raise TypeError(f"expected a DataArray or Dataset, got {type(obj).__name__}")
    return [obj], None


def _weighted_block_average_arrays(arrays, weights, y_dim, x_dim, fy, fx, sy, sx, nyo, nxo,
                                   weights_first=False):
    """sum_window(obj * w) / sum_window(w) for every array, NaN-skipping sums like xarray's
    coarsen().sum() (coarsen.py:210-213).  Arrays that share dims, shape and dtype go to the
    kernel in one launch.  `weights_first`: the product was written `weights * obj`, so the
    result's dims start with the weights' (xarray orders broadcast dims by first appearance).
    Returns [(tensor, its dims, dims of the result)] in input order."""
    results = [None] * len(arrays)
    groups = OrderedDict()
    for i, da in enumerate(arrays):
        groups.setdefault((da.dims, da.shape, str(da.dtype)), []).append(i)
    for (dims, _shape, _dt), members in groups.items():
        first = arrays[members[0]]
        a, b = (weights, first) if weights_first else (first, weights)
        sizes = OrderedDict(a.sizes)
        for d in b.dims:
            if d in sizes and sizes[d] != b.sizes[d]:
                raise ValueError(f"size of dimension {d!r} differs between the data "
                                 f"({first.sizes[d]}) and the weights ({weights.sizes[d]})")
            sizes.setdefault(d, b.sizes[d])
        for d in (y_dim, x_dim):
            if d not in sizes:
                raise ValueError(f"dimension {d!r} not found in the data or the weights")
        union = list(sizes)                                  # xarray broadcasting order
        shared = [d for d in union if d in weights.dims and d not in (y_dim, x_dim)]
        own = [d for d in union if d not in weights.dims and d not in (y_dim, x_dim)]
        order = shared + own + [y_dim, x_dim]
        n_outer = lay.numel([sizes[d] for d in shared])
        nz = lay.numel([sizes[d] for d in own])
        ny, nx = sizes[y_dim], sizes[x_dim]
        w_order = shared + [y_dim, x_dim]
        if set(w_order) == set(weights.dims):
            w = lay.device_tensor(weights, w_order)
        else:                                                # weights lacking a horizontal dim
            w = lay.broadcast_to(weights, w_order, sizes).contiguous()
        w = w.reshape(n_outer, ny, nx)
        tensors = []
        for i in members:
            da = arrays[i]
            if set(da.dims) == set(order):
                t = lay.device_tensor(da, order)
            else:
                t = lay.broadcast_to(da, order, sizes).contiguous()
            tensors.append(t.reshape(n_outer, nz, ny, nx))
        outs = ops.weighted_block_average(tensors, w2d=w, fy=fy, fx=fx, sy=sy, sx=sx, nyo=nyo, nxo=nxo)
        lead = [sizes[d] for d in shared + own]
        for i, o in zip(members, outs):
            results[i] = (o.reshape(*lead, nyo, nxo), order, union)
    return results


def weighted_block_average(obj, weights, coarsening_factor, x_dim="xaxis_1", y_dim="yaxis_2",
                           coord_func=coarsen_coords_coord_func):
    """Coarsen a DataArray or Dataset through weighted block averaging (coarsen.py:183-218).

    Args and defaults as the reference's; `weights` broadcasts against `obj` by dim name
    (2-D area against a 3-D field, or 3-D mass weights)."""
    src, wsrc = obj, weights
    obj, weights = lab.adapt(obj), lab.adapt(weights)
    f = int(coarsening_factor)
    arrays, names = _as_list(obj)
    for d in (x_dim, y_dim):
        size = weights.sizes[d] if d in weights.dims else obj.sizes.get(d)
        if size is None:
            raise ValueError(f"dimension {d!r} not found in the data or the weights")
        lay.check_exact(size, f, d)
    ny, nx = (weights.sizes.get(y_dim, obj.sizes.get(y_dim)), weights.sizes.get(x_dim, obj.sizes.get(x_dim)))
    res = _weighted_block_average_arrays(arrays, weights, y_dim, x_dim, f, f, f, f, ny // f, nx // f)
    windows = {x_dim: f, y_dim: f}
    out = []
    for da, (t, order, union) in zip(arrays, res):
        coords = _carry_coords(weights, union, windows, coord_func)
        coords.update(_carry_coords(da, union, windows, coord_func))
        data = lay.finish(t, order, union, obj)
        out.append(DataArray(data, union, coords, da.name, None))
    if names is None:
        result = out[0].rename(obj.name)
    else:
        result = Dataset(OrderedDict(zip(names, out)))
    return lab.restore(_propagate_attrs(obj, result), src)


def _edge_dims(edge, x_dim, y_dim):
    if edge == "x":
        return x_dim, y_dim
    if edge == "y":
        return y_dim, x_dim
    raise ValueError(f"'edge' most be either 'x' or 'y'; got {edge}.")


def _coarsen_downsample_coordinate(reference_obj, coarsened_obj, dim, coarsening_factor, coord_func):
    """coarsen.py:282-301: the downsampled dim's coordinate goes through coord_func by hand."""
    if dim in reference_obj.coords:
        function = _coord_function_for(coord_func, dim)
        c = reference_obj.coords[dim]
        values = _reduce_coordinate(c.values, coarsening_factor, function, pad=True)
        coarsened_obj = coarsened_obj.assign_coords({dim: ((dim,), values, c.attrs)})
    return coarsened_obj


def edge_weighted_block_average(obj, spacing, coarsening_factor, x_dim="xaxis_1", y_dim="yaxis_1",
                                edge="x", coord_func=coarsen_coords_coord_func):
    """Coarsen along a block edge (coarsen.py:221-273): weighted average over
    `coarsening_factor` consecutive cells along the edge, every `coarsening_factor`-th line of
    the other (staggered) dim kept.  Only the kept lines are read."""
    coarsen_dim, downsample_dim = _edge_dims(edge, x_dim, y_dim)
    src = obj
    obj, spacing = lab.adapt(obj), lab.adapt(spacing)
    f = int(coarsening_factor)
    arrays, names = _as_list(obj)
    sizes = dict(obj.sizes)
    sizes.update(spacing.sizes)
    lay.check_exact(sizes[coarsen_dim], f, coarsen_dim)
    n_down = -(-sizes[downsample_dim] // f)                  # isel(slice(None, None, f))
    n_coarse = sizes[coarsen_dim] // f
    if edge == "x":
        win = dict(fy=1, fx=f, sy=f, sx=f, nyo=n_down, nxo=n_coarse)
    else:
        win = dict(fy=f, fx=1, sy=f, sx=f, nyo=n_coarse, nxo=n_down)
    res = _weighted_block_average_arrays(arrays, spacing, y_dim, x_dim, weights_first=True, **win)
    windows = {coarsen_dim: f}
    out = []
    for da, (t, order, union) in zip(arrays, res):
        coords = _carry_coords(spacing, union, windows, coord_func, drop=(downsample_dim,))
        coords.update(_carry_coords(da, union, windows, coord_func, drop=(downsample_dim,)))
        data = lay.finish(t, order, union, obj)
        r = DataArray(data, union, coords, None, None)
        out.append(_coarsen_downsample_coordinate(da, r, downsample_dim, f, coord_func))
    if names is None:
        result = out[0]
        # spacing * obj keeps the name only when both operands carry the same one (xarray)
        result.name = obj.name if obj.name == spacing.name else None
    else:
        result = Dataset(OrderedDict(zip(names, out)))
    return lab.restore(_propagate_attrs(obj, result), src)


# ------------------------------------------------------------------ block reductions (a3-a6)
def _mode(arr, axis=0, nan_policy="propagate"):
    """Placeholder carrying the reference's name (coarsen.py:686): the mode over blocks runs
    on the device (cg_block_mode); this host callable exists only to be recognised."""
    raise NotImplementedError("_mode is dispatched to the CUDA block-mode kernel; it has no host path")


def _mode_reduce(arr, axis=(0,), nan_policy="propagate"):
    """coarsen.py:743-747; recognised by `xarray_block_reduce` and run as cg_block_mode."""
    raise NotImplementedError("_mode_reduce is dispatched to the CUDA block-mode kernel")


def _classify_reduction(function):
    """Reduction callable -> (kernel, skipna, kwargs).  Arbitrary Python callables cannot run
    on the device and there is no host path for data, so anything else is refused."""
    kwargs = {}
    if isinstance(function, partial):
        kwargs = dict(function.keywords or {})
        function = function.func
    table = {
        np.median: ("median", False), np.mean: ("mean", False), np.nanmean: ("mean", True),
        np.sum: ("sum", False), np.nansum: ("sum", True), np.min: ("min", False),
        np.nanmin: ("min", True), np.max: ("max", False), np.nanmax: ("max", True),
        _mode_reduce: ("mode", False),
    }
    for fn, (kernel, skipna) in table.items():
        if function is fn:
            return kernel, skipna, kwargs
    raise NotImplementedError(
        f"block reduction with {getattr(function, '__name__', function)!r} has no CUDA kernel; "
        "supported: np.median, np.mean, np.sum, np.min, np.max (and their nan- variants), "
        "and the block mode.  fv3net_b200 has no CPU fallback.")


def _result_dtype(kernel, dtype):
    """numpy's result dtype for the reduction of an array of `dtype`."""
    dtype = np.dtype(dtype)
    if dtype.kind == "f":
        return dtype
    if kernel in ("mean", "median"):
        return np.dtype(np.float64)
    if kernel == "sum":
        return np.dtype(np.int64) if dtype.kind in "ib" else np.dtype(np.uint64)
    return dtype


def _block_reduce_tensor(t, kernel, skipna, fy, fx, sy, sx, nyo, nxo, kwargs):
    """(n_outer, ny, nx) -> (n_outer, nyo, nxo) on the device."""
    if kernel == "median":
        if fy != fx or (sy, sx) != (fy, fx):
            raise NotImplementedError("block median needs square, tiling blocks")
        return ops.block_median(t, fy)
    if kernel == "mode":
        if fy != fx or (sy, sx) != (fy, fx):
            raise NotImplementedError("block mode needs square, tiling blocks")
        policy = kwargs.get("nan_policy", "propagate")
        if policy == "raise":
            if bool(torch.isnan(t).any()):
                raise ValueError("The input contains nan values")
            policy = "propagate"
        return ops.block_mode(t, fy, nan_policy=policy)
    return ops.block_reduce(t, kernel, fy, fx, sy, sx, nyo, nxo, skipna=skipna)


def _reduce_dataarray(da, windows, kernel, skipna, kwargs, cval=None):
    """Block-reduce `da` over `windows` = {dim: (window, stride)} (one or two dims).
    `cval`: pad value when a size is not a multiple of its window (skimage semantics);
    None = refuse like xarray's boundary='exact'.  Returns data with the dims of `da`."""
    active = [d for d in da.dims if d in windows and windows[d] != (1, 1)]
    if len(active) > 2:
        raise NotImplementedError("block reductions over more than two dimensions")
    if not active:
        return da.data
    others = [d for d in da.dims if d not in active]
    order = others + active
    t = lay.device_tensor(da, order)
    lead = [da.sizes[d] for d in others]
    if len(active) == 1:
        (wy, sy), (wx, sx) = (1, 1), windows[active[0]]
        ny, nx = 1, da.sizes[active[0]]
    else:
        (wy, sy), (wx, sx) = windows[active[0]], windows[active[1]]
        ny, nx = da.sizes[active[0]], da.sizes[active[1]]
    t = t.reshape(lay.numel(lead), ny, nx)
    if (wy, wx) == (sy, sx):                                  # tiling windows
        pad_y, pad_x = (-ny) % wy, (-nx) % wx
        if pad_y or pad_x:
            if cval is None:
                bad = active[-1] if pad_x else active[0]
                lay.check_exact(da.sizes[bad], windows[bad][0], bad)
            t = torch.nn.functional.pad(t, (0, pad_x, 0, pad_y), value=float(cval)).contiguous()
            ny, nx = ny + pad_y, nx + pad_x
        nyo, nxo = ny // wy, nx // wx
    else:                                                     # window along one dim, stride on both
        nyo, nxo = (ny - wy) // sy + 1, (nx - wx) // sx + 1
    out = _block_reduce_tensor(t, kernel, skipna, wy, wx, sy, sx, nyo, nxo, kwargs)
    shape = lead + ([nxo] if len(active) == 1 else [nyo, nxo])
    return lay.finish(out.reshape(shape), order, da.dims, da,
                      dtype=lay.torch_dtype(_result_dtype(kernel, da.dtype)))


def _block_reduce_dataarray_coordinates(reference_da, da, block_sizes, coord_func):
    """coarsen.py:321-347."""
    coords = OrderedDict()
    for dim, size in block_sizes.items():
        if dim in reference_da.dims and dim in reference_da.coords:
            function = _coord_function_for(coord_func, dim)
            c = reference_da.coords[dim]
            coords[dim] = DataArray(_reduce_coordinate(c.values, size, function, pad=True), (dim,),
                                    None, dim, c.attrs)
    for name, c in reference_da.coords.items():
        if name not in coords and not any(d in block_sizes for d in c.dims):
            coords[name] = c
    return da.assign_coords(coords)


def _xarray_block_reduce_dataarray(da, block_sizes, reduction_function, cval=np.nan,
                                   coord_func=coarsen_coords_coord_func):
    """coarsen.py:393-460: block reduce over arbitrary named dims; sizes that are not a
    multiple of the block are padded with `cval`; a DataArray without any of the dims is
    returned unchanged."""
    if not set(block_sizes).intersection(da.dims):
        return da
    kernel, skipna, kwargs = _classify_reduction(reduction_function)
    sizes = {d: int(s) for d, s in block_sizes.items() if d in da.dims}
    data = _reduce_dataarray(da, {d: (s, s) for d, s in sizes.items()}, kernel, skipna, kwargs, cval=cval)
    result = DataArray(data, da.dims, None, da.name, None)
    result = _block_reduce_dataarray_coordinates(da, result, sizes, coord_func)
    return _propagate_attrs(da, result)


def xarray_block_reduce(obj, block_sizes, reduction_function, cval=np.nan,
                        coord_func=coarsen_coords_coord_func):
    """A generic block reduce for labelled data (coarsen.py:463-517).  `reduction_function`
    must be one the device implements (see `_classify_reduction`)."""
    src = obj
    obj = lab.adapt(obj)
    if isinstance(obj, Dataset):
        result = Dataset(OrderedDict(
            (k, _xarray_block_reduce_dataarray(obj[k], block_sizes, reduction_function, cval=cval,
                                               coord_func=coord_func)) for k in obj.data_vars))
    else:
        result = _xarray_block_reduce_dataarray(obj, block_sizes, reduction_function, cval=cval,
                                                coord_func=coord_func)
        if result is obj:
            result = obj.copy()
    return lab.restore(_propagate_attrs(obj, result), src)


def horizontal_block_reduce(obj, coarsening_factor, reduction_function, x_dim="xaxis_1",
                            y_dim="yaxis_1", coord_func=coarsen_coords_coord_func):
    """coarsen.py:520-554."""
    block_sizes = {x_dim: coarsening_factor, y_dim: coarsening_factor}
    return xarray_block_reduce(obj, block_sizes, reduction_function, coord_func=coord_func)


def block_median(obj, coarsening_factor, x_dim="xaxis_1", y_dim="yaxis_1",
                 coord_func=coarsen_coords_coord_func):
    """Median over blocks (coarsen.py:557-588); bit-exact selection on the device."""
    return horizontal_block_reduce(obj, coarsening_factor, np.median, x_dim=x_dim, y_dim=y_dim,
                                   coord_func=coord_func)


def _coarsen_method(obj, windows, method, coord_func):
    """obj.coarsen(windows, coord_func=coord_func).<method>() of xarray (NaN-skipping,
    boundary='exact') for the methods the device implements."""
    if method not in ("sum", "mean", "min", "max"):
        raise NotImplementedError(f"coarsening method {method!r} has no CUDA kernel "
                                  "(supported: sum, mean, min, max, median, mode)")
    arrays, names = _as_list(obj)
    out = []
    for da in arrays:
        active = {d: (w, w) for d, w in windows.items() if d in da.dims}
        if not active:
            out.append(da)
            continue
        data = _reduce_dataarray(da, active, method, True, {}, cval=None)
        coords = _carry_coords(da, da.dims, {d: w for d, (w, _) in active.items()}, coord_func)
        out.append(DataArray(data, da.dims, coords, da.name, da.attrs))
    if names is None:
        return out[0]
    return Dataset(OrderedDict(zip(names, out)), attrs=obj.attrs)


def block_edge_sum(obj, coarsening_factor, x_dim="xaxis_1", y_dim="yaxis_1", edge="x",
                   coord_func=coarsen_coords_coord_func):
    """coarsen.py:591-626."""
    return block_edge_coarsen(obj, coarsening_factor, x_dim=x_dim, y_dim=y_dim, edge=edge,
                              coord_func=coord_func, method="sum")


def block_edge_coarsen(obj, coarsening_factor, x_dim="xaxis_1", y_dim="yaxis_1", edge="x",
                       coord_func=coarsen_coords_coord_func, method="sum"):
    """Reduce along a block edge, keep every `coarsening_factor`-th line of the other dim
    (coarsen.py:629-683).  One launch: the window is (1, f) with stride (f, f)."""
    coarsen_dim, downsample_dim = _edge_dims(edge, x_dim, y_dim)
    src = obj
    obj = lab.adapt(obj)
    f = int(coarsening_factor)
    if method not in ("sum", "mean", "min", "max"):
        raise NotImplementedError(f"coarsening method {method!r} has no CUDA kernel")
    arrays, names = _as_list(obj)
    out = []
    for da in arrays:
        if coarsen_dim not in da.dims:
            r = da.isel({downsample_dim: slice(None, None, f)}) if downsample_dim in da.dims else da
        else:
            lay.check_exact(da.sizes[coarsen_dim], f, coarsen_dim)
            windows = {coarsen_dim: (f, f)}
            if downsample_dim in da.dims:
                windows[downsample_dim] = (1, f)
            data = _reduce_dataarray(da, windows, method, True, {}, cval=None)
            coords = _carry_coords(da, da.dims, {coarsen_dim: f}, coord_func, drop=(downsample_dim,))
            r = DataArray(data, da.dims, coords, da.name, da.attrs)
        out.append(_coarsen_downsample_coordinate(da, r, downsample_dim, f, coord_func))
    result = out[0] if names is None else Dataset(OrderedDict(zip(names, out)))
    return lab.restore(_propagate_attrs(obj, result), src)


def _block_mode(obj, coarsening_factor, x_dim="xaxis_1", y_dim="yaxis_1",
                coord_func=coarsen_coords_coord_func, nan_policy="propagate"):
    """Mode over blocks, scipy.stats.mode semantics (coarsen.py:750-786)."""
    return horizontal_block_reduce(obj, coarsening_factor, partial(_mode_reduce, nan_policy=nan_policy),
                                   x_dim=x_dim, y_dim=y_dim, coord_func=coord_func)


CUSTOM_COARSENING_FUNCTIONS = {"median": block_median, "mode": _block_mode}


def block_coarsen(obj, coarsening_factor, x_dim="xaxis_1", y_dim="yaxis_1", method="sum",
                  coord_func=coarsen_coords_coord_func, func_kwargs=None):
    """coarsen.py:795-840."""
    if func_kwargs is None:
        func_kwargs = {}
    if method in CUSTOM_COARSENING_FUNCTIONS:
        return CUSTOM_COARSENING_FUNCTIONS[method](obj, coarsening_factor, x_dim=x_dim, y_dim=y_dim,
                                                   coord_func=coord_func, **func_kwargs)
    src = obj
    obj = lab.adapt(obj)
    f = int(coarsening_factor)
    return lab.restore(_coarsen_method(obj, {x_dim: f, y_dim: f}, method, coord_func), src)


# ------------------------------------------------------------------ upsampling (a7)
def _upsample_dataarray(da, upsampling_factor, dims):
    """Repeat every point `upsampling_factor` times along each of `dims` present in `da`; an
    odd-sized dim holds cell interfaces: all but its last point are repeated, the last is kept
    once (coarsen.py:843-866).  Like the reference's apply_ufunc-based repeat
    (xarray_utils.py:41-55) each processed dim ends up last, its coordinate and the attrs are
    dropped, the name is kept."""
    present = [d for d in dims if d in da.dims]
    if not present:
        return da
    f = int(upsampling_factor)
    out_dims = [d for d in da.dims if d not in present] + present
    sizes = OrderedDict(da.sizes)
    data_dims = list(da.dims)
    t = None
    for i in range(0, len(present), 2):
        pair = present[i:i + 2]
        others = [d for d in data_dims if d not in pair]
        order = others + pair
        if t is None:
            t = lay.device_tensor(da, order)
        else:
            t = t.permute([data_dims.index(d) for d in order]).contiguous()
        lead = [sizes[d] for d in others]
        if len(pair) == 1:
            n = sizes[pair[0]]
            o = ops.block_upsample(t.reshape(lay.numel(lead), 1, n), 1, f, False, n % 2 == 1)
            sizes[pair[0]] = o.shape[-1]
            t = o.reshape(lead + [o.shape[-1]])
        else:
            ny, nx = sizes[pair[0]], sizes[pair[1]]
            o = ops.block_upsample(t.reshape(lay.numel(lead), ny, nx), f, f, ny % 2 == 1, nx % 2 == 1)
            sizes[pair[0]], sizes[pair[1]] = o.shape[-2], o.shape[-1]
            t = o.reshape(lead + [o.shape[-2], o.shape[-1]])
        data_dims = order
    dtype = None if da.dtype.kind == "f" else lay.torch_dtype(da.dtype)
    data = lay.finish(t, data_dims, out_dims, da, dtype=dtype)
    coords = OrderedDict((k, c) for k, c in da.coords.items() if not any(d in present for d in c.dims))
    return DataArray(data, out_dims, coords, da.name, None)


def block_upsample(obj, upsampling_factor, dims):
    """Upsample by repeating values along each horizontal dim (coarsen.py:869-897); drops the
    coordinates along those dims."""
    src = obj
    obj = lab.adapt(obj)
    if isinstance(obj, Dataset):
        result = Dataset(OrderedDict((k, _upsample_dataarray(obj[k], upsampling_factor, dims))
                                     for k in obj.data_vars), attrs=obj.attrs)
    else:
        result = _upsample_dataarray(obj, upsampling_factor, dims)
    return lab.restore(result, src)


def block_upsample_like(da, reference_da, x_dim="xaxis_1", y_dim="yaxis_1"):
    """Upsample `da` back to the resolution of `reference_da` and take its horizontal
    coordinates (coarsen.py:900-938)."""
    src = da
    da, reference_da = lab.adapt(da), lab.adapt(reference_da)
    if da.sizes[x_dim] % 2 == 1:
        upsampling_factor = (reference_da.sizes[x_dim] - 1) // (da.sizes[x_dim] - 1)
    else:
        upsampling_factor = reference_da.sizes[x_dim] // da.sizes[x_dim]
    result = _upsample_dataarray(da, upsampling_factor, [x_dim, y_dim])
    coords = {d: reference_da.coords[d] for d in (x_dim, y_dim) if d in reference_da.coords}
    return lab.restore(result.assign_coords(coords), src)


# ------------------------------------------------------------------ edge -> centre (helpers)
def rename_centered_xy_coords(cell_centered_da, edge_to_center_dims=None):
    """coarsen.py:34-51."""
    edge_to_center_dims = edge_to_center_dims or {"grid_x": "grid_xt", "grid_y": "grid_yt"}
    src = cell_centered_da
    da = lab.adapt(cell_centered_da)
    for dim, centre in edge_to_center_dims.items():
        if dim in da.dims:
            da = da.assign_coords({dim: ((dim,), np.asarray(da[dim].values) - 1)})
            da = da.rename({dim: centre})
    return lab.restore(da, src)


def shift_edge_var_to_center(edge, edge_to_center_dims=None):
    """Average an edge-valued variable (u, v) onto cell centres: 0.5 * (a[i] + a[i-1]) along
    the first staggered dim found (coarsen.py:54-80)."""
    edge_to_center_dims = edge_to_center_dims or {"grid_x": "grid_xt", "grid_y": "grid_yt"}
    src = edge
    da = lab.adapt(edge)
    for staggered_dim in [d for d in edge_to_center_dims if d in da.dims]:
        axis = da.get_axis_num(staggered_dim)
        n = da.sizes[staggered_dim]
        centred = lay.finish(ops.shift_to_center(lay.device_tensor(da), axis), da.dims, da.dims, da)   # cg_shift_to_center
        coords = OrderedDict()
        for k, c in da.coords.items():
            if staggered_dim in c.dims:
                if c.ndim == 1:
                    coords[k] = DataArray(c.values[1:], c.dims, None, k, c.attrs)
            else:
                coords[k] = c
        out = DataArray(centred, da.dims, coords, da.name, None)
        return lab.restore(rename_centered_xy_coords(out, edge_to_center_dims), src)
    raise ValueError(
        "Variable to shift to center must be centered on one horizontal axis and "
        "edge-valued on the other."
    )
