"""Vertical regridding to coarse pressure levels: the reference's `vcm.cubedsphere.regridz`
API (external/vcm/vcm/cubedsphere/regridz.py) with the Fortran `mappm` remapper
(external/mappm/mappm/mappm.f90) replaced by CUDA.

Two ways in:
  * the reference's own functions (`regrid_vertical`, `regrid_to_area_weighted_pressure`,
    `regrid_to_edge_weighted_pressure`, `_mask_weights`), which return FINE-grid remapped
    fields exactly like the reference does, and
  * `pressure_level_average`, the fused form of what every caller does next (block-average
    the remapped fields with the masked / unmasked area: coarsen_restarts.py:152-225,
    fine_res_budget/budget/budgets.py:49-57): one kernel, nothing fine-grid is written.
"""
from collections import OrderedDict

import numpy as np
import torch

from .. import labeled as lab
from .. import ops
from ..labeled import DataArray, Dataset
from . import _layout as lay
from .coarsen import _as_list, _carry_coords, coarsen_coords_coord_func
from .constants import (FV_CORE_X_CENTER, FV_CORE_X_OUTER, FV_CORE_Y_CENTER, FV_CORE_Y_OUTER,
                        RESTART_Z_CENTER, RESTART_Z_OUTER, TOA_PRESSURE)


# ------------------------------------------------------------------ regrid_vertical (a9, a10)
def _n_columns(da, z_dim):
    return lay.numel([s for d, s in da.sizes.items() if d != z_dim])


def _assert_equal_number_of_columns(p_in, f_in, p_out, z_dim_center, z_dim_outer):
    """regridz.py:334-343."""
    n = _n_columns(p_in, z_dim_outer)
    if _n_columns(f_in, z_dim_center) != n or _n_columns(p_out, z_dim_outer) != n:
        raise ValueError("All dimensions except vertical must be same size for p_in, f_in and p_out")


def _assert_valid_vertical_dimension_sizes(p_in, f_in, z_dim_outer, z_dim_center):
    """regridz.py:346-350."""
    if f_in.sizes[z_dim_center] != p_in.sizes[z_dim_outer] - 1:
        raise ValueError("f_in must have a vertical dimension one shorter than p_in")


def regrid_vertical(p_in, f_in, p_out, iv=1, kord=1, z_dim_center=RESTART_Z_CENTER,
                    z_dim_outer=RESTART_Z_OUTER):
    """Remap `f_in` (layer means between the interfaces `p_in`) onto the layers between
    `p_out` with the piecewise-parabolic method of mappm.f90 (regridz.py:164-242).

    Arguments, defaults, returned dims order / attrs and the three ValueErrors are the
    reference's.  The vertical dim may sit anywhere: the kernel walks columns through strides,
    nothing is transposed to z-last.  The arithmetic and the result are float32 whatever the
    input dtype (f2py casts to Fortran `real`, test_mappm.py:14)."""
    if z_dim_center == z_dim_outer:
        raise ValueError("'z_dim_center' and 'z_dim_outer' must not be equal.")
    src = f_in
    lazy = next((a for a in (f_in, p_in, p_out) if lab.is_lazy(a)), None)
    if lazy is not None:
        # dask in -> dask out, nothing computed now (regridz.py:254-265 maps blocks lazily; test_regridz.py:63):
        # the checks run on shapes alone, the inputs are materialised and the kernel runs when the graph is computed
        mp, mf, mo = (lab.adapt(a, values=False) for a in (p_in, f_in, p_out))
        _check_regrid_vertical(mp, mf, mo, z_dim_center, z_dim_outer)
        shape = [mo.sizes[z_dim_outer] - 1 if d == z_dim_center else n for d, n in mf.sizes.items()]

        def run():
            return _regrid_vertical(lab.adapt(p_in), lab.adapt(f_in), lab.adapt(p_out), iv, kord, z_dim_center,
                                    z_dim_outer).values

        coords = OrderedDict((k, c) for k, c in mf.coords.items() if z_dim_center not in c.dims)
        meta = DataArray(np.broadcast_to(np.zeros((), np.float32), shape), mf.dims, coords, mf.name, mf.attrs)
        return lab.restore(meta, src, data=lab.defer(run, lazy, shape, np.float32))
    p_in, f_in, p_out = lab.adapt(p_in), lab.adapt(f_in), lab.adapt(p_out)
    return lab.restore(_regrid_vertical(p_in, f_in, p_out, iv, kord, z_dim_center, z_dim_outer), src)


def _check_regrid_vertical(p_in, f_in, p_out, z_dim_center, z_dim_outer):
    for name, da, zd in (("p_in", p_in, z_dim_outer), ("f_in", f_in, z_dim_center),
                         ("p_out", p_out, z_dim_outer)):
        if zd not in da.dims:
            raise ValueError(f"{name} has no dimension {zd!r}")
    _assert_equal_number_of_columns(p_in, f_in, p_out, z_dim_center, z_dim_outer)
    _assert_valid_vertical_dimension_sizes(p_in, f_in, z_dim_outer, z_dim_center)


def _regrid_vertical(p_in, f_in, p_out, iv, kord, z_dim_center, z_dim_outer):
    """The eager body of `regrid_vertical` on labelled arrays."""
    _check_regrid_vertical(p_in, f_in, p_out, z_dim_center, z_dim_outer)

    dims = f_in.dims
    axis = f_in.get_axis_num(z_dim_center)
    p_dims = tuple(z_dim_outer if d == z_dim_center else d for d in dims)
    q = lay.device_tensor(f_in).to(torch.float32)
    pe1 = lay.device_tensor(p_in, p_dims if set(p_dims) == set(p_in.dims) else None).to(torch.float32)
    pe2 = lay.device_tensor(p_out, p_dims if set(p_dims) == set(p_out.dims) else None).to(torch.float32)
    km, kn = f_in.sizes[z_dim_center], p_out.sizes[z_dim_outer] - 1
    shape = list(q.shape)
    n_outer, n_inner = lay.numel(shape[:axis]), lay.numel(shape[axis + 1:])
    shape[axis] = kn
    out = torch.empty(shape, dtype=torch.float32, device=q.device)

    def strides(nlev):
        return (nlev * n_inner, 1, n_inner)

    ops.mappm(pe1, q, pe2, n_outer, n_inner, km, kn, strides(km + 1), strides(km), strides(kn + 1),
              strides(kn), out, iv=int(iv), kord=int(kord))
    coords = OrderedDict((k, c) for k, c in f_in.coords.items() if z_dim_center not in c.dims)
    return DataArray(lay.finish(out, dims, dims, f_in), dims, coords, f_in.name, f_in.attrs)


# ------------------------------------------------------------------ composition (a14, a15)
def _canonical4(da, lead, z_dim, y_dim, x_dim):
    """(n_outer, nz, ny, nx) contiguous tensor of a (lead..., z, y, x) field."""
    t = lay.device_tensor(da, list(lead) + [z_dim, y_dim, x_dim])
    return t.reshape(lay.numel([da.sizes[d] for d in lead]), da.sizes[z_dim], da.sizes[y_dim],
                     da.sizes[x_dim])


def _weights3(weights, delp, lead, y_dim, x_dim):
    """2-D weights as a contiguous (lead..., y, x) tensor; leading dims the weights lack (the
    grid's area has no Time) are broadcast by name, like xarray does."""
    order = list(lead) + [y_dim, x_dim]
    if any(d not in order for d in weights.dims) or y_dim not in weights.dims or x_dim not in weights.dims:
        raise ValueError(f"weights dims {weights.dims} must be among {tuple(order)} and hold {y_dim!r}, {x_dim!r}")
    if set(weights.dims) == set(order):
        return lay.device_tensor(weights, order)
    return lay.broadcast_to(weights, order, delp.sizes).contiguous()


def _mask_weights(weights, phalf_coarse_on_fine, phalf_fine, dim_center=RESTART_Z_CENTER,
                  dim_outer=RESTART_Z_OUTER):
    """Weights zeroed where the coarse layer's bottom interface is not above the fine surface:
    w if p_coarse_on_fine[k+1] < p_fine[surface] else 0 (regridz.py:150-161)."""
    src = weights
    weights, pc, pf = lab.adapt(weights), lab.adapt(phalf_coarse_on_fine), lab.adapt(phalf_fine)
    hor = list(weights.dims)
    if set(hor) != set(d for d in pc.dims if d != dim_outer):
        raise ValueError(f"weights dims {weights.dims} must be the non-vertical dims of the pressures")
    # kernel layout: (n_outer, nz+1, h), z ahead of the flattened horizontal dims
    z_first = [dim_outer] + hor
    tc = lay.device_tensor(pc, z_first).unsqueeze(0)
    tf = lay.device_tensor(pf, z_first).unsqueeze(0)
    tw = lay.device_tensor(weights, hor).unsqueeze(0)
    out = ops.mask_weights(tw, tc, tf)[0]                       # (nz, *hor)
    out_dims = hor + [dim_outer]                                # where(): own dims, then the condition's
    data = lay.finish(out, z_first, out_dims, weights)
    dims = tuple(dim_center if d == dim_outer else d for d in out_dims)
    coords = OrderedDict((k, c) for k, c in weights.coords.items())
    return lab.restore(DataArray(data, dims, coords, weights.name, weights.attrs), src)


def _regrid_given_delp(ds, delp_fine, delp_coarse, weights, x_dim=FV_CORE_X_CENTER,
                       y_dim=FV_CORE_Y_CENTER, z_dim=RESTART_Z_CENTER):
    """regridz.py:115-147: upsample the coarse delp, both interface pressures, remap every
    variable on the fine grid, mask the weights.  Returns (fine-grid Dataset, masked weights)."""
    lead = [d for d in delp_fine.dims if d not in (z_dim, y_dim, x_dim)]
    dims = tuple(lead) + (z_dim, y_dim, x_dim)
    df = _canonical4(delp_fine, lead, z_dim, y_dim, x_dim)
    dc = _canonical4(delp_coarse, lead, z_dim, y_dim, x_dim)
    n_outer, nz, ny, nx = df.shape
    nyc, nxc = dc.shape[2:]
    stag_y, stag_x = nyc % 2 == 1, nxc % 2 == 1
    fy = (ny - 1) // (nyc - 1) if stag_y else ny // nyc
    fx = (nx - 1) // (nxc - 1) if stag_x else nx // nxc
    up = ops.block_upsample(dc.reshape(n_outer * nz, nyc, nxc), fy, fx, stag_y, stag_x)
    if tuple(up.shape[1:]) != (ny, nx):
        raise ValueError(f"coarse delp {tuple(dc.shape)} does not upsample to the fine grid {tuple(df.shape)}")
    pc = ops.pressure_at_interface(up.reshape(n_outer, nz, ny, nx), TOA_PRESSURE)
    pf = ops.pressure_at_interface(df, TOA_PRESSURE)
    h = ny * nx
    s_p, s_q = ((nz + 1) * h, 1, h), (nz * h, 1, h)
    pf32, pc32 = pf.to(torch.float32), pc.to(torch.float32)    # f2py hands mappm Fortran `real`s
    arrays, names = _as_list(ds)
    lead_shape = [delp_fine.sizes[d] for d in lead]
    out = OrderedDict()
    for da in arrays:
        if set(da.dims) != set(dims):
            raise ValueError(f"variable {da.name!r} has dims {da.dims}; expected {dims} in any order")
        q = _canonical4(da, lead, z_dim, y_dim, x_dim).to(torch.float32)
        res = torch.empty_like(q)
        ops.mappm(pf32, q, pc32, n_outer, h, nz, nz, s_p, s_q, s_p, s_q, res)
        data = lay.finish(res.reshape(lead_shape + [nz, ny, nx]), dims, da.dims, da)
        coords = OrderedDict((k, c) for k, c in da.coords.items())
        out[da.name] = DataArray(data, da.dims, coords, da.name, da.attrs)
    ds_regrid = Dataset(out, attrs=ds.attrs) if names is not None else out[arrays[0].name]
    tw = _weights3(weights, delp_fine, lead, y_dim, x_dim).reshape(n_outer, ny, nx)
    mw = ops.mask_weights(tw, pc, pf)
    coords = OrderedDict((k, c) for k, c in weights.coords.items())
    masked = DataArray(lay.finish(mw.reshape(lead_shape + [nz, ny, nx]), dims, dims, weights), dims,
                       coords, weights.name, weights.attrs)
    return ds_regrid, masked


def regrid_to_area_weighted_pressure(ds, delp, area, coarsening_factor, x_dim=FV_CORE_X_CENTER,
                                     y_dim=FV_CORE_Y_CENTER, z_dim=RESTART_Z_CENTER):
    """Regrid cell-centred quantities to the pressure levels of the area-weighted coarse
    `delp` (regridz.py:25-55).  Returns (fine-grid regridded data, area masked where the coarse
    layer's bottom interface is below the fine surface); dims are carried by name in the
    order (..., z, y, x)."""
    src = ds
    ds, delp, area = lab.adapt(ds), lab.adapt(delp), lab.adapt(area)
    from .coarsen import weighted_block_average

    delp_coarse = weighted_block_average(delp, area, coarsening_factor, x_dim=x_dim, y_dim=y_dim)
    regridded, masked = _regrid_given_delp(ds, delp, delp_coarse, area, x_dim=x_dim, y_dim=y_dim,
                                           z_dim=z_dim)
    return lab.restore(regridded, src), lab.restore(masked, src)


def _edge_delp(delp, edge, lead, x_dim, y_dim, z_dim, tile_dim="tile"):
    """`delp` interpolated to the cell edges across the six faces (regridz.py:85-100,
    xgcm.py:7-34): value j = (c[j-1] + c[j]) / 2 with the neighbour face supplying the halo."""
    if tile_dim not in lead or delp.sizes[tile_dim] != 6:
        raise ValueError("The input Dataset must have a `tile` coordinate.")
    other = [d for d in lead if d != tile_dim]
    # kernel layout (6, nz', n, n): tile first, the remaining leading dims folded into z
    t = lay.device_tensor(delp, [tile_dim] + other + [z_dim, y_dim, x_dim])
    n = delp.sizes[x_dim]
    nzz = lay.numel([delp.sizes[d] for d in other]) * delp.sizes[z_dim]
    interp_axis = "x" if edge == "y" else "y"
    out = ops.interp_to_edges(t.reshape(6, nzz, n, n), interp_axis)
    shape = [6] + [delp.sizes[d] for d in other] + [delp.sizes[z_dim]] + list(out.shape[2:])
    dims = [tile_dim] + other + [z_dim, y_dim, x_dim]
    return out.reshape(shape), dims, interp_axis


def staggered_delp(delp, edge, x_dim=FV_CORE_X_CENTER, y_dim=FV_CORE_Y_OUTER, z_dim=RESTART_Z_CENTER):
    """`delp` on the cell edges the winds of `edge` live on (regridz.py:85-100, the same steps as
    coarsen_restarts.py:762-790 compute_edge_delp): interpolated across the six faces along the
    other axis; the staggered dim takes the winds' name and the coordinate 1..n+1 (float32)."""
    hor_dims = {"x": x_dim, "y": y_dim}
    centre = {"x": FV_CORE_X_CENTER, "y": FV_CORE_Y_CENTER}
    lead = [d for d in delp.dims if d not in (z_dim, centre["x"], centre["y"])]
    t, dims, interp_axis = _edge_delp(delp, edge, lead, centre["x"], centre["y"], z_dim)
    rename = {centre[interp_axis]: hor_dims[interp_axis], centre[edge]: hor_dims[edge]}
    dims = [rename.get(d, d) for d in dims]
    n = delp.sizes[centre[edge]]
    coords = OrderedDict((k, c) for k, c in delp.coords.items()
                         if not any(d in (centre["x"], centre["y"]) for d in c.dims))
    if centre[edge] in delp.coords and hor_dims[edge] == centre[edge]:
        coords[centre[edge]] = delp.coords[centre[edge]]
    coords[hor_dims[interp_axis]] = ((hor_dims[interp_axis],), np.arange(1, n + 2, dtype=np.float32))
    return DataArray(lay.finish(t, dims, dims, delp), dims, coords, delp.name, None)


def regrid_to_edge_weighted_pressure(ds, delp, length, coarsening_factor, x_dim=FV_CORE_X_CENTER,
                                     y_dim=FV_CORE_Y_OUTER, z_dim=RESTART_Z_CENTER, edge="x"):
    """Regrid edge-valued quantities (u with dx, v with dy) to the pressure levels of the
    edge-weighted coarse `delp` (regridz.py:58-112).  `delp` is first interpolated to the cell
    edges; its staggered dim takes the name the winds use (`y_dim` for edge="x", `x_dim` for
    edge="y") and the coordinate 1..n+1 (float32)."""
    src = ds
    ds, delp, length = lab.adapt(ds), lab.adapt(delp), lab.adapt(length)
    from .coarsen import edge_weighted_block_average

    delp_staggered = staggered_delp(delp, edge, x_dim=x_dim, y_dim=y_dim, z_dim=z_dim)
    delp_staggered_coarse = edge_weighted_block_average(
        delp_staggered, length, coarsening_factor, x_dim=x_dim, y_dim=y_dim, edge=edge)
    regridded, masked = _regrid_given_delp(ds, delp_staggered, delp_staggered_coarse, length,
                                           x_dim=x_dim, y_dim=y_dim, z_dim=z_dim)
    return lab.restore(regridded, src), lab.restore(masked, src)


# ------------------------------------------------------------------ fused path (a14 + a1, a16)
def pressure_level_average(ds, delp, area, coarsening_factor, x_dim=FV_CORE_X_CENTER,
                           y_dim=FV_CORE_Y_CENTER, z_dim=RESTART_Z_CENTER, mask_below_surface=True,
                           coord_func=coarsen_coords_coord_func):
    """Coarse-grain cell-centred fields on surfaces of constant (coarse) pressure in one pass.

    Equivalent to the reference's two-step recipe

        ds_regrid, masked_area = regrid_to_area_weighted_pressure(ds, delp, area, f, ...)
        coarse = weighted_block_average(ds_regrid, masked_area, f, ...)      # restarts
        coarse = weighted_block_average(ds_regrid, area, f, ...)             # budget

    (coarsen_restarts.py:152-225 with `mask_below_surface=True`; fine_res_budget
    budgets.py:49-57 with False) but nothing fine-grid is materialised: each field is read
    once and only coarse values are written.  float32, iv=1, kord=1 (what those callers use).

    Returns (coarse Dataset or DataArray, coarse delp); attrs, names and the non-horizontal
    coordinates follow `weighted_block_average`."""
    src = ds
    ds, delp, area = lab.adapt(ds), lab.adapt(delp), lab.adapt(area)
    f = int(coarsening_factor)
    for d in (x_dim, y_dim):
        lay.check_exact(delp.sizes[d], f, d)
    lead = [d for d in delp.dims if d not in (z_dim, y_dim, x_dim)]
    dims = tuple(lead) + (z_dim, y_dim, x_dim)
    arrays, names = _as_list(ds)
    df = _canonical4(delp, lead, z_dim, y_dim, x_dim).to(torch.float32)
    n_outer, nz, ny, nx = df.shape
    ta = _weights3(area, delp, lead, y_dim, x_dim).reshape(n_outer, ny, nx).to(torch.float32)
    fields = []
    for da in arrays:
        if set(da.dims) != set(dims):
            raise ValueError(f"variable {da.name!r} has dims {da.dims}; expected {dims} in any order")
        fields.append(_canonical4(da, lead, z_dim, y_dim, x_dim).to(torch.float32))
    outs, dcoarse = ops.pressure_level_average(fields, df, ta, f, mask_below=bool(mask_below_surface),
                                               toa=TOA_PRESSURE)
    lead_shape = [delp.sizes[d] for d in lead]
    shape = lead_shape + [nz, ny // f, nx // f]
    windows = {x_dim: f, y_dim: f}
    result = OrderedDict()
    for da, o in zip(arrays, outs):
        coords = _carry_coords(area, dims, windows, coord_func)
        coords.update(_carry_coords(da, dims, windows, coord_func))
        result[da.name] = DataArray(lay.finish(o.reshape(shape), dims, da.dims, ds), da.dims, coords,
                                    da.name, da.attrs)
    coords = _carry_coords(area, dims, windows, coord_func)
    coords.update(_carry_coords(delp, dims, windows, coord_func))
    delp_coarse = DataArray(lay.finish(dcoarse.reshape(shape), dims, delp.dims, delp), delp.dims,
                            coords, delp.name, delp.attrs)
    coarse = Dataset(result, attrs=ds.attrs) if names is not None else result[arrays[0].name]
    return lab.restore(coarse, src), lab.restore(delp_coarse, src)
