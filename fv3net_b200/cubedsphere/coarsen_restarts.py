"""Coarse-graining of complete sets of FV3 restart files: the reference's
`vcm.cubedsphere.coarsen_restarts` recipes (external/vcm/vcm/cubedsphere/coarsen_restarts.py)
over the CUDA block / remap kernels.

Public functions, arguments and returned structure are the reference's:

    coarsen_restarts_on_sigma(coarsening_factor, grid_spec, restarts, coarsen_agrid_winds, mass_weighted)
    coarsen_restarts_on_pressure(coarsening_factor, grid_spec, restarts, coarsen_agrid_winds)
    coarsen_restarts_via_blended_method(coarsening_factor, grid_spec, restarts, coarsen_agrid_winds, mass_weighted)

`restarts` maps the four categories ("fv_core.res", "fv_srf_wnd.res", "fv_tracer.res",
"sfc_data") to Datasets; the result has the same keys, every variable in the dim order of its
input, `coarsening_factor` times coarser.  Every step that touches data is a kernel of
libcoarsen_b200 (block averages / reductions / mode / upsample, the PPM remap, the element-wise
selections of `_xops`, the column thermodynamics of `cg_hydrostatic_balance` and
`cg_blending_weights`); this module only routes variables, like the reference's does.
"""
from collections import OrderedDict

import numpy as np

from .. import labeled as lab
from .. import ops
from ..labeled import DataArray, Dataset, merge
from . import _layout as lay
from . import _xops as X
from .coarsen import (
    block_coarsen,
    block_edge_coarsen,
    block_edge_sum,
    block_median,
    block_upsample_like,
    edge_weighted_block_average,
    weighted_block_average,
)
from .constants import (
    COORD_X_CENTER,
    COORD_X_OUTER,
    COORD_Y_CENTER,
    COORD_Y_OUTER,
    FV_CORE_X_CENTER,
    FV_CORE_X_OUTER,
    FV_CORE_Y_CENTER,
    FV_CORE_Y_OUTER,
    FV_SRF_WND_X_CENTER,
    FV_SRF_WND_Y_CENTER,
    FV_TRACER_X_CENTER,
    FV_TRACER_Y_CENTER,
    RESTART_Z_CENTER,
    SFC_DATA_X_CENTER,
    SFC_DATA_Y_CENTER,
    TOA_PRESSURE,
)
from .regridz import (
    regrid_to_area_weighted_pressure,
    regrid_to_edge_weighted_pressure,
    staggered_delp,
)

FREEZING_TEMPERATURE = 273.16
SHDMIN_THRESHOLD = 0.011
STYPE_LAND_ICE = 16.0
VTYPE_LAND_ICE = 15.0
X_DIM = "xaxis_1"
Y_DIM = "yaxis_1"
SIGMA_BLEND = 0.9
FRACTION_TRACERS = ["cld_amt"]          # always area-weighted
NON_FRACTION_TRACERS = ["sphum", "liq_wat", "rainwat", "ice_wat", "snowwat", "graupel", "o3mr", "sgs_tke"]
CATEGORY_LIST = ["fv_core.res", "fv_srf_wnd.res", "fv_tracer.res", "sfc_data"]


# ------------------------------------------------------------------ entry points (:77-309)
def _grid_renamed(grid_spec, x_center, y_center):
    area = grid_spec.area.rename({COORD_X_CENTER: x_center, COORD_Y_CENTER: y_center})
    return area


def _core_grid(grid_spec):
    area = _grid_renamed(grid_spec, FV_CORE_X_CENTER, FV_CORE_Y_CENTER)
    dx = grid_spec.dx.rename({COORD_X_CENTER: FV_CORE_X_CENTER, COORD_Y_OUTER: FV_CORE_Y_OUTER})
    dy = grid_spec.dy.rename({COORD_X_OUTER: FV_CORE_X_OUTER, COORD_Y_CENTER: FV_CORE_Y_CENTER})
    return area, dx, dy


def _adapt_inputs(grid_spec, restarts):
    return lab.adapt(grid_spec), OrderedDict((k, lab.adapt(v)) for k, v in restarts.items())


def _finish(coarsened, restarts, src_restarts):
    for category in CATEGORY_LIST:
        _sync_dimension_order(coarsened[category], restarts[category])
    return OrderedDict((k, lab.restore(v, src_restarts[k])) for k, v in coarsened.items())


def coarsen_restarts_on_sigma(coarsening_factor, grid_spec, restarts, coarsen_agrid_winds=False,
                              mass_weighted=True):
    """Coarsen a complete set of restart data, averaging on model levels and using the
    'complex' surface coarsening method (coarsen_restarts.py:77-149)."""
    src = restarts
    grid_spec, restarts = _adapt_inputs(grid_spec, restarts)
    area, dx, dy = _core_grid(grid_spec)
    coarsened = OrderedDict()
    core = restarts["fv_core.res"]
    coarsened["fv_core.res"] = _coarse_grain_fv_core(
        core, core.delp, area, dx, dy, coarsening_factor, coarsen_agrid_winds, mass_weighted)
    coarsened["fv_srf_wnd.res"] = _coarse_grain_fv_srf_wnd(
        restarts["fv_srf_wnd.res"], _grid_renamed(grid_spec, FV_SRF_WND_X_CENTER, FV_SRF_WND_Y_CENTER),
        coarsening_factor)
    coarsened["fv_tracer.res"] = _coarse_grain_fv_tracer(
        restarts["fv_tracer.res"], core.delp.rename({FV_CORE_Y_CENTER: FV_TRACER_Y_CENTER}),
        _grid_renamed(grid_spec, FV_TRACER_X_CENTER, FV_TRACER_Y_CENTER), coarsening_factor, mass_weighted)
    coarsened["sfc_data"] = _coarse_grain_sfc_data_complex(
        restarts["sfc_data"], _grid_renamed(grid_spec, SFC_DATA_X_CENTER, SFC_DATA_Y_CENTER),
        coarsening_factor)
    return _finish(coarsened, restarts, src)


def coarsen_restarts_on_pressure(coarsening_factor, grid_spec, restarts, coarsen_agrid_winds=False):
    """Coarsen a complete set of restart files, averaging on pressure levels and using the
    'complex' surface coarsening method (coarsen_restarts.py:152-225)."""
    src = restarts
    grid_spec, restarts = _adapt_inputs(grid_spec, restarts)
    area, dx, dy = _core_grid(grid_spec)
    coarsened = OrderedDict()
    core = restarts["fv_core.res"]
    coarsened["fv_core.res"] = _coarse_grain_fv_core_on_pressure(
        core, core.delp, area, dx, dy, coarsening_factor, coarsen_agrid_winds)
    coarsened["fv_srf_wnd.res"] = _coarse_grain_fv_srf_wnd(
        restarts["fv_srf_wnd.res"], _grid_renamed(grid_spec, FV_SRF_WND_X_CENTER, FV_SRF_WND_Y_CENTER),
        coarsening_factor)
    coarsened["fv_tracer.res"] = _coarse_grain_fv_tracer_on_pressure(
        restarts["fv_tracer.res"], core.delp.rename({FV_CORE_Y_CENTER: FV_TRACER_Y_CENTER}),
        _grid_renamed(grid_spec, FV_TRACER_X_CENTER, FV_TRACER_Y_CENTER), coarsening_factor)
    coarsened["sfc_data"] = _coarse_grain_sfc_data_complex(
        restarts["sfc_data"], _grid_renamed(grid_spec, SFC_DATA_X_CENTER, SFC_DATA_Y_CENTER),
        coarsening_factor)
    coarsened["fv_core.res"] = _impose_hydrostatic_balance(coarsened["fv_core.res"], coarsened["fv_tracer.res"])
    return _finish(coarsened, restarts, src)


def coarsen_restarts_via_blended_method(coarsening_factor, grid_spec, restarts, coarsen_agrid_winds=False,
                                        mass_weighted=True):
    """Coarsen a complete set of restart files with the blended pressure-level / model-level
    method for 3-D fields and the 'complex' surface method (coarsen_restarts.py:228-309)."""
    src = restarts
    grid_spec, restarts = _adapt_inputs(grid_spec, restarts)
    area, dx, dy = _core_grid(grid_spec)
    coarsened = OrderedDict()
    core = restarts["fv_core.res"]
    coarsened["fv_core.res"] = _coarse_grain_fv_core_via_blended_method(
        core, core.delp, area, dx, dy, coarsening_factor, coarsen_agrid_winds, mass_weighted)
    coarsened["fv_srf_wnd.res"] = _coarse_grain_fv_srf_wnd(
        restarts["fv_srf_wnd.res"], _grid_renamed(grid_spec, FV_SRF_WND_X_CENTER, FV_SRF_WND_Y_CENTER),
        coarsening_factor)
    coarsened["fv_tracer.res"] = _coarse_grain_fv_tracer_via_blended_method(
        restarts["fv_tracer.res"], core.delp.rename({FV_CORE_Y_CENTER: FV_TRACER_Y_CENTER}),
        _grid_renamed(grid_spec, FV_TRACER_X_CENTER, FV_TRACER_Y_CENTER), coarsening_factor, mass_weighted)
    coarsened["sfc_data"] = _coarse_grain_sfc_data_complex(
        restarts["sfc_data"], _grid_renamed(grid_spec, SFC_DATA_X_CENTER, SFC_DATA_Y_CENTER),
        coarsening_factor)
    coarsened["fv_core.res"] = _impose_hydrostatic_balance(coarsened["fv_core.res"], coarsened["fv_tracer.res"])
    return _finish(coarsened, restarts, src)


# ------------------------------------------------------------------ fv_core (:316-516)
def _check_agrid_winds(ds, coarsen_agrid_winds):
    if coarsen_agrid_winds and not ("ua" in ds and "va" in ds):
        raise ValueError("If 'coarsen_agrid_winds' is active, 'ua' and 'va' "
                         "must be present in the 'fv_core.res' restart files.")


def _coarse_grain_fv_core(ds, delp, area, dx, dy, coarsening_factor, coarsen_agrid_winds=False,
                          mass_weighted=True):
    """fv_core on model levels (:316-408): phis, delp, DZ area-weighted; W, T (ua, va) by mass
    `delp * area` or area; u along x edges with dx, v along y edges with dy."""
    if mass_weighted:
        area_weighted_vars, mass_weighted_vars = ["phis", "delp", "DZ"], ["W", "T"]
    else:
        area_weighted_vars, mass_weighted_vars = ["phis", "delp", "DZ", "W", "T"], []
    _check_agrid_winds(ds, coarsen_agrid_winds)
    if coarsen_agrid_winds:
        (mass_weighted_vars if mass_weighted else area_weighted_vars).extend(["ua", "va"])
    area_weighted = weighted_block_average(ds[area_weighted_vars], area, coarsening_factor,
                                           x_dim=FV_CORE_X_CENTER, y_dim=FV_CORE_Y_CENTER)
    parts = [area_weighted]
    if mass_weighted_vars:
        mass = X.mul(delp, area)
        parts.append(weighted_block_average(ds[mass_weighted_vars], mass, coarsening_factor,
                                            x_dim=FV_CORE_X_CENTER, y_dim=FV_CORE_Y_CENTER))
    parts.append(edge_weighted_block_average(ds[["u"]], dx, coarsening_factor, x_dim=FV_CORE_X_CENTER,
                                             y_dim=FV_CORE_Y_OUTER, edge="x"))
    parts.append(edge_weighted_block_average(ds[["v"]], dy, coarsening_factor, x_dim=FV_CORE_X_OUTER,
                                             y_dim=FV_CORE_Y_CENTER, edge="y"))
    return merge(parts)


def _coarse_grain_fv_core_on_pressure(ds, delp, area, dx, dy, coarsening_factor, coarsen_agrid_winds=False):
    """fv_core on surfaces of constant pressure (:411-516), except delp, DZ and phis which stay
    on model surfaces."""
    area_weighted_vars = ["phis", "delp", "DZ"]
    masked_area_weighted_vars = ["W", "T"]
    _check_agrid_winds(ds, coarsen_agrid_winds)
    if coarsen_agrid_winds:
        masked_area_weighted_vars.extend(["ua", "va"])
    area_pressure_regridded, masked_area = regrid_to_area_weighted_pressure(
        ds[masked_area_weighted_vars], delp, area, coarsening_factor,
        x_dim=FV_CORE_X_CENTER, y_dim=FV_CORE_Y_CENTER)
    dx_pressure_regridded, masked_dx = regrid_to_edge_weighted_pressure(
        ds[["u"]], delp, dx, coarsening_factor, x_dim=FV_CORE_X_CENTER, y_dim=FV_CORE_Y_OUTER, edge="x")
    dy_pressure_regridded, masked_dy = regrid_to_edge_weighted_pressure(
        ds[["v"]], delp, dy, coarsening_factor, x_dim=FV_CORE_X_OUTER, y_dim=FV_CORE_Y_CENTER, edge="y")
    area_weighted = weighted_block_average(ds[area_weighted_vars], area, coarsening_factor,
                                           x_dim=FV_CORE_X_CENTER, y_dim=FV_CORE_Y_CENTER)
    masked_area_weighted = weighted_block_average(area_pressure_regridded, masked_area, coarsening_factor,
                                                  x_dim=FV_CORE_X_CENTER, y_dim=FV_CORE_Y_CENTER)
    edge_weighted_x = edge_weighted_block_average(dx_pressure_regridded, masked_dx, coarsening_factor,
                                                  x_dim=FV_CORE_X_CENTER, y_dim=FV_CORE_Y_OUTER, edge="x")
    edge_weighted_y = edge_weighted_block_average(dy_pressure_regridded, masked_dy, coarsening_factor,
                                                  x_dim=FV_CORE_X_OUTER, y_dim=FV_CORE_Y_CENTER, edge="y")
    return merge([area_weighted, masked_area_weighted, edge_weighted_x, edge_weighted_y])


# ------------------------------------------------------------------ blending (:519-759)
def _columns(da, z_dim, hor):
    """(tensor (n_outer, nz, h...), lead dims) with z ahead of the horizontal dims."""
    lead = [d for d in da.dims if d != z_dim and d not in hor]
    order = lead + ([z_dim] if z_dim in da.dims else []) + list(hor)
    t = lay.device_tensor(da, order)
    n_outer = lay.numel([da.sizes[d] for d in lead])
    return t.reshape((n_outer,) + tuple(t.shape[len(lead):])), lead, order


def compute_blending_weights(blending_pressure, ps_coarse, pfull_coarse):
    """(ps - pfull) / (ps - blending_pressure) below the blending pressure, 1 above (:519-536)."""
    b, ps, pf = lab.adapt(blending_pressure), lab.adapt(ps_coarse), lab.adapt(pfull_coarse)
    ratio = X.div(X.sub(ps, pf), X.sub(ps, b))
    return lab.restore(X.where(X.gt(pf, b), ratio, 1.0), pfull_coarse)


def _blending_weights_from(delp_coarse, ps_fine_min, z_dim, hor, toa_pressure):
    """Weights of one grid (cell centres or one family of edges) from its coarse delp and the
    block minimum of the fine surface pressure: `cg_blending_weights`."""
    td, lead, order = _columns(delp_coarse, z_dim, hor)
    tp, lead_p, _ = _columns(ps_fine_min, z_dim, hor)
    if lead_p != lead:
        tp = lay.device_tensor(ps_fine_min, lead + list(hor)).reshape((td.shape[0],) + tuple(td.shape[2:]))
    tp = tp.to(td.dtype)
    w = ops.blending_weights(td, tp, toa=toa_pressure, sigma=SIGMA_BLEND)
    shape = [delp_coarse.sizes[d] for d in order]
    coords = OrderedDict(delp_coarse.coords)
    data = lay.finish(w.reshape(shape), order, delp_coarse.dims, delp_coarse)
    return DataArray(data, delp_coarse.dims, coords, None, None)


def _surface_pressure(delp, z_dim, toa_pressure):
    """surface_pressure_from_delp (vertically_dependent.py:189-208) on the device."""
    hor = [d for d in delp.dims if d != z_dim][-2:]
    t, lead, order = _columns(delp, z_dim, hor)
    ps = ops.surface_pressure(t, toa=toa_pressure)
    dims = lead + hor
    shape = [delp.sizes[d] for d in dims]
    coords = OrderedDict((k, c) for k, c in delp.coords.items() if z_dim not in c.dims)
    out_dims = [d for d in delp.dims if d != z_dim]
    data = lay.finish(ps.reshape(shape), dims, out_dims, delp)
    return DataArray(data, out_dims, coords, delp.name, {"long_name": "surface pressure", "units": "Pa"})


def _compute_blending_weights_agrid(delp, area, coarsening_factor, x_dim=FV_CORE_X_CENTER,
                                    y_dim=FV_CORE_Y_CENTER, toa_pressure=TOA_PRESSURE):
    """Blending weights on the A-grid (:539-581): the blending pressure is 0.9 times the
    minimum fine surface pressure inside a coarse cell."""
    delp_coarse = weighted_block_average(delp, area, coarsening_factor, x_dim=x_dim, y_dim=y_dim)
    ps = _surface_pressure(delp, RESTART_Z_CENTER, toa_pressure)
    ps_min = block_coarsen(ps, coarsening_factor, x_dim=x_dim, y_dim=y_dim, method="min")
    return _blending_weights_from(delp_coarse, ps_min, RESTART_Z_CENTER, [y_dim, x_dim], toa_pressure)


def compute_edge_delp(delp, edge, x_dim=FV_CORE_X_CENTER, y_dim=FV_CORE_Y_CENTER):
    """Pressure thickness on grid-cell edges (:762-790)."""
    return lab.restore(staggered_delp(lab.adapt(delp), edge, x_dim=x_dim, y_dim=y_dim), delp)


def _compute_blending_weights_dgrid(delp, length, coarsening_factor, edge, x_dim, y_dim,
                                    toa_pressure=TOA_PRESSURE):
    """The same on cell edges (:584-621): minimum surface pressure along the coarse cell's edge."""
    delp_edge = staggered_delp(delp, edge, x_dim=x_dim, y_dim=y_dim)
    delp_edge_coarse = edge_weighted_block_average(delp_edge, length, coarsening_factor, x_dim=x_dim,
                                                   y_dim=y_dim, edge=edge)
    ps = _surface_pressure(delp_edge, RESTART_Z_CENTER, toa_pressure)
    ps_min = block_edge_coarsen(ps, coarsening_factor, edge=edge, x_dim=x_dim, y_dim=y_dim, method="min")
    return _blending_weights_from(delp_edge_coarse, ps_min, RESTART_Z_CENTER, [y_dim, x_dim], toa_pressure)


def blend(weights, pressure_level, model_level):
    """weights * pressure_level + (1 - weights) * model_level (:624-639)."""
    return X.blend(weights, pressure_level, model_level)


def _2d_fv_core_names_agrid(ds):
    return [v for v in ds.data_vars if RESTART_Z_CENTER not in ds[v].dims]


def _3d_fv_core_names_agrid(ds, coarsen_agrid_winds):
    ignore_vars = ["u", "v"] + ([] if coarsen_agrid_winds else ["ua", "va"])
    return [v for v in ds.data_vars if v not in ignore_vars and RESTART_Z_CENTER in ds[v].dims]


def _coarse_grain_fv_core_via_blended_method(ds, delp, area, dx, dy, coarsening_factor,
                                             coarsen_agrid_winds=False, mass_weighted=True):
    """:657-719."""
    pressure_level = _coarse_grain_fv_core_on_pressure(ds, delp, area, dx, dy, coarsening_factor,
                                                       coarsen_agrid_winds)
    model_level = _coarse_grain_fv_core(ds, delp, area, dx, dy, coarsening_factor, coarsen_agrid_winds,
                                        mass_weighted)
    weights_agrid = _compute_blending_weights_agrid(delp, area, coarsening_factor,
                                                    x_dim=FV_CORE_X_CENTER, y_dim=FV_CORE_Y_CENTER)
    weights_dgrid_u = _compute_blending_weights_dgrid(delp, dx, coarsening_factor, "x",
                                                      x_dim=FV_CORE_X_CENTER, y_dim=FV_CORE_Y_OUTER)
    weights_dgrid_v = _compute_blending_weights_dgrid(delp, dy, coarsening_factor, "y",
                                                      x_dim=FV_CORE_X_OUTER, y_dim=FV_CORE_Y_CENTER)
    names_2d_agrid = _2d_fv_core_names_agrid(ds)
    names_3d_agrid = _3d_fv_core_names_agrid(ds, coarsen_agrid_winds)
    fields_2d_agrid = model_level[names_2d_agrid]
    fields_3d_agrid = blend(weights_agrid, pressure_level[names_3d_agrid], model_level[names_3d_agrid])
    u = blend(weights_dgrid_u, pressure_level.u, model_level.u).rename("u")
    v = blend(weights_dgrid_v, pressure_level.v, model_level.v).rename("v")
    return merge([fields_2d_agrid, fields_3d_agrid, u, v])


def _coarse_grain_fv_tracer_via_blended_method(ds, delp, area, coarsening_factor, mass_weighted=True):
    """:722-759."""
    pressure_level = _coarse_grain_fv_tracer_on_pressure(ds, delp, area, coarsening_factor)
    model_level = _coarse_grain_fv_tracer(ds, delp, area, coarsening_factor, mass_weighted)
    weights = _compute_blending_weights_agrid(delp, area, coarsening_factor, x_dim=FV_TRACER_X_CENTER,
                                              y_dim=FV_TRACER_Y_CENTER)
    return blend(weights, pressure_level, model_level)


# ------------------------------------------------------------------ tracers, surface winds (:793-913)
def _coarse_grain_fv_tracer(ds, delp, area, coarsening_factor, mass_weighted=True):
    """:793-837: cld_amt by area; the mixing ratios and TKE by mass or area."""
    if mass_weighted:
        area_weighted_vars, mass_weighted_vars = FRACTION_TRACERS, NON_FRACTION_TRACERS
    else:
        area_weighted_vars, mass_weighted_vars = FRACTION_TRACERS + NON_FRACTION_TRACERS, []
    parts = [weighted_block_average(ds[area_weighted_vars], area, coarsening_factor,
                                    x_dim=FV_TRACER_X_CENTER, y_dim=FV_TRACER_Y_CENTER)]
    if mass_weighted_vars:
        mass = X.mul(delp, area)
        parts.append(weighted_block_average(ds[mass_weighted_vars], mass, coarsening_factor,
                                            x_dim=FV_TRACER_X_CENTER, y_dim=FV_TRACER_Y_CENTER))
    return merge(parts)


def _coarse_grain_fv_tracer_on_pressure(ds, delp, area, coarsening_factor):
    """:840-887: every tracer remapped to the coarse pressures, masked-area average."""
    ds_regridded, masked_area = regrid_to_area_weighted_pressure(
        ds, delp, area, coarsening_factor, x_dim=FV_TRACER_X_CENTER, y_dim=FV_TRACER_Y_CENTER)
    area_weighted = weighted_block_average(ds_regridded[FRACTION_TRACERS], masked_area, coarsening_factor,
                                           x_dim=FV_TRACER_X_CENTER, y_dim=FV_TRACER_Y_CENTER)
    masked_area_weighted = weighted_block_average(ds_regridded[NON_FRACTION_TRACERS], masked_area,
                                                  coarsening_factor, x_dim=FV_TRACER_X_CENTER,
                                                  y_dim=FV_TRACER_Y_CENTER)
    return merge([area_weighted, masked_area_weighted])


def _coarse_grain_fv_srf_wnd(ds, area, coarsening_factor):
    """:890-913."""
    return weighted_block_average(ds[["u_srf", "v_srf"]], area, coarsening_factor, x_dim="xaxis_1",
                                  y_dim="yaxis_1")


# ------------------------------------------------------------------ hydrostatic balance (:916-938)
def _impose_hydrostatic_balance(ds_fv_core, ds_fv_tracer, dim=RESTART_Z_CENTER):
    """Layer thicknesses from hydrostatic balance, surface geopotential adjusted so that the
    model-top height is kept (:916-938; height_at_interface, hydrostatic_dz and
    dz_and_top_to_phis of vcm/calc/thermo/vertically_dependent.py in one kernel)."""
    dz = ds_fv_core["DZ"]
    hor = [FV_CORE_Y_CENTER, FV_CORE_X_CENTER]
    sphum = ds_fv_tracer["sphum"].rename({FV_TRACER_Y_CENTER: FV_CORE_Y_CENTER})
    t_dz, lead, order = _columns(dz, dim, hor)
    dt = t_dz.dtype
    same = lambda da: lay.device_tensor(da, order).reshape(t_dz.shape).to(dt)        # noqa: E731
    t_phis = lay.device_tensor(ds_fv_core["phis"], lead + hor).reshape(
        (t_dz.shape[0],) + tuple(t_dz.shape[2:])).to(dt)
    dz_new, phis_new = ops.hydrostatic_balance(t_dz, t_phis, same(ds_fv_core["T"]), same(sphum),
                                               same(ds_fv_core["delp"]), toa=TOA_PRESSURE)
    shape = [dz.sizes[d] for d in order]
    coords3 = OrderedDict(dz.coords)
    phis = ds_fv_core["phis"]
    out = Dataset(attrs=ds_fv_core.attrs)
    for name in ds_fv_core.data_vars:
        if name == "DZ":
            out[name] = DataArray(lay.finish(dz_new.reshape(shape), order, dz.dims, dz), dz.dims, coords3,
                                  name, None)
        elif name == "phis":
            pshape = [phis.sizes[d] for d in lead + hor]
            out[name] = DataArray(lay.finish(phis_new.reshape(pshape), lead + hor, phis.dims, phis), phis.dims,
                                  OrderedDict(phis.coords), name, None)
        else:
            out[name] = ds_fv_core[name]
    return out


# ------------------------------------------------------------------ sfc_data (:941-1421)
def _integerize(x):
    return np.round(x).astype(x.dtype)


def _coarse_grain_sfc_data(ds, area, coarsening_factor, version="simple"):
    """:941-969: 'simple' = block median of every variable (slmsk rounded), 'complex' below."""
    if version == "simple":
        result = block_median(ds, coarsening_factor, x_dim="xaxis_1", y_dim="yaxis_1")
        slmsk = result["slmsk"]
        result["slmsk"] = slmsk.copy(_integerize(slmsk.values))
        return result
    elif version == "complex":
        return _coarse_grain_sfc_data_complex(ds, area, coarsening_factor)
    raise ValueError("Currently the only supported versions are 'simple' and 'complex'. Got {version}.")


def _coarse_grain_grid_spec(ds, coarsening_factor, x_dim_unstaggered=COORD_X_CENTER,
                            y_dim_unstaggered=COORD_Y_CENTER, x_dim_staggered=COORD_X_OUTER,
                            y_dim_staggered=COORD_Y_OUTER):
    """:972-996: dx, dy summed along their edges, area summed over blocks."""
    ds = lab.adapt(ds)
    coarse_dx = block_edge_sum(ds.dx, coarsening_factor, x_dim_unstaggered, y_dim_staggered, "x")
    coarse_dy = block_edge_sum(ds.dy, coarsening_factor, x_dim_staggered, y_dim_unstaggered, "y")
    coarse_area = block_coarsen(ds.area, coarsening_factor, x_dim_unstaggered, y_dim_unstaggered)
    return merge([coarse_dx.rename("dx"), coarse_dy.rename("dy"), coarse_area.rename("area")])


def _sync_dimension_order(a, b):
    """:999-1002."""
    for var in list(a.data_vars):
        a[var] = a[var].transpose(*b[var].dims)
    return a


def _is_float(x):
    return np.dtype(x.dtype) in (np.dtype("float32"), np.dtype("float64"))


def _doubles_to_floats(ds):
    """:1015-1029: float data and float coordinates to float32; the variables' own
    coordinates are dropped, non-float dataset coordinates (tile) are not carried."""
    coords = OrderedDict()
    for key, coord in ds.coords.items():
        if _is_float(coord):
            coords[key] = DataArray(coord.values.astype(np.float32), coord.dims, None, key, coord.attrs)
    data_vars = OrderedDict()
    for key in ds.data_vars:
        var = ds.data_vars[key]
        if _is_float(var):
            data_vars[key] = DataArray(var.astype(np.float32).data, var.dims, None, key, var.attrs)
    out = Dataset(data_vars, attrs=None)
    for key, c in coords.items():
        if all(d in out.sizes for d in c.dims):
            out.coords[key] = c
    return out


def _coarse_grain_sfc_data_complex(ds, area, coarsening_factor):
    """The 'complicated' surface method (:1032-1080): dominant surface / vegetation / soil
    types by block mode, each variable by the rule of SFC_DATA_COARSENING_METHOD, then the
    surface_chgres corrections."""
    precomputed_arguments = _compute_arguments_for_complex_sfc_coarsening(ds, coarsening_factor)
    result = Dataset()
    result["slmsk"] = precomputed_arguments["coarsened_slmsk"]
    result["vtype"] = precomputed_arguments["coarsened_vtype"]
    result["stype"] = precomputed_arguments["coarsened_stype"]
    for data_var in ds:
        if data_var not in result.data_vars:
            result[data_var] = SFC_DATA_COARSENING_METHOD[data_var](
                data_var=ds[data_var],
                coarsening_factor=coarsening_factor,
                area=area,
                is_dominant_surface_type=precomputed_arguments["is_dominant_surface_type"],
                is_dominant_vtype=precomputed_arguments["is_dominant_vtype"],
                is_dominant_stype=precomputed_arguments["is_dominant_stype"],
                vfrac=ds.vfrac,
                sncovr=ds.sncovr,
                fice=ds.fice,
                coarsened_slmsk=precomputed_arguments["coarsened_slmsk"],
            )
    result = _apply_surface_chgres_corrections(result)
    return _doubles_to_floats(result)


def _mode_omit(obj, coarsening_factor):
    return block_coarsen(obj, coarsening_factor, x_dim=X_DIM, y_dim=Y_DIM, method="mode",
                         func_kwargs={"nan_policy": "omit"})


def _compute_arguments_for_complex_sfc_coarsening(ds, coarsening_factor):
    """:1083-1120."""
    coarsened_slmsk = _mode_omit(ds.slmsk, coarsening_factor)
    upsampled_slmsk = block_upsample_like(coarsened_slmsk, ds.slmsk)
    is_dominant_surface_type = X.isclose(ds.slmsk, upsampled_slmsk)
    coarsened_vtype_and_stype = _mode_omit(X.where_method(ds[["vtype", "stype"]], is_dominant_surface_type),
                                           coarsening_factor)
    upsampled_vtype = block_upsample_like(coarsened_vtype_and_stype.vtype, ds.vtype)
    is_dominant_vtype = X.isclose(ds.vtype, upsampled_vtype)
    upsampled_stype = block_upsample_like(coarsened_vtype_and_stype.stype, ds.stype)
    is_dominant_stype = X.isclose(ds.stype, upsampled_stype)
    return {
        "coarsened_slmsk": coarsened_slmsk,
        "coarsened_vtype": coarsened_vtype_and_stype.vtype,
        "coarsened_stype": coarsened_vtype_and_stype.stype,
        "is_dominant_surface_type": is_dominant_surface_type,
        "is_dominant_vtype": is_dominant_vtype,
        "is_dominant_stype": is_dominant_stype,
    }


def _wba(data_var, weights, coarsening_factor):
    return weighted_block_average(data_var, weights, coarsening_factor, x_dim=X_DIM, y_dim=Y_DIM)


def _area_weighted_mean(data_var, coarsening_factor, area, **unused_kwargs):
    return _wba(data_var, area, coarsening_factor)


def _area_weighted_mean_over_dominant_sfc_type(data_var, coarsening_factor, area, is_dominant_surface_type,
                                               **unused_kwargs):
    return _wba(X.where_method(data_var, is_dominant_surface_type),
                X.where_method(area, is_dominant_surface_type), coarsening_factor)


def _area_and_vfrac_weighted_mean_over_dominant_sfc_and_vtype(data_var, coarsening_factor, area,
                                                              is_dominant_surface_type, is_dominant_vtype,
                                                              vfrac, **unused_kwargs):
    """:1147-1182: vfrac-weighted where the block has vegetation at all, else area-weighted."""
    mask = X.logical_and(is_dominant_surface_type, is_dominant_vtype)
    masked = X.where_method(data_var, mask)
    area_weighted_mean = _wba(masked, X.where_method(area, mask), coarsening_factor)
    area_times_vfrac = X.where_method(X.mul(area, vfrac), mask)
    area_and_vfrac_weighted_mean = _wba(masked, area_times_vfrac, coarsening_factor)
    coarsened_area_times_vfrac = block_coarsen(area_times_vfrac, coarsening_factor, x_dim=X_DIM, y_dim=Y_DIM,
                                               method="sum")
    return X.where(X.gt(coarsened_area_times_vfrac, 0.0), area_and_vfrac_weighted_mean, area_weighted_mean)


def _area_weighted_mean_over_dominant_sfc_and_stype(data_var, coarsening_factor, area,
                                                    is_dominant_surface_type, is_dominant_stype,
                                                    **unused_kwargs):
    mask = X.logical_and(is_dominant_surface_type, is_dominant_stype)
    return _wba(X.where_method(data_var, mask), X.where_method(area, mask), coarsening_factor)


def _mode(data_var, coarsening_factor, **unused_kwargs):
    return _mode_omit(data_var, coarsening_factor)


def _mode_over_dominant_sfc_type(data_var, coarsening_factor, area, is_dominant_surface_type, **unused_kwargs):
    return _mode_omit(X.where_method(data_var, is_dominant_surface_type), coarsening_factor)


def _area_and_sncovr_weighted_mean(data_var, coarsening_factor, area, sncovr, **unused_kwargs):
    return X.fillna(_wba(data_var, X.mul(area, sncovr), coarsening_factor), 0.0)


def _area_and_fice_weighted_mean(data_var, coarsening_factor, area, fice, **unused_kwargs):
    return X.fillna(_wba(data_var, X.mul(area, fice), coarsening_factor), 0.0)


def _minimum_over_dominant_sfc_type(data_var, coarsening_factor, is_dominant_surface_type, **unused_kwargs):
    return block_coarsen(X.where_method(data_var, is_dominant_surface_type), coarsening_factor, x_dim=X_DIM,
                         y_dim=Y_DIM, method="min")


def _maximum_over_dominant_sfc_type(data_var, coarsening_factor, is_dominant_surface_type, **unused_kwargs):
    return block_coarsen(X.where_method(data_var, is_dominant_surface_type), coarsening_factor, x_dim=X_DIM,
                         y_dim=Y_DIM, method="max")


def _area_or_area_and_fice_weighted_mean(data_var, coarsening_factor, area, is_dominant_surface_type, fice,
                                         coarsened_slmsk, **unused_kwargs):
    """tisfc (:1287-1316): area- and ice-fraction-weighted over sea ice, area-weighted elsewhere."""
    masked = X.where_method(data_var, is_dominant_surface_type)
    sea_ice = _wba(masked, X.where_method(X.mul(area, fice), is_dominant_surface_type), coarsening_factor)
    land_or_ocean = _wba(masked, X.where_method(area, is_dominant_surface_type), coarsening_factor)
    return X.where(X.isclose(coarsened_slmsk, 2.0), sea_ice, land_or_ocean)


SFC_DATA_COARSENING_METHOD = {
    "tsea": _area_weighted_mean,
    "alvsf": _area_weighted_mean,
    "alvwf": _area_weighted_mean,
    "alnsf": _area_weighted_mean,
    "alnwf": _area_weighted_mean,
    "facsf": _area_weighted_mean,
    "facwf": _area_weighted_mean,
    "f10m": _area_weighted_mean,
    "t2m": _area_weighted_mean,
    "q2m": _area_weighted_mean,
    "uustar": _area_weighted_mean,
    "ffmm": _area_weighted_mean,
    "ffhh": _area_weighted_mean,
    "tprcp": _area_weighted_mean,
    "snwdph": _area_weighted_mean,
    "tg3": _area_weighted_mean_over_dominant_sfc_type,
    "vfrac": _area_weighted_mean_over_dominant_sfc_type,
    "fice": _area_weighted_mean_over_dominant_sfc_type,
    "sncovr": _area_weighted_mean_over_dominant_sfc_type,
    "canopy": _area_and_vfrac_weighted_mean_over_dominant_sfc_and_vtype,
    "zorl": _area_and_vfrac_weighted_mean_over_dominant_sfc_and_vtype,
    "smc": _area_weighted_mean_over_dominant_sfc_and_stype,
    "slc": _area_weighted_mean_over_dominant_sfc_and_stype,
    "stc": _area_weighted_mean_over_dominant_sfc_and_stype,
    "srflag": _mode,
    "slope": _mode_over_dominant_sfc_type,
    "sheleg": _area_and_sncovr_weighted_mean,
    "hice": _area_and_fice_weighted_mean,
    "shdmin": _minimum_over_dominant_sfc_type,
    "shdmax": _maximum_over_dominant_sfc_type,
    "snoalb": _maximum_over_dominant_sfc_type,
    "tisfc": _area_or_area_and_fice_weighted_mean,
}


def _clip_tsea_and_t3g_at_freezing_over_ice(ds):
    """surface_chgres step 1 (:1355-1375): tsea and tg3 no warmer than 273.16 K over land ice."""
    is_land_ice = X.isclose(ds.vtype, VTYPE_LAND_ICE)
    for name in ("tsea", "tg3"):
        v = ds[name]
        clipped = X.where_method(v, X.lt(v, FREEZING_TEMPERATURE), other=FREEZING_TEMPERATURE)
        ds[name] = X.where(is_land_ice, clipped, v)
    return ds


def _ensure_stype_is_ice_if_vtype_is_ice(ds):
    """Step 2 (:1378-1386)."""
    ds["stype"] = X.where(X.isclose(ds.vtype, VTYPE_LAND_ICE), STYPE_LAND_ICE, ds.stype)
    return ds


def _zero_canopy_moisture_content_over_bare_land(ds):
    """Step 3 (:1389-1397): no vegetation (shdmin < 0.011) -> canopy = 0."""
    ds["canopy"] = X.where(X.lt(ds.shdmin, SHDMIN_THRESHOLD), 0.0, ds.canopy)
    return ds


def _zero_shdmin_over_land_ice(ds):
    """Step 4 (:1400-1408)."""
    ds["shdmin"] = X.where(X.isclose(ds.vtype, VTYPE_LAND_ICE), 0.0, ds.shdmin)
    return ds


def _apply_surface_chgres_corrections(ds):
    """:1411-1421."""
    ds = _clip_tsea_and_t3g_at_freezing_over_ice(ds)
    ds = _ensure_stype_is_ice_if_vtype_is_ice(ds)
    ds = _zero_canopy_moisture_content_over_bare_land(ds)
    return _zero_shdmin_over_land_ice(ds)
