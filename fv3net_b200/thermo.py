"""Pressure helpers on the coarse-graining path, with the reference's names and arguments
(external/vcm/vcm/calc/thermo/vertically_dependent.py:41-66, 153-179, 189-208).

`pressure_at_interface` is the sequential top-down cumulative sum from 300 Pa that feeds the
remapper (the order of the additions is part of parity).  All three run as CUDA kernels
(cg_pressure_at_interface, cg_pressure_at_midpoint_log, cg_surface_pressure_nansum); this module only
carries dims, coords and attrs.
"""
from collections import OrderedDict

from . import labeled as lab
from . import ops
from .cubedsphere import _layout as lay
from .cubedsphere.constants import COORD_Z_CENTER, COORD_Z_OUTER, TOA_PRESSURE
from .labeled import DataArray


def _interface_tensor(delp, toa_pressure, dim_center):
    """(tensor with the z axis one longer, dims) - z may sit anywhere in `delp.dims`."""
    axis = delp.get_axis_num(dim_center)
    t = lay.device_tensor(delp)
    shape = list(t.shape)
    n_outer = lay.numel(shape[:axis])
    out = ops.pressure_at_interface(t.reshape(n_outer, shape[axis], lay.numel(shape[axis + 1:])),
                                    toa=toa_pressure)
    shape[axis] += 1
    return out.reshape(shape)


def pressure_at_interface(delp, toa_pressure=TOA_PRESSURE, dim_center=COORD_Z_CENTER,
                          dim_outer=COORD_Z_OUTER):
    """Pressure at layer interfaces: cumsum([toa_pressure, delp...]) along `dim_center`
    (vertically_dependent.py:41-66).  Every coordinate except the vertical one is kept."""
    src = delp
    delp = lab.adapt(delp)
    t = _interface_tensor(delp, float(toa_pressure), dim_center)
    dims = tuple(dim_outer if d == dim_center else d for d in delp.dims)
    coords = OrderedDict((k, c) for k, c in delp.coords.items() if dim_center not in c.dims)
    data = lay.finish(t, dims, dims, delp)
    return lab.restore(DataArray(data, dims, coords, None, None), src)


def pressure_at_midpoint_log(delp, toa_pressure=TOA_PRESSURE, dim=COORD_Z_CENTER):
    """delp / diff(log(p_interface)) (vertically_dependent.py:153-179)."""
    src = delp
    delp = lab.adapt(delp)
    axis = delp.get_axis_num(dim)
    t = lay.device_tensor(delp)
    shape = list(t.shape)
    out = ops.pressure_at_midpoint_log(
        t.reshape(lay.numel(shape[:axis]), shape[axis], lay.numel(shape[axis + 1:])), toa=float(toa_pressure)).reshape(shape)
    coords = OrderedDict(delp.coords)
    return lab.restore(DataArray(lay.finish(out, delp.dims, delp.dims, delp), delp.dims, coords,
                                 delp.name, None), src)


def surface_pressure_from_delp(delp, p_toa=300.0, vertical_dim="z"):
    """sum(delp) + p_toa (vertically_dependent.py:189-208)."""
    src = delp
    delp = lab.adapt(delp)
    axis = delp.get_axis_num(vertical_dim)
    t = lay.device_tensor(delp)
    shape = list(t.shape)
    # xarray's sum skips NaN (skipna defaults to True for floats); additions top-down like numpy's reduction over
    # an outer axis
    out = ops.surface_pressure_nansum(
        t.reshape(lay.numel(shape[:axis]), shape[axis], lay.numel(shape[axis + 1:])), toa=float(p_toa))
    out = out.reshape(shape[:axis] + shape[axis + 1:])
    dims = tuple(d for d in delp.dims if d != vertical_dim)
    coords = OrderedDict((k, c) for k, c in delp.coords.items() if vertical_dim not in c.dims)
    return lab.restore(DataArray(lay.finish(out, dims, dims, delp), dims, coords, delp.name,
                                 {"long_name": "surface pressure", "units": "Pa"}), src)
