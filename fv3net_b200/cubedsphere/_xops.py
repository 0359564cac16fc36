"""Labelled element-wise steps of the restart recipes (the xarray arithmetic between the block
kernels in coarsen_restarts.py: `delp * area`, `.where(mask)`, `xr.where`, `isclose`, `&`,
`blend`, `.fillna`).  Operands broadcast by dim name like xarray (result dims in order of first
appearance), the arithmetic runs as `cg_elementwise` on the device; masks are 0 / 1 arrays in
the dtype of the data.
"""
from collections import OrderedDict

import numpy as np
import torch

from .. import ops
from ..labeled import DataArray, Dataset
from . import _layout as lay


def _union(objs):
    sizes = OrderedDict()
    for o in objs:
        for d, n in o.sizes.items():
            if d in sizes and sizes[d] != n:
                raise ValueError(f"size of dimension {d!r} differs between operands ({sizes[d]} != {n})")
            sizes.setdefault(d, n)
    return sizes


def _tensors(objs):
    """Operands expanded to the union of their dims, one dtype, contiguous."""
    sizes = _union(objs)
    dims = list(sizes)
    f64 = any(np.dtype(o.dtype) == np.float64 or np.dtype(o.dtype).kind in "iu" for o in objs)
    dt = torch.float64 if f64 else torch.float32
    out = []
    for o in objs:
        if tuple(o.dims) == tuple(dims):
            t = ops.to_device(o.data)
        else:
            t = lay.broadcast_to(o, dims, sizes)
        out.append(t.to(dt).contiguous())
    return out, dims


def _coords(objs, dims):
    coords = OrderedDict()
    for o in objs:
        for k, c in o.coords.items():
            if k not in coords and all(d in dims for d in c.dims):
                coords[k] = c
    return coords


def _name(objs):
    """xarray's result_name: the common name of the named operands, else None."""
    names = {o.name for o in objs if o.name is not None}
    return names.pop() if len(names) == 1 and all(o.name is not None for o in objs) else None


def _apply(op, arrays, scalar=None, name=None, attrs=None):
    tensors, dims = _tensors(arrays)
    a = tensors[0]
    b = tensors[1] if len(tensors) > 1 else scalar
    c = tensors[2] if len(tensors) > 2 else None
    t = ops.elementwise(op, a, b, c)
    device = any(o.is_device for o in arrays)
    data = t if device else t.cpu().numpy()
    return DataArray(data, dims, _coords(arrays, dims), name, attrs)


def mul(a, b):
    """a * b: attrs dropped, the name kept when both operands share it."""
    return _apply("mul", [a, b], name=_name([a, b]))


def sub(a, b):
    return _apply("sub", [a, b], name=_name([a, b]))


def div(a, b):
    return _apply("div", [a, b], name=_name([a, b]))


def isclose(a, b):
    """vcm.xarray_utils.isclose (np.isclose with its defaults); `b` may be a scalar."""
    if isinstance(b, DataArray):
        return _apply("isclose", [a, b], name=_name([a, b]))
    return _apply("isclose", [a], scalar=float(b), name=a.name)


def logical_and(a, b):
    return _apply("and", [a, b], name=_name([a, b]))


def gt(a, b):
    if isinstance(b, DataArray):
        return _apply("gt", [a, b], name=_name([a, b]))
    return _apply("gt", [a], scalar=float(b), name=a.name)


def lt(a, b):
    if isinstance(b, DataArray):
        return _apply("lt", [a, b], name=_name([a, b]))
    return _apply("lt", [a], scalar=float(b), name=a.name)


def where_method(da, cond, other=np.nan):
    """da.where(cond, other): name and attrs of `da` are kept."""
    if isinstance(da, Dataset):
        return Dataset(OrderedDict((k, where_method(da[k], cond, other)) for k in da.data_vars),
                       attrs=da.attrs)
    if isinstance(other, DataArray):
        out = _apply("where", [da, other, cond], name=da.name, attrs=da.attrs)
    else:
        # the scalar stands in for the second array; the condition is the third
        tensors, dims = _tensors([da, cond])
        t = ops.elementwise("where", tensors[0], float(other), tensors[1])
        data = t if (da.is_device or cond.is_device) else t.cpu().numpy()
        out = DataArray(data, dims, _coords([da, cond], dims), da.name, da.attrs)
    return out


def where(cond, x, y):
    """xr.where(cond, x, y): attrs dropped; either branch may be a scalar."""
    arrays = [o for o in (cond, x, y) if isinstance(o, DataArray)]
    sizes = _union(arrays)
    dims = list(sizes)

    def full(v, like):
        if isinstance(v, DataArray):
            return v
        t = torch.full(like.shape, float(v), dtype=like.dtype, device=like.device)
        return DataArray(t, dims)

    tensors, dims = _tensors(arrays)
    template = tensors[0]
    operands = [full(x, template), full(y, template), cond]
    named = [o for o in (cond, x, y) if isinstance(o, DataArray)]
    out = _apply("where", operands, name=_name(named))
    out.coords = _coords(named, out.dims)
    if not any(o.is_device for o in named) and out.is_device:
        out = out.copy(out.data.cpu().numpy())
    return out


def fillna(da, value):
    return _apply("fillna", [da], scalar=float(value), name=da.name, attrs=da.attrs)


def blend(weights, pressure_level, model_level):
    """weights * pressure_level + (1 - weights) * model_level (coarsen_restarts.py:624-639)."""
    if isinstance(pressure_level, Dataset):
        return Dataset(OrderedDict((k, blend(weights, pressure_level[k], model_level[k]))
                                   for k in pressure_level.data_vars))
    return _apply("blend", [weights, pressure_level, model_level])
