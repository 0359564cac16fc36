"""Minimal labelled arrays for the reference-facing API.

The reference's functions take and return xarray objects.  xarray is not a dependency
of this package (and is absent from the build image), so the API layer works on the
small `DataArray` / `Dataset` below and adapts real xarray objects at the boundary
(`adapt` / `restore`): xarray in -> xarray out, dims / coords / attrs / name carried the
way the reference's functions carry them.  `data` may be a numpy array or a torch CUDA
tensor (a device-resident field that is chained through several calls without leaving
HBM).
"""
from collections import OrderedDict

import numpy as np

try:  # (tests/test_xarray_boundary.py drives the boundary with stand-ins where xarray is absent)
    import xarray as _xr
except Exception:  # noqa: BLE001
    _xr = None


def _is_tensor(x):
    return type(x).__module__.startswith("torch") and hasattr(x, "data_ptr")


class DataArray:
    def __init__(self, data, dims=None, coords=None, name=None, attrs=None):
        if not _is_tensor(data):
            data = np.asarray(data)
        self.data = data
        if dims is None:
            dims = tuple(f"dim_{i}" for i in range(data.ndim))
        self.dims = tuple(dims)
        if len(self.dims) != data.ndim:
            raise ValueError(f"dims {self.dims} do not match data of rank {data.ndim}")
        self.coords = OrderedDict()
        if coords is not None:
            if not hasattr(coords, "keys"):           # xarray-style list of 1-D coords per dim
                coords = {d: c for d, c in zip(self.dims, coords)}
            for k, v in coords.items():
                self.coords[k] = _as_coord(k, v)
        self.name = name
        self.attrs = dict(attrs) if attrs else {}

    # ---- basic properties
    @property
    def shape(self):
        return tuple(self.data.shape)

    @property
    def ndim(self):
        return len(self.dims)

    @property
    def sizes(self):
        return OrderedDict(zip(self.dims, self.shape))

    @property
    def dtype(self):
        if _is_tensor(self.data):
            return np.dtype(str(self.data.dtype).replace("torch.", ""))
        return self.data.dtype

    @property
    def values(self):
        return self.data.detach().cpu().numpy() if _is_tensor(self.data) else self.data

    @property
    def is_device(self):
        return _is_tensor(self.data)

    def __getitem__(self, key):
        if isinstance(key, str):
            if key in self.coords:
                return self.coords[key]
            if key in self.dims:                       # default integer index, like xarray
                return DataArray(np.arange(self.sizes[key]), (key,), name=key)
            raise KeyError(key)
        raise TypeError("positional indexing is not supported; use isel")

    def __setitem__(self, key, value):
        self.coords[key] = _as_coord(key, value)

    def get_axis_num(self, dim):
        return self.dims.index(dim)

    # ---- transformations (all return new objects; inputs are never modified)
    def copy(self, data=None):
        d = self.data if data is None else data
        return DataArray(d, self.dims, self.coords, self.name, self.attrs)

    def rename(self, new=None):
        if new is None or isinstance(new, str):
            out = self.copy()
            out.name = new
            return out
        dims = tuple(new.get(d, d) for d in self.dims)
        coords = OrderedDict()
        for k, c in self.coords.items():
            nk = new.get(k, k)
            coords[nk] = DataArray(c.data, tuple(new.get(d, d) for d in c.dims), None, nk, c.attrs)
        return DataArray(self.data, dims, coords, self.name, self.attrs)

    def transpose(self, *dims):
        if tuple(dims) == self.dims:
            return self
        perm = [self.dims.index(d) for d in dims]
        data = self.data.permute(*perm) if _is_tensor(self.data) else np.transpose(self.data, perm)
        return DataArray(data, dims, self.coords, self.name, self.attrs)

    def isel(self, indexers=None, **kw):
        indexers = dict(indexers or {}, **kw)
        idx, dims = [], []
        for d in self.dims:
            sel = indexers.get(d, slice(None))
            idx.append(sel)
            if not isinstance(sel, (int, np.integer)):
                dims.append(d)
        data = self.data[tuple(idx)]
        coords = OrderedDict()
        for k, c in self.coords.items():
            if c.ndim == 1 and c.dims[0] in indexers:
                sel = indexers[c.dims[0]]
                if isinstance(sel, (int, np.integer)):
                    continue
                coords[k] = DataArray(c.data[sel], c.dims, None, k, c.attrs)
            else:
                coords[k] = c
        return DataArray(data, dims, coords, self.name, self.attrs)

    def assign_coords(self, coords=None, **kw):
        out = self.copy()
        for k, v in dict(coords or {}, **kw).items():
            out.coords[k] = _as_coord(k, v)
        return out

    def assign_attrs(self, *args, **kw):
        out = self.copy()
        for a in args:
            out.attrs.update(a)
        out.attrs.update(kw)
        return out

    def drop_vars(self, names, errors="raise"):
        names = [names] if isinstance(names, str) else list(names)
        out = self.copy()
        for n in names:
            if n in out.coords:
                del out.coords[n]
            elif errors == "raise":
                raise ValueError(f"{n} is not a coordinate")
        return out

    drop = drop_vars

    def astype(self, dtype):
        if _is_tensor(self.data):
            import torch

            return self.copy(self.data.to(getattr(torch, np.dtype(dtype).name)))
        return self.copy(self.data.astype(dtype))

    def to_dataset(self, name=None):
        name = name or self.name
        if name is None:
            raise ValueError("unable to convert unnamed DataArray to a Dataset")
        return Dataset({name: self})

    def compute(self):
        return self

    def __repr__(self):
        where = "device" if self.is_device else "host"
        return f"<fv3net_b200.DataArray {self.name!r} {dict(self.sizes)} {self.dtype} on {where}>"


def _as_coord(name, v):
    if isinstance(v, DataArray):
        return DataArray(v.values, v.dims, None, name, v.attrs)
    if isinstance(v, tuple) and len(v) in (2, 3) and isinstance(v[0], (str, tuple, list)):
        dims = (v[0],) if isinstance(v[0], str) else tuple(v[0])
        return DataArray(np.asarray(v[1]), dims, None, name, v[2] if len(v) == 3 else None)
    a = np.asarray(v)
    return DataArray(a, (name,) if a.ndim == 1 else (), None, name)


class Dataset:
    def __init__(self, data_vars=None, coords=None, attrs=None):
        self.data_vars = OrderedDict()
        self.coords = OrderedDict()
        self.attrs = dict(attrs) if attrs else {}
        for k, c in (coords or {}).items():
            self.coords[k] = _as_coord(k, c)
        for k, v in (data_vars or {}).items():
            self[k] = v

    def __setitem__(self, key, value):
        if not isinstance(value, DataArray):
            dims, data = value[0], value[1]
            value = DataArray(data, dims, attrs=value[2] if len(value) > 2 else None)
        value = value.rename(key) if value.name != key else value
        for ck, c in value.coords.items():
            self.coords.setdefault(ck, c)
        self.data_vars[key] = value

    def __getitem__(self, key):
        if isinstance(key, (list, tuple)):
            return Dataset({k: self[k] for k in key}, self.coords, self.attrs)
        if key in self.data_vars:
            v = self.data_vars[key]
            coords = OrderedDict((k, c) for k, c in self.coords.items()
                                 if all(d in v.dims for d in c.dims))
            coords.update(v.coords)
            return DataArray(v.data, v.dims, coords, key, v.attrs)
        if key in self.coords:
            return self.coords[key]
        raise KeyError(key)

    def __getattr__(self, key):
        if key.startswith("_") or key in ("data_vars", "coords", "attrs"):
            raise AttributeError(key)
        try:
            return self[key]
        except KeyError:
            raise AttributeError(key) from None

    def __iter__(self):
        return iter(self.data_vars)

    def __contains__(self, key):
        return key in self.data_vars or key in self.coords

    def __len__(self):
        return len(self.data_vars)

    def keys(self):
        return self.data_vars.keys()

    def items(self):
        return ((k, self[k]) for k in self.data_vars)

    @property
    def sizes(self):
        out = OrderedDict()
        for v in self.data_vars.values():
            out.update(v.sizes)
        return out

    dims = sizes

    def map(self, func, args=(), **kw):
        return Dataset({k: func(self[k], *args, **kw) for k in self.data_vars}, attrs=self.attrs)

    apply = map

    def rename(self, mapping):
        return Dataset({mapping.get(k, k): self[k].rename(mapping) for k in self.data_vars},
                       attrs=self.attrs)

    def assign_coords(self, coords=None, **kw):
        out = Dataset({k: self[k] for k in self.data_vars}, self.coords, self.attrs)
        for k, v in dict(coords or {}, **kw).items():
            out.coords[k] = _as_coord(k, v)
        return out

    def copy(self):
        return Dataset({k: self[k] for k in self.data_vars}, self.coords, self.attrs)

    def compute(self):
        return self

    def __repr__(self):
        return f"<fv3net_b200.Dataset {list(self.data_vars)}>"


def merge(objs):
    out = Dataset()
    for o in objs:
        if isinstance(o, DataArray):
            o = o.to_dataset()
        for k in o.data_vars:
            out[k] = o[k]
        out.attrs.update({})
    return out


# ---- comparison helpers used by the tests (xr.testing.assert_identical semantics) ----------
def identical(a, b, check_dtype=True):
    if type(a) is not type(b):
        return False
    if isinstance(a, Dataset):
        return (list(a.data_vars) == list(b.data_vars) and a.attrs == b.attrs
                and all(identical(a[k], b[k], check_dtype) for k in a.data_vars))
    if a.dims != b.dims or a.name != b.name or a.attrs != b.attrs or a.shape != b.shape:
        return False
    if check_dtype and a.dtype != b.dtype:
        return False
    if not np.array_equal(a.values, b.values, equal_nan=True):
        return False
    if set(a.coords) != set(b.coords):
        return False
    return all(identical(a.coords[k], b.coords[k], check_dtype) for k in a.coords)


def assert_identical(a, b, check_dtype=True):
    assert identical(a, b, check_dtype), f"not identical:\n{a!r}\n{b!r}\n{_dump(a)}\n{_dump(b)}"


def _dump(a):
    if isinstance(a, Dataset):
        return {k: _dump(a[k]) for k in a.data_vars}
    return dict(dims=a.dims, name=a.name, attrs=a.attrs, dtype=a.dtype, values=a.values,
                coords={k: (c.dims, c.dtype, c.values) for k, c in a.coords.items()})


# ---- xarray boundary ------------------------------------------------------------------------
def is_xarray(obj):
    return _xr is not None and isinstance(obj, (_xr.DataArray, _xr.Dataset))


def is_lazy(obj):
    """An xarray DataArray whose data are a dask array (duck-typed: chunks + compute)."""
    return (_xr is not None and isinstance(obj, _xr.DataArray) and hasattr(obj.data, "chunks")
            and hasattr(obj.data, "compute"))


def adapt(obj, values=True):
    """xarray -> labelled (dask-backed data are computed here: the kernels are eager).  `values=False`: dims,
    shape, coords, name and attrs only - the data become a zero-stride placeholder and nothing is computed."""
    if not is_xarray(obj):
        return obj
    if isinstance(obj, _xr.Dataset):
        return Dataset({k: adapt(obj[k], values) for k in obj.data_vars},
                       {k: (c.dims, np.asarray(c.values), dict(c.attrs)) for k, c in obj.coords.items()
                        if not set(c.dims) - set(obj.dims)}, dict(obj.attrs))
    coords = {k: (c.dims, np.asarray(c.values), dict(c.attrs)) for k, c in obj.coords.items()}
    if values:
        data = np.asarray(obj.values)
    else:
        data = np.broadcast_to(np.zeros((), dtype=obj.data.dtype), tuple(obj.data.shape))
    return DataArray(data, obj.dims, coords, obj.name, dict(obj.attrs))


def defer(fn, like, shape, dtype):
    """An array of `like.data`'s own library (dask.array) whose values are `fn()` when the graph is computed:
    `from_delayed(delayed(fn)(), shape, dtype)` - one chunk, the device work runs at compute time."""
    import importlib
    import sys

    mod = sys.modules[type(like.data).__module__]
    root = importlib.import_module(mod.__name__.split(".")[0])
    delayed = getattr(root, "delayed", None) or getattr(mod, "delayed")
    return mod.from_delayed(delayed(fn)(), shape=tuple(shape), dtype=np.dtype(dtype))


def restore(result, like, data=None):
    """labelled -> xarray when the caller handed us xarray; lazy (dask) in -> lazy out.  `data`: an array to put
    in place of `result.values` (the deferred array of `defer`; `result` then only carries the metadata)."""
    if not is_xarray(like) or isinstance(result, (tuple, list)):
        if isinstance(result, tuple):
            return tuple(restore(r, like) for r in result)
        return result
    if isinstance(result, Dataset):
        return _xr.Dataset({k: restore(result[k], like[k] if k in like else like) for k in result.data_vars},
                           attrs=result.attrs)
    coords = {k: _xr.Variable(c.dims, c.values, c.attrs) for k, c in result.coords.items()}
    if data is not None:
        return _xr.DataArray(data, dims=result.dims, coords=coords, name=result.name, attrs=result.attrs)
    data = result.values
    src = like if isinstance(like, _xr.DataArray) else None
    if src is not None and hasattr(src.data, "compute") and hasattr(src.data, "chunks"):
        # lazy (dask) in -> lazy out for the functions that do not defer their device work (regrid_vertical does,
        # see there): the values were computed when the call was made and are handed back as one chunk of the
        # input's own array type
        import sys

        data = sys.modules[type(src.data).__module__].from_array(data, chunks=data.shape)
    return _xr.DataArray(data, dims=result.dims, coords=coords, name=result.name, attrs=result.attrs)
