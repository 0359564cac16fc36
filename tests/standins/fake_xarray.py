"""`xarray` stand-in: DataArray / Dataset / Variable carrying dims, coords, attrs, name, .data and .values."""
from collections import OrderedDict

import numpy as np


class Variable:
    def __init__(self, dims, data, attrs=None):
        self.dims = (dims,) if isinstance(dims, str) else tuple(dims)
        self.data = data
        self.attrs = dict(attrs or {})

    @property
    def values(self):
        return np.asarray(self.data)


class DataArray:
    def __init__(self, data, dims=None, coords=None, name=None, attrs=None):
        self.data = data
        self.dims = tuple(dims)
        self.name = name
        self.attrs = dict(attrs or {})
        self.coords = OrderedDict()
        for k, c in (coords or {}).items():
            if isinstance(c, Variable):
                self.coords[k] = c
            elif isinstance(c, tuple):
                self.coords[k] = Variable(*c)
            else:
                self.coords[k] = Variable((k,), np.asarray(c))

    @property
    def values(self):
        return np.asarray(self.data)

    @property
    def shape(self):
        return tuple(self.data.shape)

    def compute(self):
        return DataArray(self.values, self.dims, self.coords, self.name, self.attrs)


class Dataset:
    def __init__(self, data_vars=None, coords=None, attrs=None):
        self.data_vars = OrderedDict(data_vars or {})
        self.attrs = dict(attrs or {})
        self.coords = OrderedDict()
        for k, c in (coords or {}).items():
            self.coords[k] = c if isinstance(c, Variable) else Variable((k,), np.asarray(c))
        for v in self.data_vars.values():
            for k, c in v.coords.items():
                self.coords.setdefault(k, c)

    @property
    def dims(self):
        sizes = OrderedDict()
        for v in self.data_vars.values():
            for d, n in zip(v.dims, v.shape):
                sizes[d] = n
        return sizes

    def __getitem__(self, k):
        return self.data_vars[k]

    def __contains__(self, k):
        return k in self.data_vars
