"""`dask.array` stand-in: an array whose values exist only after compute() (counts how often that happens)."""
import numpy as np


class Array:
    computed = 0                      # class-wide counter: the tests check WHEN the data were materialised

    def __init__(self, data, chunks):
        self._data = np.asarray(data)
        self.chunks = chunks
        self.shape = self._data.shape
        self.dtype = self._data.dtype
        self.ndim = self._data.ndim

    def compute(self):
        Array.computed += 1
        return self._data

    def __array__(self, dtype=None, copy=None):
        out = self.compute()
        return out.astype(dtype) if dtype is not None else out


def from_array(data, chunks):
    return Array(data, chunks)
