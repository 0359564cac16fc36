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


class Delayed:                        # `dask.delayed(fn)(*args)`: the call, not its result
    def __init__(self, fn, args, kwargs):
        self.fn, self.args, self.kwargs = fn, args, kwargs

    def compute(self):
        return self.fn(*self.args, **self.kwargs)


def delayed(fn):
    return lambda *args, **kwargs: Delayed(fn, args, kwargs)


class DelayedArray(Array):            # `dask.array.from_delayed`: shape and dtype now, values at compute()
    def __init__(self, value, shape, dtype):
        self._value = value
        self.shape, self.dtype, self.ndim = tuple(shape), np.dtype(dtype), len(shape)
        self.chunks = tuple((n,) for n in shape)

    def compute(self):
        Array.computed += 1
        out = np.asarray(self._value.compute())
        assert out.shape == self.shape and out.dtype == self.dtype, (out.shape, out.dtype, self.shape, self.dtype)
        return out


def from_delayed(value, shape, dtype):
    return DelayedArray(value, shape, dtype)
