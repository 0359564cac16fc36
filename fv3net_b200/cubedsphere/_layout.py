"""Host-side layout plumbing between labelled arrays and the canonical device layout.

The kernels take contiguous tensors with the horizontal dims last:
    (n_outer, ny, nx)   or   (n_outer, nz, ny, nx)
A labelled array whose dims are already in that order (the layout of the reference's
workflows: x fastest, then y, z, tile, time - fine_res_budget/budget/pipeline.py:36) goes
to the kernel without a copy; any other order is permuted once on the device.
"""
import numpy as np
import torch

from .. import ops
from ..labeled import DataArray


def numel(shape):
    return int(np.prod(shape, dtype=np.int64)) if len(shape) else 1


def device_tensor(da, order=None):
    """Data of `da` as a contiguous CUDA tensor with dims in `order` (default: its own)."""
    t = ops.to_device(da.data)
    if order is not None and tuple(order) != tuple(da.dims):
        if set(order) != set(da.dims):
            raise ValueError(f"cannot order dims {da.dims} as {tuple(order)}")
        t = t.permute([da.dims.index(d) for d in order])
    return t.contiguous()


def broadcast_to(da, dims, sizes):
    """`da` expanded (on the device, as a view) to carry every dim in `dims`, in that order."""
    t = ops.to_device(da.data)
    have = [d for d in dims if d in da.dims]
    t = t.permute([da.dims.index(d) for d in have])
    shape, expand = [], []
    for d in dims:
        if d in da.dims:
            shape.append(da.sizes[d])
            expand.append(da.sizes[d])
        else:
            shape.append(1)
            expand.append(sizes[d])
    return t.reshape(shape).expand(expand)


def finish(t, dims_now, dims_target, like, dtype=None):
    """Tensor with dims `dims_now` -> data for a labelled array with dims `dims_target`;
    host data in (numpy) gives host data out, device data stays on the device."""
    if tuple(dims_now) != tuple(dims_target):
        t = t.permute([list(dims_now).index(d) for d in dims_target])
    if dtype is not None:
        t = t.to(dtype)
    if like_is_device(like):
        return t.contiguous()
    return t.cpu().numpy()


def like_is_device(obj):
    if isinstance(obj, DataArray):
        return obj.is_device
    return any(v.is_device for v in obj.data_vars.values())


def torch_dtype(np_dtype):
    return getattr(torch, np.dtype(np_dtype).name)


def check_exact(size, window, dim):
    """xarray's coarsen(boundary='exact') refusal, same wording (xarray/core/variable.py)."""
    if size % window != 0:
        raise ValueError(
            f"Could not coarsen a dimension of size {size} with window {window} and "
            f"boundary='exact'. Try a different 'boundary' option. (dimension {dim!r})"
        )
