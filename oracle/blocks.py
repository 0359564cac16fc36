"""TEST INFRASTRUCTURE — numpy restatement of the reference's block reductions.

Plain arrays, horizontal dims are the LAST TWO axes (y, x), as in the
workflows' (tile, z, y, x) layout.  Each function cites what it restates
(paths relative to /root/reference/external/vcm/vcm).

xarray (0.19, un-vendored) semantics restated here: ``.coarsen(...).sum()``
reshapes every coarsened axis into (n/f, f) windows and applies
``nansum`` (NaN -> 0, all-NaN window -> 0) because ``skipna`` defaults to True
for floats; ``min``/``max`` likewise use nanmin/nanmax; ``boundary="exact"``
raises when the size is not a multiple of the factor.
"""
import numpy as np

__all__ = [
    "block_view", "block_sum", "weighted_block_average", "edge_weighted_block_average",
    "block_edge_coarsen", "block_coarsen", "block_median", "block_mode", "block_upsample",
    "pressure_at_interface", "pressure_at_midpoint_log", "surface_pressure_from_delp",
    "mask_weights", "coarsen_coords_coord_func", "TOA_PRESSURE",
]

TOA_PRESSURE = 300.0  # calc/thermo/constants.py:17


def _exact(n, f, what):
    if n % f != 0:
        raise ValueError(
            f"Could not coarsen a dimension of size {n} with window {f} ({what}); "
            "boundary='exact'"
        )


def block_view(a, fy, fx):
    """(..., ny, nx) -> (..., ny/fy, fy, nx/fx, fx) view."""
    ny, nx = a.shape[-2:]
    _exact(ny, fy, "y")
    _exact(nx, fx, "x")
    return a.reshape(a.shape[:-2] + (ny // fy, fy, nx // fx, fx))


def _nan_reduce(a, fy, fx, method):
    v = block_view(a, fy, fx)
    axes = (-3, -1)
    if method == "sum":
        return np.nansum(v, axis=axes)
    if method == "min":
        return np.nanmin(v, axis=axes)
    if method == "max":
        return np.nanmax(v, axis=axes)
    if method == "mean":
        return np.nanmean(v, axis=axes)
    raise ValueError(method)


def block_sum(a, f):
    """budgets.py:59 ``array.coarsen({x: f, y: f}).sum()``."""
    return _nan_reduce(np.asarray(a), f, f, "sum")


def weighted_block_average(obj, weights, f):
    """cubedsphere/coarsen.py:210-213: nansum_block(obj*w) / nansum_block(w).
    ``weights`` broadcasts against ``obj`` from the right (2-D area against a
    (tile, z, y, x) field needs shape (tile, 1, y, x))."""
    obj = np.asarray(obj)
    weights = np.asarray(weights)
    num = _nan_reduce(obj * weights, f, f, "sum")
    den = _nan_reduce(weights, f, f, "sum")
    with np.errstate(invalid="ignore", divide="ignore"):
        return num / den


def edge_weighted_block_average(obj, spacing, f, edge="x"):
    """cubedsphere/coarsen.py:250-264: 1-D block sums along the edge direction,
    then every f-th line of the other (staggered) direction."""
    obj = np.asarray(obj)
    spacing = np.asarray(spacing)
    if edge == "x":
        num = _nan_reduce(spacing * obj, 1, f, "sum")
        den = _nan_reduce(spacing, 1, f, "sum")
        with np.errstate(invalid="ignore", divide="ignore"):
            return (num / den)[..., ::f, :]
    elif edge == "y":
        num = _nan_reduce(spacing * obj, f, 1, "sum")
        den = _nan_reduce(spacing, f, 1, "sum")
        with np.errstate(invalid="ignore", divide="ignore"):
            return (num / den)[..., :, ::f]
    raise ValueError(f"'edge' most be either 'x' or 'y'; got {edge}.")


def block_edge_coarsen(obj, f, edge="x", method="sum"):
    """cubedsphere/coarsen.py:658-674 (block_edge_sum is method='sum')."""
    obj = np.asarray(obj)
    if edge == "x":
        return _nan_reduce(obj, 1, f, method)[..., ::f, :]
    elif edge == "y":
        return _nan_reduce(obj, f, 1, method)[..., :, ::f]
    raise ValueError(f"'edge' most be either 'x' or 'y'; got {edge}.")


def block_median(obj, f):
    """cubedsphere/coarsen.py:557-588 -> _skimage.py:125-202 with np.median over
    the trailing block axes: NaN in a block -> NaN; even count -> mean of the two
    middle order statistics in the input dtype."""
    obj = np.asarray(obj)
    return np.median(block_view(obj, f, f), axis=(-3, -1))


def block_mode(obj, f, nan_policy="propagate"):
    """cubedsphere/coarsen.py:686-786: scipy.stats.mode over the flattened block, as the
    reference's pinned scipy 1.7.3 / numpy 1.21.5 compute it (constraints.txt:181,250):
      'propagate': np.unique(..., return_counts) then argmax — NaN is an ordinary candidate
                   that sorts last, so it wins only when strictly more frequent than every
                   other value; ties go to the smallest value;
      'omit'     : NaNs are ignored; an all-NaN block gives NaN.
    (Pure-Python loop over blocks: the oracle is for small cases.)"""
    obj = np.asarray(obj)
    v = block_view(obj, f, f)
    v = np.moveaxis(v, -3, -2)  # (..., nyc, nxc, f, f)
    flat = v.reshape(v.shape[:-2] + (f * f,))
    out = np.empty(flat.shape[:-1], dtype=obj.dtype)
    for idx in np.ndindex(*out.shape):
        blk = flat[idx]
        if nan_policy == "omit":
            blk = blk[~np.isnan(blk)]
            if blk.size == 0:
                out[idx] = np.nan
                continue
        vals, cnts = np.unique(blk, return_counts=True)
        out[idx] = vals[cnts.argmax()]
    return out


def block_coarsen(obj, f, method="sum"):
    """cubedsphere/coarsen.py:795-840."""
    if method == "median":
        return block_median(obj, f)
    if method == "mode":
        return block_mode(obj, f)
    return _nan_reduce(np.asarray(obj), f, f, method)


def _upsample_axis(a, f, axis):
    """cubedsphere/coarsen.py:843-866: even size -> repeat all; odd (staggered)
    size -> repeat all but the last line, keep the last once."""
    n = a.shape[axis]
    if n % 2 == 1:
        idx = [slice(None)] * a.ndim
        idx[axis] = slice(None, -1)
        head = np.repeat(a[tuple(idx)], f, axis=axis)
        idx[axis] = slice(-1, None)
        return np.concatenate([head, a[tuple(idx)]], axis=axis)
    return np.repeat(a, f, axis=axis)


def block_upsample(obj, f, axes=(-1, -2)):
    """cubedsphere/coarsen.py:869-897."""
    obj = np.asarray(obj)
    for ax in axes:
        obj = _upsample_axis(obj, f, ax)
    return obj


def pressure_at_interface(delp, toa_pressure=TOA_PRESSURE, axis=-3):
    """calc/thermo/vertically_dependent.py:61-63: cumsum([toa, delp...]) in
    delp's dtype, sequentially from the top."""
    delp = np.asarray(delp)
    shp = list(delp.shape)
    shp[axis] = 1
    top = np.full(shp, toa_pressure, dtype=delp.dtype)
    return np.cumsum(np.concatenate([top, delp], axis=axis), axis=axis, dtype=delp.dtype)


def pressure_at_midpoint_log(delp, toa_pressure=TOA_PRESSURE, axis=-3):
    """calc/thermo/vertically_dependent.py:153-179: delp / diff(log(p_interface))."""
    pi = pressure_at_interface(delp, toa_pressure, axis)
    return np.asarray(delp) / np.diff(np.log(pi), axis=axis)


def surface_pressure_from_delp(delp, p_toa=300.0, axis=-3):
    """calc/thermo/vertically_dependent.py:189-208."""
    return np.nansum(np.asarray(delp), axis=axis) + p_toa


def mask_weights(weights, phalf_coarse_on_fine, phalf_fine, axis=-3):
    """cubedsphere/regridz.py:150-161: keep w where the coarse layer's bottom
    interface is strictly above the fine surface pressure, else 0."""
    pc = np.asarray(phalf_coarse_on_fine)
    pf = np.asarray(phalf_fine)
    lower = np.take(pc, range(1, pc.shape[axis]), axis=axis)
    surf = np.take(pf, [pf.shape[axis] - 1], axis=axis)
    w = np.asarray(weights)
    return np.where(lower < surf, w, np.zeros((), dtype=w.dtype)).astype(w.dtype)


def coarsen_coords_coord_func(coordinate, axis=-1):
    """cubedsphere/coarsen.py:128-132 (coordinate is the (n/f, f) window array)."""
    coordinate = np.asarray(coordinate)
    return ((coordinate[:, 0] - 1) // coordinate.shape[1] + 1).astype(int).astype(np.float32)
