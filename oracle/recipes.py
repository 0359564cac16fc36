"""TEST INFRASTRUCTURE — numpy/C restatement of the reference's compositions of
the block reductions with the vertical remap.  Arrays are (tile, z, y, x)
(2-D weights (tile, y, x)); paths are relative to /root/reference.
"""
import numpy as np

from . import blocks as B
from .mappm import mappm_strided

# external/vcm/vcm/cubedsphere/xgcm.py:7-34 — (left neighbour, right neighbour)
# of each tile along each axis, as (tile, neighbour axis).
FACE_CONNECTIONS = {
    0: {"x": ((4, "y"), (1, "x")), "y": ((5, "y"), (2, "x"))},
    1: {"x": ((0, "x"), (3, "y")), "y": ((5, "x"), (2, "y"))},
    2: {"x": ((0, "y"), (3, "x")), "y": ((1, "y"), (4, "x"))},
    3: {"x": ((2, "x"), (5, "y")), "y": ((1, "x"), (4, "y"))},
    4: {"x": ((2, "y"), (5, "x")), "y": ((3, "y"), (0, "x"))},
    5: {"x": ((4, "x"), (1, "y")), "y": ((3, "x"), (0, "y"))},
}


def regrid_vertical(p_in, f_in, p_out, iv=1, kord=1, dtype=np.float32, threads=1, zlast=True):
    """external/vcm/vcm/cubedsphere/regridz.py:164-279 on (tile, z, y, x) arrays:
    columnwise mappm, fp32 inside (f2py casts to ``real``), float32 out.

    zlast (default): the reference's access pattern - the three arrays are transposed to z-last and copied
    (regridz.py:206-242, :317-325 reshape to (ncol, nz) for f2py), the remapper walks contiguous columns, and the
    result is transposed back.  zlast=False walks the columns in place at stride ny * nx (same values)."""
    p_in, f_in, p_out = (np.ascontiguousarray(a) for a in (p_in, f_in, p_out))
    nt, km, ny, nx = f_in.shape
    kn = p_out.shape[1] - 1
    if p_in.shape != (nt, km + 1, ny, nx) or p_out.shape != (nt, kn + 1, ny, nx):
        raise ValueError("All dimensions except vertical must be same size for p_in, f_in and p_out")
    h = ny * nx
    if zlast:
        pz, fz, oz = (np.ascontiguousarray(np.moveaxis(a, 1, -1)) for a in (p_in, f_in, p_out))
        out = mappm_strided(pz, fz, oz, 1, nt * h, km, kn,
                            (0, km + 1, 1), (0, km, 1), (0, kn + 1, 1), (0, kn, 1),
                            iv=iv, kord=kord, dtype=dtype, threads=threads)
        return np.ascontiguousarray(np.moveaxis(out.reshape(nt, ny, nx, kn), -1, 1))
    out = mappm_strided(p_in, f_in, p_out, nt, h, km, kn,
                        ((km + 1) * h, 1, h), (km * h, 1, h), ((kn + 1) * h, 1, h), (kn * h, 1, h),
                        iv=iv, kord=kord, dtype=dtype, threads=threads)
    return out.reshape(nt, kn, ny, nx)


def regrid_given_delp(fields, delp_fine, delp_coarse, weights, iv=1, kord=1, threads=1):
    """regridz.py:115-147.  ``fields``: list of (tile, z, y, x) arrays.  Returns
    (list of remapped fine-grid fields, masked weights (tile, z, y, x))."""
    f = _factor(delp_fine.shape[-1], delp_coarse.shape[-1])
    dc_up = B.block_upsample(delp_coarse, f, axes=(-1, -2))          # coarsen.py:900-938
    pc = B.pressure_at_interface(dc_up, axis=1)
    pf = B.pressure_at_interface(delp_fine, axis=1)
    out = [regrid_vertical(pf, fld, pc, iv=iv, kord=kord, threads=threads) for fld in fields]
    w = np.asarray(weights)
    masked = B.mask_weights(w[:, None], pc, pf, axis=1)               # regridz.py:150-161
    return out, masked


def _factor(n_fine, n_coarse):
    if n_coarse % 2 == 1:                                             # coarsen.py:928-932
        return (n_fine - 1) // (n_coarse - 1)
    return n_fine // n_coarse


def regrid_to_area_weighted_pressure(fields, delp, area, f, threads=1):
    """regridz.py:25-55."""
    delp_coarse = B.weighted_block_average(delp, np.asarray(area)[:, None], f)
    return regrid_given_delp(fields, delp, delp_coarse, area, threads=threads)


def interp_to_edges(c, axis):
    """``xgcm.Grid.interp(delp, axis)`` with the FV3 face connections
    (regridz.py:86-100; xgcm 0.6.1 is not vendored): value j of the outer
    dimension is the mean of cells j-1 and j; the missing line at a tile face is
    the neighbour's LAST (left pad) / FIRST (right pad) line along the
    neighbour's listed axis, reversed along its transverse index when that axis
    differs from ``axis``.  Rule recovered from, and pinned by, the golden
    pressure-level u/v of test_coarsen_restarts.py."""
    c = np.asarray(c)
    nt = c.shape[0]
    assert nt == 6
    out = []
    for t in range(nt):
        (lt, la), (rt, ra) = FACE_CONNECTIONS[t][axis]

        def line(tile, naxis, last):
            i = -1 if last else 0
            ln = c[tile][..., i, :] if naxis == "y" else c[tile][..., :, i]
            return ln[..., ::-1] if naxis != axis else ln

        lo, hi = line(lt, la, True), line(rt, ra, False)
        if axis == "y":
            padded = np.concatenate([lo[..., None, :], c[t], hi[..., None, :]], axis=-2)
            out.append(0.5 * (padded[..., :-1, :] + padded[..., 1:, :]))
        else:
            padded = np.concatenate([lo[..., :, None], c[t], hi[..., :, None]], axis=-1)
            out.append(0.5 * (padded[..., :, :-1] + padded[..., :, 1:]))
    return np.stack(out)


def regrid_to_edge_weighted_pressure(fields, delp, length, f, edge="x", threads=1):
    """regridz.py:58-112."""
    interp_axis = "x" if edge == "y" else "y"
    delp_stag = interp_to_edges(delp, interp_axis)
    delp_stag_coarse = B.edge_weighted_block_average(delp_stag, np.asarray(length)[:, None], f, edge)
    return regrid_given_delp(fields, delp_stag, delp_stag_coarse, length, threads=threads)


def pressure_level_average(field, delp_fine, delp_coarse, area, f, threads=1):
    """workflows/fine_res_budget/budget/budgets.py:49-57 — remap to the
    upsampled coarse pressure, then plain (unmasked) area-weighted average."""
    pi = B.pressure_at_interface(delp_fine, axis=1)
    pi_c = B.pressure_at_interface(delp_coarse, axis=1)
    pi_c_up = B.block_upsample(pi_c, f, axes=(-1, -2))
    fg = regrid_vertical(pi, field, pi_c_up, threads=threads)
    return B.weighted_block_average(fg, np.asarray(area)[:, None], f)


def area_above_fine_surface(delp_fine, delp_coarse, area, f):
    """budgets.py:62-77."""
    p_c = B.pressure_at_midpoint_log(delp_coarse, axis=1)
    p_c_up = B.block_upsample(p_c, f, axes=(-1, -2))
    ps = B.surface_pressure_from_delp(delp_fine, axis=1)
    a = np.asarray(area)
    exposed = np.where(p_c_up <= ps[:, None], a[:, None], np.zeros((), a.dtype))
    return B.block_sum(exposed, f)
