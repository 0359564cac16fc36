"""Adjudicators for the parity tests that cannot be bit-exact (used by the -m gpu tests; numpy + the oracle only).

1. `assert_within_fp32_conditioning`: SURVEY.md §8c's rule for the PPM remap where the 3e-5 bar is too tight for
   ANY fp32 evaluation (ill-conditioned columns): the CUDA result may be at most twice as far from the fp64
   evaluation as the reference's own fp32 arithmetic is, over the case (four times column by column), or 3e-5 of
   the field scale, whichever is larger.
2. `assert_exposed_area_matches`: budgets.py:62-77 compares a fine column's surface pressure with
   delp / diff(log p) of the coarse column.  The last bit of the logarithm is the math library's, so a column may
   flip where its surface pressure sits within that rounding of a coarse mid-level pressure - and nowhere else.
"""
import numpy as np

from oracle import blocks as B, recipes


def plevel_reference(field, delp, area, f, mask_below, dtype, threads=1, dc=None):
    """The reference recipe (regridz.py:115-147 + coarsen.py:183, or budgets.py:49-57) evaluated in `dtype`;
    `dc`: a given coarse thickness instead of the block average formed here."""
    field, delp, area = (np.asarray(a).astype(dtype) for a in (field, delp, area))
    dc = B.weighted_block_average(delp, area[:, None], f) if dc is None else np.asarray(dc).astype(dtype)
    pf = B.pressure_at_interface(delp, axis=1)
    pc = B.pressure_at_interface(B.block_upsample(dc, f, axes=(-1, -2)), axis=1)
    fg = recipes.regrid_vertical(pf, field, pc, dtype=dtype, threads=threads)
    w = B.mask_weights(area[:, None], pc, pf, axis=1) if mask_below else area[:, None]
    return B.weighted_block_average(fg.astype(dtype), w, f)


def assert_within_fp32_conditioning(got, field, delp, area, f, mask_below, scale=None, floor=3e-5, threads=1, dc_gpu=None):
    """`dc_gpu`: the coarse thickness the CUDA path formed.  It is a second valid fp32 rounding of the same block
    average (numpy's pairwise order is implementation-defined too), so the reference's fp32 distance is taken over
    both: the remap amplifies that last-bit difference exactly like any other fp32 rounding."""
    e32 = plevel_reference(field, delp, area, f, mask_below, np.float32, threads)
    e64 = plevel_reference(field, delp, area, f, mask_below, np.float64, threads)
    e32b = e32 if dc_gpu is None else plevel_reference(field, delp, area, f, mask_below, np.float32, threads, dc=dc_gpu)
    ok = np.isfinite(e32) & np.isfinite(e64)
    assert np.array_equal(np.isfinite(got), np.isfinite(e32))
    scale = float(np.nanmax(np.abs(field))) if scale is None else scale
    d_gpu = np.where(ok, np.abs(got - e64), 0.0).max(axis=1)          # per coarse column: max over the levels
    d_ref = np.where(ok, np.maximum(np.abs(e32 - e64), np.abs(np.nan_to_num(e32b) - e64)), 0.0).max(axis=1)
    # over the whole case: at most twice the reference's own fp32 distance.  Column by column the two
    # fp32 evaluations round differently (the ratio of two errors of the same size scatters), so there the factor is 4
    assert d_gpu.max() <= max(2.0 * d_ref.max(), floor * scale), (
        f"max |gpu - f64| = {d_gpu.max():.3e}, fp32 reference is {d_ref.max():.3e} from f64, scale {scale:.3g}")
    allowed = np.maximum(4.0 * d_ref, floor * scale)
    worst = np.unravel_index(np.argmax(d_gpu - allowed), d_gpu.shape)
    assert (d_gpu <= allowed).all(), (
        f"column {worst}: |gpu - f64| = {d_gpu[worst]:.3e}, fp32 reference is {d_ref[worst]:.3e} from f64, scale {scale:.3g}")
    return float(d_gpu.max() / scale), float(d_ref.max() / scale)


def assert_exposed_area_matches(got, delp, delp_coarse, area, f, rtol=2e-6):
    delp, dc, area = np.asarray(delp), np.asarray(delp_coarse), np.asarray(area)
    exp = recipes.area_above_fine_surface(delp, dc, area, f)
    bad = ~np.isclose(got, exp, rtol=rtol, atol=0)
    if not bad.any():
        return 0
    # where may a column flip?  p_mid = dp / (log p1 - log p0): an error of e in each logarithm moves it by
    # 2 e / dlogp relative; e = 2 ulp of log(p) covers libm against CUDA's logf
    pi = B.pressure_at_interface(dc.astype(np.float64), axis=1)
    dlogp = np.diff(np.log(pi), axis=1)
    p_mid = B.block_upsample(B.pressure_at_midpoint_log(dc, axis=1), f, axes=(-1, -2)).astype(np.float64)
    band = B.block_upsample(4.0 * np.spacing(np.float32(np.log(pi[:, 1:].max()))) / dlogp, f, axes=(-1, -2))
    ps = B.surface_pressure_from_delp(delp, axis=1).astype(np.float64)[:, None]
    borderline = np.abs(ps - p_mid) <= band * p_mid + 2 * np.spacing(np.float32(p_mid))
    a = np.nan_to_num(area.astype(np.float64))[:, None]
    slack = B.block_sum(np.where(borderline, a, 0.0), f)              # area that may flip in each coarse cell
    excess = np.abs(got - exp) - (slack * (1 + 1e-6) + rtol * np.abs(exp))
    worst = np.unravel_index(np.argmax(np.where(bad, excess, -np.inf)), exp.shape)
    assert (excess[bad] <= 0).all(), (
        f"cell {worst}: got {got[worst]!r}, reference {exp[worst]!r}, but only {slack[worst]!r} of area sits within the "
        f"logarithm's rounding of the level's mid pressure")
    return int(bad.sum())
