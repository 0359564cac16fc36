"""TEST INFRASTRUCTURE — CPU oracle, not product code.

numpy restatement of external/mappm/mappm/interpolate_2d.f90:1-29: the Fortran's k loop is kept
(no early exit: the last matching interval wins), vectorised over the rows and targets.
Pinned by external/vcm/tests/test_interpolate.py:109-132 (scipy.interpolate.interp1d with
bounds_error=False on x = arange(10), y = x**2, xp = arange(12)); see tests/test_oracle_golden.py.
"""
import numpy as np


def interpolate_2d(xp, x, y, fill_value=np.nan):
    xp, x, y = (np.asarray(a, dtype=np.float64) for a in (xp, x, y))
    m, n_out = xp.shape
    n_in = x.shape[1]
    out = np.full((m, n_out), fill_value, dtype=np.float64)
    for k in range(n_in - 1):
        x0, x1 = x[:, k:k + 1], x[:, k + 1:k + 2]
        y0, y1 = y[:, k:k + 1], y[:, k + 1:k + 2]
        inside = (x0 <= xp) & (xp < x1)
        with np.errstate(divide="ignore", invalid="ignore"):
            weight = (xp - x0) / (x1 - x0)
            lerp = y0 * (1 - weight) + y1 * weight
        at_lo = ~inside & (x0 == xp)
        at_hi = ~inside & ~at_lo & (x1 == xp)
        out = np.where(inside, lerp, out)
        out = np.where(at_lo, np.broadcast_to(y0, out.shape), out)
        out = np.where(at_hi, np.broadcast_to(y1, out.shape), out)
    return out
