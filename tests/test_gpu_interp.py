"""cg_interpolate_2d (the f2py module's second entry point, interpolate_2d.f90) on the GPU:
bit-exact against the numpy oracle, the reference's own test through the drop-in module."""
import numpy as np
import pytest

import oracle

pytestmark = pytest.mark.gpu


def _rows(rng, m, n_in, n_out):
    x = np.cumsum(rng.uniform(0.1, 1, (m, n_in)), axis=1)
    y = rng.uniform(-5, 5, (m, n_in))
    xp = rng.uniform(-1, x.max() + 1, (m, n_out))
    xp[:, 0] = x[:, n_in // 2]                               # exact hit
    xp[:, -1] = x[:, -1]                                     # last source point
    if n_in > 4:
        x[m // 2:, 3] = x[m // 2:, 2]                        # a zero-width interval
    return xp, x, y


@pytest.mark.parametrize("m,n_in,n_out", [(1, 10, 12), (257, 79, 35), (1000, 2, 5), (33, 80, 1), (5, 1, 4)])
def test_interpolate_2d_bit_exact_vs_oracle(m, n_in, n_out):
    from fv3net_b200 import mappm

    xp, x, y = _rows(np.random.RandomState(m + n_in), m, n_in, n_out)
    got = mappm.interpolate_2d(xp, x, y, fill_value=np.nan)
    exp = oracle.interpolate_2d(xp, x, y, fill_value=np.nan)
    assert got.dtype == np.float64 and got.shape == (m, n_out)
    np.testing.assert_array_equal(got, exp)
    got = mappm.interpolate_2d(xp, x, y, fill_value=-7.5)
    np.testing.assert_array_equal(got, oracle.interpolate_2d(xp, x, y, fill_value=-7.5))


def test_interpolate_2d_reference_test():
    """external/vcm/tests/test_interpolate.py:120-132, integer inputs included."""
    import scipy.interpolate
    from fv3net_b200 import mappm

    x = np.arange(10).reshape((1, 10))
    y = (x ** 2).reshape((1, 10))
    xp = np.arange(12).reshape((1, 12))
    expected = scipy.interpolate.interp1d(x[0], y[0], bounds_error=False)(xp[0])[None]
    assert np.isnan(expected[:, -2:]).all()
    np.testing.assert_allclose(expected, mappm.interpolate_2d(xp, x, y, fill_value=np.nan))


def test_interpolate_2d_strided_layout():
    """(level, column) arrays addressed by strides: no transpose to the reference's z-last layout."""
    import torch
    from fv3net_b200 import ops

    xp, x, y = _rows(np.random.RandomState(2), 300, 40, 19)
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a.T)).cuda()          # noqa: E731  (n, m) on the device
    txp, tx, ty = dev(xp), dev(x), dev(y)
    out = ops.interpolate_2d(txp.T, tx.T, ty.T, float("nan"))
    np.testing.assert_array_equal(out.cpu().numpy(), oracle.interpolate_2d(xp, x, y))


def test_interpolate_2d_errors():
    from fv3net_b200 import mappm

    with pytest.raises(ValueError, match="shape mismatch"):
        mappm.interpolate_2d(np.zeros((2, 3)), np.zeros((2, 4)), np.zeros((2, 5)), 0.0)
    with pytest.raises(ValueError, match="2-D"):
        mappm.interpolate_2d(np.zeros(3), np.zeros((2, 4)), np.zeros((2, 4)), 0.0)
