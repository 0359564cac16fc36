"""GPU parity tests proper: every C-ABI kernel against the CPU oracle on seeded inputs.

Bars (north_star): bit-exact for indexing / selection / copies / masks; weighted sums
within rel 1e-6 (fp32) / 1e-13 (fp64) of the oracle in the same dtype; the PPM remap
(cg_mappm, compiled without FMA contraction) bit-exact against the fp32 oracle.
"""
import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

import oracle  # noqa: E402
from oracle import recipes  # noqa: E402


@pytest.fixture(scope="module")
def ops():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from fv3net_b200 import ops as _ops
    return _ops


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def host(t):
    return t.cpu().numpy()


def assert_sum_close(got, exp, scale=None, eps=None):
    """Block sums are order-dependent (numpy pairwise vs in-thread sequential), so the bar
    is on the field scale: |d| <= 16 eps * max|terms| (n = 64 terms, error ~ sqrt(n) eps
    |term| typically, n eps |term| worst case), i.e. ~2e-6 (fp32) / 4e-15 (fp64) of the
    field's magnitude, plus the same relative bound on the result itself."""
    assert got.dtype == exp.dtype, (got.dtype, exp.dtype)
    assert got.shape == exp.shape
    eps = np.finfo(exp.dtype).eps if eps is None else eps
    if scale is None:
        scale = np.nanmax(np.abs(exp)) if np.isfinite(exp).any() else 1.0
    np.testing.assert_array_equal(np.isnan(got), np.isnan(exp))
    np.testing.assert_allclose(got, exp, rtol=16 * eps, atol=16 * eps * scale, equal_nan=True)


# ------------------------------------------------------------------ weighted averages
@pytest.mark.parametrize("f,ny,nx", [(8, 48, 64), (8, 16, 24), (4, 48, 48), (2, 4, 4), (16, 32, 32)])
@pytest.mark.parametrize("dt,wdt", [("f4", "f4"), ("f8", "f4"), ("f8", "f8"), ("f4", "f8")])
def test_weighted_block_average_2d_weights(ops, f, ny, nx, dt, wdt):
    rng = np.random.RandomState(0)
    obj = rng.uniform(-1000, 1000, (3, 5, ny, nx)).astype(dt)
    area = rng.uniform(0.5, 1, (3, ny, nx)).astype(wdt)
    got = host(ops.weighted_block_average(dev(obj), w2d=dev(area), fy=f, fx=f))
    # the denominator is summed in the weights' dtype (as numpy does), so its eps sets the bar
    eps = max(np.finfo(dt).eps, np.finfo(wdt).eps)
    assert_sum_close(got, oracle.weighted_block_average(obj, area[:, None], f), scale=1000.0, eps=eps)


def test_weighted_block_average_config1(ops):
    """BASELINE config 1: C48 -> C12, f=4, 79 levels, 6 tiles, seed 0."""
    np.random.seed(0)
    obj = np.random.uniform(-1000, 1000, (6, 79, 48, 48)).astype(np.float32)
    np.random.seed(0)
    area = np.random.uniform(0.5, 1, (6, 48, 48)).astype(np.float32)
    got = host(ops.weighted_block_average(dev(obj), w2d=dev(area), fy=4, fx=4))
    assert got.shape == (6, 79, 12, 12)
    assert_sum_close(got, oracle.weighted_block_average(obj, area[:, None], 4), scale=1000.0)


@pytest.mark.parametrize("f,n", [(8, 32), (2, 6)])
def test_weighted_block_average_mass_multi_field(ops, f, n):
    rng = np.random.RandomState(1)
    fields = [rng.uniform(-10, 10, (2, 7, n, n)).astype(np.float32) for _ in range(5)]
    delp = rng.uniform(3, 5, (2, 7, n, n)).astype(np.float32)
    area = rng.uniform(0.5, 1, (2, n, n)).astype(np.float32)
    got = ops.weighted_block_average([dev(x) for x in fields], w2d=dev(area), w3d=dev(delp), fy=f, fx=f)
    mass = delp * area[:, None]
    for g, x in zip(got, fields):
        assert_sum_close(host(g), oracle.weighted_block_average(x, mass, f), scale=10.0)
    got3 = ops.weighted_block_average([dev(x) for x in fields], w3d=dev(mass), fy=f, fx=f)
    for g, x in zip(got3, fields):
        assert_sum_close(host(g), oracle.weighted_block_average(x, mass, f), scale=10.0)


@pytest.mark.parametrize("mode", ["area", "mass", "w3d"])
def test_weighted_block_average_f16_warp_path(ops, mode):
    """f = 16 at widths that are multiples of 64: a warp per four cells with shuffle reduction; weights read once per
    block for all fields; NaN products and NaN weights skipped like nansum."""
    rng = np.random.RandomState(16)
    n, f = 128, 16
    fields = [rng.uniform(-10, 10, (2, 3, n, n)).astype(np.float32) for _ in range(4)]
    delp = rng.uniform(3, 5, (2, 3, n, n)).astype(np.float32)
    area = rng.uniform(0.5, 1, (2, n, n)).astype(np.float32)
    fields[0][0, 1, 17, 33] = np.nan
    fields[1][1, 2, 64:80, 16:32] = np.nan              # an all-NaN block: 0 / den
    area[1, 100, 100] = np.nan
    if mode == "area":
        got = ops.weighted_block_average([dev(x) for x in fields], w2d=dev(area), fy=f, fx=f)
        w = np.broadcast_to(area[:, None], delp.shape)
    elif mode == "mass":
        got = ops.weighted_block_average([dev(x) for x in fields], w2d=dev(area), w3d=dev(delp), fy=f, fx=f)
        w = delp * area[:, None]
    else:
        w = delp * area[:, None]
        got = ops.weighted_block_average([dev(x) for x in fields], w3d=dev(w), fy=f, fx=f)
    for g, x in zip(got, fields):
        assert_sum_close(host(g), oracle.weighted_block_average(x, w, f), scale=10.0)
    assert host(got[1])[1, 2, 4, 1] == 0.0
    again = ops.weighted_block_average([dev(x) for x in fields], w2d=dev(area), fy=f, fx=f)
    if mode == "area":
        for a, b in zip(got, again):
            assert torch.equal(torch.nan_to_num(a), torch.nan_to_num(b))      # fixed reduction tree


@pytest.mark.parametrize("f", [8, 4])
def test_weighted_block_average_nan_skipping(ops, f):
    """xarray coarsen().sum() is nansum: NaN products are skipped, the weight stays in the
    denominator; an all-NaN block gives 0/den = 0; NaN weights drop out of both sums."""
    rng = np.random.RandomState(2)
    obj = rng.uniform(-1, 1, (1, 2, 2 * f, 2 * f)).astype(np.float32)
    w = rng.uniform(0.5, 1, (1, 2 * f, 2 * f)).astype(np.float32)
    obj[0, 0, 1, 1] = np.nan
    obj[0, 1, :f, :f] = np.nan
    w[0, f + 1, f + 1] = np.nan
    got = host(ops.weighted_block_average(dev(obj), w2d=dev(w), fy=f, fx=f))
    exp = oracle.weighted_block_average(obj, w[:, None], f)
    assert got[0, 1, 0, 0] == 0.0
    assert_sum_close(got, exp)


@pytest.mark.parametrize("f,nc", [(8, 4), (2, 2)])
@pytest.mark.parametrize("dt", ["f4", "f8"])
def test_edge_weighted_block_average(ops, f, nc, dt):
    rng = np.random.RandomState(3)
    n = f * nc
    u = rng.uniform(-1000, 1000, (6, 3, n + 1, n)).astype(dt)
    dx = rng.uniform(0.5, 1, (6, n + 1, n)).astype("f4")
    v = rng.uniform(-1000, 1000, (6, 3, n, n + 1)).astype(dt)
    dy = rng.uniform(0.5, 1, (6, n, n + 1)).astype("f4")
    gu = host(ops.weighted_block_average(dev(u), w2d=dev(dx), fy=1, fx=f, sy=f, sx=f, nyo=nc + 1, nxo=nc))
    gv = host(ops.weighted_block_average(dev(v), w2d=dev(dy), fy=f, fx=1, sy=f, sx=f, nyo=nc, nxo=nc + 1))
    eps = np.finfo("f4").eps                      # f32 spacing sums set the bar
    assert_sum_close(gu, oracle.edge_weighted_block_average(u, dx[:, None], f, "x"), scale=1000.0, eps=eps)
    assert_sum_close(gv, oracle.edge_weighted_block_average(v, dy[:, None], f, "y"), scale=1000.0, eps=eps)


# ------------------------------------------------------------------ plain reductions
@pytest.mark.parametrize("method", ["sum", "min", "max", "mean"])
@pytest.mark.parametrize("dt", ["f4", "f8"])
def test_block_reduce(ops, method, dt):
    rng = np.random.RandomState(4)
    a = rng.uniform(-5, 5, (4, 24, 16)).astype(dt)
    a[0, 0, 0] = np.nan
    got = host(ops.block_reduce(dev(a), method, 8, 8))
    exp = oracle.block_coarsen(a, 8, method)
    if method in ("min", "max"):
        np.testing.assert_array_equal(got, exp)
    else:
        assert_sum_close(got, exp, scale=5.0 * 64)


@pytest.mark.parametrize("method", ["sum", "min"])
def test_block_edge_coarsen(ops, method):
    rng = np.random.RandomState(5)
    dx = rng.uniform(0.5, 1, (6, 17, 16)).astype("f4")
    dy = rng.uniform(0.5, 1, (6, 16, 17)).astype("f4")
    gx = host(ops.block_reduce(dev(dx), method, 1, 8, 8, 8, 3, 2))
    gy = host(ops.block_reduce(dev(dy), method, 8, 1, 8, 8, 2, 3))
    ex, ey = oracle.block_edge_coarsen(dx, 8, "x", method), oracle.block_edge_coarsen(dy, 8, "y", method)
    if method == "min":
        np.testing.assert_array_equal(gx, ex)
        np.testing.assert_array_equal(gy, ey)
    else:
        assert_sum_close(gx, ex)
        assert_sum_close(gy, ey)


# ------------------------------------------------------------------ selection (bit-exact)
@pytest.mark.parametrize("f,dt", [(8, "f4"), (16, "f4"), (4, "f4"), (2, "f4"), (2, "f8"), (4, "f8"),
                                  (8, "f8"), (3, "f4"), (5, "f8")])
def test_block_median_bit_exact(ops, f, dt):
    rng = np.random.RandomState(6)
    a = rng.uniform(-5, 5, (5, 6 * f, 4 * f)).astype(dt)
    a[0, 1, 1] = np.nan                       # poisons exactly one block
    a[1, :f, :f] = 1.0                        # all ties
    a[2, f:2 * f, :f] = np.round(a[2, f:2 * f, :f])
    a[3, :f, :f] = np.inf
    a[4, 0, 0] = -np.inf
    got = host(ops.block_median(dev(a), f))
    exp = oracle.block_median(a, f)
    assert got.dtype == exp.dtype
    np.testing.assert_array_equal(got, exp)


def test_block_median_unaligned_width(ops):
    rng = np.random.RandomState(7)
    a = rng.uniform(-5, 5, (1, 16, 24)).astype("f4")[:, :, :]
    t = dev(np.concatenate([np.zeros(1, "f4"), a.ravel()]))[1:].view(1, 16, 24)  # misaligned base
    np.testing.assert_array_equal(host(ops.block_median(t.contiguous(), 8)), oracle.block_median(a, 8))


def test_block_median_golden_reference_block_reduce(ops, golden_dir):
    g = np.load(f"{golden_dir}/skimage_block_reduce.npz")
    for key in g.files:
        if key.startswith("median_in_"):
            f = int(key.split("_")[2][1:])
            got = host(ops.block_median(dev(g[key]), f))
            np.testing.assert_array_equal(got, g[key.replace("_in_", "_out_")])


@pytest.mark.parametrize("policy", ["propagate", "omit"])
@pytest.mark.parametrize("f", [2, 8])
def test_block_mode(ops, policy, f):
    rng = np.random.RandomState(8)
    a = rng.randint(0, 4, (3, 4 * f, 4 * f)).astype("f4")
    a[0, 0, 0] = np.nan
    a[1, :f, :f] = np.nan
    got = host(ops.block_mode(dev(a), f, policy))
    exp = oracle.block_mode(a, f, policy)
    np.testing.assert_array_equal(got, exp)


@pytest.mark.parametrize("policy", ["propagate", "omit"])
@pytest.mark.parametrize("f,dt", [(2, "f4"), (4, "f4"), (8, "f4"), (2, "f8"), (4, "f8"), (3, "f4"), (8, "f8"), (16, "f4")])
def test_block_mode_ties_nans_infs(ops, policy, f, dt):
    """Register-sort kernel (f = 2, 4, 8) and the warp-per-block one (other sizes, f = 8 fp64): ties go to the smallest
    value, NaN wins under 'propagate' only when strictly most frequent, +inf is an ordinary value."""
    rng = np.random.RandomState(80 + f)
    a = rng.randint(-2, 3, (5, 6 * f, 5 * f)).astype(dt)
    a[rng.uniform(size=a.shape) < 0.15] = np.nan
    a[rng.uniform(size=a.shape) < 0.10] = np.inf
    a[0, :f, :f] = np.nan                               # all NaN
    a[0, :f, f:2 * f] = np.inf                          # all +inf
    a[1, :f, :f] = np.nan
    a[1, 0, 0] = 1.0                                    # NaN strictly most frequent
    a[2, :f, :f] = np.where(np.arange(f * f).reshape(f, f) % 2 == 0, np.nan, 7.0)   # NaN ties with a value
    got = host(ops.block_mode(dev(a), f, policy))
    exp = oracle.block_mode(a, f, policy)
    np.testing.assert_array_equal(got, exp)


@pytest.mark.parametrize("stag_y,stag_x", [(False, False), (True, False), (False, True), (True, True)])
@pytest.mark.parametrize("dt", ["f4", "f8"])
def test_block_upsample(ops, stag_y, stag_x, dt):
    rng = np.random.RandomState(9)
    a = rng.uniform(-1, 1, (7, 4 + stag_y, 6 + stag_x)).astype(dt)
    got = host(ops.block_upsample(dev(a), 8, 8, stag_y, stag_x))
    np.testing.assert_array_equal(got, oracle.block_upsample(a, 8, axes=(-1, -2)))


# ------------------------------------------------------------------ pressure helpers
@pytest.mark.parametrize("dt", ["f4", "f8"])
def test_pressure_at_interface_bit_exact(ops, dt):
    rng = np.random.RandomState(10)
    delp = rng.uniform(100, 2000, (3, 79, 8, 16)).astype(dt)
    got = host(ops.pressure_at_interface(dev(delp)))
    np.testing.assert_array_equal(got, oracle.pressure_at_interface(delp, axis=1))


def test_mask_weights_bit_exact(ops):
    rng = np.random.RandomState(11)
    w = rng.uniform(0.5, 1, (2, 8, 8)).astype("f4")
    pc = np.sort(rng.uniform(300, 1e5, (2, 10, 8, 8)).astype("f4"), axis=1)
    pf = np.sort(rng.uniform(300, 1e5, (2, 10, 8, 8)).astype("f4"), axis=1)
    got = host(ops.mask_weights(dev(w), dev(pc), dev(pf)))
    np.testing.assert_array_equal(got, oracle.mask_weights(w[:, None], pc, pf, axis=1))


@pytest.mark.parametrize("axis", ["x", "y"])
@pytest.mark.parametrize("dt", ["f4", "f8"])
def test_interp_to_edges(ops, axis, dt):
    rng = np.random.RandomState(12)
    c = rng.uniform(3, 5, (6, 3, 8, 8)).astype(dt)
    got = host(ops.interp_to_edges(dev(c), axis))
    np.testing.assert_array_equal(got, recipes.interp_to_edges(c, axis).astype(dt))


def test_multicast_store_copies_16_byte_pieces(ops):
    """cg_multicast_store: every 16 bytes of a region stored once at region + offset.  On a multi-GPU box the offset
    leads to the NVLS multicast mapping of a symmetric buffer (bench.py's strong-scaling exchange checks that end
    bit for bit); here it leads to plain memory of the same device."""
    n = 4 * 100003
    whole = torch.full((2 * n + 8,), -3.0, dtype=torch.float32, device="cuda")
    whole[:n] = torch.arange(n, dtype=torch.float32, device="cuda")
    ops.multicast_store(whole[:n], 4 * (n + 4))
    w = whole.cpu().numpy()
    np.testing.assert_array_equal(w[n + 4:2 * n + 4], np.arange(n, dtype=np.float32))
    assert (w[n:n + 4] == -3.0).all() and (w[2 * n + 4:] == -3.0).all()
    with pytest.raises(ValueError):
        ops.multicast_store(whole[:n], 4)                      # not a multiple of 16
    with pytest.raises(ValueError):
        ops.multicast_store(whole[1:n + 1], 16)                # misaligned region
