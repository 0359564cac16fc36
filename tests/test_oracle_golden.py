"""Pins the CPU oracle against the reference's own known answers and golden files
(CPU only).  If these fail, no parity claim about the CUDA path means anything."""
import numpy as np
import pytest

import oracle
from oracle import recipes
from tests import synth_inputs as S

F = 2  # external/vcm/tests/test_coarsen_restarts.py:23


# ---- external/vcm/tests/test_mappm.py ------------------------------------------------
def test_mappm_interior():
    r = oracle.mappm(np.array([[0.0, 1, 2, 3, 4, 5]]), np.array([[0.0, 1, 2, 3, 4]]),
                     np.array([[0.5, 1.2, 2.4, 2.8, 3.2, 4.5]]), 1, 1.0, 1.0, 1.0, 0.0)
    assert r.dtype == np.float32
    np.testing.assert_almost_equal(r, np.array([[0.35, 1.3, 2.1, 2.5, 3.35]], np.float32), decimal=5)


def test_mappm_out_of_bounds():
    r = oracle.mappm(np.array([[1.0, 2, 3, 4, 5]]), np.array([[1.5, 2.5, 3.5, 4.5]]),
                     np.array([[0.0, 2.5, 3.5, 4.5, 50.0]]), 1, 1.0, 1.0, 1.0, 0.0)
    np.testing.assert_almost_equal(r, np.array([[1.5, 3.0, 4.0, 4.502747]], np.float32), decimal=5)


def test_mappm_nans():
    r = oracle.mappm(np.array([[1.0, 2, 3, 2, 5]]), np.full((1, 4), np.nan),
                     np.array([[0.0, 2.5, 3.5, 4.5, 50.0]]))
    assert np.isnan(r).all()


def test_mappm_f64_close_to_f32():
    rng = np.random.RandomState(1)
    dp = rng.uniform(3, 5, (50, 9))
    pe1 = np.concatenate([np.full((50, 1), 300.0), 300 + np.cumsum(dp, 1)], 1)
    pe2 = np.concatenate([np.full((50, 1), 300.0), 300 + np.cumsum(rng.uniform(3, 5, (50, 9)), 1)], 1)
    q = rng.uniform(-1000, 1000, (50, 9))
    a = oracle.mappm(pe1, q, pe2)
    b = oracle.mappm(pe1, q, pe2, dtype=np.float64)
    assert np.abs(a - b).max() < 3e-5 * 1000 * 10


# ---- kord > 7 (cs_profile): no reference test exists -> properties only ("parity unpinned")
def _two_grids(n, km, kn, seed):
    rng = np.random.RandomState(seed)
    dp = rng.uniform(1, 1.5, (n, km))
    pe1 = np.concatenate([np.full((n, 1), 3.0), 3 + np.cumsum(dp, 1)], 1)
    w = rng.uniform(1, 1.5, (n, kn))
    w = w / w.sum(1, keepdims=True) * (pe1[:, -1:] - 3)
    pe2 = np.concatenate([np.full((n, 1), 3.0), 3 + np.cumsum(w, 1)], 1)
    pe2[:, -1] = pe1[:, -1]
    pe2[:, 0] = 3.000001             # a target interface AT the source top takes q1(1) (mappm.f90:62-64)
    return pe1, pe2


@pytest.mark.parametrize("kord", [8, 9, 10, 11, 12, 13, 14, 15, 16, 17])
@pytest.mark.parametrize("iv", [-1, 0, 1, 2])
def test_cs_profile_properties(kord, iv):
    pe1, pe2 = _two_grids(64, 24, 19, kord)
    rng = np.random.RandomState(iv + 5)
    q = rng.uniform(-1, 1, (64, 24))
    if iv == 0:
        q = np.abs(q)
    out = oracle.mappm(pe1, q, pe2, iv=iv, kord=kord, dtype=np.float64)
    assert not np.isnan(out).any()
    # conservative: the column integral is kept (up to the 1e-6 sliver cut off the top)
    m1, m2 = (q * np.diff(pe1, axis=1)).sum(1), (out * np.diff(pe2, axis=1)).sum(1)
    np.testing.assert_allclose(m2, m1, atol=2e-5)
    # constants are reproduced
    c = oracle.mappm(pe1, np.full_like(q, 2.5), pe2, iv=iv, kord=kord, dtype=np.float64)
    np.testing.assert_allclose(c, 2.5, rtol=1e-12)
    if iv == 0 and kord <= 16:       # positive definite (kord > 16 is the unconstrained spline, :204-213)
        assert out.min() >= 0.0


@pytest.mark.parametrize("kord", [8, 9, 10, 11, 13, 14, 15, 16, 17])
def test_cs_profile_is_a_spline_on_smooth_data(kord):
    """Layer means of sin(p / L) remapped to another grid stay close to that grid's layer means:
    the tridiagonal solve of mappm.f90:175-201 is 4th-order, an order better than the PPM edges
    of kord <= 7 (kord = 12 flattens every extremum and is left out)."""
    pe1, pe2 = _two_grids(32, 40, 37, 3)
    L = 25.0

    def mean(pe):
        return (-L * np.cos(pe[:, 1:] / L) + L * np.cos(pe[:, :-1] / L)) / np.diff(pe, axis=1) + 2.0

    out = oracle.mappm(pe1, mean(pe1), pe2, iv=1, kord=kord, dtype=np.float64)
    ppm = oracle.mappm(pe1, mean(pe1), pe2, iv=1, kord=1, dtype=np.float64)
    err, err_ppm = np.abs(out - mean(pe2))[:, 3:-3].max(), np.abs(ppm - mean(pe2))[:, 3:-3].max()
    assert err < 2e-5 and err < 0.1 * err_ppm


def test_cs_profile_iv_minus2_is_refused():
    pe1, pe2 = _two_grids(2, 8, 8, 0)
    with pytest.raises(NotImplementedError):
        oracle.mappm(pe1, np.zeros((2, 8)), pe2, iv=-2, kord=9)


# ---- unit values of external/vcm/tests/test_cubedsphere.py ---------------------------
def test_weighted_block_average_unit():           # :206-236  -> 3.0
    data = np.array([[2.0, 6.0], [6.0, 2.0]], np.float32)
    w = np.array([[6.0, 2.0], [2.0, 6.0]], np.float32)
    np.testing.assert_allclose(oracle.weighted_block_average(data, w, 2), [[3.0]])


def test_edge_weighted_block_average_unit():      # :239-258 (test dims are (x, y); ours (y, x))
    data = np.array([[2, 6, 2], [6, 2, 6]], np.float64)          # (x=2, y=3)
    spacing = np.array([[6, 2, 6], [2, 6, 2]], np.float64)
    out = oracle.edge_weighted_block_average(data.T, spacing.T, 2, edge="x")
    np.testing.assert_array_equal(out.T, [[3.0, 3.0]])
    data = np.array([[2, 6], [6, 2], [2, 6]], np.float64)        # (x=3, y=2)
    spacing = np.array([[6, 2], [2, 6], [6, 2]], np.float64)
    out = oracle.edge_weighted_block_average(data.T, spacing.T, 2, edge="y")
    np.testing.assert_array_equal(out.T, [[3.0], [3.0]])
    with pytest.raises(ValueError, match="'edge' most be either"):
        oracle.edge_weighted_block_average(data, spacing, 2, edge="z")


def test_block_edge_sum_and_min_unit():           # :386-419
    xy = np.array([[2, 6, 2], [6, 2, 6]], np.float64)            # (x=2, y=3)
    np.testing.assert_array_equal(oracle.block_edge_coarsen(xy.T, 2, "x", "sum").T, [[8, 8]])
    np.testing.assert_array_equal(oracle.block_edge_coarsen(xy.T, 2, "x", "min").T, [[2, 2]])
    xy = np.array([[2, 6], [6, 2], [2, 6]], np.float64)          # (x=3, y=2)
    np.testing.assert_array_equal(oracle.block_edge_coarsen(xy.T, 2, "y", "sum").T, [[8], [8]])
    np.testing.assert_array_equal(oracle.block_edge_coarsen(xy.T, 2, "y", "min").T, [[2], [2]])


def test_block_upsample_staggered():              # :771-816
    a = np.array([[1.0, 2.0, 3.0]])                               # odd -> staggered
    np.testing.assert_array_equal(oracle.block_upsample(a, 2, axes=(-1,)), [[1, 1, 2, 2, 3]])
    b = np.array([[1.0, 2.0]])
    np.testing.assert_array_equal(oracle.block_upsample(b, 2, axes=(-1,)), [[1, 1, 2, 2]])


def test_coords():                                # coarsen.py:128-132
    c = np.arange(1, 9, dtype=np.float32).reshape(4, 2)
    np.testing.assert_array_equal(oracle.coarsen_coords_coord_func(c), [1, 2, 3, 4])
    c = np.arange(49, 97, dtype=np.float32).reshape(6, 8)          # 2nd subtile of 3
    np.testing.assert_array_equal(oracle.coarsen_coords_coord_func(c), [7, 8, 9, 10, 11, 12])


def test_mask_weights_truth_table():              # test_regridz.py:113-128
    w = np.arange(4, dtype=np.float32).reshape(2, 2)[None]        # (1, y, x)
    pc = np.stack([np.full((2, 2), v, np.float32) for v in (1.0, 2.0, 3.0)])[None]
    pf = pc.copy()
    pf[0, 2, 0, :] = 2.5
    pf[0, 2, 1, :] = 3.5
    m = oracle.mask_weights(w[:, None], pc, pf, axis=1)
    exp = np.broadcast_to(w[:, None], (1, 2, 2, 2)).copy()
    exp[0, 1, 0, :] = 0.0
    np.testing.assert_array_equal(m, exp)


# ---- the reference's own block_reduce outputs (tests/golden/skimage_block_reduce.npz) -
def test_block_median_matches_reference_block_reduce(golden_dir):
    g = np.load(f"{golden_dir}/skimage_block_reduce.npz")
    a = g["regtest_in"]                                           # (4,4,2), blocks (2,2,1)
    got = oracle.block_median(np.moveaxis(a, -1, 0), 2)
    np.testing.assert_array_equal(np.moveaxis(got, 0, -1), g["regtest_median"])
    for key in g.files:
        if key.startswith("median_in_"):
            f = int(key.split("_")[2][1:])
            out = oracle.block_median(g[key], f)
            exp = g[key.replace("_in_", "_out_")]
            assert out.dtype == exp.dtype
            np.testing.assert_array_equal(out, exp)               # bit-exact incl. NaN


# ---- golden JSONs of test_coarsen_restarts.py ----------------------------------------
def _core_inputs():
    delp, _ = S.generate("fv_core.res", "delp")
    area, _ = S.generate("grid", "area")
    dx, _ = S.generate("grid", "dx")
    dy, _ = S.generate("grid", "dy")
    return delp, area, dx, dy


def _close(got, tag, cat, var):
    exp, _ = S.golden(tag, cat, var)
    np.testing.assert_allclose(got, exp, rtol=1e-5, atol=0,
                               err_msg=f"{tag}/{cat}/{var}")      # xr.testing.assert_allclose default


@pytest.mark.parametrize("tag,mass", [("mass-weighted-model-level-with-agrid-winds", True),
                                      ("area-weighted-model-level-without-agrid-winds", False)])
def test_golden_model_level(tag, mass):           # coarsen_restarts.py:316-408
    delp, area, dx, dy = _core_inputs()
    for var in ("delp", "DZ"):
        x, _ = S.generate("fv_core.res", var)
        _close(oracle.weighted_block_average(x, area[:, None], F), tag, "fv_core.res", var)
    phis, _ = S.generate("fv_core.res", "phis")
    _close(oracle.weighted_block_average(phis, area, F), tag, "fv_core.res", "phis")
    w = delp * area[:, None] if mass else area[:, None]
    for var in ("W", "T"):
        x, _ = S.generate("fv_core.res", var)
        _close(oracle.weighted_block_average(x, w, F), tag, "fv_core.res", var)
    u, _ = S.generate("fv_core.res", "u")
    v, _ = S.generate("fv_core.res", "v")
    _close(oracle.edge_weighted_block_average(u, dx[:, None], F, "x"), tag, "fv_core.res", "u")
    _close(oracle.edge_weighted_block_average(v, dy[:, None], F, "y"), tag, "fv_core.res", "v")


def test_golden_pressure_level_centered():        # coarsen_restarts.py:411-516
    tag = "pressure-level-without-agrid-winds"
    delp, area, dx, dy = _core_inputs()
    _close(oracle.weighted_block_average(delp, area[:, None], F), tag, "fv_core.res", "delp")
    T, _ = S.generate("fv_core.res", "T")
    (Tr,), masked = recipes.regrid_to_area_weighted_pressure([T], delp, area, F)
    assert Tr.dtype == np.float32
    got = oracle.weighted_block_average(Tr, masked, F)
    exp, _ = S.golden(tag, "fv_core.res", "T")
    # mappm is fp32; the reference's own platform-to-platform scatter is ~1.4e-5 of
    # the field scale, so compare on the field scale (|T| <= 1000).
    assert np.abs(got - exp).max() < 3e-5 * 1000
    assert (np.abs(got - exp) <= 1e-5 * np.abs(exp)).mean() > 0.97


def test_golden_pressure_level_staggered():       # regridz.py:58-112 incl. xgcm halo
    tag = "pressure-level-without-agrid-winds"
    delp, area, dx, dy = _core_inputs()
    u, _ = S.generate("fv_core.res", "u")
    v, _ = S.generate("fv_core.res", "v")
    (ur,), mdx = recipes.regrid_to_edge_weighted_pressure([u], delp, dx, F, edge="x")
    (vr,), mdy = recipes.regrid_to_edge_weighted_pressure([v], delp, dy, F, edge="y")
    gu = oracle.edge_weighted_block_average(ur, mdx, F, "x")
    gv = oracle.edge_weighted_block_average(vr, mdy, F, "y")
    eu, _ = S.golden(tag, "fv_core.res", "u")
    ev, _ = S.golden(tag, "fv_core.res", "v")
    assert np.abs(gu - eu).max() < 3e-5 * 1000
    assert np.abs(gv - ev).max() < 3e-5 * 1000


# ---- external/vcm/tests/test_interpolate.py:109-132 (mappm.interpolate_2d) ----------------
def test_interpolate_2d_matches_reference_test():
    import scipy.interpolate

    x = np.arange(10).reshape((1, 10))
    y = (x ** 2).reshape((1, 10))
    xp = np.arange(12).reshape((1, 12))
    expected = np.zeros_like(xp, dtype=np.float64)
    for i in range(xp.shape[0]):
        expected[i] = scipy.interpolate.interp1d(x[i], y[i], bounds_error=False)(xp[i])
    assert np.isnan(expected[:, -2:]).all()
    np.testing.assert_allclose(expected, oracle.interpolate_2d(xp, x, y))


def test_interpolate_2d_random_rows_match_scipy():
    import scipy.interpolate

    rng = np.random.RandomState(0)
    x = np.cumsum(rng.uniform(0.1, 1, (20, 30)), axis=1)
    y = rng.uniform(-5, 5, (20, 30))
    xp = rng.uniform(-1, x.max() + 1, (20, 17))
    xp[:, 3] = x[:, 7]                                       # exact hits
    xp[:, 4] = x[:, -1]                                      # the last point is included (:23-24)
    got = oracle.interpolate_2d(xp, x, y, fill_value=-99.0)
    for i in range(20):
        exp = scipy.interpolate.interp1d(x[i], y[i], bounds_error=False, fill_value=-99.0)(xp[i])
        np.testing.assert_allclose(got[i], exp, rtol=1e-12, atol=1e-12)
