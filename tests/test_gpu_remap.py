"""GPU parity: PPM vertical remap (cg_mappm) and pressure-level coarse-graining.

cg_mappm is compiled without FMA contraction and follows the Fortran statement order,
so it must equal the fp32 oracle BIT FOR BIT (NaNs in the same places).  The
pressure-level averages sum 64 products in a different order than numpy, so they are
held to |d| <= 3e-5 * max|q| (the north_star tolerance for the remap) and to a relative
2e-6 where the remap itself is bit-identical (composed path).
"""
import os

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

import oracle  # noqa: E402
from oracle import recipes  # noqa: E402
from fv3net_b200 import synthetic as syn  # noqa: E402
from tests import synth_inputs as S  # noqa: E402


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


def gpu_mappm_zlast(ops, pe1, q1, pe2, iv=1, kord=1):
    ncol, km1 = pe1.shape
    km, kn = km1 - 1, pe2.shape[1] - 1
    out = torch.empty((ncol, kn), dtype=torch.float32, device="cuda")
    ops.mappm(dev(pe1.astype("f4")), dev(q1.astype("f4")), dev(pe2.astype("f4")), 1, ncol, km, kn,
              (0, km + 1, 1), (0, km, 1), (0, kn + 1, 1), (0, kn, 1), out, iv=iv, kord=kord)
    return host(out)


# ---- the reference's own known answers, through the CUDA kernel (test_mappm.py) ----------
def test_mappm_known_answers(ops):
    r = gpu_mappm_zlast(ops, np.array([[0.0, 1, 2, 3, 4, 5]]), np.array([[0.0, 1, 2, 3, 4]]),
                        np.array([[0.5, 1.2, 2.4, 2.8, 3.2, 4.5]]))
    np.testing.assert_almost_equal(r, np.array([[0.35, 1.3, 2.1, 2.5, 3.35]], "f4"), decimal=5)
    r = gpu_mappm_zlast(ops, np.array([[1.0, 2, 3, 4, 5]]), np.array([[1.5, 2.5, 3.5, 4.5]]),
                        np.array([[0.0, 2.5, 3.5, 4.5, 50.0]]))
    np.testing.assert_almost_equal(r, np.array([[1.5, 3.0, 4.0, 4.502747]], "f4"), decimal=5)
    r = gpu_mappm_zlast(ops, np.array([[1.0, 2, 3, 2, 5]]), np.full((1, 4), np.nan),
                        np.array([[0.0, 2.5, 3.5, 4.5, 50.0]]))
    assert np.isnan(r).all()


def _random_columns(rng, ncol, km, kn, spread=0.1):
    dp1 = rng.uniform(50, 1500, (ncol, km))
    pe1 = np.concatenate([np.full((ncol, 1), 300.0), 300 + np.cumsum(dp1, 1)], 1)
    scale = 1 + spread * rng.uniform(-1, 1, (ncol, 1))
    dp2 = rng.uniform(50, 1500, (ncol, kn))
    dp2 *= scale * dp1.sum(1, keepdims=True) / dp2.sum(1, keepdims=True)
    pe2 = np.concatenate([np.full((ncol, 1), 300.0), 300 + np.cumsum(dp2, 1)], 1)
    q = rng.uniform(-1000, 1000, (ncol, km))
    q[: ncol // 8, km // 3: km // 2] = 0.0          # flat stretches (dm == 0 branch)
    q[ncol // 8: ncol // 4] = np.abs(q[ncol // 8: ncol // 4])  # positive tracer-like
    return pe1.astype("f4"), q.astype("f4"), pe2.astype("f4")


@pytest.mark.parametrize("km,kn", [(79, 79), (5, 5), (4, 3), (5, 2), (32, 79), (100, 64)])
@pytest.mark.parametrize("iv,kord", [(1, 1), (0, 1), (-1, 1), (2, 1), (-2, 1), (1, 4), (0, 4),
                                     (1, 5), (1, 6), (0, 6), (1, 7), (0, 7), (-1, 7), (1, 3)])
def test_mappm_bit_exact_vs_oracle(ops, km, kn, iv, kord):
    rng = np.random.RandomState(km * 100 + kn + kord)
    pe1, q, pe2 = _random_columns(rng, 257, km, kn)
    got = gpu_mappm_zlast(ops, pe1, q, pe2, iv, kord)
    exp = oracle.mappm(pe1, q, pe2, iv=iv, kord=kord)
    np.testing.assert_array_equal(got, exp)


# ---- the streaming remapper (kord <= 6, km >= 5) against the generic kernel and the oracle ----------------------
@pytest.mark.parametrize("iv,kord", [(1, 1), (0, 2), (-1, 3), (2, 4), (-2, 5), (1, 6), (0, 5)])
@pytest.mark.parametrize("km,kn", [(79, 79), (5, 7), (6, 3), (33, 80)])
def test_mappm_stream_ties_and_ends(ops, km, kn, iv, kord):
    """Target interfaces that coincide with source interfaces, targets above the column top and below its bottom,
    one target layer that spans the whole column: the forward merge takes the Fortran's branch at every tie."""
    from fv3net_b200 import _lib
    rng = np.random.RandomState(17 * km + kn + kord)
    pe1, q, pe2 = _random_columns(rng, 300, km, kn, spread=0.3)
    for c in range(0, 300, 3):                         # ties: copy some source interfaces into the target column
        idx = rng.randint(0, km + 1, size=kn + 1)
        pe2[c] = np.sort(np.where(rng.uniform(size=kn + 1) < 0.5, pe1[c, idx], pe2[c]))
    pe2[5::7, 0] = 100.0                               # target top above the column top
    pe2[6::7, :2] = [50.0, 200.0]                      # a whole target layer above the column
    pe2[4::11, -1] = pe1[4::11, -1] + 5000.0           # target reaching far below the column
    pe2[9::13, 1:] = np.maximum(pe2[9::13, 1:], pe1[9::13, -1:] + 10.0)   # one layer spans everything
    pe2 = np.sort(pe2, axis=1)
    got = gpu_mappm_zlast(ops, pe1, q, pe2, iv, kord)
    _lib.set_option("mappm_impl", "generic")
    gen = gpu_mappm_zlast(ops, pe1, q, pe2, iv, kord)
    np.testing.assert_array_equal(got, gen)
    np.testing.assert_array_equal(got, oracle.mappm(pe1, q, pe2, iv=iv, kord=kord))


def test_mappm_stream_hands_bad_columns_to_the_generic_kernel(ops):
    """Non-increasing or NaN pressures: undefined in the Fortran, defined by the generic kernel (NaN where no layer
    brackets).  The streaming kernel flags such column groups and the generic kernel redoes them: same bits."""
    from fv3net_b200 import _lib
    rng = np.random.RandomState(99)
    pe1, q, pe2 = _random_columns(rng, 1000, 40, 37)
    pe1[3, 10] = pe1[3, 9]                              # zero thickness
    pe1[130, 20] = pe1[130, 18]                         # decreasing source pressure
    pe1[260, 5] = np.nan
    pe2[400, 8] = pe2[400, 6]                           # decreasing target pressure
    pe2[530, 3] = np.nan
    q[777, 11] = np.nan                                 # a NaN VALUE is ordinary data: no flag needed, same propagation
    got = gpu_mappm_zlast(ops, pe1, q, pe2, 1, 1)
    _lib.set_option("mappm_impl", "generic")
    gen = gpu_mappm_zlast(ops, pe1, q, pe2, 1, 1)
    np.testing.assert_array_equal(got, gen)
    assert np.isnan(got[777]).any() and np.isfinite(got[778]).all()


@pytest.mark.parametrize("km,kn", [(79, 79), (4, 3), (6, 9), (32, 79)])
@pytest.mark.parametrize("kord", [8, 9, 10, 11, 12, 13, 14, 15, 16, 17])
@pytest.mark.parametrize("iv", [1, 0, -1, 2])
def test_mappm_cs_profile_bit_exact_vs_oracle(ops, km, kn, iv, kord):
    """kord > 7: cs_profile / cs_limiters (mappm.f90:132-611).  The oracle for these orders is
    pinned only by properties (tests/test_oracle_golden.py): no reference test exists."""
    rng = np.random.RandomState(km * 100 + kn + kord)
    pe1, q, pe2 = _random_columns(rng, 129, km, kn)
    got = gpu_mappm_zlast(ops, pe1, q, pe2, iv, kord)
    exp = oracle.mappm(pe1, q, pe2, iv=iv, kord=kord)
    np.testing.assert_array_equal(got, exp)


@pytest.mark.parametrize("kord", [1, 7, 9, 13])
def test_mappm_shared_memory_variant(ops, monkeypatch, kord):
    """km <= 96 keeps the column's work arrays in local memory; deeper columns (and the option mappm_arrays = shared)
    use the shared-memory variant.  Both are the same code over two array homes: same bits."""
    rng = np.random.RandomState(40 + kord)
    pe1, q, pe2 = _random_columns(rng, 200, 79, 79)
    local = gpu_mappm_zlast(ops, pe1, q, pe2, 1, kord)
    from fv3net_b200 import _lib
    _lib.set_option("mappm_arrays", "shared")
    shared = gpu_mappm_zlast(ops, pe1, q, pe2, 1, kord)
    np.testing.assert_array_equal(local, shared)
    np.testing.assert_array_equal(shared, oracle.mappm(pe1, q, pe2, iv=1, kord=kord))


def test_mappm_tile_layout_and_shared_target(ops):
    """(tile, z, y, x) layout addressed by strides; cs_pe2 = 0 shares one target column."""
    rng = np.random.RandomState(3)
    nt, km, ny, nx = 2, 79, 8, 16
    ps = syn.surface_pressure((nt, ny, nx), seed=1)
    delp = syn.delp_from_ps(ps)
    pe1 = oracle.pressure_at_interface(delp, axis=1)
    q = rng.uniform(-50, 50, (nt, km, ny, nx)).astype("f4")
    pe2_col = pe1[:, :, 3, 5].copy()                                  # (nt, km+1)
    h = ny * nx
    out = torch.empty((nt, km, ny, nx), dtype=torch.float32, device="cuda")
    ops.mappm(dev(pe1), dev(q), dev(pe2_col), nt, h, km, km,
              ((km + 1) * h, 1, h), (km * h, 1, h), (km + 1, 0, 1), (km * h, 1, h), out)
    pe2_full = np.broadcast_to(pe2_col[:, :, None, None], pe1.shape)
    exp = recipes.regrid_vertical(pe1, q, pe2_full)
    np.testing.assert_array_equal(host(out), exp)


def test_mappm_errors(ops):
    z = torch.zeros((1, 4), dtype=torch.float32, device="cuda")
    with pytest.raises(ValueError, match="km >= 4"):
        ops.mappm(z, z, z, 1, 1, 3, 3, (0, 4, 1), (0, 3, 1), (0, 4, 1), (0, 3, 1), z)
    with pytest.raises(NotImplementedError, match="kord"):       # cs_profile would read mappm's unset `qs`
        ops.mappm(z, z, z, 1, 1, 4, 3, (0, 5, 1), (0, 4, 1), (0, 4, 1), (0, 3, 1), z, iv=-2, kord=9)


# ---- pressure-level coarse-graining --------------------------------------------------------
def _plevel_case(nt, nc, f, seed, nfields=3):
    ny = nx = nc * f
    ps = syn.surface_pressure((nt, ny, nx), seed=seed)
    delp = syn.delp_from_ps(ps)
    area = syn.area((nt, ny, nx), seed=seed + 1)
    fields = [syn.field((nt, syn.NZ, ny, nx), seed + 10 + i, scale=10.0 ** i) for i in range(nfields)]
    return fields, delp, area


_PINNED = {"impl": None}


def _check_ran(f, nz):
    """A pinned kernel must be the one that ran wherever its shape constraints hold (no silent fall-back)."""
    from fv3net_b200 import _lib
    impl = _PINNED["impl"]
    ran = _lib.lib().cg_pressure_level_last_impl().decode()
    if impl in ("pair", "pair_nansafe", "default") and f == 8 and nz <= 100:
        assert ran == "pair", ran
    elif impl in ("stream", "stream_nansafe") and f == 8:
        assert ran == "stream", ran
    elif impl == "composed":
        assert ran == "composed", ran


def _pin_impl(monkeypatch, impl):
    _PINNED["impl"] = impl
    """composed | fused | stream pin one implementation where its shape constraints hold
    (otherwise the next one runs); default = the library's own choice."""
    from fv3net_b200 import _lib
    name = "auto" if impl == "default" else impl.split("_")[0]
    _lib.set_option("plevel_impl", name)
    _lib.set_option("plevel_nansafe", "1" if impl.endswith("_nansafe") else "0")   # the NaN-skipping variant alone


@pytest.mark.parametrize("impl", ["composed", "fused", "stream", "stream_nansafe", "pair", "pair_nansafe", "default"])
@pytest.mark.parametrize("mask_below", [False, True])
@pytest.mark.parametrize("f,nc", [(8, 4), (2, 4), (16, 2)])  # even coarse sizes: odd = staggered in the reference
def test_pressure_level_average_vs_oracle(ops, monkeypatch, impl, mask_below, f, nc):
    _pin_impl(monkeypatch, impl)
    fields, delp, area = _plevel_case(2, nc, f, seed=5)
    outs, dcoarse = ops.pressure_level_average([dev(x) for x in fields], dev(delp), dev(area), f,
                                               mask_below=mask_below)
    _check_ran(f, delp.shape[1])
    dc_exp = oracle.weighted_block_average(delp, area[:, None], f)
    np.testing.assert_allclose(host(dcoarse), dc_exp, rtol=2e-6)
    if mask_below:
        remapped, masked = recipes.regrid_given_delp(fields, delp, dc_exp, area)
        exp = [oracle.weighted_block_average(r, masked, f) for r in remapped]
    else:
        exp = [recipes.pressure_level_average(x, delp, dc_exp, area, f) for x in fields]
    for g, e, x in zip(outs, exp, fields):
        g = host(g)
        scale = np.abs(x).max()
        ok = np.isfinite(e)
        assert np.array_equal(np.isfinite(g), ok)       # fully-masked blocks: 0/0 = NaN both sides
        assert np.abs(g[ok] - e[ok]).max() <= 3e-5 * scale


def _check_plevel(ops, fields, delp, area, f, mask_below, descr=None, tol=3e-5, adjudicate=False):
    """fields: numpy parents; descr: list of ("plain", i) | ("product", i, j) | ("diff", i, j, dt)."""
    d = [dev(x) for x in fields]
    descr = descr or [("plain", i) for i in range(len(fields))]
    args, cpu = [], []
    for e in descr:
        if e[0] == "plain":
            args.append(d[e[1]]); cpu.append(fields[e[1]])
        elif e[0] == "product":
            args.append(ops.product(d[e[1]], d[e[2]])); cpu.append(fields[e[1]] * fields[e[2]])
        else:
            args.append(ops.storage(d[e[1]], d[e[2]], e[3]))
            cpu.append(((fields[e[1]] - fields[e[2]]) / np.float32(e[3])).astype(np.float32))
    outs, dcoarse = ops.pressure_level_average(args, dev(delp), dev(area), f, mask_below=mask_below)
    _check_ran(f, delp.shape[1])
    dc_exp = oracle.weighted_block_average(delp, area[:, None], f)
    np.testing.assert_allclose(host(dcoarse), dc_exp, rtol=2e-6)
    if mask_below:
        remapped, masked = recipes.regrid_given_delp(cpu, delp, dc_exp, area)
        exp = [oracle.weighted_block_average(r, masked, f) for r in remapped]
    else:
        exp = [recipes.pressure_level_average(x, delp, dc_exp, area, f) for x in cpu]
    for g, e, x in zip(outs, exp, cpu):
        g = host(g)
        scale = np.nanmax(np.abs(x))
        ok = np.isfinite(e)
        assert np.array_equal(np.isfinite(g), ok)
        assert np.abs(g[ok] - e[ok]).max() <= tol * scale
        if adjudicate:          # a bar above 3e-5 has to be earned: no further from f64 than 2 x the reference's own fp32
            from tests import parity
            parity.assert_within_fp32_conditioning(g, x, delp, area, f, mask_below, scale=scale, dc_gpu=host(dcoarse))


@pytest.mark.parametrize("impl", ["fused", "stream", "stream_nansafe", "pair", "pair_nansafe"])
@pytest.mark.parametrize("mask_below", [False, True])
def test_pressure_level_budget_shaped_fields(ops, monkeypatch, impl, mask_below):
    """26 output fields like a fine_res_budget timestep: 22 first moments, 2 products and 2
    storage terms formed in the kernel from their parents (budgets.py:80, :151)."""
    _pin_impl(monkeypatch, impl)
    nt, nc, f = 1, 8, 8
    ny = nx = nc * f
    delp = syn.delp_from_ps(syn.surface_pressure((nt, ny, nx), seed=21))
    area = syn.area((nt, ny, nx), seed=22)
    parents = [syn.field((nt, syn.NZ, ny, nx), 30 + i, scale=10.0 ** (i % 4 - 2)) for i in range(22)]
    begin = [parents[1] - 0.01 * syn.field((nt, syn.NZ, ny, nx), 60 + i) for i in range(2)]
    end = [parents[1] + 0.01 * syn.field((nt, syn.NZ, ny, nx), 70 + i) for i in range(2)]
    fields = parents + begin + end
    descr = [("plain", i) for i in range(22)] + [("product", 1, 10), ("product", 0, 10),
                                                  ("diff", 24, 22, 900.0), ("diff", 25, 23, 900.0)]
    _check_plevel(ops, fields, delp, area, f, mask_below, descr)


@pytest.mark.parametrize("impl", ["composed", "stream", "stream_nansafe", "pair", "pair_nansafe"])
@pytest.mark.parametrize("mask_below", [False, True])
@pytest.mark.parametrize("case", ["rough", "steep", "shifted_top"])
def test_pressure_level_hard_columns(ops, monkeypatch, impl, mask_below, case):
    """Fine columns far from their coarse mean: many coarse interfaces inside one fine layer,
    several coarse levels below the fine surface, coarse levels above the fine top."""
    _pin_impl(monkeypatch, impl)
    nt, nc, f = 1, 4, 8
    ny = nx = nc * f
    rng = np.random.RandomState(3)
    if case == "rough":
        ps = 1e5 * (1.0 + 0.25 * rng.uniform(-1, 1, (nt, ny, nx)))
        delp = syn.delp_from_ps(ps)
    elif case == "steep":           # one thick column per block: its layers hold many coarse interfaces
        ps = np.full((nt, ny, nx), 4e4)
        ps[:, ::8, ::8] = 1.05e5
        delp = syn.delp_from_ps(ps)
    else:                           # uneven thickness profile: interfaces wander through many layers
        delp = syn.delp_from_ps(syn.surface_pressure((nt, ny, nx), seed=4))
        delp = (delp * rng.uniform(0.2, 3.0, delp.shape)).astype(np.float32)
    area = syn.area((nt, ny, nx), seed=5)
    fields = [syn.field((nt, syn.NZ, ny, nx), 80 + i, scale=10.0 ** i) for i in range(3)]
    # thin target levels inside thick fine layers divide a small difference of pressures
    # (mappm.f90:76-83): the 1e-6 summation-order difference in the coarse pressures is amplified
    # by p/dp, for the CUDA paths and for the fp32 oracle alike -> 1e-4 here instead of 3e-5
    _check_plevel(ops, fields, delp, area, f, mask_below, tol=1e-4, adjudicate=True)


@pytest.mark.parametrize("impl", ["composed", "stream", "stream_nansafe", "pair", "pair_nansafe"])
def test_pressure_level_nan_columns_are_skipped(ops, monkeypatch, impl):
    """NaN in a fine column: the block sum skips it where the remapped value is NaN
    (xarray's coarsen().sum() is nansum, coarsen.py:211), other levels keep the column."""
    _pin_impl(monkeypatch, impl)
    nt, nc, f = 1, 4, 8
    ny = nx = nc * f
    delp = syn.delp_from_ps(syn.surface_pressure((nt, ny, nx), seed=6))
    area = syn.area((nt, ny, nx), seed=7)
    fld = syn.field((nt, syn.NZ, ny, nx), 90)
    fld[0, 40, 3, 5] = np.nan
    fld[0, 10:14, 9, 17] = np.nan
    area[0, 20, 20] = np.nan
    _check_plevel(ops, [fld], delp, area, f, False)


@pytest.mark.parametrize("impl", ["stream", "stream_nansafe", "pair", "pair_nansafe", "fused"])
@pytest.mark.parametrize("nz", [6, 7, 9, 12, 33, 126])
def test_pressure_level_other_level_counts(ops, monkeypatch, impl, nz):
    """Shallow and deep columns: the end steps of the streaming walk (t = 2, t = nz, the two steps below the last
    level) sit in other places of the 4-level stages than at nz = 79, down to a single stage between them."""
    _pin_impl(monkeypatch, impl)
    nt, nc, f = 1, 4, 8
    ny = nx = nc * f
    rng = np.random.RandomState(100 + nz)
    ps = syn.surface_pressure((nt, ny, nx), seed=nz)
    frac = rng.uniform(0.5, 1.5, nz)
    frac /= frac.sum()
    delp = ((ps[:, None] - 300.0) * frac[None, :, None, None]).astype(np.float32)
    area = syn.area((nt, ny, nx), seed=nz + 1)
    fields = [syn.field((nt, nz, ny, nx), 300 + nz + i, scale=10.0 ** i) for i in range(2)]
    _check_plevel(ops, fields, delp, area, f, mask_below=bool(nz % 2))


@pytest.mark.parametrize("impl", ["composed", "stream", "stream_nansafe", "pair", "pair_nansafe"])
@pytest.mark.parametrize("mask_below", [False, True])
def test_pressure_level_nan_at_the_column_ends(ops, monkeypatch, impl, mask_below):
    """NaN in the top two levels (which only seed the streaming walk), in levels 2 and 3 (read ahead by the top end
    step; through the one-sided cubic they make the edge values of layers 0 and 1 NaN, while the top target level is
    filled from q alone, mappm.f90:62-64) and in the bottom two (one-sided cubic, extension below the surface).
    Near the model top the hybrid levels are pure pressure levels: fine and coarse interfaces coincide up to
    rounding, so whether a NaN layer touches the neighbouring target level hangs on the last bit of delp_coarse -
    the oracle is therefore given the GPU's own delp_coarse (itself checked to 2e-6)."""
    _pin_impl(monkeypatch, impl)
    nt, nc, f = 1, 4, 8
    ny = nx = nc * f
    delp = syn.delp_from_ps(syn.surface_pressure((nt, ny, nx), seed=16))
    area = syn.area((nt, ny, nx), seed=17)
    fld = syn.field((nt, syn.NZ, ny, nx), 91)
    for k, y, x in ((0, 1, 1), (1, 2, 9), (2, 11, 3), (3, 12, 20), (-1, 20, 5), (-2, 25, 30), (0, 30, 14), (1, 30, 14)):
        fld[0, k, y, x] = np.nan
    outs, dcoarse = ops.pressure_level_average([dev(fld)], dev(delp), dev(area), f, mask_below=mask_below)
    dc = host(dcoarse)
    np.testing.assert_allclose(dc, oracle.weighted_block_average(delp, area[:, None], f), rtol=2e-6)
    if mask_below:
        remapped, masked = recipes.regrid_given_delp([fld], delp, dc, area)
        exp = oracle.weighted_block_average(remapped[0], masked, f)
    else:
        exp = recipes.pressure_level_average(fld, delp, dc, area, f)
    g = host(outs[0])
    ok = np.isfinite(exp)
    assert np.array_equal(np.isfinite(g), ok)
    assert np.abs(g[ok] - exp[ok]).max() <= 3e-5 * np.nanmax(np.abs(fld))


@pytest.mark.parametrize("kernel", ["stream", "pair"])
@pytest.mark.parametrize("mask_below", [False, True])
def test_pressure_level_nonfinite_values_rerun_only_their_blocks(ops, monkeypatch, mask_below, kernel):
    """The streaming kernel's fast pass assumes finite field values and hands the CTAs that saw
    a NaN or an infinity to the NaN-skipping pass: results must equal the NaN-skipping pass alone,
    bit for bit where no CTA was rerun and everywhere else too (same arithmetic)."""
    nt, nc, f = 1, 8, 8
    ny = nx = nc * f
    delp = syn.delp_from_ps(syn.surface_pressure((nt, ny, nx), seed=16))
    area = syn.area((nt, ny, nx), seed=17)
    flds = [syn.field((nt, syn.NZ, ny, nx), 95 + i) for i in range(3)]
    flds[0][0, 2, 1, 2] = np.nan           # level 2 is read by the end step
    flds[1][0, 78, 40, 40] = np.nan
    flds[2][0, 30:33, 63, 63] = np.inf
    res = {}
    for impl in (kernel, kernel + "_nansafe"):
        _pin_impl(monkeypatch, impl)
        outs, _ = ops.pressure_level_average([dev(x) for x in flds], dev(delp), dev(area), f, mask_below=mask_below)
        res[impl] = [host(o) for o in outs]
    for a, b in zip(res[kernel], res[kernel + "_nansafe"]):
        fin = np.isfinite(b)
        assert np.array_equal(np.isfinite(a), fin)
        touched = np.zeros(a.shape, bool)      # CTAs = 4 coarse cells along x: the rerun ones must match exactly
        for (y, x) in ((0, 0), (5, 5), (7, 7)):
            touched[:, :, y, 4 * (x // 4):4 * (x // 4) + 4] = True
        np.testing.assert_array_equal(a[touched & fin], b[touched & fin])
        np.testing.assert_allclose(a[~touched & fin], b[~touched & fin], rtol=0, atol=3e-5 * 5)


@pytest.mark.parametrize("kernel", ["stream", "pair"])
def test_pressure_level_degenerate_thickness(ops, monkeypatch, kernel):
    """A source layer without thickness (0 / 0 in mappm.f90:75) or a NaN thickness makes the column's remap NaN where
    the layer is touched; the block sums skip it like nansum.  The fast pass has no NaN tests in its sums, so it must
    notice the geometry and hand the CTA to the NaN-safe variant: both give the same answer."""
    nt, nc, f = 1, 8, 8
    ny = nx = nc * f
    delp = syn.delp_from_ps(syn.surface_pressure((nt, ny, nx), seed=31))
    area = syn.area((nt, ny, nx), seed=32)
    flds = [syn.field((nt, syn.NZ, ny, nx), 33 + i, scale=10.0 ** i) for i in range(2)]
    delp[0, 30, 5, 6] = 0.0
    delp[0, 50, 40, 41] = np.nan
    res = {}
    for impl in (kernel, kernel + "_nansafe"):
        _pin_impl(monkeypatch, impl)
        outs, _ = ops.pressure_level_average([dev(x) for x in flds], dev(delp), dev(area), f)
        res[impl] = [host(o) for o in outs]
    for a, b in zip(res[kernel], res[kernel + "_nansafe"]):
        touched = np.zeros(a.shape, bool)
        touched[:, :, 0, 0:4] = True
        touched[:, :, 5, 4:8] = True
        np.testing.assert_array_equal(a[touched], b[touched])          # the rerun CTAs: the NaN-safe variant's bits
        np.testing.assert_allclose(a[~touched], b[~touched], rtol=0, atol=3e-5 * 10, equal_nan=True)


@pytest.mark.parametrize("kernel", ["pair", "stream", "fused"])
def test_pressure_level_on_a_second_device_of_the_process(ops, monkeypatch, kernel):
    """One process may drive several devices (BudgetPipeline(device=...)): the shared-memory opt-in of the fused
    kernels is a per-device function attribute, so a launch on cuda:1 after cuda:0 must still get it."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs in one process")
    nt, nc, f = 1, 8, 8
    ny = nx = nc * f
    delp = syn.delp_from_ps(syn.surface_pressure((nt, ny, nx), seed=41))
    area = syn.area((nt, ny, nx), seed=42)
    flds = [syn.field((nt, syn.NZ, ny, nx), 43 + i) for i in range(3)]
    _pin_impl(monkeypatch, kernel)
    res = []
    for d in (0, 1, 0):
        with torch.cuda.device(d):
            outs, dc = ops.pressure_level_average([torch.from_numpy(x).cuda(d) for x in flds],
                                                  torch.from_numpy(delp).cuda(d), torch.from_numpy(area).cuda(d), f)
            torch.cuda.synchronize(d)
            assert outs[0].device.index == d
            res.append([host(o) for o in outs] + [host(dc)])
    for other in res[1:]:
        for a, b in zip(res[0], other):
            np.testing.assert_array_equal(a, b)


@pytest.mark.parametrize("mask_below", [False, True])
def test_pressure_level_replicated_outputs(ops, monkeypatch, mask_below):
    """cg_pressure_level_average_replicated: every coarse value (outputs and delp_coarse) lands at its own address and
    at address + offset, bit for bit what the plain entry point gives.  The replicas of the multi-GPU path are peer
    copies of one buffer; here they are a second and third region of the same device buffer."""
    _pin_impl(monkeypatch, "default")
    nt, nc, f = 2, 8, 8
    ny = nx = nc * f
    delp = syn.delp_from_ps(syn.surface_pressure((nt, ny, nx), seed=51))
    area = syn.area((nt, ny, nx), seed=52)
    flds = [syn.field((nt, syn.NZ, ny, nx), 53 + i) for i in range(3)]
    flds[1][0, :, 3, 5] = np.nan                                        # a CTA that the NaN-safe pass redoes
    d_f, d_delp, d_area = [dev(x) for x in flds], dev(delp), dev(area)
    ref, ref_dc = ops.pressure_level_average(d_f, d_delp, d_area, f, mask_below=mask_below)
    n = ref_dc.numel()
    per = (len(flds) + 1) * n
    whole = torch.full((3 * per + 8,), -7.0, dtype=torch.float32, device="cuda")
    views = [whole[i * n:(i + 1) * n].view(ref_dc.shape) for i in range(len(flds) + 1)]
    outs, dc = ops.pressure_level_average(d_f, d_delp, d_area, f, mask_below=mask_below, outs=views[:-1],
                                          delp_coarse_out=views[-1], replica_offsets=[4 * per, 8 * per + 16])
    from fv3net_b200 import _lib
    assert _lib.lib().cg_pressure_level_last_impl() == b"pair"
    w = host(whole)
    for i, e in enumerate(ref + [ref_dc]):
        e = host(e).ravel()
        for base in (0, per, 2 * per + 4):
            np.testing.assert_array_equal(w[base + i * n:base + (i + 1) * n], e)
    assert (w[2 * per:2 * per + 4] == -7.0).all() and (w[3 * per + 4:] == -7.0).all()    # nothing else was touched
    # multicast form: ONE store per value, at address + offset only (on a multi-GPU box that address is the NVLS
    # mapping and the switch writes every copy, this GPU's included; here it is plain memory)
    whole.fill_(-7.0)
    ops.pressure_level_average(d_f, d_delp, d_area, f, mask_below=mask_below, outs=views[:-1], delp_coarse_out=views[-1],
                               multicast_offset=4 * per)
    w = host(whole)
    for i, e in enumerate(ref + [ref_dc]):
        np.testing.assert_array_equal(w[per + i * n:per + (i + 1) * n], host(e).ravel())
    assert (w[:len(flds) * n] == -7.0).all() and (w[2 * per:] == -7.0).all()
    np.testing.assert_array_equal(w[len(flds) * n:per], host(ref_dc).ravel())      # (delp_coarse's own copy is the block average's)
    with pytest.raises(ValueError):
        ops.pressure_level_average(d_f, d_delp, d_area, f, outs=views[:-1], delp_coarse_out=views[-1], replica_offsets=[6])
    with pytest.raises(ValueError):
        ops.pressure_level_average(d_f, d_delp, d_area, f, outs=views[:-1], delp_coarse_out=views[-1], replica_offsets=[16],
                                   multicast_offset=32)
    with pytest.raises(ValueError):
        ops.pressure_level_average(d_f, d_delp, d_area, f, replica_offsets=[4 * per])          # needs the caller's arrays
    d2, a2 = dev(delp[:, :, :32, :32]), dev(area[:, :32, :32])
    small = [torch.empty((nt, syn.NZ, 8, 8), device="cuda")]
    with pytest.raises(NotImplementedError):                              # f = 4: not the pair kernel's shape
        ops.pressure_level_average([dev(flds[0][:, :, :32, :32])], d2, a2, 4, outs=small,
                                   delp_coarse_out=torch.empty_like(small[0]), replica_offsets=[0])


def test_pressure_level_golden_restart_T(ops):
    """The reference's golden pressure-level T (test_coarsen_restarts.py) through CUDA."""
    tag = "pressure-level-without-agrid-winds"
    delp, _ = S.generate("fv_core.res", "delp")
    area, _ = S.generate("grid", "area")
    T, _ = S.generate("fv_core.res", "T")
    outs, dcoarse = ops.pressure_level_average([dev(T.astype("f4"))], dev(delp.astype("f4")),
                                               dev(area), 2, mask_below=True)
    exp, _ = S.golden(tag, "fv_core.res", "T")
    # This array-level entry point is fp32 end to end (the budget workflow's dtype).  The golden
    # file was produced from f64 restart fields whose interface pressures the reference sums
    # in f64 before mappm's f32 cast; with 4-Pa layers at p ~ 320 Pa the f32 cumsum alone moves
    # PL/PR by ~1e-5, hence 1e-4 of the +-1000 field scale here.  The f64 flow of the recipe
    # (fv3net_b200.cubedsphere.regridz) is tested against the same file at 3e-5.
    assert np.abs(host(outs[0]) - exp).max() < 1e-4 * 1000
    from tests import parity
    parity.assert_within_fp32_conditioning(host(outs[0]), T.astype("f4"), delp.astype("f4"), area, 2, True, scale=1000.0,
                                           dc_gpu=host(dcoarse))
    dexp, _ = S.golden(tag, "fv_core.res", "delp")
    np.testing.assert_allclose(host(dcoarse), dexp, rtol=1e-5)


@pytest.mark.parametrize("nc,f", [(4, 8), (6, 8), (8, 4)])
def test_area_above_fine_surface(ops, nc, f):
    """(4, 8): warp per strip of four cells (whole-line loads); (6, 8), (8, 4): warp per cell, any factor.
    NaN areas are skipped by both like the reference's nansum."""
    fields, delp, area = _plevel_case(2, nc, f, seed=9, nfields=0)
    area[0, 3, 5] = np.nan
    dc = oracle.weighted_block_average(delp, np.nan_to_num(area)[:, None], f)
    got = host(ops.area_above_fine_surface(dev(delp), dev(dc), dev(area), f))
    exp = recipes.area_above_fine_surface(delp, dc, area, f)
    # a coarse cell may differ from the reference only by the area of fine columns whose surface pressure sits within
    # the logarithm's rounding of that level's mid pressure (tests/parity.py); everything else to summation order
    from tests import parity
    parity.assert_exposed_area_matches(got, delp, dc, area, f)
    assert np.array_equal(np.isfinite(got), np.isfinite(exp))
    # the three formulations (warp per strip of four cells, warp per cell, three kernels with temporaries) sum the
    # same columns in the same order: same bits
    from fv3net_b200 import _lib
    forms = {}
    for impl in ("strip", "cell", "three"):
        _lib.set_option("exposed_area_impl", impl)
        forms[impl] = host(ops.area_above_fine_surface(dev(delp), dev(dc), dev(area), f))
    np.testing.assert_array_equal(forms["cell"], forms["three"])
    if f == 8 and nc % 4 == 0:
        np.testing.assert_array_equal(forms["strip"], forms["cell"])
