"""GPU parity at BASELINE.json's full sizes (configs[1]..configs[4]).

The oracle cannot finish a C384 timestep in seconds, so each case is checked two ways:

  * a SLAB of the full-size GPU result against the oracle run on just that slab (a coarse
    cell depends only on its own f x f fine columns, so the slab is a closed problem);
  * size-independent properties over the WHOLE result: power-of-two scaling (bit-exact: it
    commutes with every fp32 rounding on the path), constants, median(upsample(x)) == x,
    mirror symmetry of the selection, bitwise run-to-run reproducibility, and agreement of
    row-band shards (configs[4], C3072 by 768-row bands) with the unsharded launch.
"""
import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

import oracle  # noqa: E402
from oracle import recipes  # noqa: E402
from fv3net_b200 import sharding, synthetic as syn  # noqa: E402

F = 8
NZ = syn.NZ


@pytest.fixture(scope="module")
def ops():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from fv3net_b200 import ops as _ops
    return _ops


def host(t):
    return t.cpu().numpy()


def _gen(shape, seed, scale=1.0):
    return syn.torch_field(shape, seed, torch.device("cuda"), scale=scale)


# ---- configs[1]: C384 -> C48 area / mass weighted averages + staggered winds, 79 levels, 20 variables ----
def test_config2_block_averages_c384(ops):
    nt, n = 6, 384
    dev = torch.device("cuda")
    delp = syn.torch_delp(nt, n, n, 1, dev)
    area = syn.torch_area((nt, n, n), 2, dev)
    rows = slice(40, 56)                        # two coarse rows of tile 3, every level
    tile = 3
    # 3 area-weighted + 13 mass-weighted fields, in two launches sharing their weight reads
    area_w = [delp] + [_gen((nt, NZ, n, n), 10 + i, 10.0 ** i) for i in range(2)]
    out_a = ops.weighted_block_average(area_w, w2d=area, fy=F, fx=F)
    mass_w = [_gen((nt, NZ, n, n), 20 + i, 10.0 ** (i % 4 - 1)) for i in range(13)]
    out_m = ops.weighted_block_average(mass_w, w2d=area, w3d=delp, fy=F, fx=F)
    a_h, d_h = host(area[tile, rows]), host(delp[tile, :, rows])
    for src, got in zip(area_w, out_a):
        exp = oracle.weighted_block_average(host(src[tile, :, rows])[None], a_h[None, None], F)[0]
        g = host(got[tile, :, 5:7])
        assert np.abs(g - exp).max() <= 16 * np.finfo("f4").eps * np.abs(host(src[tile, :, rows])).max()
    for src, got in zip(mass_w[:3], out_m[:3]):
        s_h = host(src[tile, :, rows])
        exp = oracle.weighted_block_average(s_h[None], (d_h * a_h[None])[None], F)[0]
        assert np.abs(host(got[tile, :, 5:7]) - exp).max() <= 16 * np.finfo("f4").eps * np.abs(s_h).max()
    # properties over the whole result: x4 is exact; a constant field averages to itself within the summation tolerance
    again = ops.weighted_block_average([4.0 * t for t in mass_w[:2]], w2d=area, w3d=delp, fy=F, fx=F)
    for a, b in zip(again, out_m[:2]):
        assert torch.equal(a, 4.0 * b)
    const = torch.full((nt, NZ, n, n), 7.25, device=dev)
    c = ops.weighted_block_average(const, w2d=area, w3d=delp, fy=F, fx=F)
    assert (c - 7.25).abs().max().item() <= 16 * 7.25 * np.finfo("f4").eps
    del area_w, mass_w, const, again
    # staggered winds: u (ny + 1, nx) with dx, v (ny, nx + 1) with dy; only the kept lines are read
    u = _gen((nt, NZ, n + 1, n), 40, 30.0)
    dx = syn.torch_area((nt, n + 1, n), 41, dev)
    v = _gen((nt, NZ, n, n + 1), 42, 30.0)
    dy = syn.torch_area((nt, n, n + 1), 43, dev)
    uc = ops.weighted_block_average(u, w2d=dx, fy=1, fx=F, sy=F, sx=F, nyo=n // F + 1, nxo=n // F)
    vc = ops.weighted_block_average(v, w2d=dy, fy=F, fx=1, sy=F, sx=F, nyo=n // F, nxo=n // F + 1)
    assert tuple(uc.shape) == (nt, NZ, 49, 48) and tuple(vc.shape) == (nt, NZ, 48, 49)
    exp_u = oracle.edge_weighted_block_average(host(u[tile])[None], host(dx[tile])[None, None], F, edge="x")[0]
    exp_v = oracle.edge_weighted_block_average(host(v[tile])[None], host(dy[tile])[None, None], F, edge="y")[0]
    assert np.abs(host(uc[tile]) - exp_u).max() <= 16 * np.finfo("f4").eps * 30.0 * 6
    assert np.abs(host(vc[tile]) - exp_v).max() <= 16 * np.finfo("f4").eps * 30.0 * 6


# ---- configs[2]: C384 -> C48 pressure-level coarse-graining of a budget-shaped timestep ----
def _budget_fields(nt, n, nfields, seed0):
    return [_gen((nt, NZ, n, n), seed0 + i, 10.0 ** (i % 5 - 2)) for i in range(nfields)]


@pytest.mark.parametrize("mask_below", [False, True])
def test_config3_pressure_level_c384(ops, mask_below):
    nt, n = 6, 384
    dev = torch.device("cuda")
    delp = syn.torch_delp(nt, n, n, 3, dev)
    area = syn.torch_area((nt, n, n), 4, dev)
    fields = _budget_fields(nt, n, 8, 50)
    args = fields + [ops.product(fields[0], fields[1]), ops.storage(fields[2], fields[3], 900.0)]
    outs, dc = ops.pressure_level_average(args, delp, area, F, mask_below=mask_below)
    torch.cuda.synchronize()
    # (1) bitwise reproducible: the 64 columns of a cell are summed in a fixed order
    outs2, dc2 = ops.pressure_level_average(args, delp, area, F, mask_below=mask_below)
    for a, b in zip(outs, outs2):
        assert torch.equal(a, b) or torch.equal(torch.isnan(a), torch.isnan(b)) and torch.equal(
            torch.nan_to_num(a), torch.nan_to_num(b))
    assert torch.equal(dc, dc2)
    # (2) slabs against the oracle: two coarse rows (16 fine rows; an odd coarse size would read as a staggered
    # dim in block_upsample, coarsen.py:847-864) of two tiles, every level and column
    R = 2 * F
    for tile, r0 in ((0, 0), (4, 368)):
        rows = slice(r0, r0 + R)
        d_h, a_h = host(delp[tile:tile + 1, :, rows]), host(area[tile:tile + 1, rows])
        dc_exp = oracle.weighted_block_average(d_h, a_h[:, None], F)
        np.testing.assert_allclose(host(dc[tile:tile + 1, :, r0 // F:r0 // F + 2]), dc_exp, rtol=2e-6)
        cpu = [host(t[tile:tile + 1, :, rows]) for t in fields]
        cpu += [cpu[0] * cpu[1], ((cpu[2] - cpu[3]) / np.float32(900.0)).astype(np.float32)]
        if mask_below:
            remapped, masked = recipes.regrid_given_delp(cpu, d_h, dc_exp, a_h, threads=8)
            exp = [oracle.weighted_block_average(r, masked, F) for r in remapped]
        else:
            exp = [recipes.pressure_level_average(x, d_h, dc_exp, a_h, F, threads=8) for x in cpu]
        for g, e, x in zip(outs, exp, cpu):
            g = host(g[tile:tile + 1, :, r0 // F:r0 // F + 2])
            ok = np.isfinite(e)
            assert np.array_equal(np.isfinite(g), ok)
            assert np.abs(g[ok] - e[ok]).max() <= 3e-5 * np.abs(x).max()
    # (3) x 2^k is exact through PPM edges, limiter, integrals and the block sum
    scaled, _ = ops.pressure_level_average([16.0 * t for t in fields[:4]], delp, area, F, mask_below=mask_below,
                                           want_delp_coarse=False)
    for a, b in zip(scaled, outs[:4]):
        assert torch.equal(torch.nan_to_num(a), torch.nan_to_num(16.0 * b))
    # (4) a constant comes back as the constant (zero slopes, flat parabolas) wherever a level has weight
    const, _ = ops.pressure_level_average([torch.full_like(delp, -3.5)], delp, area, F, mask_below=mask_below,
                                          want_delp_coarse=False)
    c = const[0]
    fin = torch.isfinite(c)
    assert fin.float().mean().item() > 0.9
    assert (c[fin] + 3.5).abs().max().item() <= 3.5 * 32 * np.finfo("f4").eps


@pytest.mark.parametrize("mask_below", [False, True])
def test_config3_deterministic_under_load(ops, mask_below):
    """The streaming kernel hands buffers between warps through mbarriers and named barriers; a missed hand-over
    would show as run-to-run differences.  Twelve runs of a 26-field timestep with a rough surface (many layers cut by
    several coarse interfaces, many coarse levels below the fine surface: the rare paths), while a second stream keeps
    the memory system busy, must agree bit for bit."""
    nt, n = 6, 384
    dev = torch.device("cuda")
    g = torch.Generator(device=dev).manual_seed(77)
    base = syn.torch_delp(nt, n, n, 7, dev)
    rough = 1.0 + 0.25 * (torch.rand((nt, 1, n, n), generator=g, device=dev) * 2 - 1) * (
        torch.rand((nt, 1, n, n), generator=g, device=dev) < 0.3)
    delp = (base * rough).contiguous()
    area = syn.torch_area((nt, n, n), 8, dev)
    fields = _budget_fields(nt, n, 22, 90)
    args = fields + [ops.product(fields[0], fields[1]), ops.product(fields[2], fields[1]),
                     ops.storage(fields[3], fields[4], 900.0), ops.storage(fields[5], fields[6], 900.0)]
    ref, dc = ops.pressure_level_average(args, delp, area, F, mask_below=mask_below)
    ref = [t.clone() for t in ref]
    noise_a = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    noise_b = torch.empty_like(noise_a)
    side = torch.cuda.Stream()
    for i in range(12):
        if i % 2:
            with torch.cuda.stream(side):
                for _ in range(8):
                    noise_b.copy_(noise_a)
        outs, _ = ops.pressure_level_average(args, delp, area, F, mask_below=mask_below, want_delp_coarse=False)
        for a, b in zip(outs, ref):
            assert torch.equal(torch.nan_to_num(a), torch.nan_to_num(b)) and torch.equal(torch.isnan(a), torch.isnan(b)), i
    torch.cuda.synchronize()
    # the rare paths against the oracle on a slab of this rough timestep (ill-conditioned columns: 1e-4 as in
    # test_gpu_remap.test_pressure_level_hard_columns), with the GPU's own delp_coarse as the target pressures
    tile, r0 = 5, 192
    rows = slice(r0, r0 + 2 * F)
    d_h, a_h = host(delp[tile:tile + 1, :, rows]), host(area[tile:tile + 1, rows])
    dc_h = host(dc[tile:tile + 1, :, r0 // F:r0 // F + 2])
    np.testing.assert_allclose(dc_h, oracle.weighted_block_average(d_h, a_h[:, None], F), rtol=2e-6)
    for v in (0, 7, 21):
        x = host(fields[v][tile:tile + 1, :, rows])
        if mask_below:
            remapped, masked = recipes.regrid_given_delp([x], d_h, dc_h, a_h, threads=8)
            e = oracle.weighted_block_average(remapped[0], masked, F)
        else:
            e = recipes.pressure_level_average(x, d_h, dc_h, a_h, F, threads=8)
        gg = host(ref[v][tile:tile + 1, :, r0 // F:r0 // F + 2])
        ok = np.isfinite(e)
        assert np.array_equal(np.isfinite(gg), ok)
        assert np.abs(gg[ok] - e[ok]).max() <= 1e-4 * np.abs(x).max()
        # the bar above 3e-5 is adjudicated by the f64 evaluation (tests/parity.py); delp_coarse as the oracle forms it
        from tests import parity
        parity.assert_within_fp32_conditioning(gg, x, d_h, a_h, F, mask_below, threads=8, dc_gpu=dc_h)


def test_config3_exposed_area_c384(ops):
    nt, n = 6, 384
    dev = torch.device("cuda")
    delp = syn.torch_delp(nt, n, n, 5, dev)
    area = syn.torch_area((nt, n, n), 6, dev)
    dc = ops.weighted_block_average(delp, w2d=area, fy=F, fx=F)
    got = ops.area_above_fine_surface(delp, dc, area, F)
    tile, r0 = 2, 128
    exp = recipes.area_above_fine_surface(host(delp[tile:tile + 1, :, r0:r0 + 2 * F]), host(dc[tile:tile + 1, :, r0 // F:r0 // F + 2]),
                                          host(area[tile:tile + 1, r0:r0 + 2 * F]), F)
    g = host(got[tile:tile + 1, :, r0 // F:r0 // F + 2])
    # a cell may differ only by the area of columns whose surface pressure sits within the logarithm's rounding of
    # the level's mid pressure (tests/parity.py)
    from tests import parity
    nflip = parity.assert_exposed_area_matches(g, host(delp[tile:tile + 1, :, r0:r0 + 2 * F]), host(dc[tile:tile + 1, :, r0 // F:r0 // F + 2]),
                                               host(area[tile:tile + 1, r0:r0 + 2 * F]), F)
    assert nflip <= 0.01 * g.size
    # the exposed area never exceeds the block's area and grows monotonically upward (lower pressure)
    tot = ops.block_reduce(area, "sum", F, F)
    assert (got <= tot[:, None] * (1 + 1e-6)).all()
    assert (got[:, :-1] >= got[:, 1:] * (1 - 1e-6)).all()


# ---- configs[3]: 2-D sweep, 96 timesteps x 6 tiles x 384 x 384: block_median and block_upsample ----
def test_config4_median_upsample_sweep(ops):
    nt, n = 96 * 6, 384
    dev = torch.device("cuda")
    g = torch.Generator(device=dev).manual_seed(7)
    fine = torch.rand((nt, n, n), generator=g, device=dev)
    med = ops.block_median(fine, F)
    assert tuple(med.shape) == (nt, n // F, n // F)
    # slab against numpy's median through the oracle (4 of 576 images, first / last included)
    for i in (0, 1, 300, nt - 1):
        np.testing.assert_array_equal(host(med[i:i + 1]), oracle.block_median(host(fine[i:i + 1]), F))
    # selection commutes with mirroring the grid (blocks map onto blocks) and with x -> -x for an even count
    assert torch.equal(ops.block_median(torch.flip(fine, (1, 2)).contiguous(), F), torch.flip(med, (1, 2)))
    assert torch.equal(ops.block_median(-fine, F), -med)
    # upsample: every fine cell of a block holds the coarse value; median(upsample(x)) == x bit for bit
    up = ops.block_upsample(med, F, F)
    assert tuple(up.shape) == (nt, n, n)
    assert torch.equal(up.view(nt, n // F, F, n // F, F), med[:, :, None, :, None].expand(nt, n // F, F, n // F, F))
    assert torch.equal(ops.block_median(up, F), med)
    # NaN poisons only its own block
    fine[5, 17, 200] = float("nan")
    m2 = ops.block_median(fine, F)
    assert torch.isnan(m2[5, 2, 25]) and torch.isnan(m2).sum().item() == 1


# ---- configs[4]: C3072 -> C384 by (tile, 768-row band): 24 shards, 3 per GPU ----
def test_config5_c3072_row_band(ops):
    n, band = 3072, 768
    shards = sharding.plan(1, 8, n_tile=6, ny=n, factor=F)
    assert len(shards) == 8 and sum(len(s) for s in shards) >= 8
    dev = torch.device("cuda")
    # one (tile, band) shard, generated on the device: (1, 79, 768, 3072) fp32 = 745 MB per field
    delp = syn.torch_delp(1, band, n, 9, dev)
    area = syn.torch_area((1, band, n), 10, dev)
    fields = [_gen((1, NZ, band, n), 70 + i, 10.0 ** i) for i in range(3)]
    outs, dc = ops.pressure_level_average(fields, delp, area, F)
    assert tuple(outs[0].shape) == (1, NZ, band // F, n // F)
    # a narrower band of the same data gives the same coarse rows bit for bit (no cross-row coupling)
    sub = slice(256, 512)
    outs_b, dc_b = ops.pressure_level_average([t[:, :, sub].contiguous() for t in fields],
                                              delp[:, :, sub].contiguous(), area[:, sub].contiguous(), F)
    for a, b in zip(outs, outs_b):
        assert torch.equal(torch.nan_to_num(a[:, :, 32:64]), torch.nan_to_num(b))
    assert torch.equal(dc[:, :, 32:64], dc_b)
    # slab against the oracle: two coarse rows across all 3072 columns
    r0 = 40 * F
    d_h, a_h = host(delp[:, :, r0:r0 + 2 * F]), host(area[:, r0:r0 + 2 * F])
    dc_exp = oracle.weighted_block_average(d_h, a_h[:, None], F)
    for g, t in zip(outs, fields):
        x = host(t[:, :, r0:r0 + 2 * F])
        e = recipes.pressure_level_average(x, d_h, dc_exp, a_h, F, threads=8)
        ok = np.isfinite(e)
        gg = host(g[:, :, 40:42])
        assert np.array_equal(np.isfinite(gg), ok)
        assert np.abs(gg[ok] - e[ok]).max() <= 3e-5 * np.abs(x).max()


# ---- N > 1 on one box: shards by tile / row band, one NCCL gather of the coarse outputs ----
def test_nccl_gather_matches_unsharded():
    """Runs tools/multi_gpu_gather.py on every GPU of the box (skipped on a one-GPU box; the gloo world-2 twin of
    the host logic is tests/test_sharding.py)."""
    import os
    import subprocess
    import sys
    n = torch.cuda.device_count() if torch.cuda.is_available() else 0
    if n < 2:
        pytest.skip("needs at least two GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}",
                        "--master-addr", "127.0.0.1", "--master-port", "29533", os.path.join(root, "tools", "multi_gpu_gather.py")],
                       capture_output=True, text=True, timeout=600, cwd=root)
    assert r.returncode == 0, r.stderr[-2000:]
    assert '"bit_identical_to_unsharded": true' in r.stdout
