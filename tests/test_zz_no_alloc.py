"""GPU test of the allocation-free entry point cg_area_above_fine_surface_static.

No counterpart in the reference: the same three
kernels as cg_area_above_fine_surface's general form, with the two temporaries in a file-scope
device array instead of cudaMallocAsync.  Called straight through the C ABI (ctypes), compared
with the existing entry point on two calls in a row with DIFFERENT inputs (the array keeps the
first call's contents) and sizes, and refused beyond the array's size.

Named test_zz_* so that pytest collects it last.
"""
import ctypes

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

from fv3net_b200 import synthetic as syn  # noqa: E402


@pytest.fixture(scope="module")
def lib():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from fv3net_b200 import _lib
    return _lib


def _case(nt, nc, f, seed):
    ny = nx = nc * f
    delp = syn.delp_from_ps(syn.surface_pressure((nt, ny, nx), seed=seed))
    area = syn.area((nt, ny, nx), seed=seed + 1)
    h = (delp * area[:, None]).reshape(nt, syn.NZ, nc, f, nc, f).sum((3, 5))
    dc = (h / area.reshape(nt, nc, f, nc, f).sum((2, 4))[:, None]).astype(np.float32)
    return delp, dc, area


def _call(lib, name, delp, dc, area, f):
    nt, nz, ny, nx = delp.shape
    d, c, a = (torch.from_numpy(np.ascontiguousarray(x)).cuda() for x in (delp, dc, area))
    out = torch.full((nt, nz, ny // f, nx // f), float("nan"), dtype=torch.float32, device="cuda")
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = getattr(lib.lib(), name)(d.data_ptr(), c.data_ptr(), a.data_ptr(), out.data_ptr(), nt, nz, ny, nx, f,
                                  300.0, stream)
    torch.cuda.synchronize()
    return rc, out.cpu().numpy()


def test_static_scratch_matches_allocating_entry_point_twice(lib):
    assert lib.lib().cg_static_scratch_bytes() == 16 << 20
    # two calls in a row: the second is smaller and has other values, so stale scratch would show
    for nt, nc, f, seed in ((2, 6, 8, 40), (1, 8, 4, 50), (2, 6, 8, 40)):
        delp, dc, area = _case(nt, nc, f, seed)
        area[0, 3, 5] = np.nan                      # NaN areas are skipped like the reference's nansum
        rc0, exp = _call(lib, "cg_area_above_fine_surface", delp, dc, area, f)
        rc1, got = _call(lib, "cg_area_above_fine_surface_static", delp, dc, area, f)
        assert rc0 == 0 and rc1 == 0, lib.lib().cg_last_error()
        assert np.isfinite(got).all()
        # same tolerance as tests/test_gpu_remap.py::test_area_above_fine_surface (summation order)
        np.testing.assert_allclose(got, exp, rtol=2e-6, atol=0)


def test_static_scratch_rejects_sizes_beyond_the_array(lib):
    # 24 tiles of 384 x 384: 14.2 MiB of surface pressure + 2.8 MiB of coarse interfaces > 16 MiB.
    # Arguments are validated before anything is launched, so small dummy buffers are enough.
    t = torch.zeros(16, dtype=torch.float32, device="cuda")
    rc = lib.lib().cg_area_above_fine_surface_static(t.data_ptr(), t.data_ptr(), t.data_ptr(), t.data_ptr(),
                                                     24, 79, 384, 384, 8, 300.0, None)
    assert rc == lib.CG_EINVAL
    assert b"static array holds" in lib.lib().cg_last_error()
