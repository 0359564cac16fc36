"""fv3net_b200.staging: host arrays in, host arrays out, chunked by tile through two device
buffer sets - the result must be the device-resident step's, bit for bit (a coarse cell only
needs its own fine columns, so cutting the timestep along the tile dimension changes nothing)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _timestep(nt, nz, n, seed):
    from fv3net_b200 import budget as B, synthetic as syn

    dev = torch.device("cuda")
    shape3 = (nt, nz, n, n)
    delp = syn.torch_delp(nt, n, n, seed, dev) if nz == 79 else None
    if delp is None:
        delp = (torch.rand(shape3, device=dev) * 400 + 100).float()
    area = syn.torch_area((nt, n, n), seed + 1, dev)
    names3 = B.FIRST_MOMENTS_3D
    middle3d = {k: syn.torch_field(shape3, seed + 10 + i, dev) for i, k in enumerate(names3)}
    middle2d = {k: syn.torch_field((nt, n, n), seed + 100 + i, dev) for i, k in enumerate(B.GFSPHYSICS_2D[:3])}
    begin = {k: middle3d[k] - 0.01 for k in B.STORAGE_TERMS}
    end = {k: middle3d[k] + 0.02 for k in B.STORAGE_TERMS}
    return dict(middle3d=middle3d, middle2d=middle2d, delp=delp, area=area, exposed_area_fine=area * 0.5,
                begin=begin, end=end)


@pytest.mark.parametrize("nt,nz,n,chunk", [(6, 79, 64, 1), (6, 79, 32, 2), (4, 12, 16, 1), (3, 79, 32, 3)])
def test_staged_step_equals_device_step(nt, nz, n, chunk):
    from fv3net_b200 import budget as B, staging

    inp = _timestep(nt, nz, n, seed=nt + n)
    direct = B.recoarsen_budget_inputs(**inp)
    host = staging.pin({k: ({m: t.cpu() for m, t in v.items()} if isinstance(v, dict) else v.cpu())
                        for k, v in inp.items()})
    pipe = staging.BudgetPipeline(chunk=chunk)
    out = pipe.run(host)
    assert list(out) == list(direct)
    for k in direct:
        assert not out[k].is_cuda and out[k].is_pinned()
        np.testing.assert_array_equal(out[k].numpy(), direct[k].cpu().numpy(), err_msg=k)
    assert pipe.h2d_bytes == sum(t.numel() * 4 for t in staging._flatten(host).values())
    assert pipe.d2h_bytes == sum(t.numel() * 4 for t in direct.values())
    # a second timestep through the same pipeline reuses the buffers and the output arrays
    again = pipe.run(host, out)
    assert again is out
    for k in direct:
        np.testing.assert_array_equal(out[k].numpy(), direct[k].cpu().numpy(), err_msg=k)


def test_staging_refuses_device_inputs_and_ragged_chunks():
    from fv3net_b200 import staging

    inp = _timestep(3, 12, 16, seed=1)
    with pytest.raises(ValueError, match="host arrays"):
        staging.BudgetPipeline().run(inp)
    host = staging.pin({k: ({m: t.cpu() for m, t in v.items()} if isinstance(v, dict) else v.cpu())
                        for k, v in inp.items()})
    with pytest.raises(ValueError, match="do not split"):
        staging.BudgetPipeline(chunk=2).run(host)
