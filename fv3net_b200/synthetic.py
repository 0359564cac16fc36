"""Synthetic inputs of the shapes BASELINE.json names (there is no network for data).

The 79-level hybrid coefficients are the model's own vertical grid: the `a79`/`b79`
data tables of external/fv3gfs-fortran/FV3/atmos_cubed_sphere/tools/fv_eta.F90:989-1014
(top interface 300 Pa = vcm's _TOA_PRESSURE).  Interface pressure = a + b * ps.
"""
import numpy as np

A79 = np.array([
    300, 646.7159, 1045.222, 1469.188, 1897.829, 2325.385, 2754.396, 3191.294,
    3648.332, 4135.675, 4668.282, 5247.94, 5876.271, 6554.716, 7284.521,
    8066.738, 8902.188, 9791.482, 10734.99, 11626.25, 12372.12, 12990.41,
    13496.29, 13902.77, 14220.98, 14460.58, 14629.93, 14736.33, 14786.17,
    14785.11, 14738.12, 14649.66, 14523.7, 14363.82, 14173.24, 13954.91,
    13711.48, 13445.4, 13158.9, 12854.07, 12532.8, 12196.85, 11847.88,
    11487.39, 11116.82, 10737.48, 10350.62, 9957.395, 9558.875, 9156.069,
    8749.922, 8341.315, 7931.065, 7519.942, 7108.648, 6698.281, 6290.007,
    5884.984, 5484.372, 5089.319, 4700.96, 4320.421, 3948.807, 3587.201,
    3236.666, 2898.237, 2572.912, 2261.667, 1965.424, 1685.079, 1421.479,
    1175.419, 947.6516, 738.8688, 549.713, 380.7626, 232.5417, 105.481,
    -0.0008381903, 0], dtype=np.float64)

B79 = np.array([
    0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0.001065947,
    0.004128662, 0.009006631, 0.01554263, 0.02359921, 0.03305481, 0.0438012,
    0.05574095, 0.06878554, 0.08285347, 0.09786981, 0.1137643, 0.130471,
    0.1479275, 0.1660746, 0.1848558, 0.2042166, 0.2241053, 0.2444716,
    0.2652672, 0.286445, 0.3079604, 0.3297701, 0.351832, 0.3741062,
    0.3965532, 0.4191364, 0.4418194, 0.4645682, 0.48735, 0.5101338,
    0.5328897, 0.5555894, 0.5782067, 0.6007158, 0.6230936, 0.6452944,
    0.6672683, 0.6889648, 0.7103333, 0.7313231, 0.7518838, 0.7719651,
    0.7915173, 0.8104913, 0.828839, 0.846513, 0.8634676, 0.8796583,
    0.8950421, 0.9095779, 0.9232264, 0.9359506, 0.9477157, 0.9584892,
    0.9682413, 0.9769447, 0.9845753, 0.9911126, 0.9965372, 1], dtype=np.float64)

NZ = 79


def surface_pressure(shape, seed=0, smooth_amp=0.03, rough_frac=0.05, rough_amp=0.10):
    """ps = 1e5 * (1 + noise): gentle noise everywhere plus ~5 % of columns perturbed by up
    to +-10 % (orography-like), so coarse target levels fall inside and below fine columns."""
    rng = np.random.RandomState(seed)
    ps = 1e5 * (1.0 + smooth_amp * rng.uniform(-1, 1, shape))
    rough = rng.uniform(0, 1, shape) < rough_frac
    ps = np.where(rough, ps * (1.0 + rough_amp * rng.uniform(-1, 1, shape)), ps)
    return ps


def delp_from_ps(ps, dtype=np.float32):
    """(..., ny, nx) surface pressure -> (..., 79, ny, nx) layer thickness."""
    da = np.diff(A79)[:, None, None]
    db = np.diff(B79)[:, None, None]
    ps = np.asarray(ps)
    return (da + db * ps[..., None, :, :]).astype(dtype)


def field(shape, seed, scale=1.0, dtype=np.float32):
    rng = np.random.RandomState(seed)
    return (scale * rng.standard_normal(shape)).astype(dtype)


def area(shape, seed=12345, dtype=np.float32):
    rng = np.random.RandomState(seed)
    return rng.uniform(0.5, 1.0, shape).astype(dtype)


# --- device-side generators (bench: inputs are created directly in HBM) ----------------
def torch_delp(n_tile, ny, nx, seed, device):
    import torch

    g = torch.Generator(device=device).manual_seed(seed)
    u = lambda: torch.rand((n_tile, ny, nx), generator=g, device=device, dtype=torch.float64) * 2 - 1  # noqa: E731
    ps = 1e5 * (1.0 + 0.03 * u())
    rough = torch.rand((n_tile, ny, nx), generator=g, device=device) < 0.05
    ps = torch.where(rough, ps * (1.0 + 0.10 * u()), ps)
    da = torch.tensor(np.diff(A79), device=device)[None, :, None, None]
    db = torch.tensor(np.diff(B79), device=device)[None, :, None, None]
    return (da + db * ps[:, None]).to(torch.float32).contiguous()


def torch_field(shape, seed, device, scale=1.0):
    import torch

    g = torch.Generator(device=device).manual_seed(seed)
    return (scale * torch.randn(shape, generator=g, device=device, dtype=torch.float32)).contiguous()


def torch_area(shape, seed, device):
    import torch

    g = torch.Generator(device=device).manual_seed(seed)
    return (0.5 + 0.5 * torch.rand(shape, generator=g, device=device, dtype=torch.float32)).contiguous()
