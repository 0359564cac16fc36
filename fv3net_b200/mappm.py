"""Drop-in for the f2py module `mappm` (external/mappm: mappm.f90:10-31), the one native
interface on the reference's path: `import fv3net_b200.mappm as mappm` makes
`vcm/cubedsphere/regridz.py:273-275` run the CUDA remapper unchanged.

    mappm(pe1[n, km+1], q1[n, km], pe2[n, kn+1], i1, i2, iv, kord, ptop) -> q2[n, kn] float32

Host numpy arrays in, host float32 array out (f2py casts every input to Fortran `real`);
`i1`, `i2`, `iv`, `kord` may arrive as floats (external/vcm/tests/test_mappm.py:9-13);
columns i1..i2 (1-based, inclusive) are remapped, the others come back as zeros like the
Fortran's untouched intent(out) array; `ptop` is unused there and here.
"""
import numpy as np
import torch

from . import ops


def mappm(pe1, q1, pe2, i1, i2, iv, kord, ptop=0.0):
    pe1 = np.ascontiguousarray(pe1, dtype=np.float32)
    q1 = np.ascontiguousarray(q1, dtype=np.float32)
    pe2 = np.ascontiguousarray(pe2, dtype=np.float32)
    if pe1.ndim != 2 or q1.ndim != 2 or pe2.ndim != 2:
        raise ValueError("mappm expects 2-D arrays (columns, levels)")
    n, km = q1.shape
    kn = pe2.shape[1] - 1
    if pe1.shape != (n, km + 1) or pe2.shape[0] != n:
        raise ValueError(f"shape mismatch: pe1 {pe1.shape}, q1 {q1.shape}, pe2 {pe2.shape}")
    i1, i2, iv, kord = int(i1), int(i2), int(iv), int(kord)
    if not (1 <= i1 <= i2 <= n):
        raise ValueError(f"column range i1={i1}, i2={i2} outside 1..{n}")
    lo, cnt = i1 - 1, i2 - i1 + 1
    d_pe1, d_q1, d_pe2 = (ops.to_device(a[lo:lo + cnt]) for a in (pe1, q1, pe2))
    out = torch.empty((cnt, kn), dtype=torch.float32, device=d_q1.device)
    # columns are the outer index, levels contiguous: strides (outer, column, level)
    ops.mappm(d_pe1, d_q1, d_pe2, cnt, 1, km, kn, (km + 1, 0, 1), (km, 0, 1), (kn + 1, 0, 1),
              (kn, 0, 1), out, iv=iv, kord=kord)
    q2 = np.zeros((n, kn), dtype=np.float32)
    q2[lo:lo + cnt] = out.cpu().numpy()
    return q2


def interpolate_2d(xp, x, y, fill_value):
    """`mappm.interpolate_2d(xp[m, n_out], x[m, n_in], y[m, n_in], fill_value) -> y_out[m, n_out]`
    float64 (interpolate_2d.f90:1-29; called by vcm/interpolate.py:176-180 with
    fill_value=np.nan).  Host arrays in, host array out, any input dtype (f2py casts to real(8))."""
    xp, x, y = (np.ascontiguousarray(a, dtype=np.float64) for a in (xp, x, y))
    if xp.ndim != 2 or x.ndim != 2 or y.ndim != 2:
        raise ValueError("interpolate_2d expects 2-D arrays (rows, levels)")
    if x.shape != y.shape or xp.shape[0] != x.shape[0]:
        raise ValueError(f"shape mismatch: xp {xp.shape}, x {x.shape}, y {y.shape}")
    out = ops.interpolate_2d(ops.to_device(xp), ops.to_device(x), ops.to_device(y), float(fill_value))
    return out.cpu().numpy()
