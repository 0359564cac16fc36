"""Array-level entry points: torch CUDA tensors (or numpy arrays, staged through
pinned memory) in canonical layout, horizontal dims last.

  3-D fields  : (n_outer, nz, ny, nx)      2-D fields / weights : (n_outer, ny, nx)

Every function launches hand-written CUDA through the C ABI (include/coarsen_b200.h)
on torch's current stream.  torch is only the allocator / stream provider.
"""
import ctypes

import numpy as np
import torch

from . import _lib as L

_DT = {torch.float32: L.CG_F32, torch.float64: L.CG_F64}


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def device():
    if not torch.cuda.is_available():
        raise L.CoarsenLibraryError("fv3net_b200 needs a CUDA device; there is no CPU fallback")
    return torch.device("cuda", torch.cuda.current_device())


def to_device(a, dtype=None):
    """numpy / torch -> contiguous CUDA tensor (float32 or float64)."""
    if isinstance(a, torch.Tensor):
        t = a
    else:
        a = np.asarray(a)
        if a.dtype.kind != "f" or a.dtype.itemsize < 4:
            a = a.astype(np.float64 if a.dtype.kind in "iub" and a.dtype.itemsize >= 4 else np.float32)
        t = torch.from_numpy(np.ascontiguousarray(a))
    if dtype is not None and t.dtype != dtype:
        t = t.to(dtype)
    if t.dtype not in _DT:
        t = t.to(torch.float64 if t.dtype in (torch.int32, torch.int64) else torch.float32)
    if not t.is_cuda:
        t = t.to(device(), non_blocking=False)
    return t.contiguous()


def to_host(t):
    return t.detach().cpu().numpy()


def _p(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else None


def _promote(a, b):
    return torch.float64 if torch.float64 in (a, b) else torch.float32


def weighted_block_average(objs, w2d=None, w3d=None, fy=8, fx=8, sy=None, sx=None, nyo=None, nxo=None):
    """coarsen.py:183-218 / :221-273.  objs: list of (n_outer, nz, ny, nx) tensors of one
    dtype sharing the weights; w2d (n_outer, ny, nx) and/or w3d (n_outer, nz, ny, nx); both
    given = mass weighting w3d*w2d formed in the kernel.  Returns a list of coarse tensors."""
    single = isinstance(objs, torch.Tensor)
    objs = [objs] if single else list(objs)
    n_outer, nz, ny, nx = objs[0].shape
    sy = fy if sy is None else sy
    sx = fx if sx is None else sx
    if nyo is None:
        if ny % sy or (fy != sy):
            raise ValueError("nyo must be given for non-tiling windows")
        nyo = ny // sy
    if nxo is None:
        if nx % sx or (fx != sx):
            raise ValueError("nxo must be given for non-tiling windows")
        nxo = nx // sx
    if w2d is not None and w3d is not None:
        mode, wdt = L.CG_WMASS, w3d.dtype
        if w2d.dtype != w3d.dtype:
            wdt = _promote(w2d.dtype, w3d.dtype)
            w2d, w3d = w2d.to(wdt), w3d.to(wdt)
    elif w3d is not None:
        mode, wdt = L.CG_W3D, w3d.dtype
    elif w2d is not None:
        mode, wdt = L.CG_W2D, w2d.dtype
    else:
        raise ValueError("weights are required")
    if w2d is not None and tuple(w2d.shape) != (n_outer, ny, nx):
        raise ValueError(f"2-D weights have shape {tuple(w2d.shape)}, expected {(n_outer, ny, nx)}")
    if w3d is not None and tuple(w3d.shape) != (n_outer, nz, ny, nx):
        raise ValueError(f"3-D weights have shape {tuple(w3d.shape)}, expected {(n_outer, nz, ny, nx)}")
    dt = objs[0].dtype
    out_dt = _promote(dt, wdt)
    outs = []
    for i in range(0, len(objs), L.CG_MAX_FIELDS):
        chunk = objs[i:i + L.CG_MAX_FIELDS]
        for o in chunk:
            if o.dtype != dt or tuple(o.shape) != (n_outer, nz, ny, nx) or not o.is_contiguous():
                raise ValueError("all fields of one call must share dtype, shape and be contiguous")
        res = [torch.empty((n_outer, nz, nyo, nxo), dtype=out_dt, device=o.device) for o in chunk]
        pin, _k1 = L.ptr_array([o.data_ptr() for o in chunk])
        pout, _k2 = L.ptr_array([r.data_ptr() for r in res])
        L.check(L.lib().cg_weighted_block_average(
            pin, pout, len(chunk), _p(w3d), _p(w2d), mode, n_outer, nz, ny, nx,
            fy, fx, sy, sx, nyo, nxo, _DT[dt], _DT[wdt], _stream()))
        outs.extend(res)
    return outs[0] if single else outs


def block_reduce(a, method, fy, fx, sy=None, sx=None, nyo=None, nxo=None, skipna=True):
    """Window reduction of (n_outer, ny, nx); method in sum/min/max/mean.  skipna=True is
    xarray's coarsen (nansum...), False is numpy's plain reduction (NaN propagates)."""
    code = {"sum": L.CG_SUM, "min": L.CG_MIN, "max": L.CG_MAX, "mean": L.CG_MEAN}[method]
    if not skipna:
        code |= L.CG_NAN_PROPAGATE
    n_outer, ny, nx = a.shape
    sy = fy if sy is None else sy
    sx = fx if sx is None else sx
    nyo = (ny - fy) // sy + 1 if nyo is None else nyo
    nxo = (nx - fx) // sx + 1 if nxo is None else nxo
    out = torch.empty((n_outer, nyo, nxo), dtype=a.dtype, device=a.device)
    L.check(L.lib().cg_block_reduce(_p(a), _p(out), code, n_outer, ny, nx, fy, fx, sy, sx, nyo, nxo,
                                    _DT[a.dtype], _stream()))
    return out


def block_median(a, f):
    n_outer, ny, nx = a.shape
    if ny % f or nx % f:
        raise ValueError(f"dimension sizes ({ny}, {nx}) are not multiples of the block size {f}")
    out = torch.empty((n_outer, ny // f, nx // f), dtype=a.dtype, device=a.device)
    L.check(L.lib().cg_block_median(_p(a), _p(out), n_outer, ny, nx, f, _DT[a.dtype], _stream()))
    return out


def block_mode(a, f, nan_policy="propagate"):
    if nan_policy not in ("propagate", "omit"):
        raise ValueError("nan_policy must be 'propagate' or 'omit' on the device path")
    n_outer, ny, nx = a.shape
    if ny % f or nx % f:
        raise ValueError(f"dimension sizes ({ny}, {nx}) are not multiples of the block size {f}")
    out = torch.empty((n_outer, ny // f, nx // f), dtype=a.dtype, device=a.device)
    L.check(L.lib().cg_block_mode(_p(a), _p(out), n_outer, ny, nx, f, int(nan_policy == "omit"),
                                  _DT[a.dtype], _stream()))
    return out


def block_upsample(a, fy, fx, stag_y=False, stag_x=False):
    n_outer, nyc, nxc = a.shape
    ny = (nyc - 1) * fy + 1 if stag_y else nyc * fy
    nx = (nxc - 1) * fx + 1 if stag_x else nxc * fx
    out = torch.empty((n_outer, ny, nx), dtype=a.dtype, device=a.device)
    L.check(L.lib().cg_block_upsample(_p(a), _p(out), n_outer, nyc, nxc, fy, fx, int(stag_y),
                                      int(stag_x), _DT[a.dtype], _stream()))
    return out


def pressure_at_interface(delp, toa=300.0):
    """(n_outer, nz, *h) -> (n_outer, nz+1, *h)."""
    n_outer, nz = delp.shape[:2]
    h = int(np.prod(delp.shape[2:], dtype=np.int64))
    out = torch.empty((n_outer, nz + 1) + tuple(delp.shape[2:]), dtype=delp.dtype, device=delp.device)
    L.check(L.lib().cg_pressure_at_interface(_p(delp), _p(out), float(toa), n_outer, nz, h,
                                             _DT[delp.dtype], _stream()))
    return out


def mask_weights(w2d, pc, pf):
    """regridz.py:150-161: w2d (n_outer,*h), pc (n_outer,nz+1,*h), pf (n_outer,nzf+1,*h)."""
    n_outer, nz1 = pc.shape[:2]
    h = int(np.prod(pc.shape[2:], dtype=np.int64))
    if pf.dtype != pc.dtype:
        dt = _promote(pf.dtype, pc.dtype)
        pc, pf = pc.to(dt), pf.to(dt)
    out = torch.empty((n_outer, nz1 - 1) + tuple(pc.shape[2:]), dtype=w2d.dtype, device=pc.device)
    L.check(L.lib().cg_mask_weights(_p(w2d), _p(pc), _p(pf), _p(out), n_outer, nz1 - 1,
                                    pf.shape[1] - 1, h, _DT[pc.dtype], _DT[w2d.dtype], _stream()))
    return out


def interp_to_edges(c, axis):
    """(6, nz, n, n) cell-centred -> y-edges (axis='y') or x-edges (axis='x')."""
    t, nz, n, n2 = c.shape
    if t != 6 or n != n2:
        raise ValueError("interp_to_edges needs (6, nz, n, n) data on all six tiles")
    ax = {"y": 0, "x": 1}[axis]
    shape = (6, nz, n + 1, n) if ax == 0 else (6, nz, n, n + 1)
    out = torch.empty(shape, dtype=c.dtype, device=c.device)
    L.check(L.lib().cg_interp_to_edges(_p(c), _p(out), nz, n, ax, _DT[c.dtype], _stream()))
    return out


def mappm(pe1, q1, pe2, n_outer, n_inner, km, kn, s_pe1, s_q1, s_pe2, s_q2, out, iv=1, kord=1):
    """Strided PPM remap (fp32); s_* = (outer, column, level) element strides."""
    for t in (pe1, q1, pe2, out):
        if t.dtype != torch.float32:
            raise ValueError("mappm runs in float32 (the Fortran `real`)")
    strides = (ctypes.c_long * 12)(*s_pe1, *s_q1, *s_pe2, *s_q2)
    L.check(L.lib().cg_mappm(_p(pe1), _p(q1), _p(pe2), _p(out), n_outer, n_inner, km, kn, strides,
                             int(iv), int(kord), _stream()))
    return out


_WORKSPACE = {}


def _workspace(nbytes, dev):
    key = (dev.index, torch.cuda.current_stream().cuda_stream)
    ws = _WORKSPACE.get(key)
    if ws is None or ws.numel() < nbytes:
        ws = torch.empty(max(nbytes, 1), dtype=torch.uint8, device=dev)
        _WORKSPACE[key] = ws
    return ws


def release_workspaces():
    """Drop the cached workspaces (one per device and stream; they only grow otherwise)."""
    _WORKSPACE.clear()


def product(a, b):
    """Field descriptor: q = a * b, formed inside the kernel (budgets.py:151)."""
    return ("product", a, b, 1.0)


def storage(end, begin, dt):
    """Field descriptor: q = (end - begin) / dt, formed inside the kernel (budgets.py:80)."""
    return ("diff", end, begin, float(dt))


def _describe(fields):
    descs = (L.CgField * max(len(fields), 1))()
    keep = []
    for i, fld in enumerate(fields):
        if isinstance(fld, tuple):
            kind, a, b, scale = fld
            op = {"product": L.CG_FIELD_PRODUCT, "diff": L.CG_FIELD_DIFF}[kind]
        else:
            a, b, op, scale = fld, None, L.CG_FIELD_PLAIN, 1.0
        for t in (a, b):
            if t is not None and (t.dtype != torch.float32 or not t.is_contiguous()):
                raise ValueError("pressure_level_average needs contiguous float32 tensors")
        descs[i].a = a.data_ptr()
        descs[i].b = b.data_ptr() if b is not None else None
        descs[i].op = op
        descs[i].scale = scale
        keep.append((a, b))
    return descs, keep


def pressure_level_average(fields, delp, area, f, mask_below=False, toa=300.0, want_delp_coarse=True,
                           outs=None, delp_coarse_out=None, replica_offsets=None, multicast_offset=None):
    """Fused regrid-to-coarse-pressure + (masked) area-weighted block average.
    fields: list of (n_outer, nz, ny, nx) float32 tensors or descriptors from product() /
    storage(); delp same shape; area (n_outer, ny, nx).
    outs / delp_coarse_out: arrays the kernels write their results into (contiguous float32 of the coarse shape,
    e.g. views of a collective's send buffer: sharding.CoarseBuffers) instead of new ones.
    replica_offsets: byte offsets at which every coarse value is stored once more by the kernel itself
    (cg_pressure_level_average_replicated: peer copies of the output buffer over NVLink; needs `outs`).
    multicast_offset: instead, the byte offset from the output buffer to its NVLS multicast mapping - every value is
    stored once with multimem.st and the NVSwitch writes all copies.
    Returns (list of coarse fields, delp_coarse or None)."""
    fields = list(fields)
    n_outer, nz, ny, nx = delp.shape
    if ny % f or nx % f:
        raise ValueError(f"dimension sizes ({ny}, {nx}) are not multiples of the factor {f}")
    for t in (delp, area):
        if t.dtype != torch.float32 or not t.is_contiguous():
            raise ValueError("pressure_level_average needs contiguous float32 tensors")
    dev = delp.device
    cshape = (n_outer, nz, ny // f, nx // f)

    def given(t, what):
        if tuple(t.shape) != cshape or t.dtype != torch.float32 or not t.is_contiguous() or t.device != dev:
            raise ValueError(f"{what} must be a contiguous float32 array of shape {cshape} on {dev}")
        return t

    if delp_coarse_out is not None:
        dcoarse = given(delp_coarse_out, "delp_coarse_out")
    else:
        dcoarse = torch.empty(cshape, dtype=torch.float32, device=dev) if want_delp_coarse else None
    outs_given = outs is not None
    if outs is not None:
        outs = [given(t, f"outs[{i}]") for i, t in enumerate(outs)]
        if len(outs) != len(fields):
            raise ValueError(f"{len(outs)} output arrays for {len(fields)} fields")
    else:
        outs = [torch.empty(cshape, dtype=torch.float32, device=dev) for _ in fields]
    reps = [int(x) for x in (replica_offsets if replica_offsets is not None else [])]
    kind = 0
    if multicast_offset is not None:
        if reps:
            raise ValueError("either replica_offsets or multicast_offset")
        reps, kind = [int(multicast_offset)], 1
    if reps and outs_given is False:
        raise ValueError("replica_offsets need the caller's own output arrays (outs=, delp_coarse_out=)")
    rep_arr = (ctypes.c_longlong * max(len(reps), 1))(*reps)
    step = L.CG_MAX_FIELDS
    first = True
    for i in range(0, max(len(fields), 1), step):
        chunk, res = fields[i:i + step], outs[i:i + step]
        nbytes = L.lib().cg_pressure_level_workspace_bytes(len(chunk), n_outer, nz, ny, nx, f)
        ws = _workspace(nbytes, dev)
        descs, _keep = _describe(chunk)
        pout, _k2 = L.ptr_array([t.data_ptr() for t in res])
        args = (ctypes.cast(descs, ctypes.c_void_p), pout, len(chunk), _p(delp), _p(area),
                _p(dcoarse) if first else None, n_outer, nz, ny, nx, f, int(mask_below), float(toa))
        if reps:                            # (no fallback: only the pair-thread kernel stores replicas)
            L.check(L.lib().cg_pressure_level_average_replicated(*args, kind, len(reps), rep_arr, _p(ws), ws.numel(), _stream()))
            first = False
            continue
        rc = L.lib().cg_pressure_level_average(*args, _p(ws), ws.numel(), _stream())
        if rc == L.CG_ENOTSUP:
            # arguments the streaming kernels do not take (e.g. views that are not 16-byte aligned): the
            # composed recipe needs its fine-grid temporaries
            ws = _workspace(L.lib().cg_pressure_level_fallback_workspace_bytes(len(chunk), n_outer, nz, ny, nx, f), dev)
            rc = L.lib().cg_pressure_level_average(*args, _p(ws), ws.numel(), _stream())
        L.check(rc)
        first = False
    return outs, dcoarse


def multicast_store(t, mc_offset):
    """Broadcast the contiguous tensor `t` (a region of a symmetric buffer) into every GPU's copy of that buffer:
    one multimem.st per 16 bytes to `t`'s address + mc_offset (sharding.CoarseBuffers.multicast_offset())."""
    if not t.is_contiguous():
        raise ValueError("multicast_store needs a contiguous tensor")
    L.check(L.lib().cg_multicast_store(_p(t), int(mc_offset), t.numel() * t.element_size(), _stream()))


def area_above_fine_surface(delp, delp_coarse, area, f, toa=300.0):
    n_outer, nz, ny, nx = delp.shape
    out = torch.empty((n_outer, nz, ny // f, nx // f), dtype=torch.float32, device=delp.device)
    L.check(L.lib().cg_area_above_fine_surface(_p(delp), _p(delp_coarse), _p(area), _p(out), n_outer,
                                               nz, ny, nx, f, float(toa), _stream()))
    return out


# ---- restart-coarsening recipes (coarsen_restarts.py): element-wise steps and column thermodynamics
_EW = {"mul": L.CG_EW_MUL, "isclose": L.CG_EW_ISCLOSE, "and": L.CG_EW_AND, "where": L.CG_EW_WHERE,
       "gt": L.CG_EW_GT, "lt": L.CG_EW_LT, "blend": L.CG_EW_BLEND, "fillna": L.CG_EW_FILLNA, "sub": L.CG_EW_SUB, "div": L.CG_EW_DIV}


def elementwise(op, a, b=None, c=None):
    """out = op(a, b, c) on same-shape tensors of one dtype; `b` may be a Python scalar.
    Masks are 0 / 1 in the data dtype.  op: mul, isclose, and, where (c ? a : b), gt, lt,
    blend (a*b + (1-a)*c), fillna."""
    scalar = 0.0
    tb = b
    if not isinstance(b, torch.Tensor):
        scalar, tb = (0.0 if b is None else float(b)), None
    for t in (tb, c):
        if t is not None and (t.dtype != a.dtype or tuple(t.shape) != tuple(a.shape)):
            raise ValueError("elementwise operands must share shape and dtype")
    a = a.contiguous()
    tb = tb.contiguous() if tb is not None else None
    c = c.contiguous() if c is not None else None
    out = torch.empty_like(a)
    L.check(L.lib().cg_elementwise(_EW[op], _p(a), _p(tb), _p(c), scalar, _p(out), a.numel(),
                                   _DT[a.dtype], _stream()))
    return out


def _column_shape(t):
    n_outer, nz = t.shape[:2]
    return n_outer, nz, int(np.prod(t.shape[2:], dtype=np.int64))


def surface_pressure(delp, toa=300.0):
    """(n_outer, nz, *h) -> (n_outer, *h): sum over z + toa (vertically_dependent.py:189-208)."""
    n_outer, nz, h = _column_shape(delp)
    out = torch.empty((n_outer,) + tuple(delp.shape[2:]), dtype=delp.dtype, device=delp.device)
    L.check(L.lib().cg_surface_pressure(_p(delp), _p(out), float(toa), n_outer, nz, h,
                                        _DT[delp.dtype], _stream()))
    return out


def surface_pressure_nansum(delp, toa=300.0):
    """(n_outer, nz, *h) -> (n_outer, *h): NaN-skipping sum over z + toa (xarray's DataArray.sum)."""
    n_outer, nz, h = _column_shape(delp)
    out = torch.empty((n_outer,) + tuple(delp.shape[2:]), dtype=delp.dtype, device=delp.device)
    L.check(L.lib().cg_surface_pressure_nansum(_p(delp), _p(out), float(toa), n_outer, nz, h,
                                               _DT[delp.dtype], _stream()))
    return out


def pressure_at_midpoint_log(delp, toa=300.0):
    """(n_outer, nz, *h) -> same shape: delp / diff(log(p_interface)) (vertically_dependent.py:153-179)."""
    n_outer, nz, h = _column_shape(delp)
    out = torch.empty_like(delp)
    L.check(L.lib().cg_pressure_at_midpoint_log(_p(delp), _p(out), float(toa), n_outer, nz, h,
                                                _DT[delp.dtype], _stream()))
    return out


def shift_to_center(a, axis):
    """0.5 * (a[i + 1] + a[i]) along `axis` (coarsen.py:54-80): that dim shrinks by one."""
    a = a.contiguous()
    shape = list(a.shape)
    n_outer = int(np.prod(shape[:axis], dtype=np.int64)) if axis else 1
    inner = int(np.prod(shape[axis + 1:], dtype=np.int64)) if axis + 1 < len(shape) else 1
    shape[axis] -= 1
    out = torch.empty(shape, dtype=a.dtype, device=a.device)
    L.check(L.lib().cg_shift_to_center(_p(a), _p(out), n_outer, a.shape[axis], inner, _DT[a.dtype], _stream()))
    return out


def blending_weights(delp_coarse, ps_min, toa=300.0, sigma=0.9):
    """coarsen_restarts.py:519-581: delp_coarse (n_outer, nz, *h), ps_min (n_outer, *h) the block
    minimum of the fine surface pressure -> weights (n_outer, nz, *h)."""
    n_outer, nz, h = _column_shape(delp_coarse)
    if ps_min.dtype != delp_coarse.dtype or ps_min.numel() != n_outer * h:
        raise ValueError("ps_min must be (n_outer, *h) in the dtype of delp_coarse")
    out = torch.empty_like(delp_coarse)
    L.check(L.lib().cg_blending_weights(_p(delp_coarse), _p(ps_min), _p(out), float(toa), float(sigma),
                                        n_outer, nz, h, _DT[delp_coarse.dtype], _stream()))
    return out


def hydrostatic_balance(dz, phis, temp, sphum, delp, toa=300.0):
    """coarsen_restarts.py:916-938: returns (hydrostatic DZ, adjusted phis) keeping the model-top
    height.  3-D arguments (n_outer, nz, *h), phis (n_outer, *h), one dtype."""
    n_outer, nz, h = _column_shape(dz)
    for t in (temp, sphum, delp):
        if t.dtype != dz.dtype or tuple(t.shape) != tuple(dz.shape):
            raise ValueError("hydrostatic_balance operands must share shape and dtype")
    if phis.dtype != dz.dtype or phis.numel() != n_outer * h:
        raise ValueError("phis must be (n_outer, *h) in the dtype of DZ")
    dz_out, phis_out = torch.empty_like(dz), torch.empty_like(phis)
    L.check(L.lib().cg_hydrostatic_balance(_p(dz), _p(phis), _p(temp), _p(sphum), _p(delp), _p(dz_out),
                                           _p(phis_out), float(toa), n_outer, nz, h, _DT[dz.dtype],
                                           _stream()))
    return dz_out, phis_out


def interpolate_2d(xp, x, y, fill_value=float("nan"), strides=None):
    """interpolate_2d.f90: rows of y given at x, linearly interpolated to xp; float64 tensors
    (m, n_out), (m, n_in), (m, n_in) -> (m, n_out).  `strides` = ((row, level) element strides of
    xp, x, y, out) for layouts other than row-major."""
    for t in (xp, x, y):
        if t.dtype != torch.float64:
            raise ValueError("interpolate_2d runs in float64 (the Fortran's real(8))")
    m, n_out = xp.shape
    n_in = x.shape[1]
    if tuple(x.shape) != (m, n_in) or tuple(y.shape) != (m, n_in):
        raise ValueError(f"shape mismatch: xp {tuple(xp.shape)}, x {tuple(x.shape)}, y {tuple(y.shape)}")
    out = torch.empty((m, n_out), dtype=torch.float64, device=xp.device)
    if strides is None:
        strides = tuple(t.stride() for t in (xp, x, y, out))
    flat = (ctypes.c_long * 8)(*[int(v) for pair in strides for v in pair])
    L.check(L.lib().cg_interpolate_2d(_p(xp), _p(x), _p(y), _p(out), float(fill_value), m, n_in, n_out,
                                      flat, _stream()))
    return out
