"""Achieved HBM bandwidth of every kernel on the path beside the dominant one, at BASELINE.json's sizes.

    python tools/kernel_rooflines.py > gpurun_out/kernel_rooflines.txt        (on the GPU box)

Each line: algorithmic bytes (every input element read once, every output element written once), the average
launch time over 20 launches after 3 warm-ups (CUDA events on the launching stream; every case moves more
than the 126 MB L2 or is flushed by a 512 MB write between launches) and the fraction of the measured copy
bandwidth (MEASURED_PEAKS.json hbm_gbs).  The dominant kernel is bench.py's business.
"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fv3net_b200 import ops, synthetic as syn  # noqa: E402

dev = torch.device("cuda")
try:
    PEAK = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    PEAK = 6400.0
flush_buf = torch.empty(512 << 20, dtype=torch.uint8, device=dev)


def timeit(fn, flush, n=20, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    tot = 0.0
    for _ in range(n):
        if flush:
            flush_buf.fill_(1)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        b.synchronize()
        tot += a.elapsed_time(b)
    return tot / n


def report(name, nbytes, fn, flush=None):
    flush = nbytes < (256 << 20) if flush is None else flush
    ms = timeit(fn, flush)
    gbs = nbytes / ms / 1e6
    print(f"{name:66s} {nbytes / 1e6:10.1f} MB {ms * 1e3:10.1f} us {gbs:8.0f} GB/s {100 * gbs / PEAK:6.1f} % of {PEAK:.0f}"
          + ("  (L2 flushed)" if flush else ""))


nt, nz, n, F = 6, syn.NZ, 384, 8
delp = syn.torch_delp(nt, n, n, 1, dev)
area = syn.torch_area((nt, n, n), 2, dev)
fields = [syn.torch_field((nt, nz, n, n), 10 + i, dev) for i in range(13)]
f3, f2, c3 = 4 * nt * nz * n * n, 4 * nt * n * n, 4 * nt * nz * (n // F) ** 2

report("weighted_block_average, 1 field, area weights (C384, 79 levels)", f3 + f2 + c3,
       lambda: ops.weighted_block_average(fields[0], w2d=area, fy=F, fx=F))
report("weighted_block_average, 13 fields, mass weights delp*area", 14 * f3 + f2 + 13 * c3,
       lambda: ops.weighted_block_average(fields, w2d=area, w3d=delp, fy=F, fx=F))
c3_16 = 4 * nt * nz * (n // 16) ** 2
report("weighted_block_average f=16 (C384 -> C24), 13 fields, mass weights", 14 * f3 + f2 + 13 * c3_16,
       lambda: ops.weighted_block_average(fields, w2d=area, w3d=delp, fy=16, fx=16))
report("weighted_block_average f=16, 1 field, area weights", f3 + f2 + c3_16,
       lambda: ops.weighted_block_average(fields[0], w2d=area, fy=16, fx=16))
u = syn.torch_field((nt, nz, n + 1, n), 40, dev)
dx = syn.torch_area((nt, n + 1, n), 41, dev)
report("edge_weighted_block_average u/dx (reads the 49 kept rows only)", (f3 + f2) // 8 + c3,
       lambda: ops.weighted_block_average(u, w2d=dx, fy=1, fx=F, sy=F, sx=F, nyo=n // F + 1, nxo=n // F))
report("pressure_at_interface (79 -> 80 levels)", f3 + f3 // nz * (nz + 1), lambda: ops.pressure_at_interface(delp))
dc = ops.weighted_block_average(delp, w2d=area, fy=F, fx=F)
report("area_above_fine_surface (exposed area, delp re-read)", f3 + f2 + 2 * c3,
       lambda: ops.area_above_fine_surface(delp, dc, area, F))
sweep = torch.rand((96 * 6, n, n), device=dev)
med = ops.block_median(sweep, F)
report("block_median f=8, 96 x 6 x 384 x 384 (configs[3])", sweep.numel() * 4 + med.numel() * 4, lambda: ops.block_median(sweep, F))
report("block_upsample f=8 back to 384 x 384 (configs[3])", sweep.numel() * 4 + med.numel() * 4, lambda: ops.block_upsample(med, F, F))
report("block_reduce sum f=8, 576 images", sweep.numel() * 4 + med.numel() * 4, lambda: ops.block_reduce(sweep, "sum", F, F))
ints = torch.randint(0, 5, (96 * 6, n, n), device=dev).float()
report("block_mode f=8, 576 images (5 surface types)", ints.numel() * 4 + med.numel() * 4, lambda: ops.block_mode(ints, F))
# the unfused remapper (regrid_vertical): fine-grid result, like the reference's mappm call
pe1 = ops.pressure_at_interface(delp)
dcu = torch.repeat_interleave(torch.repeat_interleave(dc, F, dim=2), F, dim=3)
pe2 = ops.pressure_at_interface(dcu)
out = torch.empty_like(delp)
h = n * n
from fv3net_b200 import _lib  # noqa: E402
report("cg_mappm kord=1, streaming remapper (one C384 field, fine-grid result)", 4 * f3 + 2 * f3 // nz,
       lambda: ops.mappm(pe1, fields[0], pe2, nt, h, nz, nz, ((nz + 1) * h, 1, h), (nz * h, 1, h), ((nz + 1) * h, 1, h),
                         (nz * h, 1, h), out, iv=1, kord=1))
_lib.set_option("mappm_impl", "generic")
report("cg_mappm kord=1, generic kernel (column work arrays in local memory)", 4 * f3 + 2 * f3 // nz,
       lambda: ops.mappm(pe1, fields[0], pe2, nt, h, nz, nz, ((nz + 1) * h, 1, h), (nz * h, 1, h), ((nz + 1) * h, 1, h),
                         (nz * h, 1, h), out, iv=1, kord=1))
_lib.set_option("mappm_impl", "auto")
report("cg_mappm kord=9 (cubic-spline profile)", 4 * f3 + 2 * f3 // nz,
       lambda: ops.mappm(pe1, fields[0], pe2, nt, h, nz, nz, ((nz + 1) * h, 1, h), (nz * h, 1, h), ((nz + 1) * h, 1, h),
                         (nz * h, 1, h), out, iv=1, kord=9))
