"""cg_mappm on one C384 field (fine-grid result), streaming and generic: python tools/mappm_bench.py [n_launches]"""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fv3net_b200 import _lib, ops, synthetic as syn
dev = torch.device("cuda"); nt, nz, n, F = 6, syn.NZ, 384, 8
delp = syn.torch_delp(nt, n, n, 1, dev); area = syn.torch_area((nt, n, n), 2, dev)
q = syn.torch_field((nt, nz, n, n), 10, dev)
dc = ops.weighted_block_average(delp, w2d=area, fy=F, fx=F)
pe1 = ops.pressure_at_interface(delp)
pe2 = ops.pressure_at_interface(torch.repeat_interleave(torch.repeat_interleave(dc, F, dim=2), F, dim=3))
out = torch.empty_like(delp); h = n * n
def run():
    ops.mappm(pe1, q, pe2, nt, h, nz, nz, ((nz + 1) * h, 1, h), (nz * h, 1, h), ((nz + 1) * h, 1, h), (nz * h, 1, h), out, iv=1, kord=1)
for impl in ("auto", "generic"):
    _lib.set_option("mappm_impl", impl)
    for _ in range(3): run()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    N = int(sys.argv[1]) if len(sys.argv) > 1 else 10
    a.record()
    for _ in range(N): run()
    b.record(); b.synchronize()
    print(impl, a.elapsed_time(b) / N, "ms per launch")
