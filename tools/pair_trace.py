"""Timeline of one CTA of the pair-thread kernel (library built with EXTRA=-DCG_PAIR_TRACE):
python tools/pair_trace.py  -> per-stage table of who waits for whom, in cycles."""
import ctypes, os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from fv3net_b200 import _lib, budget as B

class A: tiles = 6; n = 384
spec = bench.workload(A)
inp = bench.make_inputs(spec, torch.device("cuda", 0), seed=0)
for _ in range(2):
    B.recoarsen_budget_inputs(**inp)
torch.cuda.synchronize()
lib = ctypes.CDLL(_lib.LIB_PATH)
S, T, V = 10, 256, 6
buf = np.zeros(S * T * V, dtype=np.int64)
lib.cg_debug_pair_trace.argtypes = [ctypes.c_void_p, ctypes.c_long]
assert lib.cg_debug_pair_trace(buf.ctypes.data, buf.size) == 0
tr = buf.reshape(S, T, V)
t0 = tr[tr > 0].min()
tr = np.where(tr > 0, tr - t0, -1)
nch, total = 28, 224
print("walker rows (start, after cempty wait, done):")
for r in range(8): print("  row", r, tr[9, r, :3])
print("it  chunk | geom: start  wait_handback  work  done_at | field0: wait_rec  wait_tma  work  done_at | fieldN done_at")
prev_f = 0
for it in range(total):
    g = it % 2 * 3          # first warp of the set (sub = 0); worst of the three below
    gs = [tr[(it % 2) * 3 + k, it] for k in range(3)]
    gdone = max(x[3] for x in gs); gstart = min(x[0] for x in gs)
    gwait = max(x[2] - x[1] for x in gs); gwork = max(x[3] - x[2] for x in gs)
    gtma = max(x[4] - x[2] for x in gs); gtop = max(x[1] - x[0] for x in gs)
    f, fn = tr[7, it], tr[8, it]
    print(f"{it:3d} {it % nch:3d} | {gstart:8d} {gwait:6d} {gwork:6d} {gdone:8d} top {gtop:5d} tma {gtma:5d} | {f[1]-f[0]:6d} {f[2]-f[1]:6d} {f[3]-f[2]:6d} {f[3]:8d} | {fn[3]:8d}  period {f[3]-prev_f}")
    prev_f = f[3]
print("total cycles", tr[7, total - 1, 3])
