"""Launch count and time of the restart-coarsening recipes (coarsen_restarts.py) at the reference's own test size
(48 x 48 fine cells per tile, factor 2, 6 tiles x 79 levels: launch-bound) - what VERDICT r01 weak 13 asked to be timed:
    python tools/restart_recipes_timing.py"""
import os, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fv3net_b200 import _lib
from fv3net_b200.cubedsphere import coarsen_restarts as R
from tests import synth_inputs as S

CATS = ["fv_core.res", "fv_tracer.res", "fv_srf_wnd.res", "sfc_data"]
restarts = {c: S.dataset(c) for c in CATS}
grid = S.dataset("grid")
cases = [("coarsen_restarts_on_sigma (mass-weighted, agrid winds)", lambda: R.coarsen_restarts_on_sigma(2, grid, restarts, coarsen_agrid_winds=True, mass_weighted=True)),
         ("coarsen_restarts_on_pressure (agrid winds)", lambda: R.coarsen_restarts_on_pressure(2, grid, restarts, coarsen_agrid_winds=True)),
         ("coarsen_restarts_via_blended_method (agrid winds)", lambda: R.coarsen_restarts_via_blended_method(2, grid, restarts, coarsen_agrid_winds=True))]
sizes = {c: sum(int(restarts[c][v].values.size) for v in restarts[c].data_vars) for c in CATS}
print("# inputs (elements per category):", sizes)
print(f"{'recipe':62s} {'launches':>9s} {'ms per call (host arrays in and out)':>40s}")
for name, fn in cases:
    fn(); torch.cuda.synchronize()
    n0 = _lib.lib().cg_launch_count()
    t0 = time.perf_counter()
    reps = 3
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / reps
    n = (_lib.lib().cg_launch_count() - n0) // reps
    print(f"{name:62s} {n:9d} {1e3 * dt:40.1f}")
