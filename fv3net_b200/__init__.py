"""fv3net_b200 — B200-native (sm_100a) coarse-graining hot path of fv3net's `vcm`.

Layers (top to bottom):
  fv3net_b200.cubedsphere   the reference's Python API (same names and signatures as
                            vcm.cubedsphere.{coarsen,regridz}) on labelled arrays
  fv3net_b200.budget        the fine_res_budget composition (workflows/fine_res_budget)
  fv3net_b200.ops           array-level entry points on torch CUDA tensors
  fv3net_b200._lib          ctypes binding of lib/libcoarsen_b200.so (include/coarsen_b200.h)
  fv3net_b200/csrc          hand-written CUDA kernels + the C ABI

There is no CPU fallback: without the built CUDA library every compute call raises.
"""
__version__ = "0.1.0"
