"""ctypes binding of libcoarsen_b200.so (the C ABI in include/coarsen_b200.h).

The product has no CPU fallback: if the CUDA library is missing, importing a
compute entry point raises, loudly.
"""
import ctypes
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libcoarsen_b200.so")
CSRC = os.path.join(_HERE, "csrc")

CG_OK, CG_EINVAL, CG_ENOTSUP, CG_ECUDA = 0, -1, -2, -3
CG_F32, CG_F64 = 0, 1
CG_SUM, CG_MIN, CG_MAX, CG_MEAN = 0, 1, 2, 3
CG_NAN_PROPAGATE = 16
CG_W2D, CG_W3D, CG_WMASS = 0, 1, 2
CG_MAX_FIELDS = 32

_c = ctypes
_vp, _i, _l, _d = _c.c_void_p, _c.c_int, _c.c_long, _c.c_double
_pp = _c.POINTER(_c.c_void_p)

# name -> (restype, argtypes); mirrors include/coarsen_b200.h one to one
SIGNATURES = {
    "cg_last_error": (_c.c_char_p, []),
    "cg_version": (_i, []),
    "cg_set_option": (_i, [_c.c_char_p, _c.c_char_p]),
    "cg_launch_count": (_l, []),
    "cg_device_info": (_i, [_c.POINTER(_i), _c.POINTER(_i), _c.POINTER(_l)]),
    "cg_weighted_block_average": (_i, [_pp, _pp, _i, _vp, _vp, _i, _l, _i, _i, _i,
                                       _i, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "cg_block_reduce": (_i, [_vp, _vp, _i, _l, _i, _i, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "cg_block_median": (_i, [_vp, _vp, _l, _i, _i, _i, _i, _vp]),
    "cg_block_mode": (_i, [_vp, _vp, _l, _i, _i, _i, _i, _i, _vp]),
    "cg_block_upsample": (_i, [_vp, _vp, _l, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "cg_pressure_at_interface": (_i, [_vp, _vp, _d, _l, _i, _l, _i, _vp]),
    "cg_mask_weights": (_i, [_vp, _vp, _vp, _vp, _l, _i, _i, _l, _i, _i, _vp]),
    "cg_interp_to_edges": (_i, [_vp, _vp, _i, _i, _i, _i, _vp]),
    "cg_mappm": (_i, [_vp, _vp, _vp, _vp, _l, _l, _i, _i, _c.POINTER(_l), _i, _i, _vp]),
    "cg_pressure_level_last_impl": (_c.c_char_p, []),
    "cg_pressure_level_workspace_bytes": (_l, [_i, _l, _i, _i, _i, _i]),
    "cg_pressure_level_fallback_workspace_bytes": (_l, [_i, _l, _i, _i, _i, _i]),
    "cg_pressure_level_average": (_i, [_vp, _pp, _i, _vp, _vp, _vp, _l, _i, _i, _i, _i, _i, _d,
                                       _vp, _l, _vp]),
    "cg_pressure_level_average_replicated": (_i, [_vp, _pp, _i, _vp, _vp, _vp, _l, _i, _i, _i, _i, _i, _d,
                                                  _i, _i, _c.POINTER(_c.c_longlong), _vp, _l, _vp]),
    "cg_multicast_store": (_i, [_vp, _c.c_longlong, _l, _vp]),
    "cg_area_above_fine_surface": (_i, [_vp, _vp, _vp, _vp, _l, _i, _i, _i, _i, _d, _vp]),
    "cg_static_scratch_bytes": (_l, []),
    "cg_area_above_fine_surface_static": (_i, [_vp, _vp, _vp, _vp, _l, _i, _i, _i, _i, _d, _vp]),
    "cg_interpolate_2d": (_i, [_vp, _vp, _vp, _vp, _d, _l, _i, _i, _c.POINTER(_l), _vp]),
    "cg_elementwise": (_i, [_i, _vp, _vp, _vp, _d, _vp, _l, _i, _vp]),
    "cg_surface_pressure": (_i, [_vp, _vp, _d, _l, _i, _l, _i, _vp]),
    "cg_surface_pressure_nansum": (_i, [_vp, _vp, _d, _l, _i, _l, _i, _vp]),
    "cg_pressure_at_midpoint_log": (_i, [_vp, _vp, _d, _l, _i, _l, _i, _vp]),
    "cg_shift_to_center": (_i, [_vp, _vp, _l, _i, _l, _i, _vp]),
    "cg_blending_weights": (_i, [_vp, _vp, _vp, _d, _d, _l, _i, _l, _i, _vp]),
    "cg_hydrostatic_balance": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _d, _l, _i, _l, _i, _vp]),
}
(CG_EW_MUL, CG_EW_ISCLOSE, CG_EW_AND, CG_EW_WHERE, CG_EW_GT, CG_EW_LT, CG_EW_BLEND,
 CG_EW_FILLNA, CG_EW_SUB, CG_EW_DIV) = range(10)

CG_FIELD_PLAIN, CG_FIELD_PRODUCT, CG_FIELD_DIFF = 0, 1, 2


class CgField(ctypes.Structure):
    _fields_ = [("a", ctypes.c_void_p), ("b", ctypes.c_void_p), ("op", ctypes.c_int),
                ("scale", ctypes.c_float)]


_LIB = None


class CoarsenLibraryError(RuntimeError):
    pass


def build(verbose=False):
    """Compile the CUDA library in-tree for sm_100a (nvcc cross-compiles without a GPU)."""
    cmd = ["make", "-C", CSRC, "-j8"]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if verbose or res.returncode != 0:
        print(res.stdout)
        print(res.stderr)
    if res.returncode != 0:
        raise CoarsenLibraryError("building libcoarsen_b200.so failed (see output above)")
    return LIB_PATH


def lib():
    """The loaded library.  Raises if it has not been built: there is no fallback."""
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise CoarsenLibraryError(
                f"{LIB_PATH} is missing. Build it with `python -c 'import __graft_entry__ as g; "
                "g.build()'` (or `make -C fv3net_b200/csrc`). fv3net_b200 has no CPU fallback."
            )
        handle = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)  # AttributeError if the header and the .so diverge
            fn.restype = res
            fn.argtypes = args
        _LIB = handle
    return _LIB


def check(rc):
    """Map a C status to the exception the reference's Python raises for the same mistake."""
    if rc == CG_OK:
        return
    msg = lib().cg_last_error().decode("utf-8", "replace")
    if rc == CG_EINVAL:
        raise ValueError(msg)
    if rc == CG_ENOTSUP:
        raise NotImplementedError(msg)
    raise CoarsenLibraryError(msg)


def set_option(name, value):
    """Per-thread behaviour switch of the library (include/coarsen_b200.h, cg_set_option)."""
    check(lib().cg_set_option(name.encode(), str(value).encode()))


def ptr_array(ptrs):
    arr = (ctypes.c_void_p * len(ptrs))(*ptrs)
    return ctypes.cast(arr, _pp), arr
