"""TEST INFRASTRUCTURE — ctypes binding of oracle/libcgref.so.

``mappm(pe1, q1, pe2, i1, i2, iv, kord, ptop)`` has the argument list of the
f2py extension the reference calls (regridz.py:273-275; mappm.f90:10): arrays
are (ncol, nz) with z last, inputs are cast to float32 the way f2py casts to
Fortran ``real``, the result is float32 (test_mappm.py:14).
"""
import ctypes
import os
import subprocess
from concurrent.futures import ThreadPoolExecutor

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

_STRIDES = [ctypes.c_long] * 12


def build_library(force=False):
    """Compile oracle/libcgref.so with gcc (building the checker is not using it)."""
    so = os.path.join(_HERE, "libcgref.so")
    srcs = [os.path.join(_HERE, f) for f in ("cgref.c", "cgref_mappm.inc", "Makefile")]
    if force or not os.path.exists(so) or any(
        os.path.getmtime(s) > os.path.getmtime(so) for s in srcs
    ):
        subprocess.check_call(["make", "-C", _HERE, "-s", "libcgref.so"])
    return so


def _lib():
    global _LIB
    if _LIB is None:
        lib = ctypes.CDLL(build_library())
        for sfx in ("f32", "f64"):
            fn = getattr(lib, f"cgref_mappm_{sfx}")
            fn.restype = ctypes.c_int
            fn.argtypes = (
                [ctypes.c_void_p] * 4
                + [ctypes.c_long, ctypes.c_long, ctypes.c_int, ctypes.c_int]
                + _STRIDES
                + [ctypes.c_int, ctypes.c_int, ctypes.c_int]
            )
        lib.cgref_ppm_profile_f32.restype = ctypes.c_int
        lib.cgref_ppm_profile_f32.argtypes = (
            [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int]
            + [ctypes.c_void_p] * 3
        )
        _LIB = lib
    return _LIB


def _check(rc):
    if rc == -1:
        raise ValueError("cgref: invalid argument (km >= 4 and kn >= 1 required)")
    if rc == -2:
        raise NotImplementedError("cgref: kord > 7 with iv = -2 reads an uninitialised array in the reference (mappm.f90:35, :163)")
    if rc != 0:
        raise RuntimeError(f"cgref: error {rc}")


def mappm_strided(pe1, q1, pe2, n_outer, n_inner, km, kn, s_pe1, s_q1, s_pe2, s_q2,
                  iv=1, kord=1, dtype=np.float32, out=None, threads=1):
    """Strided batch remap on flat arrays.  ``s_*`` = (outer, column, level)
    strides in elements.  Returns the flat output array ``out``."""
    dtype = np.dtype(dtype)
    fn = _lib().cgref_mappm_f32 if dtype == np.float32 else _lib().cgref_mappm_f64
    pe1 = np.ascontiguousarray(pe1, dtype=dtype).ravel()
    q1 = np.ascontiguousarray(q1, dtype=dtype).ravel()
    pe2 = np.ascontiguousarray(pe2, dtype=dtype).ravel()
    if out is None:
        n_out = (n_outer - 1) * s_q2[0] + (n_inner - 1) * s_q2[1] + (kn - 1) * s_q2[2] + 1
        out = np.empty(n_out, dtype=dtype)
    isz = dtype.itemsize

    def run(o0, o1, i0, i1):
        def off(a, s):
            return ctypes.c_void_p(a.ctypes.data + (o0 * s[0] + i0 * s[1]) * isz)

        _check(fn(off(pe1, s_pe1), off(q1, s_q1), off(pe2, s_pe2), off(out, s_q2),
                  o1 - o0, i1 - i0, km, kn, *s_pe1, *s_q1, *s_pe2, *s_q2, iv, kord, 1))

    if threads <= 1:
        run(0, n_outer, 0, n_inner)
    else:
        # split the inner (column) range; ctypes releases the GIL during the call
        edges = np.linspace(0, n_inner, threads + 1).astype(int)
        with ThreadPoolExecutor(threads) as ex:
            futs = [ex.submit(run, 0, n_outer, int(a), int(b))
                    for a, b in zip(edges[:-1], edges[1:]) if b > a]
            for f in futs:
                f.result()
    return out


def mappm(pe1, q1, pe2, i1=1, i2=None, iv=1, kord=1, ptop=0.0, dtype=np.float32, threads=1):
    """Drop-in for ``mappm.mappm`` (f2py): (ncol, km+1), (ncol, km), (ncol, kn+1)
    -> (ncol, kn) float32.  ``i1``, ``i2`` and ``ptop`` are accepted and ignored
    exactly as the reference call site uses them (regridz.py:270-275)."""
    pe1 = np.asarray(pe1)
    q1 = np.asarray(q1)
    pe2 = np.asarray(pe2)
    ncol, km1 = pe1.shape
    km, kn = km1 - 1, pe2.shape[1] - 1
    if q1.shape != (ncol, km) or pe2.shape[0] != ncol:
        raise ValueError("mappm: shape mismatch")
    out = mappm_strided(pe1, q1, pe2, 1, ncol, km, kn,
                        (0, km + 1, 1), (0, km, 1), (0, kn + 1, 1), (0, kn, 1),
                        iv=int(iv), kord=int(kord), dtype=dtype, threads=threads)
    return out.reshape(ncol, kn)


def ppm_profile(q, dp, iv=1, kord=1):
    """(al, ar, a6) of one fp32 column — for unit tests of the CUDA profile stage."""
    q = np.ascontiguousarray(q, dtype=np.float32)
    dp = np.ascontiguousarray(dp, dtype=np.float32)
    km = q.shape[0]
    al, ar, a6 = (np.empty(km, np.float32) for _ in range(3))
    _check(_lib().cgref_ppm_profile_f32(q.ctypes.data, dp.ctypes.data, km, iv, kord,
                                        al.ctypes.data, ar.ctypes.data, a6.ctypes.data))
    return al, ar, a6
