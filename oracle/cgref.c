/* TEST INFRASTRUCTURE — CPU oracle, not product code.
 *
 * libcgref.so: C ABI over the column restatement of the reference's Fortran
 * remapper (see cgref_mappm.inc for the file:line map).  Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
 * may load this library.  The product (fv3net_b200) never does.
 *
 * Replaces, for checking purposes, the f2py entry point
 *   mappm.mappm(pe1[n,km+1], q1[n,km], pe2[n,kn+1], i1, i2, iv, kord, ptop)
 * (mappm.f90:10,23-31; call site vcm/cubedsphere/regridz.py:273-275).
 *
 * Build: see oracle/Makefile (gcc -O2 -ffp-contract=off; no -ffast-math).
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>

#define CAT_(a, b) a##b
#define CAT(a, b) CAT_(a, b)

#define REAL float
#define NAME(x) CAT(x, _f32)
#define FMAX fmaxf
#define FMIN fminf
#define FABS fabsf
#define COPYSIGN copysignf
#include "cgref_mappm.inc"
#undef REAL
#undef NAME
#undef FMAX
#undef FMIN
#undef FABS
#undef COPYSIGN

#define REAL double
#define NAME(x) CAT(x, _f64)
#define FMAX fmax
#define FMIN fmin
#define FABS fabs
#define COPYSIGN copysign
#include "cgref_mappm.inc"
#undef REAL
#undef NAME
#undef FMAX
#undef FMIN
#undef FABS
#undef COPYSIGN

/* Error codes shared with include/coarsen_b200.h */
#define CG_OK 0
#define CG_EINVAL -1
#define CG_ENOTSUP -2

/* Strided batch driver.  Column c = (o, i), o in [0,n_outer), i in [0,n_inner):
 *   element k of column c of array A lives at A[o*os_A + i*cs_A + k*ls_A]
 * so both the reference's z-last (ncol, nz) layout (n_outer=1, cs=nz, ls=1) and
 * the workflows' (tile, z, y, x) layout (n_outer=tile, os=nz*ny*nx, cs=1,
 * ls=ny*nx) are addressed without a transpose.  pe2 may be shared by groups of
 * columns through its own strides (e.g. coarse pressure, cs_pe2 = 0).
 */
#define DEFINE_BATCH(SFX, T)                                                              \
    int cgref_mappm_##SFX(const T *pe1, const T *q1, const T *pe2, T *q2,                 \
                          long n_outer, long n_inner, int km, int kn,                     \
                          long os_pe1, long cs_pe1, long ls_pe1,                          \
                          long os_q1, long cs_q1, long ls_q1,                             \
                          long os_pe2, long cs_pe2, long ls_pe2,                          \
                          long os_q2, long cs_q2, long ls_q2,                             \
                          int iv, int kord, int nthreads)                                 \
    {                                                                                     \
        if (km < 4 || kn < 1) return CG_EINVAL;                                           \
        if (kord > 7 && iv == -2) return CG_ENOTSUP; /* reads an uninitialised qs there */ \
        long ncol = n_outer * n_inner;                                                    \
        (void)nthreads; /* threading is done by the caller (ctypes releases the GIL) */  \
        {                                                                                 \
            T *buf = (T *)malloc(sizeof(T) * (size_t)(2 * (km + 1) + 2 * (kn + 1) + 9 * km + 8)); \
            T *cpe1 = buf, *cq1 = cpe1 + km + 1, *cpe2 = cq1 + km + 1;                    \
            T *cq2 = cpe2 + kn + 1, *wk = cq2 + kn + 1;                                   \
            for (long c = 0; c < ncol; ++c) {                                             \
                long o = c / n_inner, i = c % n_inner;                                    \
                for (int k = 0; k <= km; ++k) cpe1[k] = pe1[o * os_pe1 + i * cs_pe1 + k * ls_pe1]; \
                for (int k = 0; k < km; ++k) cq1[k] = q1[o * os_q1 + i * cs_q1 + k * ls_q1];      \
                for (int k = 0; k <= kn; ++k) cpe2[k] = pe2[o * os_pe2 + i * cs_pe2 + k * ls_pe2]; \
                remap_column_##SFX(km, cpe1, cq1, kn, cpe2, cq2, iv, kord, wk);           \
                for (int k = 0; k < kn; ++k) q2[o * os_q2 + i * cs_q2 + k * ls_q2] = cq2[k];      \
            }                                                                             \
            free(buf);                                                                    \
        }                                                                                 \
        return CG_OK;                                                                     \
    }

DEFINE_BATCH(f32, float)
DEFINE_BATCH(f64, double)

/* PPM coefficients of one column, exposed so tests can compare the CUDA
 * kernel's profile stage on its own. */
int cgref_ppm_profile_f32(const float *q, const float *dp, int km, int iv, int kord,
                          float *al, float *ar, float *a6)
{
    if (km < 4) return CG_EINVAL;
    if (kord > 7 && iv == -2) return CG_ENOTSUP;
    float *wk = (float *)malloc(sizeof(float) * (size_t)(5 * km + 8));
    if (kord > 7)
        cs_column_f32(km, q, dp, al, ar, a6, iv, kord, wk);
    else
        ppm_column_f32(km, q, dp, al, ar, a6, iv, kord, wk);
    free(wk);
    return CG_OK;
}
