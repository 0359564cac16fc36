// cg_mappm — general PPM vertical remap (mappm.f90:10-126; ppm_profile :614-851 and
// ppm_limiters :854-931 for kord <= 7, cs_profile :132-532 and cs_limiters :535-611 for
// kord > 7), any iv, any km/kn, strided columns.  This is the drop-in for the f2py entry point the reference calls from
// regridz.py:273; the fused pressure-level kernel (plevel_fused.cu) is the fast
// path for the iv=1/kord=1 coarse-graining case.
//
// One thread per column; the column's profile (pe1, q, AL, AR, A6, dc) lives in
// shared memory laid out [level][thread] so that every access — including the
// data-dependent source-layer index of the overlap loop — is bank-conflict free.
// Compiled with --fmad=false: every statement keeps the Fortran's operation order
// and rounding, so results are bit-identical to the CPU oracle (gcc
// -ffp-contract=off), which is what gfortran produces without FMA contraction.
#include "mappm_common.cuh"
#include <cstdlib>
#include <cmath>
#include <mutex>
#include <vector>

namespace cg {

constexpr int TPB = 32;            // shared-memory variant: columns per block
constexpr int TPB_LOCAL = 128;     // local-memory variant
constexpr int KMAX_LOCAL = 96;     // deepest column the local-memory variant takes

struct ColumnViews {
    float *pe, *q, *al, *ar, *a6, *dc, *h2;
    int tid, stride;                 // element k of an array sits at k * stride + tid
    __device__ __forceinline__ float &PE(int k) const { return pe[k * stride + tid]; }
    __device__ __forceinline__ float &Q(int k) const { return q[k * stride + tid]; }
    __device__ __forceinline__ float &AL(int k) const { return al[k * stride + tid]; }
    __device__ __forceinline__ float &AR(int k) const { return ar[k * stride + tid]; }
    __device__ __forceinline__ float &A6(int k) const { return a6[k * stride + tid]; }
    __device__ __forceinline__ float &DC(int k) const { return dc[k * stride + tid]; }
    __device__ __forceinline__ float &H2(int k) const { return h2[k * stride + tid]; }
    __device__ __forceinline__ float DP(int k) const { return PE(k + 1) - PE(k); }
    __device__ __forceinline__ float D4(int k) const { return DP(k - 1) + DP(k); }
    __device__ __forceinline__ float DELQ(int k) const { return Q(k + 1) - Q(k); }
};

__device__ __forceinline__ void limit_at(const ColumnViews &c, int j, int lmt)
{
    float al = c.AL(j), ar = c.AR(j), a6 = c.A6(j);
    limit_layer(c.DC(j), c.Q(j), al, ar, a6, lmt);
    c.AL(j) = al; c.AR(j) = ar; c.A6(j) = a6;
}

// ppm_profile for the thread's column (mappm.f90:614-851)
__device__ void ppm_column(const ColumnViews &c, int km, int iv, int kord)
{
    for (int j = 1; j <= km - 2; ++j) {                                   // :658-668
        const float c1 = (c.DP(j - 1) + 0.5f * c.DP(j)) / c.D4(j + 1);
        const float c2 = (c.DP(j + 1) + 0.5f * c.DP(j)) / c.D4(j);
        const float df2 = c.DP(j) * (c1 * c.DELQ(j) + c2 * c.DELQ(j - 1)) / (c.D4(j) + c.DP(j + 1));
        const float hi = max3f(c.Q(j - 1), c.Q(j), c.Q(j + 1)) - c.Q(j);
        const float lo = c.Q(j) - min3f(c.Q(j - 1), c.Q(j), c.Q(j + 1));
        c.DC(j) = copysignf(fminf(fminf(fabsf(df2), hi), lo), df2);
    }
    for (int j = 2; j <= km - 2; ++j) {                                   // :674-683
        const float c1 = c.DELQ(j - 1) * c.DP(j - 1) / c.D4(j);
        const float a1 = c.D4(j - 1) / (c.D4(j) + c.DP(j - 1));
        const float a2 = c.D4(j + 1) / (c.D4(j) + c.DP(j));
        c.AL(j) = c.Q(j - 1) + c1 + 2.f / (c.D4(j - 1) + c.D4(j + 1)) *
                  (c.DP(j) * (c1 * (a1 - a2) + a2 * c.DC(j - 1)) - c.DP(j - 1) * a1 * c.DC(j));
    }
    {                                                                     // top :689-707
        const float d1 = c.DP(0), d2 = c.DP(1);
        const float q0 = c.Q(0), q1 = c.Q(1);
        const float qm = (d2 * q0 + d1 * q1) / (d1 + d2);
        const float dq = 2.f * (q1 - q0) / (d1 + d2);
        const float c1 = 4.f * (c.AL(2) - qm - d2 * dq) / (d2 * (2.f * d2 * d2 + d1 * (d2 + 3.f * d1)));
        const float c3 = dq - 0.5f * c1 * (d2 * (5.f * d1 + d2) - 3.f * d1 * d1);
        float al1 = qm - 0.25f * c1 * d1 * d2 * (d2 + 3.f * d1);
        float al0 = d1 * (2.f * c1 * (d1 * d1) - c3) + al1;
        al1 = fmaxf(al1, fminf(q0, q1));
        al1 = fminf(al1, fmaxf(q0, q1));
        c.DC(0) = 0.5f * (al1 - q0);
        if (iv == 0) {                                                    // :711-725
            al0 = fmaxf(0.f, al0);
            al1 = fmaxf(0.f, al1);
        } else if (iv == -1) {
            if (al0 * q0 <= 0.f) al0 = 0.f;
        } else if (iv == 2 || iv == -2) {
            al0 = q0;
        }
        c.AL(0) = al0;
        c.AL(1) = al1;
    }
    {                                                                     // bottom :729-761
        const int b = km - 1, e = km - 2;
        const float d1 = c.DP(b), d2 = c.DP(e);
        const float qb = c.Q(b), qe = c.Q(e);
        const float qm = (d2 * qb + d1 * qe) / (d1 + d2);
        const float dq = 2.f * (qe - qb) / (d1 + d2);
        const float c1 = (c.AL(e) - qm - d2 * dq) / (d2 * (2.f * d2 * d2 + d1 * (d2 + 3.f * d1)));
        const float c3 = dq - 2.f * c1 * (d2 * (5.f * d1 + d2) - 3.f * d1 * d1);
        float alb = qm - c1 * d1 * d2 * (d2 + 3.f * d1);
        float arb = d1 * (8.f * c1 * (d1 * d1) - c3) + alb;
        alb = fmaxf(alb, fminf(qb, qe));
        alb = fminf(alb, fmaxf(qb, qe));
        c.DC(b) = 0.5f * (qb - alb);
        if (iv == 0) {
            alb = fmaxf(0.f, alb);
            arb = fmaxf(0.f, arb);
        } else if (iv < 0) {
            if (qb * arb <= 0.f) arb = 0.f;
        }
        c.AL(b) = alb;
        c.AR(b) = arb;
    }
    for (int j = 0; j <= km - 2; ++j) c.AR(j) = c.AL(j + 1);              // :763-767

    for (int j = 0; j < 2; ++j) {                                         // :773-778
        c.A6(j) = curv(c.Q(j), c.AL(j), c.AR(j));
        limit_at(c, j, 0);
    }
    if (kord >= 7) {                                                      // :780-826
        for (int j = 1; j <= km - 2; ++j)
            c.H2(j) = 2.f * (c.DC(j + 1) / c.DP(j + 1) - c.DC(j - 1) / c.DP(j - 1)) /
                      (c.DP(j) + 0.5f * (c.DP(j - 1) + c.DP(j + 1))) * (c.DP(j) * c.DP(j));
        const float fac = 1.5f;
        for (int j = 2; j <= km - 3; ++j) {
            const float q = c.Q(j), dcj = c.DC(j);
            const float pmp = 2.f * dcj;
            float qmp = q + pmp;
            float lac = q + fac * c.H2(j - 1) + dcj;
            const float ar = fminf(fmaxf(c.AR(j), min3f(q, qmp, lac)), max3f(q, qmp, lac));
            qmp = q - pmp;
            lac = q + fac * c.H2(j + 1) - dcj;
            const float al = fminf(fmaxf(c.AL(j), min3f(q, qmp, lac)), max3f(q, qmp, lac));
            c.AR(j) = ar;
            c.AL(j) = al;
            c.A6(j) = curv(q, al, ar);
            if (iv == 0 && kord >= 6) limit_at(c, j, 2);
        }
    } else {                                                              // :828-842
        int lmt = kord - 3;
        if (lmt < 0) lmt = 0;
        if (iv == 0 && lmt > 2) lmt = 2;
        for (int j = 2; j <= km - 3; ++j) {
            c.A6(j) = kord != 4 ? curv(c.Q(j), c.AL(j), c.AR(j)) : 0.f;
            if (kord != 6) limit_at(c, j, lmt);
        }
    }
    for (int j = km - 2; j < km; ++j) {                                   // :844-849
        c.A6(j) = curv(c.Q(j), c.AL(j), c.AR(j));
        limit_at(c, j, 0);
    }
}

// ---- kord > 7: cs_profile (mappm.f90:132-532) and cs_limiters (:535-611) -----------------------------
// Cubic-spline edge values from a tridiagonal solve along the column, then the kord 8..16 constraint
// variants.  1-based accessors below keep each statement next to its Fortran line.  Edge values live in
// H2 (km + 1 entries), gam in DC, the three extremum flags as bits of FL.
// cs_limiters for one layer; `mode` is its `iv` argument (0: positive definite, 1, else: standard PPM)
__device__ __forceinline__ void cs_limit(bool extm, float q, float &al, float &ar, float &a6, int mode)
{
    const float r12 = 1.f / 12.f;
    if (mode == 0) {
        if (q <= 0.f) {
            al = q; ar = q; a6 = 0.f;
        } else if (fabsf(ar - al) < -a6) {
            const float d = ar - al;
            if ((q + 0.25f * (d * d) / a6 + a6 * r12) < 0.f) {
                if (q < ar && q < al) {
                    ar = q; al = q; a6 = 0.f;
                } else if (ar > al) {
                    a6 = 3.f * (al - q);
                    ar = al - a6;
                } else {
                    a6 = 3.f * (ar - q);
                    al = ar - a6;
                }
            }
        }
        return;
    }
    const bool flat = mode == 1 ? ((q - al) * (q - ar) >= 0.f) : extm;
    if (flat) {
        al = q; ar = q; a6 = 0.f;
    } else {
        const float da1 = ar - al;
        const float da2 = da1 * da1;
        const float a6da = a6 * da1;
        if (a6da < -da2) {
            a6 = 3.f * (al - q);
            ar = al - a6;
        } else if (a6da > da2) {
            a6 = 3.f * (ar - q);
            al = ar - a6;
        }
    }
}

struct CsColumn {
    const ColumnViews &c;
    float *fl;
    __device__ __forceinline__ float A1(int k) const { return c.Q(k - 1); }
    __device__ __forceinline__ float DELP(int k) const { return c.DP(k - 1); }
    __device__ __forceinline__ float &AL(int k) const { return c.AL(k - 1); }
    __device__ __forceinline__ float &AR(int k) const { return c.AR(k - 1); }
    __device__ __forceinline__ float &A6(int k) const { return c.A6(k - 1); }
    __device__ __forceinline__ float &QE(int k) const { return c.H2(k - 1); }
    __device__ __forceinline__ float &GAM(int k) const { return c.DC(k - 1); }
    __device__ __forceinline__ int FLAGS(int k) const { return __float_as_int(fl[(k - 1) * c.stride + c.tid]); }
    __device__ __forceinline__ void set_flags(int k, int v) const { fl[(k - 1) * c.stride + c.tid] = __int_as_float(v); }
    __device__ __forceinline__ bool extm(int k) const { return FLAGS(k) & 1; }
    __device__ __forceinline__ bool ext5(int k) const { return FLAGS(k) & 2; }
    __device__ __forceinline__ bool ext6(int k) const { return FLAGS(k) & 4; }
    __device__ __forceinline__ float curvature(int k) const { return 3.f * (2.f * A1(k) - (AL(k) + AR(k))); }
    __device__ __forceinline__ void flatten(int k) const { AL(k) = A1(k); AR(k) = A1(k); }
    // Huynh's 2nd constraint on both edges (:331-340 and its repeats)
    __device__ __forceinline__ void huynh(int k) const
    {
        const float a = A1(k);
        const float pmp_1 = a - 2.f * GAM(k + 1);
        const float lac_1 = pmp_1 + 1.5f * GAM(k + 2);
        AL(k) = fminf(fmaxf(AL(k), min3f(a, pmp_1, lac_1)), max3f(a, pmp_1, lac_1));
        const float pmp_2 = a + 2.f * GAM(k);
        const float lac_2 = pmp_2 - 1.5f * GAM(k - 1);
        AR(k) = fminf(fmaxf(AR(k), min3f(a, pmp_2, lac_2)), max3f(a, pmp_2, lac_2));
    }
    __device__ __forceinline__ void limit(int k, int mode) const
    {
        float al = AL(k), ar = AR(k), a6 = A6(k);
        cs_limit(extm(k), A1(k), al, ar, a6, mode);
        AL(k) = al; AR(k) = ar; A6(k) = a6;
    }
};

__device__ void cs_column(const ColumnViews &cv, float *fl, int km, int iv, int kord)
{
    const CsColumn c{cv, fl};
    const int ak = kord < 0 ? -kord : kord;
    float d4 = 0.f;
    {                                                                     // :175-201 spline solve
        const float grat = c.DELP(2) / c.DELP(1);
        float bet = grat * (grat + 0.5f);
        c.QE(1) = ((grat + grat) * (grat + 1.f) * c.A1(1) + c.A1(2)) / bet;
        c.GAM(1) = (1.f + grat * (grat + 1.5f)) / bet;
        for (int k = 2; k <= km; ++k) {
            d4 = c.DELP(k - 1) / c.DELP(k);
            bet = 2.f + d4 + d4 - c.GAM(k - 1);
            c.QE(k) = (3.f * (c.A1(k - 1) + d4 * c.A1(k)) - c.QE(k - 1)) / bet;
            c.GAM(k) = d4 / bet;
        }
        const float a_bot = 1.f + d4 * (d4 + 1.5f);
        c.QE(km + 1) = (2.f * d4 * (d4 + 1.f) * c.A1(km) + c.A1(km - 1) - a_bot * c.QE(km)) /
                       (d4 * (d4 + 0.5f) - a_bot * c.GAM(km));
        for (int k = km; k >= 1; --k) c.QE(k) = c.QE(k) - c.GAM(k) * c.QE(k + 1);
    }
    if (ak > 16) {                                                        // :204-213
        for (int k = 1; k <= km; ++k) {
            c.AL(k) = c.QE(k);
            c.AR(k) = c.QE(k + 1);
            c.A6(k) = c.curvature(k);
        }
        return;
    }
    // large-scale constraints :222-258
    c.QE(2) = fminf(c.QE(2), fmaxf(c.A1(1), c.A1(2)));
    c.QE(2) = fmaxf(c.QE(2), fminf(c.A1(1), c.A1(2)));
    for (int k = 2; k <= km; ++k) c.GAM(k) = c.A1(k) - c.A1(k - 1);
    for (int k = 3; k <= km - 1; ++k) {
        const float lo = fminf(c.A1(k - 1), c.A1(k)), hi = fmaxf(c.A1(k - 1), c.A1(k));
        float q = c.QE(k);
        if (c.GAM(k - 1) * c.GAM(k + 1) > 0.f) {
            q = fminf(q, hi);
            q = fmaxf(q, lo);
        } else if (c.GAM(k - 1) > 0.f) {
            q = fmaxf(q, lo);
        } else {
            q = fminf(q, hi);
            if (iv == 0) q = fmaxf(0.f, q);
        }
        c.QE(k) = q;
    }
    c.QE(km) = fminf(c.QE(km), fmaxf(c.A1(km - 1), c.A1(km)));
    c.QE(km) = fmaxf(c.QE(km), fminf(c.A1(km - 1), c.A1(km)));
    for (int k = 1; k <= km; ++k) {                                       // :260-287
        c.AL(k) = c.QE(k);
        c.AR(k) = c.QE(k + 1);
    }
    for (int k = 1; k <= km; ++k) {
        int f;
        if (k == 1 || k == km)
            f = (c.AL(k) - c.A1(k)) * (c.AR(k) - c.A1(k)) > 0.f;
        else
            f = c.GAM(k) * c.GAM(k + 1) < 0.f;
        if (ak > 9) {
            const float x0 = 2.f * c.A1(k) - (c.AL(k) + c.AR(k));
            const float x1 = fabsf(c.AL(k) - c.AR(k));
            c.A6(k) = 3.f * x0;
            f |= (fabsf(x0) > x1) ? 2 : 0;
            f |= (fabsf(c.A6(k)) > x1) ? 4 : 0;
        }
        c.set_flags(k, f);
    }
    // top two layers :296-324
    if (iv == 0) {
        c.AL(1) = fmaxf(0.f, c.AL(1));
    } else if (iv == -1) {
        if (c.AL(1) * c.A1(1) <= 0.f) c.AL(1) = 0.f;
    } else if (iv == 2) {
        c.AL(1) = c.A1(1); c.AR(1) = c.A1(1); c.A6(1) = 0.f;
    }
    if (iv != 2) {
        c.A6(1) = c.curvature(1);
        c.limit(1, 1);
    }
    c.A6(2) = c.curvature(2);
    c.limit(2, 2);
    for (int k = 3; k <= km - 2; ++k) {                                   // interior :329-497
        if (ak < 9) {
            c.huynh(k);
            c.A6(k) = c.curvature(k);
        } else if (ak == 9 || ak == 12) {
            const bool wave = ak == 12 ? c.extm(k) : ((c.extm(k) && c.extm(k - 1)) || (c.extm(k) && c.extm(k + 1)));
            if (wave) {
                c.flatten(k);
                c.A6(k) = 0.f;
            } else {
                c.A6(k) = 6.f * c.A1(k) - 3.f * (c.AL(k) + c.AR(k));
                if (fabsf(c.A6(k)) > fabsf(c.AL(k) - c.AR(k))) {
                    c.huynh(k);
                    c.A6(k) = 6.f * c.A1(k) - 3.f * (c.AL(k) + c.AR(k));
                }
            }
        } else if (ak == 10) {
            if (c.ext5(k)) {
                if (c.ext5(k - 1) || c.ext5(k + 1)) c.flatten(k);
                else if (c.ext6(k - 1) || c.ext6(k + 1)) c.huynh(k);
            } else if (c.ext6(k)) {
                if (c.ext5(k - 1) || c.ext5(k + 1)) c.huynh(k);
            }
            c.A6(k) = c.curvature(k);
        } else if (ak == 13) {
            if (c.ext6(k) && c.ext6(k - 1) && c.ext6(k + 1)) c.flatten(k);
            c.A6(k) = c.curvature(k);
        } else if (ak == 14) {
            c.A6(k) = c.curvature(k);
        } else if (ak == 15) {
            if (c.ext5(k)) {
                if (c.ext5(k - 1) || c.ext5(k + 1)) c.flatten(k);
            } else if (c.ext6(k)) {
                c.huynh(k);
            }
            c.A6(k) = c.curvature(k);
        } else if (ak == 16) {
            if (c.ext5(k)) {
                if (c.ext5(k - 1) || c.ext5(k + 1)) c.flatten(k);
                else if (c.ext6(k - 1) || c.ext6(k + 1)) c.huynh(k);
            }
            c.A6(k) = c.curvature(k);
        } else {                                                          // kord = 11
            if (c.ext5(k) && (c.ext5(k - 1) || c.ext5(k + 1))) {
                c.flatten(k);
                c.A6(k) = 0.f;
            } else {
                c.A6(k) = c.curvature(k);
            }
        }
        if (iv == 0) c.limit(k, 0);
    }
    // bottom two layers :502-519
    if (iv == 0) {
        c.AR(km) = fmaxf(0.f, c.AR(km));
    } else if (iv == -1) {
        if (c.AR(km) * c.A1(km) <= 0.f) c.AR(km) = 0.f;
    }
    c.A6(km - 1) = c.curvature(km - 1);
    c.limit(km - 1, 2);
    c.A6(km) = c.curvature(km);
    c.limit(km, 1);
}

// Two homes for the column's work arrays (pe, q, AL, AR, A6, dc, h2 [, flags]: 7-8 values per level).  SHARED:
// interleaved in shared memory, 32 columns per block - 71 KB per warp at km = 79, i.e. 3 warps per SM, for any
// km that fits 227 KB.  LOCAL (km <= KMAX_LOCAL): thread-private arrays in local memory, which the hardware
// interleaves across the lanes the same way and serves from L1 / L2; occupancy is then bounded by registers only
// (one C384 field: 7.5 ms shared, see DESIGN.md for the local figure).
template <bool LOCAL>
__global__ void __launch_bounds__(LOCAL ? TPB_LOCAL : TPB)
mappm_generic_kernel(const float *__restrict__ pe1, const float *__restrict__ q1,
                     const float *__restrict__ pe2, float *__restrict__ q2,
                     long ncol, long n_inner, int km, int kn, Strides s, Pe2Block pb,
                     int iv, int kord, const int *flags)
{
    extern __shared__ float smem[];
    float local_buf[LOCAL ? 8 * KMAX_LOCAL + 2 : 1];
    const long col = blockIdx.x * (long)(LOCAL ? TPB_LOCAL : TPB) + threadIdx.x;
    if (col >= ncol) return;  // no block-level synchronisation below
    // second pass behind the streaming remapper: only the column groups it could not take
    if (flags && flags[col / STREAM_TPB] == 0) return;
    ColumnViews c;
    c.tid = LOCAL ? 0 : threadIdx.x;
    c.stride = LOCAL ? 1 : TPB;
    c.pe = LOCAL ? local_buf : smem;
    c.q = c.pe + (km + 1) * c.stride;
    c.al = c.q + km * c.stride;
    c.ar = c.al + km * c.stride;
    c.a6 = c.ar + km * c.stride;
    c.dc = c.a6 + km * c.stride;
    c.h2 = c.dc + km * c.stride;
    float *fl = c.h2 + (km + 1) * c.stride;                               // kord > 7 only

    const long o = col / n_inner, i = col % n_inner;
    const float *p1 = pe1 + o * s.v[0] + i * s.v[1];
    const float *f1 = q1 + o * s.v[3] + i * s.v[4];
    long i2 = i;
    if (pb.f > 0) i2 = (long)((i / pb.nx) / pb.f) * pb.nxc + (i % pb.nx) / pb.f;
    const float *p2 = pe2 + o * s.v[6] + i2 * s.v[7];
    float *f2 = q2 + o * s.v[9] + i * s.v[10];
    const long l_pe1 = s.v[2], l_q1 = s.v[5], l_pe2 = s.v[8], l_q2 = s.v[11];

    for (int k = 0; k <= km; ++k) c.PE(k) = p1[k * l_pe1];
    for (int k = 0; k < km; ++k) c.Q(k) = f1[k * l_q1];
    if (kord > 7)
        cs_column(c, fl, km, iv, kord);
    else
        ppm_column(c, km, iv, kord);

    const float r3 = 1.f / 3.f, r23 = 2.f / 3.f;
    const float pe_top = c.PE(0), pe_bot = c.PE(km), q_top = c.Q(0), q_bot = c.Q(km - 1);
    int k0 = 0;
    float top = p2[0];
    for (int k = 0; k < kn; ++k) {                                        // mappm.f90:58-124
        const float bot = p2[(long)(k + 1) * l_pe2];
        float res;
        if (top <= pe_top) {
            res = q_top;
        } else if (top >= pe_bot) {
            res = q_bot;
        } else {
            int found = 0, k1 = 0;
            float qsum = 0.f, dpsum = 0.f;
            res = NAN;  // no bracketing layer: undefined in the Fortran, NaN here
            for (int L = k0; L < km; ++L) {
                const float pl0 = c.PE(L), pl1 = c.PE(L + 1);
                if (top >= pl0 && top <= pl1) {
                    k0 = L;
                    const float dpl = pl1 - pl0;
                    const float pl = (top - pl0) / dpl;
                    const float al = c.AL(L), ar = c.AR(L), a6 = c.A6(L);
                    if (bot <= pl1) {
                        const float pr = (bot - pl0) / dpl;
                        const float tt = r3 * (pr * (pr + pl) + pl * pl);
                        res = al + 0.5f * (a6 + ar - al) * (pr + pl) - a6 * tt;
                        found = 2;
                    } else {
                        const float delp = pl1 - top;
                        const float tt = r3 * (1.f + pl * (1.f + pl));
                        qsum = delp * (al + 0.5f * (a6 + ar - al) * (1.f + pl) - a6 * tt);
                        dpsum = delp;
                        k1 = L + 1;
                        found = 1;
                    }
                    break;
                }
            }
            if (found == 1) {
                bool closed = false;
                for (int L = k1; L < km; ++L) {
                    const float pl0 = c.PE(L), pl1 = c.PE(L + 1);
                    const float dpl = pl1 - pl0;
                    if (bot > pl1) {
                        qsum = qsum + dpl * c.Q(L);
                        dpsum = dpsum + dpl;
                    } else {
                        const float delp = bot - pl0;
                        const float esl = delp / dpl;
                        const float al = c.AL(L), ar = c.AR(L), a6 = c.A6(L);
                        qsum = qsum + delp * (al + 0.5f * esl * (ar - al + a6 * (1.f - r23 * esl)));
                        dpsum = dpsum + delp;
                        k0 = L;
                        closed = true;
                        break;
                    }
                }
                if (!closed) {
                    const float delp = bot - pe_bot;
                    if (delp > 0.f) {
                        qsum = qsum + delp * q_bot;
                        dpsum = dpsum + delp;
                    }
                }
                res = qsum / dpsum;
            }
        }
        f2[(long)k * l_q2] = res;
        top = bot;
    }
}

int launch_mappm_stream(const float *pe1, const float *q1, const float *pe2, float *q2, long ncol, long n_inner,
                        int km, int kn, const Strides &s, const Pe2Block &pb, int iv, int kord, int *flags,
                        cudaStream_t st);

static int launch_generic(const float *pe1, const float *q1, const float *pe2, float *q2, long ncol, long n_inner,
                          int km, int kn, const Strides &s, const Pe2Block &pb, int iv, int kord, const int *flags,
                          cudaStream_t st);

// One int per 128 columns between the streaming pass and the generic pass behind it.  Kept per (device, stream) and
// grown on demand: a stream-ordered allocation per call would go back to the driver at every synchronisation
// (5 ms per call, measured), and calls on different streams must not share flags.  A few KB each (1.7 MB at C3072).
static int *flag_buffer(size_t bytes, cudaStream_t st)
{
    struct Entry { int dev; cudaStream_t st; int *ptr; size_t bytes; };
    static std::mutex mu;
    static std::vector<Entry> cache;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) { set_error("cudaGetDevice failed"); return nullptr; }
    std::lock_guard<std::mutex> lock(mu);
    for (Entry &e : cache)
        if (e.dev == dev && e.st == st) {
            if (e.bytes >= bytes) return e.ptr;
            cudaStreamSynchronize(st);                                     // (earlier calls on the stream still read it)
            cudaFree(e.ptr);
            e.ptr = nullptr; e.bytes = 0;
            const cudaError_t err = cudaMalloc((void **)&e.ptr, bytes);
            if (err != cudaSuccess) { cuda_fail(err, "cudaMalloc(mappm flags)"); return nullptr; }
            e.bytes = bytes;
            return e.ptr;
        }
    Entry e{dev, st, nullptr, bytes};
    const cudaError_t err = cudaMalloc((void **)&e.ptr, bytes);
    if (err != cudaSuccess) { cuda_fail(err, "cudaMalloc(mappm flags)"); return nullptr; }
    cache.push_back(e);
    return e.ptr;
}

int launch_mappm(const float *pe1, const float *q1, const float *pe2, float *q2,
                 long n_outer, long n_inner, int km, int kn, const long *strides,
                 int pe2_f, int pe2_nx, int pe2_nxc, int iv, int kord, cudaStream_t st)
{
    CG_REQUIRE(pe1 && q1 && pe2 && q2 && strides, "null pointer");
    CG_REQUIRE(n_outer >= 0 && n_inner >= 0, "bad column counts");
    CG_REQUIRE(km >= 4 && kn >= 1, "mappm needs km >= 4 and kn >= 1 (got %d, %d)", km, kn);
    CG_REQUIRE(iv >= -2 && iv <= 2, "iv must be in [-2, 2]");
    if (kord > 7 && iv == -2) {
        set_error("kord=%d with iv=-2: cs_profile reads an array mappm never sets (mappm.f90:35, :163)", kord);
        return CG_ENOTSUP;
    }
    const long ncol = n_outer * n_inner;
    if (ncol == 0) return CG_OK;
    Strides s;
    for (int i = 0; i < 12; ++i) s.v[i] = strides[i];
    Pe2Block pb{pe2_f, pe2_nx, pe2_nxc};
    if (kord <= 6 && km >= 5 && !options().mappm_impl && !options().mappm_arrays) {
        // the streaming remapper; column groups it flags (pressures that do not increase downward, NaN) are redone
        // by the generic kernel, whose other blocks return at once
        const size_t fbytes = sizeof(int) * (size_t)((ncol + STREAM_TPB - 1) / STREAM_TPB);
        int *flags = flag_buffer(fbytes, st);
        if (!flags) return CG_ECUDA;
        cudaError_t e = cudaMemsetAsync(flags, 0, fbytes, st);
        int rc = e == cudaSuccess ? CG_OK : cuda_fail(e, "cudaMemsetAsync(mappm flags)");
        if (rc == CG_OK) rc = launch_mappm_stream(pe1, q1, pe2, q2, ncol, n_inner, km, kn, s, pb, iv, kord, flags, st);
        if (rc == CG_OK) rc = launch_generic(pe1, q1, pe2, q2, ncol, n_inner, km, kn, s, pb, iv, kord, flags, st);
        return rc;
    }
    return launch_generic(pe1, q1, pe2, q2, ncol, n_inner, km, kn, s, pb, iv, kord, nullptr, st);
}

static int launch_generic(const float *pe1, const float *q1, const float *pe2, float *q2, long ncol, long n_inner,
                          int km, int kn, const Strides &s, const Pe2Block &pb, int iv, int kord, const int *flags,
                          cudaStream_t st)
{
    if (km <= KMAX_LOCAL && !options().mappm_arrays) {                  // (option mappm_arrays = shared: tests)
        mappm_generic_kernel<true><<<cdiv(ncol, TPB_LOCAL), TPB_LOCAL, 0, st>>>(pe1, q1, pe2, q2, ncol, n_inner, km, kn, s, pb, iv, kord, flags);
        CG_LAUNCH_CHECK("cg_mappm");
        return CG_OK;
    }
    const size_t smem = sizeof(float) * TPB * (size_t)(kord > 7 ? 8 * km + 2 : 7 * km + 1);
    CG_REQUIRE(smem <= 227 * 1024, "km=%d needs %zu B of shared memory per block", km, smem);
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(mappm_generic_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(mappm)");
    }
    mappm_generic_kernel<false><<<cdiv(ncol, TPB), TPB, smem, st>>>(pe1, q1, pe2, q2, ncol, n_inner, km, kn, s, pb, iv, kord, flags);
    CG_LAUNCH_CHECK("cg_mappm");
    return CG_OK;
}

}  // namespace cg

using namespace cg;

extern "C" int cg_mappm(const float *pe1, const float *q1, const float *pe2, float *q2,
                        long n_outer, long n_inner, int km, int kn, const long *strides,
                        int iv, int kord, void *stream)
{
    return launch_mappm(pe1, q1, pe2, q2, n_outer, n_inner, km, kn, strides, 0, 0, 0, iv, kord,
                        (cudaStream_t)stream);
}
