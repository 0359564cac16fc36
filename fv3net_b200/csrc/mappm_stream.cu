// Streaming PPM remapper: mappm (mappm.f90:10-126) with ppm_profile (:614-851) and ppm_limiters (:854-931) for
// kord <= 6, any iv, km >= 5 - the common case of regrid_vertical (regridz.py:273-275 calls it with iv = 1, kord = 1)
// and of the edge-weighted pressure recipes.  Same results as mappm_generic.cu bit for bit (both are compiled with
// --fmad=false and keep the Fortran's statements), without its whole-column work arrays:
//   * one thread per column, lanes over adjacent columns (coalesced for the (tile, z, y, x) layout);
//   * the source column is walked ONCE, top down, with a rolling window of four levels in registers: the limited
//     slope of a layer needs its two neighbours, its lower edge the slope below, so the parabola of layer L is
//     final when level L + 2 has arrived;
//   * the overlap integrals of mappm.f90:58-124 are formed in the same walk by a two-pointer merge over the
//     target interfaces, with the additions of every target layer in the Fortran's order.
// The merge is the Fortran's forward search only for columns whose pressures increase downward.  A column group
// (128 columns) that holds a non-finite or non-increasing source thickness, or a decreasing target pressure, is
// flagged and redone by the generic kernel (launched behind this one; its other blocks return at once).
#include "mappm_common.cuh"
#include <cmath>

namespace cg {

namespace {

// limited slope of an interior layer (mappm.f90:658-668): q- q q+, thicknesses d- d d+
__device__ __forceinline__ float slope_dc(float qm, float q, float qp, float dm, float d, float dp)
{
    const float d4 = dm + d, d4p = d + dp;
    const float c1 = (dm + 0.5f * d) / d4p;
    const float c2 = (dp + 0.5f * d) / d4;
    const float df2 = d * (c1 * (qp - q) + c2 * (q - qm)) / (d4 + dp);
    const float hi = max3f(qm, q, qp) - q;
    const float lo = q - min3f(qm, q, qp);
    return copysignf(fminf(fminf(fabsf(df2), hi), lo), df2);
}

// upper edge of layer j, 2 <= j <= km - 2 (mappm.f90:674-683): q(j-1), q(j); thicknesses of j-2, j-1, j, j+1;
// slopes of j-1 and j
__device__ __forceinline__ float edge_al(float qm, float q, float dmm, float dm, float d, float dp, float dcm, float dc)
{
    const float d4m = dmm + dm, d4 = dm + d, d4p = d + dp;
    const float c1 = (q - qm) * dm / d4;
    const float a1 = d4m / (d4 + dm);
    const float a2 = d4p / (d4 + d);
    return qm + c1 + 2.f / (d4m + d4p) * (d * (c1 * (a1 - a2) + a2 * dcm) - dm * a1 * dc);
}

struct Target {                      // the walk over the target layers of the column
    const float *p2n;                // running pointers: interface k + 2, layer k
    float *f2;
    long l_pe2, l_q2;
    int k, kn, mode;                 // mode 1: target layer k is open (its top lies in a source layer above)
    float top, bot, botn;            // pe2[k], pe2[k+1], pe2[k+2] (read ahead)
    float qsum, dpsum;
    bool bad;

    __device__ __forceinline__ void emit(float v)
    {
        *f2 = v;
        f2 += l_q2;
        ++k;
        bad = bad || !(bot >= top);
        top = bot;
        bot = botn;
        p2n += l_pe2;
        if (k + 2 <= kn) botn = *p2n;
    }
};

// everything source layer L = [pl0, pl1] contributes (mappm.f90:58-124, statement for statement)
__device__ __forceinline__ void consume(Target &t, float pl0, float pl1, float q, float al, float ar, float a6,
                                        float pe_top, float pe_bot, float q_top, float q_bot)
{
    const float r3 = 1.f / 3.f, r23 = 2.f / 3.f;
    const float dpl = pl1 - pl0;
    while (t.k < t.kn) {
        if (t.mode == 0) {
            if (t.top <= pe_top) { t.emit(q_top); continue; }             // :62-64
            if (t.top >= pe_bot) { t.emit(q_bot); continue; }             // :65-67
            if (!(t.top >= pl0 && t.top <= pl1)) return;                  // :70-75: the search goes on below
            const float pl = (t.top - pl0) / dpl;
            if (t.bot <= pl1) {                                           // :76-83
                const float pr = (t.bot - pl0) / dpl;
                const float tt = r3 * (pr * (pr + pl) + pl * pl);
                t.emit(al + 0.5f * (a6 + ar - al) * (pr + pl) - a6 * tt);
                continue;
            }
            const float delp = pl1 - t.top;                               // :85-92
            const float tt = r3 * (1.f + pl * (1.f + pl));
            t.qsum = delp * (al + 0.5f * (a6 + ar - al) * (1.f + pl) - a6 * tt);
            t.dpsum = delp;
            t.mode = 1;
            return;
        }
        if (t.bot > pl1) {                                                // :98-102
            t.qsum = t.qsum + dpl * q;
            t.dpsum = t.dpsum + dpl;
            return;
        }
        const float delp = t.bot - pl0;                                   // :103-112
        const float esl = delp / dpl;
        t.qsum = t.qsum + delp * (al + 0.5f * esl * (ar - al + a6 * (1.f - r23 * esl)));
        t.dpsum = t.dpsum + delp;
        t.mode = 0;
        t.emit(t.qsum / t.dpsum);
    }
}

__global__ void __launch_bounds__(STREAM_TPB)
mappm_stream_kernel(const float *__restrict__ pe1, const float *__restrict__ q1, const float *__restrict__ pe2,
                    float *__restrict__ q2, long ncol, long n_inner, int km, int kn, Strides s, Pe2Block pb,
                    int iv, int kord, int *__restrict__ flags)
{
    const long col = blockIdx.x * (long)STREAM_TPB + threadIdx.x;
    if (col >= ncol) return;
    const long o = col / n_inner, i = col % n_inner;
    const float *p1 = pe1 + o * s.v[0] + i * s.v[1];
    const float *f1 = q1 + o * s.v[3] + i * s.v[4];
    long i2 = i;
    if (pb.f > 0) i2 = (long)((i / pb.nx) / pb.f) * pb.nxc + (i % pb.nx) / pb.f;
    const long l_pe1 = s.v[2], l_q1 = s.v[5];

    Target t;
    const float *p2 = pe2 + o * s.v[6] + i2 * s.v[7];
    t.f2 = q2 + o * s.v[9] + i * s.v[10];
    t.l_pe2 = s.v[8];
    t.l_q2 = s.v[11];
    t.k = 0; t.kn = kn; t.mode = 0; t.qsum = 0.f; t.dpsum = 0.f; t.bad = false;
    t.top = p2[0];
    t.bot = p2[t.l_pe2];
    t.botn = kn >= 2 ? p2[2 * t.l_pe2] : t.bot;
    t.p2n = p2 + 2 * t.l_pe2;

    const float pe_top = p1[0], pe_bot = p1[(long)km * l_pe1];
    const float q_top = f1[0], q_bot = f1[(long)(km - 1) * l_q1];

    // parabola limiter of the interior layers (mappm.f90:828-842)
    int lmt = kord - 3;
    if (lmt < 0) lmt = 0;
    if (iv == 0 && lmt > 2) lmt = 2;

    // levels 0 .. 4 (km >= 5); window at step L: q(L) q(L+1) q(L+2) | interfaces L .. L+3 | one level read ahead
    float pa = pe_top, pb_ = p1[l_pe1], pc = p1[2 * l_pe1], pd = p1[3 * l_pe1], pe = p1[4 * l_pe1];
    float qa = q_top, qb = f1[l_q1], qc = f1[2 * l_q1], qd = f1[3 * l_q1];
    float pnext = p1[5 * l_pe1], qnext = f1[4 * l_q1];                   // interface 5, level 4
    float d0 = pb_ - pa, d1 = pc - pb_, d2 = pd - pc, d3 = pe - pd;      // mappm.f90:41-46
    bool bad = !(d0 > 0.f) || !(d1 > 0.f) || !(d2 > 0.f) || !(d3 > 0.f);
    const float dc1 = slope_dc(qa, qb, qc, d0, d1, d2);
    float dc2 = slope_dc(qb, qc, qd, d1, d2, d3);
    float al2 = edge_al(qb, qc, d0, d1, d2, d3, dc1, dc2);
    float al0, al1, dc0;
    {                                                                     // top :689-725
        const float qm = (d1 * qa + d0 * qb) / (d0 + d1);
        const float dq = 2.f * (qb - qa) / (d0 + d1);
        const float c1 = 4.f * (al2 - qm - d1 * dq) / (d1 * (2.f * d1 * d1 + d0 * (d1 + 3.f * d0)));
        const float c3 = dq - 0.5f * c1 * (d1 * (5.f * d0 + d1) - 3.f * d0 * d0);
        al1 = qm - 0.25f * c1 * d0 * d1 * (d1 + 3.f * d0);
        al0 = d0 * (2.f * c1 * (d0 * d0) - c3) + al1;
        al1 = fmaxf(al1, fminf(qa, qb));
        al1 = fminf(al1, fmaxf(qa, qb));
        dc0 = 0.5f * (al1 - qa);
        if (iv == 0) {
            al0 = fmaxf(0.f, al0);
            al1 = fmaxf(0.f, al1);
        } else if (iv == -1) {
            if (al0 * qa <= 0.f) al0 = 0.f;
        } else if (iv == 2 || iv == -2) {
            al0 = qa;
        }
    }
    {                                                                     // layers 0 and 1 (:773-778)
        float al = al0, ar = al1, a6 = curv(qa, al0, al1);
        limit_layer(dc0, qa, al, ar, a6, 0);
        consume(t, pa, pb_, qa, al, ar, a6, pe_top, pe_bot, q_top, q_bot);
        al = al1; ar = al2; a6 = curv(qb, al1, al2);
        limit_layer(dc1, qb, al, ar, a6, 0);
        consume(t, pb_, pc, qb, al, ar, a6, pe_top, pe_bot, q_top, q_bot);
    }
    // window for L = 2: levels 2, 3, 4
    float dm = d1;                                                        // thickness of level L - 1
    qa = qc; qb = qd; qc = qnext;
    pa = pc; pb_ = pd; pc = pe; pd = pnext;
    float dL = d2, dN = d3, dNN = pd - pc;                               // levels L, L + 1, L + 2
    bad = bad || !(dNN > 0.f);
    float dcL = dc2, alL = al2;
    const float *f1n = f1 + 5 * l_q1, *p1n = p1 + 6 * l_pe1;            // level L + 3, interface L + 4 at L = 2
    if (km > 5) { qnext = *f1n; pnext = *p1n; }

    for (int L = 2; L <= km - 3; ++L) {
        // level L + 2 has arrived: slope of L + 1, upper edge of L + 1 = lower edge of L
        const float dcN = slope_dc(qa, qb, qc, dL, dN, dNN);
        const float alN = edge_al(qa, qb, dm, dL, dN, dNN, dcL, dcN);
        float al = alL, ar = alN, a6 = kord != 4 ? curv(qa, alL, alN) : 0.f;
        if (kord != 6) limit_layer(dcL, qa, al, ar, a6, lmt);
        consume(t, pa, pb_, qa, al, ar, a6, pe_top, pe_bot, q_top, q_bot);
        // on to L + 1; level L + 3 was read one step ahead, level L + 4 is requested now
        dm = dL; dL = dN; dN = dNN;
        qa = qb; qb = qc; pa = pb_; pb_ = pc; pc = pd;
        dcL = dcN; alL = alN;
        if (L + 3 <= km - 1) {
            qc = qnext; pd = pnext;
            dNN = pd - pc;
            bad = bad || !(dNN > 0.f);
            f1n += l_q1; p1n += l_pe1;
            if (L + 4 <= km - 1) { qnext = *f1n; pnext = *p1n; }
        }
    }
    {                                                                     // bottom :729-761; now qa = q(km-2), qb = q(km-1)
        const float e1 = dN, e2 = dL;                                     // thickness of the bottom layer and the one above
        const float qm = (e2 * qb + e1 * qa) / (e1 + e2);
        const float dq = 2.f * (qa - qb) / (e1 + e2);
        const float c1 = (alL - qm - e2 * dq) / (e2 * (2.f * e2 * e2 + e1 * (e2 + 3.f * e1)));
        const float c3 = dq - 2.f * c1 * (e2 * (5.f * e1 + e2) - 3.f * e1 * e1);
        float alb = qm - c1 * e1 * e2 * (e2 + 3.f * e1);
        float arb = e1 * (8.f * c1 * (e1 * e1) - c3) + alb;
        alb = fmaxf(alb, fminf(qb, qa));
        alb = fminf(alb, fmaxf(qb, qa));
        const float dcb = 0.5f * (qb - alb);
        if (iv == 0) {
            alb = fmaxf(0.f, alb);
            arb = fmaxf(0.f, arb);
        } else if (iv < 0) {
            if (qb * arb <= 0.f) arb = 0.f;
        }
        float al = alL, ar = alb, a6 = curv(qa, alL, alb);                // layers km - 2 and km - 1 (:844-849)
        limit_layer(dcL, qa, al, ar, a6, 0);
        consume(t, pa, pb_, qa, al, ar, a6, pe_top, pe_bot, q_top, q_bot);
        al = alb; ar = arb; a6 = curv(qb, alb, arb);
        limit_layer(dcb, qb, al, ar, a6, 0);
        consume(t, pb_, pc, qb, al, ar, a6, pe_top, pe_bot, q_top, q_bot);
    }
    // target layers that reach below the column (:115-121) or lie under it (:65-67)
    while (t.k < t.kn) {
        if (t.mode == 1) {
            const float delp = t.bot - pe_bot;
            if (delp > 0.f) {
                t.qsum = t.qsum + delp * q_bot;
                t.dpsum = t.dpsum + delp;
            }
            t.mode = 0;
            t.emit(t.qsum / t.dpsum);
        } else if (t.top <= pe_top) {
            t.emit(q_top);
        } else if (t.top >= pe_bot) {
            t.emit(q_bot);
        } else {
            t.emit(NAN);                                                  // no bracketing layer (never for increasing pressures)
        }
    }
    if (bad || t.bad || !(pe_bot == pe_bot)) flags[col / STREAM_TPB] = 1;
}

}  // namespace

// kord <= 6 and km >= 5; `flags`: one int per 128 columns, zeroed by the caller (stream-ordered)
int launch_mappm_stream(const float *pe1, const float *q1, const float *pe2, float *q2, long ncol, long n_inner,
                        int km, int kn, const Strides &s, const Pe2Block &pb, int iv, int kord, int *flags,
                        cudaStream_t st)
{
    mappm_stream_kernel<<<cdiv(ncol, STREAM_TPB), STREAM_TPB, 0, st>>>(pe1, q1, pe2, q2, ncol, n_inner, km, kn, s, pb,
                                                                      iv, kord, flags);
    CG_LAUNCH_CHECK("cg_mappm(stream)");
    return CG_OK;
}

}  // namespace cg
