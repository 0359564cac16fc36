// Pair-thread streaming pressure-level coarse-graining for sm_100a: f = 8, iv = 1, kord = 1, nx % 32 == 0.
// (regridz.py:115-147 + coarsen.py:183, budgets.py:49-57; mappm.f90:10-126, :614-851, :854-931)
//
// Same tandem walk as plevel_stream.cu (source layers in order, coarse interfaces met on the way close one
// target level and open the next; the pressure-only geometry is formed once per column and level and shared by
// all fields through shared-memory records), re-cut so that every lane works:
//   * A CTA owns 4 x-adjacent coarse cells = 32 fine columns = 16 column PAIRS x 8 rows x nz levels; one CTA
//     of 16 warps per SM.
//   * A field thread is one work item (column pair, field): 16 V items flattened over the field warps
//     (V = 26 -> 13 full warps).  It carries the rolling PPM state of its two columns as f32x2 operands (one
//     instruction, two columns) and owns a PRIVATE accumulator column acc[target level][item] in shared memory:
//     its read-modify-writes never meet another thread's, and bank = lane whatever the level.  The four pairs of
//     a cell are summed once at the end in a fixed order (bitwise reproducible).
//   * Three geometry warps (lanes over the 32 columns, warp g takes level 3 chunk + g) write the records and
//     issue the TMA loads: 13 + 3 warps = 4 per scheduler.
//   * The walk is unrolled by the 3 levels of a pipeline stage with the rolling state in 3-slot rings indexed
//     at compile time: no register moves between steps.
//   * Slope and parabola limiters are written for the instruction mix of sm_100a: |.|-minimum with the sign
//     spliced back (FMNMX, LOP3) and 0 / 1 masks (FSET) instead of chains of half-rate 3-input min / max.
// Every fine input element is read from HBM exactly once per launch.
#include "plevel_common.cuh"
#include <cmath>
#include <cstring>
#include <mutex>

namespace cg {
namespace pairk {
using namespace plv;

constexpr int NCELL = 4;             // coarse cells per CTA
constexpr int NCOL = 32;             // fine columns per CTA row = lanes of a geometry warp
constexpr int NPAIR = NCOL / 2;      // column pairs per CTA row
constexpr int KC = 3;                // levels per pipeline stage
constexpr int NGW = 2 * KC;          // geometry warps: one per level of a stage, two stages in flight
#ifndef CG_NSTAGE
#define CG_NSTAGE 3
#endif
constexpr int NSTAGE = CG_NSTAGE;    // stages of the field ring = geometry record buffers (3 or 4)
constexpr int BAR_STAGE0 = 2;        // named barriers BAR_STAGE0 + b: stage / buffer b is handed back
constexpr int BAR_OUT = BAR_STAGE0 + NSTAGE, BAR_WALK = BAR_OUT + 1;
constexpr int BOX_BYTES = KC * 128;  // one source, one stage: 3 rows of 32 columns (odd row count: 8 swizzle phases over 8 sources)
constexpr int MAXFW = 13;            // field warps: 20 warps with the geometry warps and the column walker = 5 per scheduler, 96 registers
constexpr int MAXV = MAXFW * 32 / NPAIR;     // 26 output fields per launch
constexpr int MAXSRC = 2 * MAXV;
// geometry record of one step, structure of arrays over the 16 pairs: group g of pair p at g * GROUP_BYTES + 16 p
//   0: E1a E1b E2a E2b | 1: Fa Fb Ga Gb | 2: -Ha -Hb dpwa dpwb | 3: exa exb eya eyb | 4: eza ezb -ewa -ewb
//   5: offa offb infos flag   (off = accumulator byte offset of the level the first event closes; infos = info_a |
//      info_b << 8, info = number of coarse interfaces in the layer; flag = some column has further events)
constexpr int GROUP_BYTES = NPAIR * 16;
constexpr int NGROUP = 6;
constexpr int REC_LEVEL = NGROUP * GROUP_BYTES;
constexpr int REC_BYTES = KC * REC_LEVEL;
constexpr int MAXE = 2;              // listed further events per (column, layer)
constexpr int INFO_SLOW = 255;
constexpr int OVA_BYTES = KC * NCOL * MAXE * 16;     // (x, y, z, -w) per event
constexpr int OVB_BYTES = KC * NCOL * MAXE * 4;      // accumulator byte offset per event
constexpr int TAILMAX = NSTAGE > 3 ? 8 : 16;         // listed further events of the layer below the surface: (x, offset) per event
constexpr int TAIL_BYTES = NCOL * TAILMAX * 8;
constexpr int EDGE_WORDS = 24;       // per column: one-sided cubic coefficients, top (0..8) and bottom (12..20)
enum { CTX_NZ, CTX_MASK, CTX_TOA, CTX_REC, CTX_OVA, CTX_OVB, CTX_EDGE, CTX_TAIL, CTX_CELL0, CTX_CELLB, CTX_ACC, CTX_PE2,
       CTX_RDP2, CTX_ASTRIDE, CTX_WORDS };

struct Params {
    const float *src[MAXSRC];        // distinct parent arrays, (n_outer, nz, ny, nx)
    float *outs[MAXV];
    int fa[MAXV], fb[MAXV];          // TMA sources of the parents (fb = fa for a plain field)
    const float *area;               // (n_outer, ny, nx)
    const float *psurf;              // (n_outer, ny, nx) sequential-cumsum surface pressure, mask_below only
    const float *delp_coarse;        // (n_outer, nz, ny/8, nx/8)
    const float *delp;               // (n_outer, nz, ny, nx), read by the geometry warps
    float fc1[MAXV], fc2[MAXV], fc3[MAXV];   // q = (a - c1 b) (c2 b + c3): plain (0,0,1), product (0,1,0), diff (1,0,1/scale)
    int any_derived;
    int nsrc, nfields, nz, ny, nx, mask_below;
    int nfw;                         // field warps = ceil(16 nfields / 32)
    float toa;
    int *flags;                      // per CTA: the fast variant met a non-finite field value (NaN-safe variant reruns the CTA)
    // cg_pressure_level_average_replicated: every coarse value is also stored at its address + rep[r] bytes (peer
    // copies of the output buffer over NVLink: the stores are the all-gather); rep_dc: delp_coarse as well
    int nrep, rep_dc, rep_kind;      // rep_kind CG_REPLICA_MULTICAST: rep[0] leads to a multicast mapping (one multimem.st)
    long long rep[CG_MAX_REPLICAS];
};

// ---- shared-memory carve-up --------------------------------------------------------------------
struct Layout {
    int ring, stage_bytes, bars, cell0, cell_bytes;
    int pe2, rdp2, den, ratio;       // within a cell block
    int rec, ova, ovb, tail, ctx, edge, colw, colps, cnt, cnt_bytes, srow, acc, acc_stride, total;
};

__host__ __device__ inline Layout make_layout(int nsrc, int nz, int nfw)
{
    Layout L;
    int o = 0;
    L.stage_bytes = (nsrc * BOX_BYTES + 1023) & ~1023;       // stage bases keep the swizzle phase of a row
    L.ring = o;  o += NSTAGE * L.stage_bytes;    o = (o + 127) & ~127;
    L.bars = o;  o += 128;
    L.cell0 = o;
    int w = 0;
    L.pe2 = w;   w += (nz + 2) * 4;              w = (w + 15) & ~15;
    L.rdp2 = w;  w += (nz + 1) * 4;              w = (w + 15) & ~15;
    L.den = w;   w += (nz + 1) * 4;              w = (w + 15) & ~15;
    L.ratio = w; w += (nz + 1) * 4;              // thickness of level k - 1 over thickness of level k
    L.cell_bytes = (w + 127) & ~127;
    o += NCELL * L.cell_bytes;
    L.rec = o;   o += NSTAGE * REC_BYTES;
    L.ova = o;   o += NSTAGE * OVA_BYTES;
    L.ovb = o;   o += NSTAGE * OVB_BYTES;
    L.tail = o;  o += TAIL_BYTES;
    L.ctx = o;   o += 128;
    L.edge = o;  o += NCOL * EDGE_WORDS * 4;
    L.colw = o;  o += 8 * NCOL * 4;              // [row][column of the CTA]
    L.colps = o; o += 8 * NCOL * 4;
    L.cnt_bytes = ((nz + 2) * NCOL + 127) & ~127;    // per row: coarse interfaces at or above every fine interface
    L.cnt = o;   o += 2 * L.cnt_bytes;
    L.srow = o;  o += ((nsrc * KC * 8) + 127) & ~127;    // per ring row (source, level of the stage): address of its level 0 plane | level
    L.acc_stride = nfw * 32 * 4;                 // bytes per accumulator slot: one float per work item
    L.acc = o;   o += (nz + 1) * L.acc_stride;
    L.total = o;
    return L;
}

// 1 / x for the pressure differences of the geometry (never denormal or huge): one MUFU.RCP, none of the range
// tests that div.approx carries
__device__ __forceinline__ float frcp(float x)
{
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
// NVLS: one store to a multicast address, the NVSwitch writes the value into every GPU's copy of the buffer
__device__ __forceinline__ void multimem_st(float *mc, float v)
{
    asm volatile("multimem.st.relaxed.sys.global.f32 [%0], %1;" ::"l"(mc), "f"(v) : "memory");
}
__device__ __forceinline__ void multimem_st4(float *mc, float4 v)
{
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
                 : "memory");
}
__device__ __forceinline__ int ldsu8(uint32_t addr)
{
    int v;
    asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(addr));
    return v;
}
__device__ __forceinline__ void stsu8(uint32_t addr, int v)
{
    asm volatile("st.shared.u8 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}

#ifdef CG_PAIR_TRACE
// timeline of one CTA (debug builds only): clock64 at the hand-over points of every stage
constexpr int TRACE_CTA = 700, TRACE_SLOTS = 10, TRACE_STAGES = 256, TRACE_VALS = 6;
__device__ long long g_trace[TRACE_SLOTS * TRACE_STAGES * TRACE_VALS];
#define TRACE(slot, it, k) do { if (!NS && cta == TRACE_CTA && lane == 0 && (it) < TRACE_STAGES) g_trace[((slot) * TRACE_STAGES + (it)) * TRACE_VALS + (k)] = clock64(); } while (0)
#else
#define TRACE(slot, it, k) do { } while (0)
#endif

// ---- geometry warps ---------------------------------------------------------------------------------
// Step t of column x (geometry warp g handles t = 3 chunk + g): level t arrives.  P(t-3) .. P(t+1) are the
// column's interface pressures; cnt = the column's entry in the walker's table of the row.
// Writes the column's half of the pair record:
//   E1, E2, F-1, G, H: PPM stencil of layer t-1 (mappm.f90:651-683, delp-only factors)
//   dpw: weight of the whole layer t-2 in the target level open at its bottom
//   ex, ey, ez, ew, off: the first event of layer t-2 (a no-op when it has none); further events go to the lists
template <bool NS>
__device__ __forceinline__ void geometry_step(uint32_t cnt, float pm3, float pm2, float pm1, float p0, float pp1,
                                              const CellG &g, const ColG &col, int t, int kk, int x, uint32_t rec,
                                              uint32_t ova, uint32_t ovb, uint32_t tail, uint32_t edge, int nz,
                                              int mask_below, float toa, int astride)
{
    // (approximate division, 2 ulp: these factors multiply differences of q)
    float E1 = 0.f, E2 = 0.f, Fm = 0.f, G = 0.f, H = 0.f;
    const int j = t - 1;
    if (j >= 1 && j <= nz - 2) {
        const float dm1 = pm1 - pm2, d0 = p0 - pm1, dp1 = pp1 - p0;     // mappm.f90:43 (from the rounded sums)
        const float d4 = dm1 + d0, d4p = d0 + dp1;
        const float r4 = __fdividef(1.f, d4), inv = __fdividef(d0, d4 + dp1);
        E1 = __fdividef(dm1 + 0.5f * d0, d4p) * inv;
        E2 = (dp1 + 0.5f * d0) * r4 * inv;
        if (j >= 2) {
            const float dm2 = pm2 - pm3, d4m = dm2 + dm1;
            const float cq = dm1 * r4, a1 = __fdividef(d4m, d4 + dm1), a2 = __fdividef(d4p, d4 + d0);
            const float gg = __fdividef(2.f, d4m + d4p);
            Fm = cq * (1.f + gg * d0 * (a1 - a2)) - 1.f;
            G = gg * d0 * a2;
            H = gg * dm1 * a1;
        }
    }
    // one-sided cubics at the column ends (mappm.f90:689-747), once per column
    if (t == 2 || t == nz) {
        const bool top = (t == 2);
        const float q0 = pm2, q1 = pm1, q2 = p0;                        // P(m), P(m+1), P(m+2), m = 0 or nz - 2
        const float d1 = top ? q1 - q0 : q2 - q1, d2 = top ? q2 - q1 : q1 - q0;
        const float D = d2 * (2.f * d2 * d2 + d1 * (d2 + 3.f * d1));
        const float Pq = d2 * (5.f * d1 + d2) - 3.f * d1 * d1;
        const float Rr = d1 * d2 * (d2 + 3.f * d1);
        const float r12 = 1.f / (d1 + d2);
        // pair-interleaved like the records: word i of column x sits at pair base + 8 i + 4 (x & 1)
        const uint32_t e = edge + 4 * ((x >> 1) * (2 * EDGE_WORDS) + (x & 1)) + 8 * (top ? 0 : 12);
        sts32(e, d2 * r12);             sts32(e + 8, d1 * r12);             sts32(e + 16, 2.f * r12);
        sts32(e + 24, (top ? 4.f : 1.f) / D);
        sts32(e + 32, d2);              sts32(e + 40, (top ? 0.5f : 2.f) * Pq);
        sts32(e + 48, (top ? 0.25f : 1.f) * Rr);
        sts32(e + 56, (top ? 2.f : 8.f) * d1 * d1 * d1);
        sts32(e + 64, d1);
    }
    // events of layer L = t - 2: the coarse interfaces in (top, bottom] of the layer (the two-pointer merge of
    // mappm.f90:70-75 is the column walker's); L == nz is everything below the surface
    const int L = t - 2, kn = nz;
    const float pt = pm2, pb = pm1;
    // cnt[L] = number of coarse interfaces at or above the top of layer L, left by the column walker
    int i0 = 0, i1 = 0, open_slot = DUMMY_SLOT;
    float dpw = 0.f;
    if (L >= 0 && L <= nz) {
        i0 = L == 0 ? 0 : ldsu8(cnt + NCOL * L);
        if (L < nz) {
            i1 = ldsu8(cnt + NCOL * (L + 1));
            const int ko = i1 - 1;                                      // level open at the bottom of L
            if (ko >= 0 && ko < kn && lds32(g.pe2 + 4 * ko) > toa) {
                dpw = level_scale(g, col, ko, mask_below) * (pb - pt);
                open_slot = ko + 1;
            }
        } else {
            i1 = kn + 1;
        }
    }
    const int n = i1 - i0;
    const bool below = L == nz;
    const int cap = below ? TAILMAX : MAXE;
    const uint32_t oa = ova + (kk * NCOL + x) * (MAXE * 16), ob = ovb + (kk * NCOL + x) * (MAXE * 4);
    // no coarse interface in the layer: the record is a no-op that keeps the running sum (info = 0); the fast
    // variant has no such select and instead moves the running sum to the open level's slot early (I = 0)
    Event e0 = {0.f, 0.f, 0.f, 0.f, (!NS && n == 0) ? open_slot : DUMMY_SLOT};
    int info = n;
    const int nlist = n - 1 <= cap ? n : 1;
    for (int e = 0; e < nlist; ++e) {
        const Event ev = make_event(g, col, pt, pb, L, i0 + e, nz, mask_below, toa);
        if (e == 0) e0 = ev;
        else if (below) {
            const uint32_t a = tail + (x * TAILMAX + e - 1) * 8;
            sts32(a, ev.x);
            stsi32(a + 4, ev.slot * astride);
        } else {
            sts128(oa + 16 * (e - 1), make_float4(ev.x, ev.y, ev.z, -ev.w));
            stsi32(ob + 4 * (e - 1), ev.slot * astride);
        }
    }
    if (n - 1 > cap) {                                                  // park what the field threads need to form them
        sts128(oa, make_float4(pt, pb, col.w, col.ps));
        stsi32(ob, i0);
        stsi32(ob + 4, n);
        info = INFO_SLOW;
    }
    const int other = __shfl_xor_sync(0xffffffffu, info, 1);
    const uint32_t r = rec + kk * REC_LEVEL + (x >> 1) * 16 + 4 * (x & 1);
    sts32(r, E1);                           sts32(r + 8, E2);
    sts32(r + GROUP_BYTES, Fm);             sts32(r + GROUP_BYTES + 8, G);
    sts32(r + 2 * GROUP_BYTES, -H);         sts32(r + 2 * GROUP_BYTES + 8, dpw);
    sts32(r + 3 * GROUP_BYTES, e0.x);       sts32(r + 3 * GROUP_BYTES + 8, e0.y);
    sts32(r + 4 * GROUP_BYTES, e0.z);       sts32(r + 4 * GROUP_BYTES + 8, -e0.w);
    stsi32(r + 5 * GROUP_BYTES, e0.slot * astride);
    stsi32(r + 5 * GROUP_BYTES + 8, (x & 1) ? ((info > 1 || other > 1) ? 1 : 0) : (info | other << 8));
}

// The same for an interior step, 3 <= t <= nz + 1, t != nz: layer L = t - 2 lies inside the column (1 <= L < nz) and
// every coarse interface it meets lies below the column top, so make_event reduces to one straight line.
//   interface i in the layer, s = (pe2[i] - pt) / dpl:  a = scale * dpl * s,  I = a (q + (1-s)^2 x' - s (1-s) y')
//   closes level i - 1 (i >= 2; level 0 is filled from the top layer's mean, mappm.f90:62-64) and restarts the
//   running sum of level i at -(thickness(i-1) / thickness(i)) I
template <bool NS>
__device__ __forceinline__ void geometry_interior(uint32_t cnt, float pm3, float pm2, float pm1, float p0, float pp1,
                                                  const CellG &g, uint32_t gratio, float cw, float cps, int t, int kk,
                                                  int x, uint32_t rec, uint32_t ova, uint32_t ovb, uint32_t edge,
                                                  int nz, int mask_below, int astride)
{
    float E1 = 0.f, E2 = 0.f, Fm = 0.f, G = 0.f, H = 0.f;
    if (t <= nz - 1) {                                                  // layer j = t - 1 in 2 .. nz - 2
        const float dm2 = pm2 - pm3, dm1 = pm1 - pm2, d0 = p0 - pm1, dp1 = pp1 - p0;    // mappm.f90:43 (from the rounded sums)
        const float d4 = dm1 + d0, d4p = d0 + dp1, d4m = dm2 + dm1;
        const float r4 = frcp(d4), inv = d0 * frcp(d4 + dp1);
        E1 = (dm1 + 0.5f * d0) * frcp(d4p) * inv;
        E2 = (dp1 + 0.5f * d0) * r4 * inv;
        const float cq = dm1 * r4, a1 = d4m * frcp(d4 + dm1), a2 = d4p * frcp(d4 + d0);
        const float gg = 2.f * frcp(d4m + d4p);
        Fm = cq * (1.f + gg * d0 * (a1 - a2)) - 1.f;
        G = gg * d0 * a2;
        H = gg * dm1 * a1;
    }
    const int L = t - 2;
    const float pt = pm2, dpl = pm1 - pm2;
    // the column walker's table: coarse interfaces at or above the top / the bottom of the layer (>= 1)
    const int i0 = ldsu8(cnt + NCOL * L), i1 = ldsu8(cnt + NCOL * (L + 1));
    const int n = i1 - i0, ko = i1 - 1;                                 // ko = level open at the bottom of the layer
    auto scale = [&](int k) {                                           // weight / thickness of level k, 0 <= k < nz
        float v = cw * lds32(g.rdp2 + 4 * k);
        if (mask_below) v = lds32(g.pe2 + 4 * (k + 1)) < cps ? v : 0.f; // regridz.py:150-161
        return v;
    };
    auto restart = [&](int i) {                                         // scale(i) / scale(i - 1), 2 <= i <= nz
        float v = lds32(gratio + 4 * i);                                // (0 at i == nz)
        if (mask_below) v = lds32(g.pe2 + 4 * (i + 1)) < cps ? v : 0.f; // (the mask only switches off downwards)
        return v;
    };
    const bool open = ko >= 1 && ko < nz;
    const float dpw = open ? scale(ko < nz ? ko : nz - 1) * dpl : 0.f;
    const int open_slot = open ? ko + 1 : DUMMY_SLOT;
    const float rdpl = frcp(dpl);
    const uint32_t oa = ova + (kk * NCOL + x) * (MAXE * 16), ob = ovb + (kk * NCOL + x) * (MAXE * 4);
    float ex = 0.f, ey = 0.f, ez = 0.f, ew = 0.f;
    int slot = NS ? DUMMY_SLOT : open_slot;                             // no interface in the layer: see geometry_step
    int info = n;
    {
        const int i = i0 < nz ? i0 : nz;
        const bool end_ok = i0 >= 2;
        const int ks = end_ok ? i - 1 : (i < nz ? i : nz - 1);
        const float s = (lds32(g.pe2 + 4 * i) - pt) * rdpl, oms = 1.f - s;
        const float a = scale(ks) * dpl * s;
        if (n > 0) {
            ex = a; ey = a * oms * oms; ez = -a * s * oms;
            ew = end_ok ? restart(i) : 1.f;
            slot = end_ok ? i : DUMMY_SLOT;
        }
    }
    if (n > 1) {
        if (n - 1 <= MAXE) {
            for (int e = 1; e < n; ++e) {
                const int i = i0 + e;                                   // 2 <= i <= nz
                const float s = (lds32(g.pe2 + 4 * i) - pt) * rdpl, oms = 1.f - s;
                const float a = scale(i - 1) * dpl * s;
                sts128(oa + 16 * (e - 1), make_float4(a, a * oms * oms, -a * s * oms, -restart(i)));
                stsi32(ob + 4 * (e - 1), i * astride);
            }
        } else {                                                        // park what the field threads need to form them
            sts128(oa, make_float4(pt, pm1, cw, cps));
            stsi32(ob, i0);
            stsi32(ob + 4, n);
            info = INFO_SLOW;
        }
    }
    const int other = __shfl_xor_sync(0xffffffffu, info, 1);
    const uint32_t r = rec + kk * REC_LEVEL + (x >> 1) * 16 + 4 * (x & 1);
    sts32(r, E1);                           sts32(r + 8, E2);
    sts32(r + GROUP_BYTES, Fm);             sts32(r + GROUP_BYTES + 8, G);
    sts32(r + 2 * GROUP_BYTES, -H);         sts32(r + 2 * GROUP_BYTES + 8, dpw);
    sts32(r + 3 * GROUP_BYTES, ex);         sts32(r + 3 * GROUP_BYTES + 8, ey);
    sts32(r + 4 * GROUP_BYTES, ez);         sts32(r + 4 * GROUP_BYTES + 8, -ew);
    stsi32(r + 5 * GROUP_BYTES, slot * astride);
    stsi32(r + 5 * GROUP_BYTES + 8, (x & 1) ? ((info > 1 || other > 1) ? 1 : 0) : (info | other << 8));
}

// ---- field threads: rare paths (out of line) -------------------------------------------------------
__device__ __forceinline__ uint32_t ctx_u32(uint32_t sctx, int i) { return (uint32_t)ldsi32(sctx + 4 * i); }

// close a target level with `tot` (NaN-skipping block sum, coarsen.py:211)
template <bool NS>
__device__ __forceinline__ void flush(uint32_t acc_item, int off, float tot)
{
    red_add(acc_item + off, (!NS || tot == tot) ? tot : 0.f);
}

// Events 1 .. n-1 of both columns of pair p at step t; r2 = restart values left by the first events.
// code = kk | buffer << 2 | p << 4 | t << 8 | infos << 16.  Everything the events need is recomputed from the
// context block so that the call site stays a handful of instructions (3 % of the pair steps of a budget timestep).
template <bool NS>
__device__ __noinline__ u64 extras_pair(uint32_t sctx, int code, u64 q2, u64 x2, u64 y2, u64 r2)
{
    const int kk = code & 3, b = (code >> 2) & 3, p = (code >> 4) & 15, t = (code >> 8) & 255;
    const int info2[2] = {(code >> 16) & 255, (code >> 24) & 255};
    const int nz = ldsi32(sctx + 4 * CTX_NZ);
    const uint32_t acc_item = ctx_u32(sctx, CTX_ACC) + 4 * threadIdx.x;
    float qv[2], xv[2], yv[2], rv[2];
    upk(q2, qv[0], qv[1]); upk(x2, xv[0], xv[1]); upk(y2, yv[0], yv[1]); upk(r2, rv[0], rv[1]);
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const int info = info2[h];
        if (info <= 1) continue;
        const int xc = 2 * p + h;
        const float q = qv[h], x = xv[h], y = yv[h];
        float r = rv[h];
        if (info != INFO_SLOW && t == nz + 2) {                         // below the surface: I = x * q, restart at 0
            const uint32_t a = ctx_u32(sctx, CTX_TAIL) + xc * (TAILMAX * 8);
            for (int e = 0; e + 2 <= info; ++e) {
                const float I = lds32(a + 8 * e) * q;
                flush<NS>(acc_item, ldsi32(a + 8 * e + 4), r + I);
                r = 0.f * I;
            }
            rv[h] = r;
            continue;
        }
        const uint32_t oa = ctx_u32(sctx, CTX_OVA) + b * OVA_BYTES + (kk * NCOL + xc) * (MAXE * 16);
        const uint32_t ob = ctx_u32(sctx, CTX_OVB) + b * OVB_BYTES + (kk * NCOL + xc) * (MAXE * 4);
        if (info != INFO_SLOW) {
            for (int e = 0; e + 2 <= info; ++e) {
                const float4 ev = lds128(oa + 16 * e);
                const float I = event_sum<NS>(ev.x, ev.y, ev.z, q, x, y);
                flush<NS>(acc_item, ldsi32(ob + 4 * e), r + I);
                r = ev.w * I;                                           // stored negated
            }
            rv[h] = r;
            continue;
        }
        // more than the lists hold (rare): formed on the fly from the column's pressures parked in the list
        const float4 pp = lds128(oa);                                   // pt, pb, weight, surface pressure
        const int i0 = ldsi32(ob), n = ldsi32(ob + 4);
        const int mask_below = ldsi32(sctx + 4 * CTX_MASK), astride = ldsi32(sctx + 4 * CTX_ASTRIDE);
        const float toa = lds32(sctx + 4 * CTX_TOA);
        const uint32_t cb = ctx_u32(sctx, CTX_CELL0) + (p >> 2) * ctx_u32(sctx, CTX_CELLB);
        CellG g;
        g.pe2 = cb + ctx_u32(sctx, CTX_PE2);
        g.rdp2 = cb + ctx_u32(sctx, CTX_RDP2);
        ColG col;
        col.w = pp.z;
        col.ps = pp.w;
        for (int e = 1; e < n; ++e) {
            const Event ev = make_event(g, col, pp.x, pp.y, t - 2, i0 + e, nz, mask_below, toa);
            const float I = event_sum<NS>(ev.x, ev.y, ev.z, q, x, y);
            flush<NS>(acc_item, ev.slot * astride, r + I);
            r = -(ev.w * I);
        }
        rv[h] = r;
    }
    return pk(rv[0], rv[1]);
}

// Loads of the hot step the compiler may schedule freely (no volatile, no memory clobber): the records and the field
// boxes are read-only while a stage is being consumed.  Their addresses derive from opaque() values formed after
// the stage's barrier waits, and the state they feed is pinned before the stage is handed back (pin_state).
__device__ __forceinline__ void lds2x64_s(uint32_t addr, u64 &a, u64 &b)
{
    asm("ld.shared.v2.b64 {%0,%1}, [%2];" : "=l"(a), "=l"(b) : "r"(addr));
}
__device__ __forceinline__ u64 lds64_s(uint32_t addr)
{
    u64 a;
    asm("ld.shared.b64 %0, [%1];" : "=l"(a) : "r"(addr));
    return a;
}
__device__ __forceinline__ int4 lds128i_s(uint32_t addr)
{
    int4 v;
    asm("ld.shared.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
    return v;
}

// one pair's record of one step
struct PairRec {
    u64 E1, E2, Fm, G, NH, DPW, EX, EY, EZ, NW;
    int4 oi;                         // offa, offb, infos, flag
};
__device__ __forceinline__ void load_pair_rec(PairRec &r, uint32_t rb)      // rb = record of the level + 16 p
{
    lds2x64(rb, r.E1, r.E2);
    lds2x64(rb + GROUP_BYTES, r.Fm, r.G);
    lds2x64(rb + 2 * GROUP_BYTES, r.NH, r.DPW);
    lds2x64(rb + 3 * GROUP_BYTES, r.EX, r.EY);
    lds2x64(rb + 4 * GROUP_BYTES, r.EZ, r.NW);
    r.oi = lds128i(rb + 5 * GROUP_BYTES);
}
__device__ __forceinline__ void load_pair_rec_s(PairRec &r, uint32_t rb)
{
    lds2x64_s(rb, r.E1, r.E2);
    lds2x64_s(rb + GROUP_BYTES, r.Fm, r.G);
    lds2x64_s(rb + 2 * GROUP_BYTES, r.NH, r.DPW);
    lds2x64_s(rb + 3 * GROUP_BYTES, r.EX, r.EY);
    lds2x64_s(rb + 4 * GROUP_BYTES, r.EZ, r.NW);
    r.oi = lds128i_s(rb + 5 * GROUP_BYTES);
}

// Both columns of a pair at one of the end steps of the walk (t == 2: nothing above / one-sided cubic at the top,
// mappm.f90:689-707; t >= nz: one-sided cubic at the bottom, :729-747, the bottom layer, the layer below the
// surface).  State of the pair: A = q(t-1), B = q(t-2), dq1 = delq(t-1), dcp = dc(t-2), xp = AL(t-2) - q(t-2), run.
// Returns delq(t), dc(t-1), AL(t-1) - q(t-1), run and the 0 / 1 mask "layer t-1 is not flat".
// The parabola limiter is the compare form of ppm_limiters lmt = 0 (:875-894): the edges of the top and bottom
// layers may lie on the same side of the mean, which the sign-splice form of the hot step does not cover.
// code = kk | buffer << 2 | p << 4.  Out of line: 4 of nz + 1 steps.
struct EndOut { u64 dq, dc, x, run, mk; };
template <bool NS>
__device__ __noinline__ EndOut end_pair(uint32_t sctx, int t, int code, u64 qn, u64 q3, u64 A, u64 B, u64 dq1, u64 dcp,
                                        u64 xp, u64 run)
{
    const int kk = code & 3, b = (code >> 2) & 3, p = (code >> 4) & 15;
    const int nz = ldsi32(sctx + 4 * CTX_NZ);
    const uint32_t rec = ctx_u32(sctx, CTX_REC) + b * REC_BYTES;
    PairRec r;
    load_pair_rec(r, rec + kk * REC_LEVEL + 16 * p);
    const u64 half = pk(0.5f, 0.5f);
    u64 dqn = sub2(qn, A);                                                  // delq(t)
    const u64 df2 = fma2(r.E1, dqn, mul2(r.E2, dq1));                       // layer t-1, mappm.f90:661
    u64 dcn = slope2(df2, dq1, dqn);                                        // dc(t-1)
    u64 xn = fma2(r.NH, dcn, fma2(r.G, dcp, mul2(r.Fm, dq1)));              // AL(t-1) - q(t-1), :674-683
    u64 x = xp, y = add2(xn, dq1), dc = dcp;                                // layer t-2
    const uint32_t edge = ctx_u32(sctx, CTX_EDGE) + 4 * (p * (2 * EDGE_WORDS));
    if (t == 2) {
        // level 3 is visible (q3; same or next stage): AL of layer 2 from the record of step 3, then the cubic
        // through the top two layers and that edge.  Step 3 is level 0 of the next stage's records.
        const uint32_t r3 = ctx_u32(sctx, CTX_REC) + ((b + 1) % NSTAGE) * REC_BYTES + 16 * p;
        u64 e1, e2, fm, gg, nh, unused;
        lds2x64(r3, e1, e2); lds2x64(r3 + GROUP_BYTES, fm, gg); lds2x64(r3 + 2 * GROUP_BYTES, nh, unused);
        const u64 dq3 = sub2(q3, qn);
        const u64 dc2 = slope2(fma2(e1, dq3, mul2(e2, dqn)), dqn, dq3);
        const u64 al2 = add2(qn, fma2(nh, dc2, fma2(gg, dcn, mul2(fm, dqn))));
        u64 c0, c1_, c2, c3_, c4, c5, c6, c7, c8, pad;
        lds2x64(edge, c0, c1_); lds2x64(edge + 16, c2, c3_); lds2x64(edge + 32, c4, c5); lds2x64(edge + 48, c6, c7);
        lds2x64(edge + 64, c8, pad);
        const u64 q0 = B, q1 = A;
        const u64 qm = fma2(c0, q0, mul2(c1_, q1)), dq = mul2(c2, sub2(q1, q0));
        const u64 cc = mul2(sub2(sub2(al2, qm), mul2(c4, dq)), c3_);
        const u64 c3 = sub2(dq, mul2(cc, c5));
        const u64 inner = sub2(qm, mul2(cc, c6));
        const u64 top = add2(sub2(mul2(cc, c7), mul2(c3, c8)), inner);
        const u64 al1 = clamp2(inner, q0, q1);
        x = sub2(top, q0); y = sub2(al1, q0); dc = mul2(half, sub2(al1, q0)); xn = sub2(al1, q1);
    } else if (t == nz) {
        const uint32_t eb = edge + 8 * 12;
        u64 c0, c1_, c2, c3_, c4, c5, c6, c7, c8, pad;
        lds2x64(eb, c0, c1_); lds2x64(eb + 16, c2, c3_); lds2x64(eb + 32, c4, c5); lds2x64(eb + 48, c6, c7);
        lds2x64(eb + 64, c8, pad);
        const u64 qe = A, qi = B;
        const u64 ali = add2(qi, xp);
        const u64 qm = fma2(c0, qe, mul2(c1_, qi)), dq = mul2(c2, sub2(qi, qe));
        const u64 cc = mul2(sub2(sub2(ali, qm), mul2(c4, dq)), c3_);
        const u64 c3 = sub2(dq, mul2(cc, c5));
        const u64 inner = sub2(qm, mul2(cc, c6));
        const u64 bot = add2(sub2(mul2(cc, c7), mul2(c3, c8)), inner);
        const u64 alb = clamp2(inner, qe, qi);
        y = sub2(alb, qi); xn = sub2(alb, qe); dcn = mul2(half, sub2(qe, alb)); dqn = sub2(bot, qe);   // layer nz-1 is finished next step
    } else if (t == nz + 1) {
        y = dq1;                                                            // (dq1 holds bottom edge - q(nz-1))
    } else if (t > nz + 1) {                                                // below the surface
        x = 0; y = 0; dc = 0;
    }
    // limited parabola of layer t-2 in deviation form, x = AL - q, y = AR - q, A6 = -3 (x + y)
    // (ppm_limiters lmt = 0, mappm.f90:875-894), then the events of the layer and its whole-layer term
    const u64 q = B;
    {
        const u64 da1 = sub2(y, x), sm = add2(x, y);
        const u64 tt = mul2(sm, da1), da2 = mul2(da1, mul2(da1, pk(1.f / 3.f, 1.f / 3.f)));
        float xa, xb, ya, yb, ta, tb, ga, gb, ca, cb;
        upk(x, xa, xb); upk(y, ya, yb); upk(tt, ta, tb); upk(da2, ga, gb); upk(dc, ca, cb);
        const float mxa = -2.f * xa, mxb = -2.f * xb, mya = -2.f * ya, myb = -2.f * yb;
        const bool k1a = ta > ga, k2a = -ta > ga, k1b = tb > gb, k2b = -tb > gb;    // k1 and k2 exclude each other (da2 >= 0)
        ya = k1a ? mxa : ya;    xa = k2a ? mya : xa;                    // a6da < -da2: AR = AL - 3 (AL - q); > da2: AL = AR - 3 (AR - q)
        yb = k1b ? mxb : yb;    xb = k2b ? myb : xb;
        xa = ca == 0.f ? 0.f : xa;  ya = ca == 0.f ? 0.f : ya;
        xb = cb == 0.f ? 0.f : xb;  yb = cb == 0.f ? 0.f : yb;
        x = pk(xa, xb); y = pk(ya, yb);
    }
    const uint32_t acc_item = ctx_u32(sctx, CTX_ACC) + 4 * threadIdx.x;
    const u64 I = event_sum2<NS>(r.EX, r.EY, r.EZ, q, x, y);
    float ta_, tb_, ra, rb_, na, nb;
    upk(add2(run, I), ta_, tb_); upk(run, ra, rb_); upk(mul2(r.NW, I), na, nb);     // -w is what the record holds
    flush<NS>(acc_item, r.oi.x, ta_);
    flush<NS>(acc_item, r.oi.y, tb_);
    const int ia = r.oi.z & 255, ib = (r.oi.z >> 8) & 255;
    u64 rs = pk((NS && ia == 0) ? ra : na, (NS && ib == 0) ? rb_ : nb);
    if (r.oi.w) rs = extras_pair<NS>(sctx, kk | b << 2 | p << 4 | t << 8 | r.oi.z << 16, q, x, y, rs);
    float ca, cb;
    upk(dcn, ca, cb);
    EndOut o;
    o.dq = dqn; o.dc = dcn; o.x = xn;
    o.mk = pk(ca != 0.f ? 1.f : 0.f, cb != 0.f ? 1.f : 0.f);
    o.run = fma2(r.DPW, q, rs);
    return o;
}

// ---- field threads: the hot step -----------------------------------------------------------------
// Rolling state of a work item (two adjacent columns, one field): q(t-1), q(t-2) in a 3-slot ring indexed by
// t mod 3 = level within the stage (compile time), delq(t-1), dc(t-2), AL(t-2) - q(t-2), the mask "layer t-2 is
// not flat", the running integral of the target level that is open.
struct Pair {
    u64 q[3], dq, dc, xp, mk, run;
};
// the per-thread description of a field: byte offsets of its two columns in the parents' boxes for the three
// levels of a stage (128-B swizzle folded in: stage bases are multiples of 1024), q = (a - c1 b) (c2 b + c3)
struct Field {
    int offa[KC], offb[KC];
    u64 n1, m2, a3;
};
template <int KK>
__device__ __forceinline__ u64 load_q(uint32_t stage, const Field &f, int any_derived)
{
    u64 q = lds64_s(stage + f.offa[KK]);
    if (any_derived) {                                   // uniform across the launch (budgets.py:80, :151)
        const u64 b = lds64_s(stage + f.offb[KK]);
        q = mul2(fma2(b, f.n1, q), fma2(b, f.m2, f.a3));
    }
    return q;
}
__device__ __forceinline__ float absmin3_sign(float a, float b, float c, float s)
{
    return copysignf(fminf(fminf(fabsf(a), fabsf(b)), fabsf(c)), s);
}
__device__ __forceinline__ float absmin2_sign(float a, float b, float s) { return copysignf(fminf(fabsf(a), fabsf(b)), s); }

// One step of the tandem walk for a pair: level t = 3 chunk + S arrives (qn), layer t-2 is finished.
// recp = record of the step + 16 p; code = buffer << 2 | p << 4.
// Fast variant (finite values):
//   dc = sign(df2) min(|df2|, |delq(t-1)|, |delq(t)|) where the two differences agree in sign, else 0
//        (mappm.f90:663-665: max(q-,q,q+) - q and q - min(q-,q,q+) are |delq(t)|, |delq(t-1)| or 0; df2 is a
//        positive combination of the two differences, so its sign is theirs)
//   limiter lmt = 0 (:875-894) in deviation form x = AL - q, y = AR - q: the edges of an interior layer that is
//        not flat lie on opposite sides of the mean, where "a6da < -da2 -> AR = AL - 3 (AL - q)" and its mirror
//        are y = sign(y) min(|y|, 2 |x|), x = sign(x) min(|x|, 2 |y|); flat layers through the 0 / 1 mask
// NaN-safe variant: the compare forms, NaN behaving as in the Fortran.
template <bool NS, int S>
__device__ __forceinline__ void hot_step(Pair &s, uint32_t recp, u64 qn, uint32_t acc_item, uint32_t sctx, int code,
                                         int t, u64 &chk)
{
    constexpr int a = (S + 2) % 3, b = (S + 1) % 3;                     // slots of q(t-1), q(t-2)
    PairRec r;
    load_pair_rec_s(r, recp);
    const u64 A = s.q[a], B = s.q[b], dq1 = s.dq;
    const u64 dqn = sub2(qn, A);                                        // delq(t)
    const u64 df2 = fma2(r.E1, dqn, mul2(r.E2, dq1));                   // layer t-1, mappm.f90:661
    u64 dcn, mkn, xl, yl;
    if (!NS) {
        float dfa, dfb, d1a, d1b, dna, dnb, saa, sab;
        upk(df2, dfa, dfb); upk(dq1, d1a, d1b); upk(dqn, dna, dnb);
        // a b > 0 from the product's sign (one packed multiply instead of a packed sum and difference; a product that
        // underflows - both differences below 1e-19 - gives slope 0 where the compare form gives one of that size)
        upk(mul2(dq1, dqn), saa, sab);
        mkn = pk(gt01(saa, 0.f), gt01(sab, 0.f));
        dcn = mul2(pk(absmin3_sign(dfa, d1a, dna, d1a), absmin3_sign(dfb, d1b, dnb, d1b)), mkn);
    } else {
        dcn = slope2(df2, dq1, dqn);
        mkn = 0;
    }
    const u64 xn = fma2(r.NH, dcn, fma2(r.G, s.dc, mul2(r.Fm, dq1)));   // AL(t-1) - q(t-1), :674-683
    if (!NS) {
        const u64 x = s.xp, y = mul2(add2(xn, dq1), s.mk);
        float xa, xb, ya, yb, x2a, x2b, y2a, y2b;
        upk(x, xa, xb); upk(y, ya, yb); upk(add2(x, x), x2a, x2b); upk(add2(y, y), y2a, y2b);
        xl = pk(absmin2_sign(xa, y2a, xa), absmin2_sign(xb, y2b, xb));
        yl = pk(absmin2_sign(ya, x2a, ya), absmin2_sign(yb, x2b, yb));
    } else {
        const u64 x = s.xp, y = add2(xn, dq1);
        const u64 da1 = sub2(y, x), sm = add2(x, y);
        const u64 tt = mul2(sm, da1), da2 = mul2(da1, mul2(da1, pk(1.f / 3.f, 1.f / 3.f)));
        float xa, xb, ya, yb, ta, tb, ga, gb, ca, cb;
        upk(x, xa, xb); upk(y, ya, yb); upk(tt, ta, tb); upk(da2, ga, gb); upk(s.dc, ca, cb);
        const float mxa = -2.f * xa, mxb = -2.f * xb, mya = -2.f * ya, myb = -2.f * yb;
        ya = ta > ga ? mxa : ya;    xa = -ta > ga ? mya : xa;
        yb = tb > gb ? mxb : yb;    xb = -tb > gb ? myb : xb;
        xa = ca == 0.f ? 0.f : xa;  ya = ca == 0.f ? 0.f : ya;
        xb = cb == 0.f ? 0.f : xb;  yb = cb == 0.f ? 0.f : yb;
        xl = pk(xa, xb); yl = pk(ya, yb);
    }
    // the first event of each column closes a target level and restarts the next
    const u64 I = event_sum2<NS>(r.EX, r.EY, r.EZ, B, xl, yl);
    float ta_, tb_;
    upk(add2(s.run, I), ta_, tb_);
    flush<NS>(acc_item, r.oi.x, ta_);
    flush<NS>(acc_item, r.oi.y, tb_);
    u64 rs = mul2(r.NW, I);
    if (NS) {                                                           // no event: the running sum is kept
        float ra, rb_, na, nb;
        upk(s.run, ra, rb_); upk(rs, na, nb);
        rs = pk((r.oi.z & 255) == 0 ? ra : na, ((r.oi.z >> 8) & 255) == 0 ? rb_ : nb);
    }
    if (r.oi.w) rs = extras_pair<NS>(sctx, S | code | t << 8 | r.oi.z << 16, B, xl, yl, rs);
    s.run = fma2(r.DPW, B, rs);
    if (!NS) chk = fma2(qn, 0ull, chk);
    s.q[S] = qn; s.dq = dqn; s.dc = dcn; s.xp = xn; s.mk = mkn;
}

// one value of delp_coarse to every GPU's copy (cg_pressure_level_average_replicated); out of line like the outputs
__device__ __noinline__ void replicate_value(const Params &p, float *dst, float v)
{
    if (p.rep_kind == CG_REPLICA_MULTICAST) multimem_st((float *)((char *)dst + p.rep[0]), v);
    else for (int r = 0; r < p.nrep; ++r) *(float *)((char *)dst + p.rep[r]) = v;
}

// cg_pressure_level_average_replicated: the CTA's four cells of a (field, level) as one 16-byte store, sent to
// every GPU's copy of the output buffer - through the multicast mapping (one multimem.st, the NVSwitch
// replicates) or with one store per peer over NVLink.  Same sums in the same order as the plain output loop.
// Out of line: the single-GPU kernel's code and registers stay as they are.
__device__ __noinline__ void replicated_outputs(const Params &p, uint32_t acc0, int acc_stride, uint32_t den0, int cell_bytes,
                                                int nz, int V, int nfw, int tid, long o, long hc, long cell_off)
{
    for (int idx = tid; idx < nz * V; idx += nfw * 32) {
        const int vv = idx % V, k = idx / V;
        float r4[NCELL];
#pragma unroll
        for (int c = 0; c < NCELL; ++c) {
            const uint32_t a = acc0 + (k + 1) * acc_stride + 4 * ((NCELL * c) * V + vv);
            float sum = lds32(a);
            sum += lds32(a + 4 * V);
            sum += lds32(a + 8 * V);
            sum += lds32(a + 12 * V);
            r4[c] = sum / lds32(den0 + c * cell_bytes + 4 * k);
        }
        float *dst = p.outs[vv] + (o * nz + k) * hc + cell_off;
        const float4 val = make_float4(r4[0], r4[1], r4[2], r4[3]);
        if (p.rep_kind == CG_REPLICA_MULTICAST) {
            multimem_st4((float *)((char *)dst + p.rep[0]), val);       // lands in this GPU's copy as well
        } else {
            *(float4 *)dst = val;
            for (int r = 0; r < p.nrep; ++r) *(float4 *)((char *)dst + p.rep[r]) = val;
        }
    }
}

// ---- the kernel ------------------------------------------------------------------------------------
// (nfw + 4) warps: field warps 0 .. nfw-1, the three geometry warps, the column walker.  120 registers per
// thread at 544 threads.
// Two variants.  NS = false (fast): assumes every field value is finite, which lets the block sums drop their
// NaN tests and selects; it checks that assumption (0 * q summed over everything it reads) and flags the CTA
// otherwise.  NS = true: NaN-skipping sums exactly like the reference's nansum; launched after the fast one, its
// CTAs return at once unless flagged (p.flags == nullptr: runs everywhere).
template <bool NS>
__global__ void __launch_bounds__((MAXFW + NGW + 1) * 32, 1)
plevel_pair_kernel(const __grid_constant__ Params p)
{
    const int cta = (blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
    if (NS && p.flags && p.flags[cta] == 0) return;
    extern __shared__ __align__(1024) unsigned char smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int nz = p.nz, V = p.nfields, nsrc = p.nsrc, nfw = p.nfw;
    const Layout lay = make_layout(nsrc, nz, nfw);
    const uint32_t sbase = smem_u32(smem);
    const uint32_t ring = sbase + lay.ring;
    // mbarriers: full[NSTAGE] (field ring, TMA), gfull[NSTAGE] (records of geometry buffer b are written): the
    // field warps wait on them one by one, so that they are not tied to each other.  Named barrier 2 + b = stage /
    // buffer b is handed back (bar.arrive by the field warps, bar.sync by the geometry warps)
    const uint32_t bars = sbase + lay.bars;
    const uint32_t gfull = bars + 8 * NSTAGE;
    // count tables of the column walker, one per row parity: cfull (walker -> geometry warps), cempty (back)
    const uint32_t cfull = gfull + 8 * NSTAGE, cempty = cfull + 16;
    const uint32_t sctx = sbase + lay.ctx;
    const int nch = (nz + 3 + KC - 1) / KC;                     // steps t = 0 .. nz + 2
    const int total = 8 * nch;                                  // stages of the CTA
    const int bx = blockIdx.x, by = blockIdx.y;
    const long o = blockIdx.z;
    const float toa = p.toa;
    const int mask_below = p.mask_below;
    const uint32_t rec0 = sbase + lay.rec, ova0 = sbase + lay.ova, ovb0 = sbase + lay.ovb, edge = sbase + lay.edge;
    const uint32_t tail = sbase + lay.tail;
    const int nthreads = (nfw + KC) * 32;                       // field warps + the geometry warps of one stage

    for (int R = tid; R < nsrc * KC; R += blockDim.x) {
        const int kk = R % KC;
        const u64 a = (u64)(uintptr_t)(p.src[R / KC] + (long)kk * p.ny * p.nx) | (u64)kk;     // (16-byte aligned planes)
        asm volatile("st.shared.b64 [%0], %1;" ::"r"(sbase + lay.srow + 8 * R), "l"(a) : "memory");
    }
    if (tid == 0) {
        if (!NS) p.flags[cta] = 0;
        for (int s = 0; s < NSTAGE; ++s) {
            mbar_init(bars + 8 * s, KC * 32);               // full: every thread of the stage's geometry warps, when its copies have landed
            mbar_init(gfull + 8 * s, KC);                   // records: one arrival per geometry warp of the stage
        }
        for (int s = 0; s < 2; ++s) {
            mbar_init(cfull + 8 * s, 1);
            mbar_init(cempty + 8 * s, NGW);
        }
        asm volatile("fence.mbarrier_init.release.cluster;");
        stsi32(sctx + 4 * CTX_NZ, nz);              stsi32(sctx + 4 * CTX_MASK, mask_below);
        sts32(sctx + 4 * CTX_TOA, toa);             stsi32(sctx + 4 * CTX_REC, (int)rec0);
        stsi32(sctx + 4 * CTX_OVA, (int)ova0);      stsi32(sctx + 4 * CTX_OVB, (int)ovb0);
        stsi32(sctx + 4 * CTX_EDGE, (int)edge);     stsi32(sctx + 4 * CTX_TAIL, (int)tail);
        stsi32(sctx + 4 * CTX_CELL0, (int)(sbase + lay.cell0));
        stsi32(sctx + 4 * CTX_CELLB, lay.cell_bytes);
        stsi32(sctx + 4 * CTX_ACC, (int)(sbase + lay.acc));
        stsi32(sctx + 4 * CTX_PE2, lay.pe2);        stsi32(sctx + 4 * CTX_RDP2, lay.rdp2);
        stsi32(sctx + 4 * CTX_ASTRIDE, lay.acc_stride);
    }
    __syncthreads();

    const int ny = p.ny, nx = p.nx, nyc = ny >> 3, nxc = nx >> 3;
    const long h = (long)ny * nx, hc = (long)nyc * nxc;

    // roles of the last seven warps: the column walker sits in the middle (with 13 field warps it then shares its
    // scheduler with four field warps and no geometry warp; the other three schedulers get three field warps and two
    // geometry warps each)
    if (warp == nfw + KC) {
        // ---------------- column walker: lane = fine column, one row ahead of the geometry warps ----------
        // The two-pointer merge of mappm.f90:70-75: walking the column down with the reference's sequential
        // cumsum of interface pressures (same additions as the geometry warps: bit-identical), count the coarse
        // interfaces at or above every fine interface.  cnt[L][x] for L = 1 .. nz; one byte each.
        const int x = lane;
        const uint32_t pe2 = sbase + lay.cell0 + (x >> 3) * lay.cell_bytes + lay.pe2;
        asm volatile("bar.sync %0, %1;" ::"n"(BAR_WALK), "n"((NGW + 1) * 32) : "memory");   // the cells' coarse interfaces are in place
        const float *dcol = p.delp + o * nz * h + (long)(8 * by) * nx + NCOL * bx + x;
        constexpr int WB = 8;                                            // levels per batch of loads
        for (int row = 0; row < 8; ++row) {
            const float *d = dcol + (long)row * nx;
            float cur[WB], nxt[WB];
#pragma unroll
            for (int i = 0; i < WB; ++i) cur[i] = i < nz ? ldg_stream(d + (long)i * h) : 0.f;
            TRACE(9, row, 0);
            if (row >= 2) mbar_wait(cempty + 8 * (row & 1), ((row >> 1) - 1) & 1);
            TRACE(9, row, 1);
            const uint32_t crow = sbase + lay.cnt + (row & 1) * lay.cnt_bytes + x;
            float P = toa, pnext = lds32(pe2);
            int i = 0;
            for (int l0 = 0; l0 < nz; l0 += WB) {
#pragma unroll
                for (int j = 0; j < WB; ++j) nxt[j] = l0 + WB + j < nz ? ldg_stream(d + (long)(l0 + WB + j) * h) : 0.f;
#pragma unroll
                for (int j = 0; j < WB; ++j) {
                    const int L = l0 + j + 1;                            // interface below level l0 + j
                    if (L <= nz) {
                        P = P + cur[j];
                        while (i <= nz && pnext <= P) { ++i; pnext = lds32(pe2 + 4 * (i <= nz ? i : nz)); }
                        stsu8(crow + NCOL * L, i);
                    }
                }
#pragma unroll
                for (int j = 0; j < WB; ++j) cur[j] = nxt[j];
            }
            __syncwarp();
            TRACE(9, row, 2);
            if (lane == 0) mbar_arrive(cfull + 8 * (row & 1));
        }
        return;
    }
    if (warp >= nfw) {
        // ---------------- geometry warp gw: lane = fine column x of the CTA's 32, level 3 chunk + gw ----
        const int gw = warp - nfw < KC ? warp - nfw : warp - nfw - 1;
        const int x = lane, cell = x >> 3;
        const uint32_t colw = sbase + lay.colw, colps = sbase + lay.colps;
        // CTA prologue: coarse interfaces (sequential cumsum, vertically_dependent.py:61-63), 1 / thickness,
        // sum of the 64 columns' weights per target level (coarsen.py:212); cells and rows dealt over the warps
        for (int r = gw; r < 8; r += NGW) {
            const long col = o * h + (long)(8 * by + r) * nx + NCOL * bx + x;
            const float a = p.area[col];
            sts32(colw + 4 * (r * NCOL + x), (a == a) ? a : 0.f);
            sts32(colps + 4 * (r * NCOL + x), mask_below ? p.psurf[col] : 0.f);
        }
        for (int c = gw; c < NCELL; c += NGW) {
            const uint32_t cbw = sbase + lay.cell0 + c * lay.cell_bytes;
            for (int k = lane; k < nz; k += 32) {
                const float *src = p.delp_coarse + (o * nz + k) * hc + (long)by * nxc + NCELL * bx + c;
                const float dcv = *src;
                sts32(cbw + lay.rdp2 + 4 * k, dcv);
                if (p.rep_dc) replicate_value(p, const_cast<float *>(src), dcv);
            }
            __syncwarp();
            if (lane == 0) {
                float pe = toa;
                sts32(cbw + lay.pe2, pe);
                for (int k = 0; k < nz; ++k) { pe = pe + lds32(cbw + lay.rdp2 + 4 * k); sts32(cbw + lay.pe2 + 4 * (k + 1), pe); }
            }
            __syncwarp();
            for (int k = lane; k < nz; k += 32)
                sts32(cbw + lay.rdp2 + 4 * k, 1.f / (lds32(cbw + lay.pe2 + 4 * (k + 1)) - lds32(cbw + lay.pe2 + 4 * k)));
            __syncwarp();
            for (int k = lane; k <= nz; k += 32)
                sts32(cbw + lay.ratio + 4 * k, (k >= 1 && k < nz) ? lds32(cbw + lay.rdp2 + 4 * k) / lds32(cbw + lay.rdp2 + 4 * (k - 1)) : 0.f);
        }
        asm volatile("bar.sync 1, %0;" ::"n"(NGW * 32) : "memory");     // the geometry warps: weights and cell arrays are in place
        for (int c = gw; c < NCELL; c += NGW) {
            const uint32_t cbw = sbase + lay.cell0 + c * lay.cell_bytes;
            // the cell's 64 columns: rows lane / 8 and 4 + lane / 8, column lane % 8
            const int xc = 8 * c + (lane & 7), r0 = lane >> 3;
            const float w0 = lds32(colw + 4 * (r0 * NCOL + xc)), w1 = lds32(colw + 4 * ((r0 + 4) * NCOL + xc));
            const float s0 = lds32(colps + 4 * (r0 * NCOL + xc)), s1 = lds32(colps + 4 * ((r0 + 4) * NCOL + xc));
            if (!mask_below) {
                float d = w0 + w1;
#pragma unroll
                for (int m = 1; m < 32; m <<= 1) d += __shfl_xor_sync(0xffffffffu, d, m);
                for (int k = lane; k < nz; k += 32) sts32(cbw + lay.den + 4 * k, d);
            } else {
                for (int k = 0; k < nz; ++k) {
                    const float lower = lds32(cbw + lay.pe2 + 4 * (k + 1));
                    float d = (lower < s0 ? w0 : 0.f) + (lower < s1 ? w1 : 0.f);
#pragma unroll
                    for (int m = 1; m < 32; m <<= 1) d += __shfl_xor_sync(0xffffffffu, d, m);
                    if (lane == 0) sts32(cbw + lay.den + 4 * k, d);
                }
            }
        }
        asm volatile("bar.sync %0, %1;" ::"n"(BAR_WALK), "n"((NGW + 1) * 32) : "memory");   // with the column walker
        CellG g;
        g.pe2 = sbase + lay.cell0 + cell * lay.cell_bytes + lay.pe2;
        g.rdp2 = sbase + lay.cell0 + cell * lay.cell_bytes + lay.rdp2;
        const uint32_t gratio = sbase + lay.cell0 + cell * lay.cell_bytes + lay.ratio;
        // warp gw takes level sub of the stages it = par, par + 2, ...: two stages are in the making at any time
        const int sub = gw % KC, par = gw / KC;
        const float *dcol = p.delp + o * nz * h + (long)(8 * by) * nx + NCOL * bx + x;
        // Interface pressures by the reference's sequential cumsum (vertically_dependent.py:61-63): every warp
        // carries the running sum of its own column walk, same order of additions.  A step at level t needs
        // P(t-3) .. P(t+1); the warp's next step of the row is at t + 6.  dn = thicknesses of levels t-3 .. t+2,
        // loaded one step ahead; e0, e1 = levels t-5, t-4 at the first step of a row, where t <= 5.
        float dn[2 * KC], e0 = 0.f, e1 = 0.f, pbase = toa;
        // (addresses are carried from step to step: the next thicknesses at dnext, the stage's planes at snext bytes)
        const long dstep = 2L * KC * h, dwrap = (long)nx - (long)nch * KC * h;
        const float *dnext = dcol + (long)(par * KC + sub - 3) * h;
        long snext = 4 * (((o * nz + par * KC) * (long)ny + 8 * by) * nx);
        auto load_delp = [&](int chunk2) {
            const int t2 = chunk2 * KC + sub;
            const float *d = dnext;
            if (t2 >= 2 * KC && t2 + 2 < nz) {                       // inside the column
#pragma unroll
                for (int i = 0; i < 2 * KC; ++i) dn[i] = __ldg(d + i * h);
            } else {
#pragma unroll
                for (int i = 0; i < 2 * KC; ++i) {
                    const int lvl = t2 - 3 + i;
                    dn[i] = (lvl >= 0 && lvl < nz) ? __ldg(d + i * h) : 0.f;
                }
                if (t2 < 2 * KC) {                                   // first step of the warp in this row
                    e0 = t2 >= 5 ? __ldg(d - 2 * h) : 0.f;
                    e1 = t2 >= 4 ? __ldg(d - h) : 0.f;
                }
            }
        };
        int row_n = 0, chunk_n = par, b = par;                       // (nch >= 3: the first stages are in row 0)
        load_delp(chunk_n);
        // The field boxes of a stage, [source][level][32 columns] with the 16-byte chunks of a row xor-swizzled by
        // the row index (the field threads' 8-byte reads of a column pair are then conflict-free), are filled by
        // asynchronous 16-byte copies: an instruction moves four rows; warp sub takes instructions sub, sub + 3, ...
        const uint32_t srow = sbase + lay.srow;
        const int nrows = nsrc * KC;
        const long lane_off = 4 * (long)(NCOL * bx + 4 * (lane & 7));        // bytes
        // the first NCP of this lane's rows (all of them for up to 28 sources) with their addresses in registers
        constexpr int NCP = 7;
        const char *csrc[NCP];
        uint32_t cdst[NCP];                                              // byte offset in the stage | level of the stage; ~0: none
#pragma unroll
        for (int i = 0; i < NCP; ++i) {
            const int R = 4 * sub + (lane >> 3) + 4 * KC * i;
            csrc[i] = nullptr;
            cdst[i] = ~0u;
            if (R < nrows) {
                const u64 e = lds64(srow + 8 * R);
                csrc[i] = (const char *)(uintptr_t)(e & ~15ull) + lane_off;
                cdst[i] = (uint32_t)(R * 128 + ((((uint32_t)lane ^ (uint32_t)R) & 7u) << 4)) | ((uint32_t)e & 3u);
            }
        }
        int row = -1;
        float cw = 0.f, cps = 0.f;
        uint32_t cnt = 0;
        bool bad_geometry = false;                                   // a thickness that is not positive and finite
        for (int it = par; it < total; it += 2) {
            const int chunk = chunk_n, r2 = row_n;
            TRACE(gw, it, 0);
            const long stage_off = snext;
            chunk_n += 2;
            dnext += dstep;
            snext += 4 * dstep;
            if (chunk_n >= nch) { chunk_n -= nch; ++row_n; dnext += dwrap; snext += 4 * dwrap; }
            if (r2 != row) {
                if (row >= 0) { __syncwarp(); if (lane == 0) mbar_arrive(cempty + 8 * (row & 1)); }
                row = r2;
                cw = lds32(colw + 4 * (row * NCOL + x));
                cps = lds32(colps + 4 * (row * NCOL + x));
                mbar_wait(cfull + 8 * (row & 1), (row >> 1) & 1);
                cnt = sbase + lay.cnt + (row & 1) * lay.cnt_bytes + x;
                pbase = (toa + e0) + e1;
            }
            const float pm3 = pbase, pm2 = pm3 + dn[0], pm1 = pm2 + dn[1], p0 = pm1 + dn[2], pp1 = p0 + dn[3];
            pbase = (pp1 + dn[4]) + dn[5];
            if (!NS) {
                // the fast variant's block sums have no NaN tests: a source layer without thickness (0 / 0 in
                // mappm.f90:75) sends the CTA to the NaN-safe variant, like a non-finite field value does
                const int t1 = chunk * KC + sub;
                if (t1 >= 1 && t1 <= nz) bad_geometry = bad_geometry || !(p0 - pm1 > 0.f) || !(p0 - pm1 < 3.0e38f);
            }
            if (it + 2 < total) load_delp(chunk_n);
            // stage b of the field ring and geometry buffer b are released together by the field warps
            TRACE(gw, it, 1);
            if (it >= NSTAGE) asm volatile("bar.sync %0, %1;" ::"r"(BAR_STAGE0 + b), "r"(nthreads) : "memory");
            TRACE(gw, it, 2);
            {
                const uint32_t dst0 = ring + b * lay.stage_bytes;
                const int lv0 = chunk * KC;
                if (lv0 + KC <= nz) {                                        // every level of the stage is in the column
#pragma unroll
                    for (int i = 0; i < NCP; ++i)
                        if (cdst[i] != ~0u)
                            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst0 + (cdst[i] & ~15u)),
                                         "l"(csrc[i] + stage_off) : "memory");
                } else {
#pragma unroll
                    for (int i = 0; i < NCP; ++i)
                        if (cdst[i] != ~0u) {
                            const int sz = lv0 + (int)(cdst[i] & 3u) < nz ? 16 : 0;   // levels past the column bottom: zeros
                            asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst0 + (cdst[i] & ~15u)),
                                         "l"(csrc[i] + stage_off), "r"(sz) : "memory");
                        }
                }
                for (int R = 4 * sub + (lane >> 3) + 4 * KC * NCP; R < nrows; R += 4 * KC) {     // more than 28 sources
                    const u64 e = lds64(srow + 8 * R);
                    const int kk = (int)e & 3;
                    const char *src = (const char *)(uintptr_t)(e & ~15ull) + stage_off + lane_off;
                    const uint32_t dst = dst0 + R * 128 + ((((uint32_t)lane ^ (uint32_t)R) & 7u) << 4);
                    const int sz = lv0 + kk < nz ? 16 : 0;
                    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(sz) : "memory");
                }
                asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(bars + 8 * b) : "memory");
            }
            TRACE(gw, it, 4);
            const int t = chunk * KC + sub;
            if (t >= 3 && t <= nz + 1 && t != nz) {
                geometry_interior<NS>(cnt, pm3, pm2, pm1, p0, pp1, g, gratio, cw, cps, t, sub, x, rec0 + b * REC_BYTES,
                                      ova0 + b * OVA_BYTES, ovb0 + b * OVB_BYTES, edge, nz, mask_below, lay.acc_stride);
            } else if (t >= 2) {
                ColG col;
                col.w = cw;
                col.ps = cps;
                geometry_step<NS>(cnt, pm3, pm2, pm1, p0, pp1, g, col, t, sub, x, rec0 + b * REC_BYTES,
                                  ova0 + b * OVA_BYTES, ovb0 + b * OVB_BYTES, tail, edge, nz, mask_below, toa, lay.acc_stride);
            }
            __syncwarp();                                            // (mbarrier.arrive releases the warp's writes)
            TRACE(gw, it, 3);
            if (lane == 0) mbar_arrive(gfull + 8 * b);
            b = (b + 2) % NSTAGE;
        }
        __syncwarp();
        if (lane == 0 && row >= 0) mbar_arrive(cempty + 8 * (row & 1));
        if (!NS && bad_geometry) p.flags[cta] = 1;
        return;
    }

    // ---------------- field threads: work item = (column pair, field) ----------------------------------
    const int item = tid, nitems = NPAIR * V;
    const bool active = item < nitems;
    const int pr = active ? item / V : NPAIR - 1;                    // pair of the CTA row (idle lanes shadow the last one)
    const int v = active ? item - pr * V : 0;
    const uint32_t acc_item = sbase + lay.acc + 4 * item;            // every thread owns a column of the accumulator
    for (int k = 0; k <= nz; ++k) sts32(acc_item + k * lay.acc_stride, 0.f);
    Field fld;
    {
        const int fa = p.fa[v], fb = p.fb[v];
        const float c1 = p.fc1[v], c2 = p.fc2[v], c3 = p.fc3[v];
#pragma unroll
        for (int kk = 0; kk < KC; ++kk) {
            fld.offa[kk] = fa * BOX_BYTES + kk * 128 + ((((fa * KC + kk) ^ (pr >> 1)) & 7) << 4) + 8 * (pr & 1);
            fld.offb[kk] = fb * BOX_BYTES + kk * 128 + ((((fb * KC + kk) ^ (pr >> 1)) & 7) << 4) + 8 * (pr & 1);
        }
        fld.n1 = pk(-c1, -c1); fld.m2 = pk(c2, c2); fld.a3 = pk(c3, c3);
    }
    const int any_derived = p.any_derived;
    const uint32_t recp = rec0 + 16 * pr;
    const int hot_end = nz / KC;                                     // chunks 1 .. hot_end - 1 hold levels 3 .. nz - 1 only

    u64 chk = 0;
    for (int row = 0; row < 8; ++row) {
        Pair st;
        st.q[0] = st.q[1] = st.q[2] = 0; st.dq = 0; st.dc = 0; st.xp = 0; st.mk = 0; st.run = 0;
        for (int chunk = 0; chunk < nch; ++chunk) {
            const int it = row * nch + chunk, b = it % NSTAGE;
            if (warp == 0 || warp == nfw - 1) TRACE(warp == 0 ? 7 : 8, it, 0);
            mbar_wait(gfull + 8 * b, (it / NSTAGE) & 1);
            if (warp == 0 || warp == nfw - 1) TRACE(warp == 0 ? 7 : 8, it, 1);
            mbar_wait(bars + 8 * b, (it / NSTAGE) & 1);
            if (warp == 0 || warp == nfw - 1) TRACE(warp == 0 ? 7 : 8, it, 2);
            const uint32_t stage = opaque(ring + b * lay.stage_bytes);
            const uint32_t rec = opaque(recp + b * REC_BYTES);
            const int code = b << 2 | pr << 4;
            if (chunk == 0) {
                // the walk starts at t = 2: levels 0 and 1 only seed the rolling state; the one-sided cubic at the
                // top needs level 3 and the record of step 3, which are the head of the next stage
                const int b1 = (it + 1) % NSTAGE;
                mbar_wait(gfull + 8 * b1, ((it + 1) / NSTAGE) & 1);
                mbar_wait(bars + 8 * b1, ((it + 1) / NSTAGE) & 1);
                const u64 l0 = load_q<0>(stage, fld, any_derived), l1 = load_q<1>(stage, fld, any_derived);
                const u64 l2 = load_q<2>(stage, fld, any_derived);
                const u64 l3 = load_q<0>(opaque(ring + b1 * lay.stage_bytes), fld, any_derived);
                if (!NS) chk = fma2(l2, 0ull, fma2(l1, 0ull, fma2(l0, 0ull, chk)));
                const EndOut e = end_pair<NS>(sctx, 2, 2 | code, l2, l3, l1, l0, sub2(l1, l0), 0ull, 0ull, 0ull);
                st.q[0] = l0; st.q[1] = l1; st.q[2] = l2;
                st.dq = e.dq; st.dc = e.dc; st.xp = e.x; st.mk = e.mk; st.run = e.run;
            } else if (chunk < hot_end) {
                const int t0 = chunk * KC;
                hot_step<NS, 0>(st, rec, load_q<0>(stage, fld, any_derived), acc_item, sctx, code, t0, chk);
                hot_step<NS, 1>(st, rec + REC_LEVEL, load_q<1>(stage, fld, any_derived), acc_item, sctx, code, t0 + 1, chk);
                hot_step<NS, 2>(st, rec + 2 * REC_LEVEL, load_q<2>(stage, fld, any_derived), acc_item, sctx, code, t0 + 2, chk);
            } else {
                // the last chunks: levels up to nz - 1, then t = nz, nz + 1 (the bottom layers) and nz + 2 (below the surface)
#pragma unroll
                for (int kk = 0; kk < KC; ++kk) {
                    const int t = chunk * KC + kk;
                    if (t > nz + 2) continue;
                    const u64 A = st.q[(kk + 2) % 3], B = st.q[(kk + 1) % 3];
                    u64 qn = A;
                    if (t < nz) qn = kk == 0 ? load_q<0>(stage, fld, any_derived) : kk == 1 ? load_q<1>(stage, fld, any_derived) : load_q<2>(stage, fld, any_derived);
                    if (t < nz) {
                        if (kk == 0) hot_step<NS, 0>(st, rec, qn, acc_item, sctx, code, t, chk);
                        else if (kk == 1) hot_step<NS, 1>(st, rec + REC_LEVEL, qn, acc_item, sctx, code, t, chk);
                        else hot_step<NS, 2>(st, rec + 2 * REC_LEVEL, qn, acc_item, sctx, code, t, chk);
                    } else {
                        const EndOut e = end_pair<NS>(sctx, t, kk | code, qn, 0ull, A, B, st.dq, st.dc, st.xp, st.run);
                        st.q[kk] = qn;
                        st.dq = e.dq; st.dc = e.dc; st.xp = e.x; st.mk = e.mk; st.run = e.run;
                    }
                }
            }
            // every load of the stage has been consumed before it is handed back
            asm volatile("" ::"l"(st.q[0]), "l"(st.q[1]), "l"(st.q[2]), "l"(st.dq), "l"(st.dc), "l"(st.xp), "l"(st.mk),
                         "l"(st.run), "l"(chk) : "memory");
            __syncwarp();
            if (warp == 0 || warp == nfw - 1) TRACE(warp == 0 ? 7 : 8, it, 3);
            if (it + NSTAGE < total) asm volatile("bar.arrive %0, %1;" ::"r"(BAR_STAGE0 + b), "r"(nthreads) : "memory");
        }
    }
    if (!NS) {
        float c0, c1;
        upk(chk, c0, c1);
        if (active && (c0 != c0 || c1 != c1)) p.flags[cta] = 1;
    }

    // coarse outputs: the four pairs of a cell in a fixed order, block sum / weight sum (coarsen.py:213);
    // 0/0 -> NaN like the reference
    asm volatile("bar.sync %0, %1;" ::"n"(BAR_OUT), "r"(nfw * 32) : "memory");
    const uint32_t acc0 = sbase + lay.acc;
    if (p.nrep == 0) {
        for (int idx = tid; idx < nz * NCELL * V; idx += nfw * 32) {
            const int c = idx % NCELL, vv = (idx / NCELL) % V, k = idx / (NCELL * V);
            const uint32_t a = acc0 + (k + 1) * lay.acc_stride + 4 * ((NCELL * c) * V + vv);
            float sum = lds32(a);
            sum += lds32(a + 4 * V);
            sum += lds32(a + 8 * V);
            sum += lds32(a + 12 * V);
            const float den = lds32(sbase + lay.cell0 + c * lay.cell_bytes + lay.den + 4 * k);
            p.outs[vv][(o * nz + k) * hc + (long)by * nxc + NCELL * bx + c] = sum / den;
        }
        return;
    }
    replicated_outputs(p, acc0, lay.acc_stride, sbase + lay.cell0 + lay.den, lay.cell_bytes, nz, V, nfw, tid, o, hc,
                       (long)by * nxc + NCELL * bx);
}

// ---- host side ----------------------------------------------------------------------------------
constexpr int SMEM_LIMIT = 227 * 1024 - 1024;

// one launch: nfields <= MAXV output fields
static int launch_group(const cg_field *fields, float *const *outs, int nfields, const float *delp, const float *area,
                        const float *delp_coarse, const float *psurf, int *flags, long n_outer, int nz, int ny, int nx,
                        int mask_below, float toa, int nansafe_only, int nrep, const long long *rep, int rep_dc,
                        int rep_kind, cudaStream_t st)
{
    static Params p;                       // ~8 KB: keep it off the stack; filled under a lock
    static std::mutex mu;
    std::lock_guard<std::mutex> lock(mu);
    const float *srcs[MAXSRC];
    int nsrc = 0, any_derived = 0;
    auto source_index = [&](const float *ptr) {
        for (int i = 0; i < nsrc; ++i)
            if (srcs[i] == ptr) return i;
        srcs[nsrc] = ptr;
        return nsrc++;
    };
    for (int v = 0; v < nfields; ++v) {
        p.fa[v] = source_index(fields[v].a);
        p.fb[v] = fields[v].op != CG_FIELD_PLAIN ? source_index(fields[v].b) : p.fa[v];
        p.fc1[v] = fields[v].op == CG_FIELD_DIFF ? 1.f : 0.f;
        p.fc2[v] = fields[v].op == CG_FIELD_PRODUCT ? 1.f : 0.f;
        p.fc3[v] = fields[v].op == CG_FIELD_DIFF ? 1.f / fields[v].scale : (fields[v].op == CG_FIELD_PRODUCT ? 0.f : 1.f);
        any_derived |= fields[v].op != CG_FIELD_PLAIN;
        p.outs[v] = outs[v];
    }
    p.any_derived = any_derived;
    const int nfw = (NPAIR * nfields + 31) / 32;
    const Layout lay = make_layout(nsrc, nz, nfw);
    if (lay.total > SMEM_LIMIT) return CG_ENOTSUP;
    for (int i = 0; i < nsrc; ++i) {
        if (((uintptr_t)srcs[i]) & 15) return CG_ENOTSUP;
        p.src[i] = srcs[i];
    }
    p.area = area; p.psurf = psurf; p.delp_coarse = delp_coarse; p.delp = delp;
    p.nsrc = nsrc; p.nfields = nfields; p.nz = nz; p.ny = ny; p.nx = nx; p.mask_below = mask_below; p.toa = toa;
    p.nfw = nfw;
    p.nrep = nrep; p.rep_dc = rep_dc; p.rep_kind = rep_kind;
    if (nrep > 0)                           // 16-byte stores of the four cells of a CTA
        for (int v = 0; v < nfields; ++v)
            if (((uintptr_t)outs[v]) & 15) return CG_ENOTSUP;
    for (int r = 0; r < CG_MAX_REPLICAS; ++r) p.rep[r] = r < nrep ? rep[r] : 0;
    // per device attributes: cheap, so set on every call (a process may drive several devices)
    for (int v = 0; v < 2; ++v) {
        const void *fn = v ? (const void *)plevel_pair_kernel<true> : (const void *)plevel_pair_kernel<false>;
        cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT);
        if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(plevel_pair)");
        e = cudaFuncSetAttribute(fn, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
        if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(carveout)");
    }
    dim3 grid(nx / 32, ny / 8, (unsigned)n_outer);
    const int nthreads = (nfw + NGW + 1) * 32;
    if (!nansafe_only) {
        p.flags = flags;
        plevel_pair_kernel<false><<<grid, nthreads, lay.total, st>>>(p);
        CG_LAUNCH_CHECK("cg_pressure_level_average(pair)");
    }
    p.flags = nansafe_only ? nullptr : flags;
    plevel_pair_kernel<true><<<grid, nthreads, lay.total, st>>>(p);
    CG_LAUNCH_CHECK("cg_pressure_level_average(pair, NaN-safe pass)");
    return CG_OK;
}

}  // namespace pairk

#ifdef CG_PAIR_TRACE
extern "C" int cg_debug_pair_trace(long long *dst, long n)
{
    const long have = pairk::TRACE_SLOTS * pairk::TRACE_STAGES * pairk::TRACE_VALS;
    return (int)cudaMemcpyFromSymbol(dst, pairk::g_trace, 8 * (n < have ? n : have));
}
#endif

// groups of at most 26 output fields, each as one launch of 16 V work items per CTA row
static void pair_groups(int nfields, int &ngroups, int &per)
{
    ngroups = (nfields + pairk::MAXV - 1) / pairk::MAXV;
    per = (nfields + ngroups - 1) / ngroups;
}

// shape constraints alone, with the smallest possible number of sources; launch_plevel_pair checks the real one
int pair_supported(int nfields, int nz, int ny, int nx, int f)
{
    if (!(f == 8 && nx % 32 == 0 && ny % 8 == 0 && nz >= 6 && nz <= 250 && nfields >= 1)) return 0;
    int ng, per;
    pair_groups(nfields, ng, per);
    const pairk::Layout lay = pairk::make_layout(1, nz, (pairk::NPAIR * per + 31) / 32);
    return lay.total <= pairk::SMEM_LIMIT;
}

static int distinct_sources(const cg_field *fields, int n)
{
    const float *srcs[pairk::MAXSRC];
    int nsrc = 0;
    for (int v = 0; v < n; ++v)
        for (int h = 0; h < (fields[v].op != CG_FIELD_PLAIN ? 2 : 1); ++h) {
            const float *ptr = h ? fields[v].b : fields[v].a;
            int i = 0;
            while (i < nsrc && srcs[i] != ptr) ++i;
            if (i == nsrc) srcs[nsrc++] = ptr;
        }
    return nsrc;
}

// fields/outs as in cg_pressure_level_average; delp_coarse (and psurf when mask_below) already computed
int launch_plevel_pair(const cg_field *fields, float *const *outs, int nfields, const float *delp,
                       const float *area, const float *delp_coarse, const float *psurf, int *flags, long n_outer,
                       int nz, int ny, int nx, int mask_below, float toa, int nansafe_only, int nrep, const long long *rep,
                       int rep_dc, int rep_kind, cudaStream_t st)
{
    if (n_outer > 65535 || n_outer * nz > 0x7fffffffL) return CG_ENOTSUP;
    if (((uintptr_t)delp) & 3) return CG_ENOTSUP;
    for (int v = 0; v < nfields; ++v)
        if ((((uintptr_t)fields[v].a) & 15) || (fields[v].op != CG_FIELD_PLAIN && (((uintptr_t)fields[v].b) & 15))) return CG_ENOTSUP;
    int ng, per;
    pair_groups(nfields, ng, per);
    for (int v0 = 0; v0 < nfields; v0 += per) {          // every group must fit before anything is launched
        const int n = nfields - v0 < per ? nfields - v0 : per;
        const pairk::Layout lay = pairk::make_layout(distinct_sources(fields + v0, n), nz, (pairk::NPAIR * n + 31) / 32);
        if (lay.total > pairk::SMEM_LIMIT) return CG_ENOTSUP;
    }
    for (int v0 = 0; v0 < nfields; v0 += per) {
        const int n = nfields - v0 < per ? nfields - v0 : per;
        // (delp_coarse goes to the replicas with the first group only)
        const int rc = pairk::launch_group(fields + v0, outs + v0, n, delp, area, delp_coarse, psurf, flags, n_outer, nz, ny,
                                           nx, mask_below, toa, nansafe_only, nrep, rep, rep_dc && v0 == 0, rep_kind, st);
        if (rc) return rc;
    }
    return CG_OK;
}

}  // namespace cg
