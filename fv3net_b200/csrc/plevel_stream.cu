// Streaming pressure-level coarse-graining for sm_100a: f = 8, iv = 1, kord = 1, nx % 32 == 0.
// (regridz.py:115-147 + coarsen.py:183, budgets.py:49-57; mappm.f90:10-126, :614-851, :854-931)
//
// Mapping ("lanes over fields").  A CTA owns 4 x-adjacent coarse cells; consumer warp w owns
// cell w for the whole launch and walks its 8 rows one after the other.  Inside a warp,
// LANE v IS FIELD v and every thread carries the 8 fine columns of the current row, so
//   * everything that depends only on pressure (the geometry: PPM stencil coefficients of
//     every layer, where each coarse interface cuts the fine column and with which weight)
//     is the same for all lanes of a warp.  It is built once per (column, level) by the
//     warp itself (lane = (column, level of the chunk)), parked in shared memory and read
//     back by all lanes with broadcast loads: no per-field integer work, no divergence;
//   * field data is fetched by TMA (cp.async.bulk.tensor, one [4 levels][32 columns] box
//     per source array and stage, 128B-swizzled) into a 2-stage ring shared by the 4
//     warps; a lane reads its own field's 8 columns with two conflict-free LDS.128;
//   * the vertical walk is a tandem walk: source layers in order (rolling PPM state in
//     registers), coarse interfaces met on the way close one target layer and open the
//     next ("events"); per-target sums live in a per-warp shared-memory accumulator
//     [target level][lane], so the 64 columns of a cell are summed by one warp in a
//     fixed order (bitwise reproducible) and no cross-warp reduction exists.
// Every fine input element is read from HBM exactly once per launch; products / storage
// terms of the budget workflow are formed from their parents in registers.
// The prepass (plevel.cu) supplies delp_coarse and, for mask_below, the fine surface pressure.
#include "common.cuh"
#include <cuda.h>
#include <cmath>
#include <cstring>
#include <mutex>
#include <vector>

namespace cg {
namespace stream {

constexpr int C = 8;                 // fine columns per thread (one row of a coarse cell)
constexpr int NCELL = 4;             // coarse cells per CTA = consumer warps
constexpr int KC = 4;                // levels per pipeline stage
constexpr int BOX_ROWS = KC + 1;     // 128-B rows per source box: odd stride -> 8 distinct swizzle phases over 8 lanes
constexpr int BOX_BYTES = BOX_ROWS * 128;
constexpr int NSTAGE = 2;            // == geometry buffers: both are released by the same barrier
constexpr int MAXF = CG_MAX_FIELDS;  // output fields per launch (lanes)
constexpr int MAXSRC = 2 * MAXF;     // distinct parent arrays
constexpr int EVCAP = 4;             // precomputed events per (column, layer); more are formed on the fly
constexpr int PRING = 16;            // levels kept in the interface-pressure ring
constexpr int NTHREADS = 2 * NCELL * 32;         // consumer warps + geometry warps (which also issue the TMA loads)
constexpr int CONSUMER_REGS = 168, HELPER_REGS = 88;   // setmaxnreg targets; launch value 128 x 256 threads
constexpr int REC_BYTES = C * KC * 32, EV_BYTES = C * KC * EVCAP * 16;
constexpr int DUMMY_SLOT = 0;        // accumulator slot k + 1 belongs to target level k

struct Params {
    CUtensorMap maps[MAXSRC];
    float *outs[MAXF];
    int fa[MAXF], fb[MAXF], fop[MAXF];
    float finv[MAXF];                // 1 / scale of CG_FIELD_DIFF
    const float *area;               // (n_outer, ny, nx)
    const float *psurf;              // (n_outer, ny, nx) sequential-cumsum surface pressure, mask_below only
    const float *delp_coarse;        // (n_outer, nz, ny/8, nx/8)
    const float *delp;               // (n_outer, nz, ny, nx), read by the geometry warps
    int nsrc;                        // TMA sources (field parents)
    int nfields, nz, ny, nx, mask_below;
    float toa;
};

// ---- shared-memory carve-up --------------------------------------------------------------------
struct Layout {
    int ring, bars, warp0, warp_bytes;
    int acc, rec, ev, pe1, pe2, rdp2, den, edge, colw, colps, nmax, total;
};

__host__ __device__ inline Layout make_layout(int nsrc, int nz, int V)
{
    Layout L;
    int o = 0;
    L.ring = o;  o += NSTAGE * nsrc * BOX_BYTES;
    L.bars = o;  o += 256;
    L.warp0 = o;
    int w = 0;
    L.acc = w;   w += (nz + 1) * V * 4;          w = (w + 15) & ~15;
    L.rec = w;   w += 2 * REC_BYTES;             // double-buffered between a geometry warp and its consumer
    L.ev = w;    w += 2 * EV_BYTES;
    L.nmax = w;  w += 2 * KC * 4;
    L.pe1 = w;   w += PRING * C * 4;
    L.pe2 = w;   w += (nz + 2) * 4;              w = (w + 15) & ~15;
    L.rdp2 = w;  w += (nz + 1) * 4;              w = (w + 15) & ~15;
    L.den = w;   w += (nz + 1) * 4;              w = (w + 15) & ~15;
    L.edge = w;  w += C * 24 * 4;
    L.colw = w;  w += 64 * 4;                    // 64 columns at cell start; [row parity][8] afterwards
    L.colps = w; w += 64 * 4;
    L.warp_bytes = (w + 127) & ~127;
    L.total = L.warp0 + NCELL * L.warp_bytes;
    return L;
}

// ---- PTX helpers ---------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, int count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, int bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, int parity)
{
    uint32_t done = 0;
    while (!done)
        asm volatile("{.reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, 0x989680; selp.u32 %0, 1, 0, p;}"
                     : "=r"(done) : "r"(bar), "r"(parity) : "memory");
}
// wait of a warp that is ahead of its partner: sleep between polls instead of spinning on the issue port
__device__ __forceinline__ void mbar_wait_relaxed(uint32_t bar, int parity)
{
    uint32_t done = 0;
    for (;;) {
        asm volatile("{.reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p;}"
                     : "=r"(done) : "r"(bar), "r"(parity) : "memory");
        if (done) break;
        __nanosleep(1500);
    }
}
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap *map, int x, int y, int z, uint32_t bar)
{
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
        ::"r"(dst), "l"(map), "r"(x), "r"(y), "r"(z), "r"(bar) : "memory");
}
__device__ __forceinline__ float4 lds128(uint32_t addr)
{
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
    return v;
}
__device__ __forceinline__ float lds32(uint32_t addr)
{
    float v;
    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr));
    return v;
}
__device__ __forceinline__ float max3(float a, float b, float c)
{
    float r;
    asm("max.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
    return r;
}
__device__ __forceinline__ float min3(float a, float b, float c)
{
    float r;
    asm("min.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
    return r;
}
// byte address of 16-B chunk `chunk` of the 128-B row starting at absolute shared address `row`
// under CU_TENSOR_MAP_SWIZZLE_128B (chunk index xor row index mod 8; verified by tools/tma_probe.cu)
__device__ __forceinline__ uint32_t swz(uint32_t row, int chunk)
{
    return row + ((((row >> 7) ^ (uint32_t)chunk) & 7u) << 4);
}

__device__ __forceinline__ void sts32(uint32_t addr, float v)
{
    asm volatile("st.shared.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory");
}
__device__ __forceinline__ void sts128(uint32_t addr, float4 v)
{
    asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

// ---- geometry: what the whole warp shares ------------------------------------------------------
// All of it is a function of pressure only (mappm.f90:41-46, :58-124, :651-683).  Shared-window
// byte addresses of the warp's private arrays (uniform across the warp):
struct WG {
    uint32_t acc, rec, ev, pe1, pe2, rdp2, den, edge, colw, colps, nmax;
};

__device__ __forceinline__ int count_le(uint32_t pe2, int n, float p)
{   // number of target interfaces i in [0, n) with pe2[i] <= p (pe2 ascending)
    int lo = 0, hi = n;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (lds32(pe2 + 4 * mid) <= p) lo = mid + 1; else hi = mid;
    }
    return lo;
}

// weight of fine column c in target level k (regridz.py:150-161 when mask_below)
__device__ __forceinline__ float level_weight(const WG &g, int c, int k, int mask_below)
{
    const float w = lds32(g.colw + 4 * c);
    if (!mask_below) return w;
    return lds32(g.pe2 + 4 * (k + 1)) < lds32(g.colps + 4 * c) ? w : 0.f;
}
// ... times 1 / thickness of that level
__device__ __forceinline__ float level_scale(const WG &g, int c, int k, int mask_below)
{
    return level_weight(g, c, k, mask_below) * lds32(g.rdp2 + 4 * k);
}

// Event of target interface i met in source layer L of column c (L == nz: below the surface).
// Record (x, y, z, w):  I = x*q + y*x' + z*y'  closes a target level: run + I goes to the accumulator
// slot held in the low 7 bits of z; the running sum of the next level restarts at -w * I.
__device__ __forceinline__ float4 event_record(const WG &g, int c, int L, int i, int nz, int mask_below, float toa)
{
    const int kn = nz;
    float4 e = make_float4(0.f, 0.f, 0.f, 0.f);
    int slot = DUMMY_SLOT;
    if (L < nz) {
        const float pt = lds32(g.pe1 + 4 * ((L & (PRING - 1)) * C + c));
        const float pb = lds32(g.pe1 + 4 * (((L + 1) & (PRING - 1)) * C + c));
        const float pi = lds32(g.pe2 + 4 * i);
        if (pi <= toa) {                                                // mappm.f90:62-64, only met in L == 0
            if (i < kn) { e.x = level_weight(g, c, i, mask_below); slot = i + 1; }
        } else {
            const float dpl = pb - pt;
            const float s = __fdividef(pi - pt, dpl);
            const int ke = i - 1;
            const bool end_ok = ke >= 0 && lds32(g.pe2 + 4 * (ke < 0 ? 0 : ke)) > toa;
            const bool start_ok = i < kn;
            const float sc_e = end_ok ? level_scale(g, c, ke, mask_below) : 0.f;
            const float sc_s = start_ok ? level_scale(g, c, i, mask_below) : 0.f;
            const float sc = end_ok ? sc_e : sc_s;
            const float a = sc * dpl * s, oms = 1.f - s;
            e.x = a;
            e.y = a * oms * oms;
            e.z = -a * s * oms;
            e.w = end_ok ? (sc_e != 0.f ? __fdividef(sc_s, sc_e) : 0.f) : 1.f;
            slot = end_ok ? ke + 1 : DUMMY_SLOT;
        }
    } else {                                                            // mappm.f90:65-67, :115-120
        const float psf = lds32(g.pe1 + 4 * ((nz & (PRING - 1)) * C + c));
        const int ke = i - 1;                                           // level whose bottom is interface i
        if (ke >= 0 && ke < kn) {
            const float top = lds32(g.pe2 + 4 * ke);
            if (top >= psf) { e.x = level_weight(g, c, ke, mask_below); slot = ke + 1; }
            else if (top > toa) { e.x = level_scale(g, c, ke, mask_below) * (lds32(g.pe2 + 4 * i) - psf); slot = ke + 1; }
        }
    }
    e.z = __int_as_float((__float_as_int(e.z) & ~127) | slot);
    return e;
}
// Build the records of steps t0 .. t0+KC-1; lane = (column c, step kk).  Record of step t:
//   r0 = (E1, E2, F-1, G) and r1.x = H: PPM stencil of layer t-1 (mappm.f90:651-683, delp-only factors)
//   r1.y = weight of the whole layer t-2 in the target level open at its bottom
//   r1.z = first interface | count << 8 of the events of layer t-2;  r1.w = 1 if it has none
// plus up to EVCAP event records of layer t-2 (the first one is a no-op when the layer has none).
__device__ __forceinline__ void build_chunk(const WG &g, float d /* delp(t0 + kk, column c) */, int lane, int t0, int nz,
                                            int mask_below, float toa)
{
    const int c = lane & 7, kk = lane >> 3, t = t0 + kk;
    // interface pressures by the reference's sequential cumsum (vertically_dependent.py:61-63):
    // the running sum is handed from level to level, same order of additions
    {
        if (t0 == 0 && kk == 0) sts32(g.pe1 + 4 * c, toa);
        const float pe0 = (t0 == 0) ? toa : lds32(g.pe1 + 4 * ((t0 & (PRING - 1)) * C + c));
        float pe = 0.f;
#pragma unroll
        for (int j = 0; j < KC; ++j) {
            const float prev = __shfl_up_sync(0xffffffffu, pe, 8);
            if (kk == j) pe = (j == 0 ? pe0 : prev) + d;
        }
        if (t < nz) sts32(g.pe1 + 4 * (((t + 1) & (PRING - 1)) * C + c), pe);
    }
    __syncwarp();
    auto P = [&](int m) { return lds32(g.pe1 + 4 * ((m & (PRING - 1)) * C + c)); };

    // (approximate division, 2 ulp: these factors multiply differences of q)
    float E1 = 0.f, E2 = 0.f, Fm = 0.f, G = 0.f, H = 0.f;
    const int j = t - 1;
    if (j >= 1 && j <= nz - 2) {
        const float p0 = P(j - 1), p1 = P(j), p2 = P(j + 1), p3 = P(j + 2);
        const float dm1 = p1 - p0, d0 = p2 - p1, dp1 = p3 - p2;        // mappm.f90:43 (from the rounded sums)
        const float d4 = dm1 + d0, d4p = d0 + dp1;
        const float r4 = __fdividef(1.f, d4), inv = __fdividef(d0, d4 + dp1);
        E1 = __fdividef(dm1 + 0.5f * d0, d4p) * inv;
        E2 = (dp1 + 0.5f * d0) * r4 * inv;
        if (j >= 2) {
            const float dm2 = p0 - P(j - 2), d4m = dm2 + dm1;
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
        const int m = top ? 0 : nz - 2;
        const float q0 = P(m), q1 = P(m + 1), q2 = P(m + 2);
        const float d1 = top ? q1 - q0 : q2 - q1, d2 = top ? q2 - q1 : q1 - q0;
        const float D = d2 * (2.f * d2 * d2 + d1 * (d2 + 3.f * d1));
        const float Pq = d2 * (5.f * d1 + d2) - 3.f * d1 * d1;
        const float Rr = d1 * d2 * (d2 + 3.f * d1);
        const float r12 = 1.f / (d1 + d2);
        const uint32_t e = g.edge + 4 * (c * 24 + (top ? 0 : 12));
        sts128(e, make_float4(d2 * r12, d1 * r12, 2.f * r12, (top ? 4.f : 1.f) / D));
        sts128(e + 16, make_float4(d2, (top ? 0.5f : 2.f) * Pq, (top ? 0.25f : 1.f) * Rr, (top ? 2.f : 8.f) * d1 * d1 * d1));
        sts32(e + 32, d1);
    }
    // events of layer L = t - 2 and its whole-layer weight.  Interfaces at or above the bottom of a
    // layer are counted once (binary search); the count of the layer above comes from the lane above.
    const int L = t - 2, kn = nz;
    float dpw = 0.f, pt = 0.f, pb = 0.f;
    int i1 = 0;
    if (L >= 0 && L < nz) { pt = P(L); pb = P(L + 1); i1 = count_le(g.pe2, kn + 1, pb); }
    else if (L == nz) i1 = kn + 1;
    int i0 = __shfl_up_sync(0xffffffffu, i1, 8);
    if (kk == 0) i0 = __float_as_int(lds32(g.edge + 4 * (c * 24 + 22)));       // last layer of the previous chunk
    if (L <= 0) i0 = 0;
    if (kk == KC - 1) sts32(g.edge + 4 * (c * 24 + 22), __int_as_float(i1));
    int n = (L >= 0 && L <= nz) ? i1 - i0 : 0;
    if (n < 0) n = 0;                                                   // NaN / non-monotone pressures
    if (L >= 0 && L < nz) {
        const int ko = i1 - 1;                                          // level open at the bottom of L
        if (ko >= 0 && ko < kn && lds32(g.pe2 + 4 * ko) > toa) dpw = level_scale(g, c, ko, mask_below) * (pb - pt);
    }
    const uint32_t evp = g.ev + 16 * ((c * KC + kk) * EVCAP);
    if (n == 0) sts128(evp, make_float4(0.f, 0.f, __int_as_float(DUMMY_SLOT), 0.f));
    const int nw = n < EVCAP ? n : EVCAP;
    for (int e = 0; e < nw; ++e) sts128(evp + 16 * e, event_record(g, c, L, i0 + e, nz, mask_below, toa));
    const uint32_t r = g.rec + 32 * (c * KC + kk);
    sts128(r, make_float4(E1, E2, Fm, G));
    sts128(r + 16, make_float4(H, dpw, __int_as_float(i0 | (n << 8)), n == 0 ? 1.f : 0.f));
    int nm = n;                                                         // most events of any column at this step
    nm = max(nm, __shfl_xor_sync(0xffffffffu, nm, 1));
    nm = max(nm, __shfl_xor_sync(0xffffffffu, nm, 2));
    nm = max(nm, __shfl_xor_sync(0xffffffffu, nm, 4));
    if (c == 0) sts32(g.nmax + 4 * kk, (float)nm);
    __syncwarp();
}

// ---- per-field work -----------------------------------------------------------------------------
// one event: close a target level with run + I (NaN-skipping block sum, coarsen.py:211), restart the
// running sum of the next one.  acc_lane = &acc[0][lane]; slot 0 collects what belongs to no level.
__device__ __forceinline__ void apply_event(const float4 ev, float q, float x, float y, float keep, float &run,
                                            uint32_t acc_lane, int V4, bool active)
{
    const float I = ev.x * q + ev.y * x + ev.z * y;
    const float tot = run + I;
    const uint32_t a = acc_lane + (__float_as_int(ev.z) & 127) * V4;
    const float upd = lds32(a) + (tot == tot ? tot : 0.f);
    asm volatile("{.reg .pred p; setp.ne.s32 p, %2, 0; @p st.shared.f32 [%0], %1;}" ::"r"(a), "f"(upd), "r"((int)active) : "memory");
    run = keep != 0.f ? run : -(ev.w * I);
}

// events 1 .. n-1 of a layer cut by several coarse interfaces (out of line: rare)
__device__ __noinline__ float extra_events(WG g, int c, int kk, int L, int evinfo, float q, float x, float y, float run,
                                           uint32_t acc_lane, int V4, bool active, int nz, int mask_below, float toa)
{
    const int n = evinfo >> 8, i0 = evinfo & 255;
    const uint32_t evp = g.ev + 16 * ((c * KC + kk) * EVCAP);
    for (int e = 1; e < n; ++e) {
        const float4 ev = e < EVCAP ? lds128(evp + 16 * e) : event_record(g, c, L, i0 + e, nz, mask_below, toa);
        apply_event(ev, q, x, y, 0.f, run, acc_lane, V4, active);
    }
    return run;
}

// Limited parabola of layer L in deviation form, x = AL - q, y = AR - q, A6 = -3 (x + y)
// (ppm_limiters lmt = 0, mappm.f90:875-894), then the events of the layer and its whole-layer term.
template <bool GENERAL>
__device__ __forceinline__ void finish_layer(const WG &g, int c, int kk, int L, float q, float x, float y, float dc,
                                             float &run, const float4 r1, const float4 ev0, uint32_t acc_lane, int V4,
                                             bool active, int nz, int mask_below, float toa)
{
    const float da1 = y - x, s = x + y;
    const float t = s * da1, da2 = da1 * (da1 * (1.f / 3.f));
    const float m2x = -2.f * x, m2y = -2.f * y;
    const bool c1 = t > da2, c2 = -t > da2, flat = dc == 0.f;
    y = c1 ? m2x : y;                  // a6da < -da2: AR = AL - 3 (AL - q)
    x = (c2 && !c1) ? m2y : x;         // a6da >  da2: AL = AR - 3 (AR - q)
    x = flat ? 0.f : x;
    y = flat ? 0.f : y;
    apply_event(ev0, q, x, y, r1.w, run, acc_lane, V4, active);
    if (GENERAL) {
        const int evinfo = __float_as_int(r1.z);
        if (evinfo >= (2 << 8))        // a layer cut by several coarse interfaces: the others out of line
            run = extra_events(g, c, kk, L, evinfo, q, x, y, run, acc_lane, V4, active, nz, mask_below, toa);
    }
    run = fmaf(r1.y, q, run);
}

// one-sided cubic at the column top (mappm.f90:689-707), evaluated at step t = 2 once level 3 is
// visible: returns (x, y, dc) of layer 0 and x of layer 1
__device__ __noinline__ float4 top_fix(uint32_t rec3, uint32_t edge, float qn, float dqn, float dcn, float q1, float q0, float q3)
{
    const float4 r0p = lds128(rec3);
    const float Hp = lds32(rec3 + 16);
    const float dq3 = q3 - qn;
    const float df2 = r0p.x * dq3 + r0p.y * dqn;
    const float na = -dqn;
    const float hi3 = max3(na, dq3, 0.f), nlo = min3(na, dq3, 0.f);
    const float dc2 = max3(min3(df2, hi3, -nlo), -hi3, nlo);
    const float al2 = qn + (r0p.z * dqn + r0p.w * dcn - Hp * dc2);
    const float4 ea = lds128(edge), eb = lds128(edge + 16);
    const float sd1 = lds32(edge + 32);
    const float qm = ea.x * q0 + ea.y * q1, dq = ea.z * (q1 - q0);
    const float c1 = (al2 - qm - eb.x * dq) * ea.w;
    const float c3 = dq - c1 * eb.y;
    const float inner = qm - c1 * eb.z;
    const float top = c1 * eb.w - c3 * sd1 + inner;
    const float al1 = fminf(fmaxf(inner, fminf(q0, q1)), fmaxf(q0, q1));
    return make_float4(top - q0, al1 - q0, 0.5f * (al1 - q0), al1 - q1);
}

// ... and at the bottom (:729-747), at step t = nz: returns y of layer nz-2 and (x, dc, y) of layer nz-1
__device__ __noinline__ float4 bottom_fix(uint32_t edge, float qe, float qi, float xi)
{
    const float4 ea = lds128(edge), eb = lds128(edge + 16);
    const float sd1 = lds32(edge + 32);
    const float ali = qi + xi;
    const float qm = ea.x * qe + ea.y * qi, dq = ea.z * (qi - qe);
    const float c1 = (ali - qm - eb.x * dq) * ea.w;
    const float c3 = dq - c1 * eb.y;
    const float inner = qm - c1 * eb.z;
    const float bot = c1 * eb.w - c3 * sd1 + inner;
    const float alb = fminf(fmaxf(inner, fminf(qe, qi)), fmaxf(qe, qi));
    return make_float4(alb - qi, alb - qe, 0.5f * (qe - alb), bot - qe);
}

// one 128-B row of a lane's own field: its 8 columns of the warp's cell
__device__ __forceinline__ void load_row(float (&qv)[C], uint32_t stage, int kk, int w, int fa, int fb, int fop, float finv)
{
    const uint32_t ra = stage + fa * BOX_BYTES + kk * 128;
    const float4 lo = lds128(swz(ra, 2 * w)), hi = lds128(swz(ra, 2 * w + 1));
    qv[0] = lo.x; qv[1] = lo.y; qv[2] = lo.z; qv[3] = lo.w;
    qv[4] = hi.x; qv[5] = hi.y; qv[6] = hi.z; qv[7] = hi.w;
    if (fop != CG_FIELD_PLAIN) {                  // budgets.py:151 products, :80 storage terms
        const uint32_t rb = stage + fb * BOX_BYTES + kk * 128;
        const float4 l2 = lds128(swz(rb, 2 * w)), h2 = lds128(swz(rb, 2 * w + 1));
        const float bv[C] = {l2.x, l2.y, l2.z, l2.w, h2.x, h2.y, h2.z, h2.w};
#pragma unroll
        for (int c = 0; c < C; ++c) qv[c] = (fop == CG_FIELD_PRODUCT) ? qv[c] * bv[c] : (qv[c] - bv[c]) * finv;
    }
}

__device__ __forceinline__ float limited_slope(float df2, float dq_lo, float dq_hi)
{   // mappm.f90:663-665: sign(min(|df2|, max(q-,q,q+) - q, q - min(q-,q,q+)), df2)
    const float na = -dq_lo;
    const float hi3 = max3(na, dq_hi, 0.f), nlo = min3(na, dq_hi, 0.f);
    return max3(min3(df2, hi3, -nlo), -hi3, nlo);
}


// One step of the tandem walk for the 8 columns of a row: level t arrives, layer t-2 is finished.
// Steps t < 2 finish nothing (their records hold a no-op event and zero weights); the steps at the ends
// of the column (t == 2, t >= nz) patch in the one-sided cubics.  GENERAL = false is the straight-line body
// for interior levels where no column is cut by more than one coarse interface (the 8 columns interleave).
template <bool GENERAL>
__device__ __forceinline__ void level_step(const WG &g, uint32_t stage, int t, int kk, int w, int fa, int fb, int fop,
                                           float finv, float (&A)[C], float (&B)[C], float (&dq1)[C], float (&dcp)[C],
                                           float (&xp)[C], float (&run)[C], uint32_t acc_lane, int V4,
                                           bool active, int nz, int mask_below, float toa)
{
    float qv[C], q3[C];
    const bool ends = GENERAL && ((t == 2) || (t >= nz));
    if (!GENERAL || t < nz) load_row(qv, stage, kk, w, fa, fb, fop, finv);
    else {
#pragma unroll
        for (int c = 0; c < C; ++c) qv[c] = A[c];
    }
    if (GENERAL && t == 2) load_row(q3, stage, 3, w, fa, fb, fop, finv);           // level 3 is in the same stage (KC = 4)
    const uint32_t rk = g.rec + 32 * kk, ek = g.ev + 16 * EVCAP * kk;
    float4 r0 = lds128(rk), r1 = lds128(rk + 16), e0 = lds128(ek);
#pragma unroll
    for (int c = 0; c < C; ++c) {
        float4 r0n = r0, r1n = r1, e0n = e0;
        if (c + 1 < C) {
            r0n = lds128(rk + 32 * KC * (c + 1));
            r1n = lds128(rk + 32 * KC * (c + 1) + 16);
            e0n = lds128(ek + 16 * EVCAP * KC * (c + 1));
        }
        const float qn = qv[c];
        float dqn = qn - A[c];                                              // delq(t)
        const float df2 = r0.x * dqn + r0.y * dq1[c];                       // layer t-1, mappm.f90:661
        float dcn = limited_slope(df2, dq1[c], dqn);                        // dc(t-1)
        float xn = r0.z * dq1[c] + r0.w * dcp[c] - r1.x * dcn;              // AL(t-1) - q(t-1), :674-683
        float x = xp[c], y = xn + dq1[c], dc = dcp[c];                      // layer t-2
        if (GENERAL && ends) {
            if (t == 2) {
                const float4 r = top_fix(g.rec + 32 * (c * KC + 3), g.edge + 4 * (c * 24), qn, dqn, dcn, A[c], B[c], q3[c]);
                x = r.x; y = r.y; dc = r.z; xn = r.w;
            } else if (t == nz) {
                const float4 r = bottom_fix(g.edge + 4 * (c * 24 + 12), A[c], B[c], xp[c]);
                y = r.x; xn = r.y; dcn = r.z; dqn = r.w;                    // layer nz-1 is finished next step
            } else if (t == nz + 1) {
                y = dq1[c];
            } else {                                                        // below the surface
                x = 0.f; y = 0.f; dc = 0.f;
            }
        }
        finish_layer<GENERAL>(g, c, kk, t - 2, B[c], x, y, dc, run[c], r1, e0, acc_lane, V4, active, nz, mask_below, toa);
        B[c] = A[c]; A[c] = qn; dq1[c] = dqn; dcp[c] = dcn; xp[c] = xn;
        r0 = r0n; r1 = r1n; e0 = e0n;
    }
}

// 128 registers at launch (2 CTAs of 256 threads per SM); the consumer warpgroup then takes 168 and the
// geometry warpgroup drops to 88 (setmaxnreg)
__global__ void __maxnreg__(128)
plevel_stream_kernel(const __grid_constant__ Params p)
{
    extern __shared__ __align__(1024) unsigned char smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int nz = p.nz, V = p.nfields, nsrc = p.nsrc;
    const Layout lay = make_layout(nsrc, nz, V);
    const uint32_t sbase = smem_u32(smem);
    const uint32_t ring = sbase + lay.ring;
    // barriers: full[NSTAGE], empty[NSTAGE] (field ring); gfull[NCELL][2], gempty[NCELL][2] (geometry buffers)
    const uint32_t bars = sbase + lay.bars;
    const uint32_t gbars = bars + 8 * 2 * NSTAGE;
    const int stage_bytes = nsrc * BOX_BYTES;
    const int nch = (nz + 3 + KC - 1) / KC;                     // steps t = 0 .. nz + 2
    const int bx = blockIdx.x, by = blockIdx.y;
    const long o = blockIdx.z;

    if (tid == 0) {
        for (int s = 0; s < NSTAGE; ++s) {
            mbar_init(bars + 8 * s, NCELL);                 // full: one expect_tx arrival per geometry warp
            mbar_init(bars + 8 * (NSTAGE + s), NCELL);      // empty: one arrival per consumer warp
        }
        for (int i = 0; i < 2 * NCELL; ++i) mbar_init(gbars + 8 * i, 1);
        asm volatile("fence.mbarrier_init.release.cluster;");
    }
    __syncthreads();

    const int w = warp & (NCELL - 1);                           // the cell this warp works for
    const uint32_t wb = sbase + lay.warp0 + w * lay.warp_bytes;
    WG g;
    g.acc = wb + lay.acc;   g.rec = wb + lay.rec;   g.ev = wb + lay.ev;     g.pe1 = wb + lay.pe1;
    g.pe2 = wb + lay.pe2;   g.rdp2 = wb + lay.rdp2; g.den = wb + lay.den;   g.edge = wb + lay.edge;
    g.colw = wb + lay.colw; g.colps = wb + lay.colps; g.nmax = wb + lay.nmax;
    const uint32_t gfull = gbars + 8 * (2 * w);
    const int ny = p.ny, nx = p.nx, nyc = ny >> 3, nxc = nx >> 3;
    const long h = (long)ny * nx, hc = (long)nyc * nxc;
    const int cx = 4 * bx + w;
    const float toa = p.toa;
    const int mask_below = p.mask_below;

    if (warp >= NCELL) {
        // ---------------- geometry warp of cell w: everything that depends on pressure only ---------
        asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(HELPER_REGS));
        // cell prologue: coarse interfaces (sequential cumsum, vertically_dependent.py:61-63), 1/thickness,
        // sum of the 64 columns' weights per target level (coarsen.py:212)
        for (int k = lane; k < nz; k += 32) sts32(g.rdp2 + 4 * k, p.delp_coarse[(o * nz + k) * hc + (long)by * nxc + cx]);
        for (int i = lane; i < 64; i += 32) {
            const long col = o * h + (long)(8 * by + (i >> 3)) * nx + 8 * cx + (i & 7);
            const float a = p.area[col];
            sts32(g.colw + 4 * i, (a == a) ? a : 0.f);
            sts32(g.colps + 4 * i, mask_below ? p.psurf[col] : 0.f);
        }
        __syncwarp();
        if (lane == 0) {
            float pe = toa;
            sts32(g.pe2, pe);
            for (int k = 0; k < nz; ++k) { pe = pe + lds32(g.rdp2 + 4 * k); sts32(g.pe2 + 4 * (k + 1), pe); }
        }
        __syncwarp();
        for (int k = lane; k < nz; k += 32) {
            const float lower = lds32(g.pe2 + 4 * (k + 1));
            sts32(g.rdp2 + 4 * k, 1.f / (lower - lds32(g.pe2 + 4 * k)));
            float d = 0.f;
            for (int i = 0; i < 64; ++i) d += (!mask_below || lower < lds32(g.colps + 4 * i)) ? lds32(g.colw + 4 * i) : 0.f;
            sts32(g.den + 4 * k, d);
        }
        __syncwarp();

        const int c = lane & 7, kk = lane >> 3;
        const float *dcol = p.delp + o * nz * h + (long)(8 * by) * nx + 8 * cx + c;
        auto load_delp = [&](int row, int chunk) {
            const int lvl = chunk * KC + kk;
            return lvl < nz ? ldg_stream(dcol + (long)lvl * h + (long)row * nx) : 0.f;
        };
        float dnext = load_delp(0, 0);
        for (int row = 0; row < 8; ++row) {
            // this row's 8 columns, double-buffered by row parity (the consumer may still be on the previous row)
            WG gr = g;
            gr.colw = g.colw + 32 * (row & 1);
            gr.colps = g.colps + 32 * (row & 1);
            if (lane < C) {
                const long col = o * h + (long)(8 * by + row) * nx + 8 * cx + lane;
                const float a = p.area[col];
                sts32(gr.colw + 4 * lane, (a == a) ? a : 0.f);
                sts32(gr.colps + 4 * lane, mask_below ? p.psurf[col] : 0.f);
            }
            __syncwarp();
            for (int chunk = 0; chunk < nch; ++chunk) {
                const int it = row * nch + chunk, b = it & 1;
                const float d = dnext;
                if (chunk + 1 < nch) dnext = load_delp(row, chunk + 1);
                else if (row + 1 < 8) dnext = load_delp(row + 1, 0);
                // stage b of the field ring and geometry buffer b are released together by the consumers
                if (it >= 2) mbar_wait_relaxed(bars + 8 * (NSTAGE + b), ((it >> 1) - 1) & 1);
                {   // this warp's share of the TMA boxes: parents w, w + 4, ...
                    const int mine = (nsrc - w + NCELL - 1) / NCELL;
                    if (lane == 0) mbar_expect_tx(bars + 8 * b, mine * KC * 128);
                    __syncwarp();
                    const int src = w + NCELL * lane;
                    if (lane < mine)
                        tma_load_3d(ring + b * stage_bytes + src * BOX_BYTES, &p.maps[src], 32 * bx, 8 * by + row,
                                    (int)(o * nz) + chunk * KC, bars + 8 * b);
                }
                WG gb = gr;
                gb.rec = g.rec + b * REC_BYTES;
                gb.ev = g.ev + b * EV_BYTES;
                gb.nmax = g.nmax + b * KC * 4;
                build_chunk(gb, d, lane, chunk * KC, nz, mask_below, toa);
                if (lane == 0) mbar_arrive(gfull + 8 * b);
            }
        }
        return;
    }

    // ---------------- consumer warp of cell (by, 4 bx + w): lanes = fields ----------------------------
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(CONSUMER_REGS));
    const bool active = lane < V;
    const int V4 = 4 * V;
    const uint32_t acc_lane = g.acc + 4 * lane;
    for (int i = lane; i < (nz + 1) * V; i += 32) sts32(g.acc + 4 * i, 0.f);
    __syncwarp();
    const int fa = active ? p.fa[lane] : 0, fb = active ? p.fb[lane] : 0, fop = active ? p.fop[lane] : 0;
    const float finv = active ? p.finv[lane] : 1.f;

    for (int row = 0; row < 8; ++row) {
        // rolling state of the 8 columns: A = q(t-1), B = q(t-2), dq1 = delq(t-1), dcp = dc(t-2), xp = x(t-2)
        float A[C], B[C], dq1[C], dcp[C], xp[C], run[C];
#pragma unroll
        for (int c = 0; c < C; ++c) { A[c] = 0.f; B[c] = 0.f; dq1[c] = 0.f; dcp[c] = 0.f; xp[c] = 0.f; run[c] = 0.f; }
        WG gr = g;
        gr.colw = g.colw + 32 * (row & 1);
        gr.colps = g.colps + 32 * (row & 1);

        for (int chunk = 0; chunk < nch; ++chunk) {
            const int it = row * nch + chunk, s = it % NSTAGE, b = it & 1, t0 = chunk * KC;
            mbar_wait(gfull + 8 * b, (it >> 1) & 1);
            mbar_wait(bars + 8 * s, (it / NSTAGE) & 1);
            const uint32_t stage = ring + s * stage_bytes;
            WG gb = gr;
            gb.rec = g.rec + b * REC_BYTES;
            gb.ev = g.ev + b * EV_BYTES;
            gb.nmax = g.nmax + b * KC * 4;
#pragma unroll 1
            for (int kk = 0; kk < KC; ++kk) {
                const int t = t0 + kk;
                if (t > nz + 2) break;
                if (t >= 3 && t < nz && lds32(gb.nmax + 4 * kk) <= 1.f)
                    level_step<false>(gb, stage, t, kk, w, fa, fb, fop, finv, A, B, dq1, dcp, xp, run, acc_lane, V4,
                                      active, nz, mask_below, toa);
                else
                    level_step<true>(gb, stage, t, kk, w, fa, fb, fop, finv, A, B, dq1, dcp, xp, run, acc_lane, V4,
                                     active, nz, mask_below, toa);
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(bars + 8 * (NSTAGE + s));
        }
    }

    // coarse outputs: block sum / weight sum (coarsen.py:213); 0/0 -> NaN like the reference
    __syncwarp();
    if (active) {
        float *out = p.outs[lane];
        for (int k = 0; k < nz; ++k)
            out[(o * nz + k) * hc + (long)by * nxc + cx] = lds32(acc_lane + (k + 1) * V4) / lds32(g.den + 4 * k);
    }
}

// fine surface pressure as regridz.py:127-147 sees it: the last interface of the sequential
// cumsum from toa (vertically_dependent.py:61-63), bitwise
__global__ void __launch_bounds__(256)
surface_pressure_seq_kernel(const float *__restrict__ delp, float *__restrict__ ps, float toa,
                            long n_outer, int nz, long h)
{
    const long total = n_outer * h;
    for (long idx = blockIdx.x * (long)blockDim.x + threadIdx.x; idx < total;
         idx += (long)gridDim.x * blockDim.x) {
        const long o = idx / h, i = idx - o * h;
        const float *d = delp + o * nz * h + i;
        float pe = toa;
        for (int k = 0; k < nz; ++k) pe = pe + ldg_stream(d + (long)k * h);
        ps[idx] = pe;
    }
}

// ---- host side ----------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn()
{
    static EncodeTiledFn fn = [] {
        void *f = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) != cudaSuccess) f = nullptr;
        return (EncodeTiledFn)f;
    }();
    return fn;
}

struct MapKey {
    const void *ptr;
    long n_outer;
    int nz, ny, nx;
    bool operator==(const MapKey &o) const
    { return ptr == o.ptr && n_outer == o.n_outer && nz == o.nz && ny == o.ny && nx == o.nx; }
};

// (n_outer, nz, ny, nx) fp32 as a 3-D tensor (x, y, outer*z); box = [KC][1][32], 128B swizzle
static int tensor_map_for(CUtensorMap *out, const float *ptr, long n_outer, int nz, int ny, int nx)
{
    static thread_local std::vector<std::pair<MapKey, CUtensorMap>> cache;
    const MapKey key{ptr, n_outer, nz, ny, nx};
    for (auto &e : cache)
        if (e.first == key) { *out = e.second; return CG_OK; }
    EncodeTiledFn enc = encode_fn();
    if (!enc) { set_error("cuTensorMapEncodeTiled is not available from this driver"); return CG_ECUDA; }
    const cuuint64_t dims[3] = {(cuuint64_t)nx, (cuuint64_t)ny, (cuuint64_t)(n_outer * nz)};
    const cuuint64_t strides[2] = {(cuuint64_t)nx * 4, (cuuint64_t)nx * ny * 4};
    const cuuint32_t box[3] = {32, 1, KC};
    const cuuint32_t es[3] = {1, 1, 1};
    CUtensorMap m;
    const CUresult r = enc(&m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, (void *)ptr, dims, strides, box, es,
                           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                           CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled failed with CUresult %d", (int)r); return CG_ECUDA; }
    if (cache.size() >= 256) cache.erase(cache.begin(), cache.begin() + 128);
    cache.emplace_back(key, m);
    *out = m;
    return CG_OK;
}

}  // namespace stream

int stream_supported(int nfields, int nz, int ny, int nx, int f)
{
    if (!(f == 8 && nx % 32 == 0 && ny % 8 == 0 && nz >= 6 && nz <= 126 && nfields >= 1)) return 0;
    // worst case: every field has two parents
    const stream::Layout lay = stream::make_layout(2 * nfields, nz, nfields);
    return lay.total <= 227 * 1024 - 1024;
}

long stream_workspace_bytes(long n_outer, int nz, int ny, int nx)
{
    const long h = (long)ny * nx, hc = (long)(ny / 8) * (nx / 8);
    return 4 * n_outer * nz * hc + 256 + 4 * n_outer * h + 256;
}

int launch_surface_pressure_seq(const float *delp, float *ps, float toa, long n_outer, int nz, long h, cudaStream_t st)
{
    long b = (n_outer * h + 255) / 256;
    const long cap = (long)sm_count() * 16;
    stream::surface_pressure_seq_kernel<<<(unsigned)(b > cap ? cap : (b < 1 ? 1 : b)), 256, 0, st>>>(delp, ps, toa, n_outer, nz, h);
    CG_LAUNCH_CHECK("surface_pressure_seq");
    return CG_OK;
}

// fields/outs as in cg_pressure_level_average; delp_coarse (and psurf when mask_below) already computed
int launch_plevel_stream(const cg_field *fields, float *const *outs, int nfields, const float *delp,
                         const float *area, const float *delp_coarse, const float *psurf, long n_outer,
                         int nz, int ny, int nx, int mask_below, float toa, cudaStream_t st)
{
    using namespace stream;
    if (n_outer > 65535 || n_outer * nz > 0x7fffffffL) return CG_ENOTSUP;
    static Params p;                       // ~10 KB: keep it off the stack; filled under a lock
    static std::mutex mu;
    std::lock_guard<std::mutex> lock(mu);
    const float *srcs[MAXSRC];
    int nsrc = 0;
    auto source_index = [&](const float *ptr) {
        for (int i = 0; i < nsrc; ++i)
            if (srcs[i] == ptr) return i;
        srcs[nsrc] = ptr;
        return nsrc++;
    };
    for (int v = 0; v < nfields; ++v) {
        p.fa[v] = source_index(fields[v].a);
        p.fb[v] = fields[v].op != CG_FIELD_PLAIN ? source_index(fields[v].b) : p.fa[v];
        p.fop[v] = fields[v].op;
        p.finv[v] = fields[v].op == CG_FIELD_DIFF ? 1.f / fields[v].scale : 1.f;
        p.outs[v] = outs[v];
    }
    if (((uintptr_t)delp) & 3) return CG_ENOTSUP;
    for (int i = 0; i < nsrc; ++i) {
        if (((uintptr_t)srcs[i]) & 15) return CG_ENOTSUP;
        const int rc = tensor_map_for(&p.maps[i], srcs[i], n_outer, nz, ny, nx);
        if (rc) return rc;
    }
    const Layout lay = make_layout(nsrc, nz, nfields);
    if (lay.total > 227 * 1024 - 1024) return CG_ENOTSUP;
    p.area = area; p.psurf = psurf; p.delp_coarse = delp_coarse; p.delp = delp;
    p.nsrc = nsrc; p.nfields = nfields; p.nz = nz; p.ny = ny; p.nx = nx; p.mask_below = mask_below; p.toa = toa;
    static int max_set = 0;
    if (lay.total > max_set) {
        cudaError_t e = cudaFuncSetAttribute(plevel_stream_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - 1024);
        if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(plevel_stream)");
        e = cudaFuncSetAttribute(plevel_stream_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
        if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(carveout)");
        max_set = 227 * 1024 - 1024;
    }
    dim3 grid(nx / 32, ny / 8, (unsigned)n_outer);
    plevel_stream_kernel<<<grid, NTHREADS, lay.total, st>>>(p);
    CG_LAUNCH_CHECK("cg_pressure_level_average(stream)");
    return CG_OK;
}

}  // namespace cg
