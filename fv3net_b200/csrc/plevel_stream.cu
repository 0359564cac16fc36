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
constexpr int PRING = 16;            // levels kept in the interface-pressure ring
constexpr int NTHREADS = 2 * NCELL * 32;         // consumer warps + geometry warps (which also issue the TMA loads)
constexpr int CONSUMER_REGS = 184, HELPER_REGS = 72;   // setmaxnreg targets; launch value 128 x 256 threads
// geometry record of one step and one PAIR of adjacent columns (a, b): 6 x 16 B, pair-interleaved so that
// the field warps load every coefficient as an f32x2 operand
//   +0  E1a E1b E2a E2b | +16 Fa Fb Ga Gb | +32 -Ha -Hb dpwa dpwb | +48 exa exb eya eyb
//   +64 eza ezb -ewa -ewb | +80 offa offb infoa infob   (off = accumulator byte offset, info = i0 | n << 8)
constexpr int REC_BLOCK = 96;
constexpr int REC_BYTES = KC * (C / 2) * REC_BLOCK;
constexpr int ACC_STRIDE = 128;      // bytes per accumulator slot: one float per lane
constexpr int DUMMY_SLOT = 0;        // accumulator slot k + 1 belongs to target level k

struct Params {
    CUtensorMap maps[MAXSRC];
    float *outs[MAXF];
    int fa[MAXF], fb[MAXF];          // TMA sources of the parents (fb = fa for a plain field)
    const float *area;               // (n_outer, ny, nx)
    const float *psurf;              // (n_outer, ny, nx) sequential-cumsum surface pressure, mask_below only
    const float *delp_coarse;        // (n_outer, nz, ny/8, nx/8)
    const float *delp;               // (n_outer, nz, ny, nx), read by the geometry warps
    float fc1[MAXF], fc2[MAXF], fc3[MAXF];   // q = (a - c1 b) (c2 b + c3): plain (0,0,1), product (0,1,0), diff (1,0,1/scale)
    int any_derived;                 // some field is a product or a difference
    int nsrc;                        // TMA sources (field parents)
    int nfields, nz, ny, nx, mask_below;
    float toa;
};

// ---- shared-memory carve-up --------------------------------------------------------------------
struct Layout {
    int ring, bars, warp0, warp_bytes;
    int acc, rec, pe1, pe2, rdp2, den, edge, colw, colps, nmax, total;
};

__host__ __device__ inline Layout make_layout(int nsrc, int nz, int V)
{
    Layout L;
    (void)V;
    int o = 0;
    L.ring = o;  o += NSTAGE * nsrc * BOX_BYTES;
    L.bars = o;  o += 256;
    L.warp0 = o;
    int w = 0;
    L.acc = w;   w += (nz + 1) * ACC_STRIDE;
    L.rec = w;   w += 2 * REC_BYTES;             // double-buffered between a geometry warp and its consumer
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

// ---- f32x2: two adjacent columns per instruction (FFMA2 / FADD2 / FMUL2: one issue slot, two results) ----
typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float lo, float hi)
{
    u64 r;
    asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void upk(u64 v, float &lo, float &hi)
{
    asm("mov.b64 {%0,%1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c)
{
    u64 r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
    return r;
}
__device__ __forceinline__ u64 add2(u64 a, u64 b)
{
    u64 r;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ u64 sub2(u64 a, u64 b)
{
    u64 r;
    asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ u64 mul2(u64 a, u64 b)
{
    u64 r;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ void lds2x64(uint32_t addr, u64 &a, u64 &b)
{
    asm volatile("ld.shared.v2.b64 {%0,%1}, [%2];" : "=l"(a), "=l"(b) : "r"(addr));
}
__device__ __forceinline__ int4 lds128i(uint32_t addr)
{
    int4 v;
    asm volatile("ld.shared.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
    return v;
}

// ---- geometry: what the whole warp shares ------------------------------------------------------
// All of it is a function of pressure only (mappm.f90:41-46, :58-124, :651-683).  Shared-window
// byte addresses of the warp's private arrays (uniform across the warp):
struct WG {
    uint32_t acc, rec, pe1, pe2, rdp2, den, edge, colw, colps, nmax;
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
// Record (x, y, z, w):  I = x*q + y*x' + z*y'  closes a target level: run + I goes to accumulator
// slot `slot`; the running sum of the next level restarts at -w * I.
__device__ __forceinline__ float4 event_record(const WG &g, int c, int L, int i, int nz, int mask_below, float toa,
                                               int &slot)
{
    const int kn = nz;
    float4 e = make_float4(0.f, 0.f, 0.f, 0.f);
    slot = DUMMY_SLOT;
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
    return e;
}
// Build the records of steps t0 .. t0+KC-1; lane = (column c, step kk).  Record of step t (REC_BLOCK):
//   E1, E2, F-1, G, H: PPM stencil of layer t-1 (mappm.f90:651-683, delp-only factors)
//   dpw: weight of the whole layer t-2 in the target level open at its bottom
//   ex, ey, ez, ew, off: the first event of layer t-2 (a no-op when it has none)
//   info = first interface | count << 8 of the events of layer t-2 (further events are formed on the fly)
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
    float4 e0 = make_float4(0.f, 0.f, 0.f, 0.f);
    int slot = DUMMY_SLOT;
    if (n > 0) e0 = event_record(g, c, L, i0, nz, mask_below, toa, slot);
    const uint32_t r = g.rec + (kk * (C / 2) + (c >> 1)) * REC_BLOCK + 4 * (c & 1);
    sts32(r, E1);           sts32(r + 8, E2);
    sts32(r + 16, Fm);      sts32(r + 24, G);
    sts32(r + 32, -H);      sts32(r + 40, dpw);
    sts32(r + 48, e0.x);    sts32(r + 56, e0.y);
    sts32(r + 64, e0.z);    sts32(r + 72, -e0.w);
    sts32(r + 80, __int_as_float(slot * ACC_STRIDE));
    sts32(r + 88, __int_as_float(i0 | (n << 8)));
    int nm = n;                                                         // most events of any column at this step
    nm = max(nm, __shfl_xor_sync(0xffffffffu, nm, 1));
    nm = max(nm, __shfl_xor_sync(0xffffffffu, nm, 2));
    nm = max(nm, __shfl_xor_sync(0xffffffffu, nm, 4));
    if (c == 0) sts32(g.nmax + 4 * kk, (float)nm);
    __syncwarp();
}

// ---- per-field work -----------------------------------------------------------------------------
// the record of one column at one step, gathered from its pair block (general step only)
struct Rec {
    float E1, E2, Fm, G, H, dpw;
    float4 ev;
    int off, info;
};
__device__ __forceinline__ Rec load_rec(uint32_t rec, int kk, int c)
{
    const uint32_t b = rec + (kk * (C / 2) + (c >> 1)) * REC_BLOCK + 4 * (c & 1);
    Rec r;
    r.E1 = lds32(b);        r.E2 = lds32(b + 8);
    r.Fm = lds32(b + 16);   r.G = lds32(b + 24);
    r.H = -lds32(b + 32);   r.dpw = lds32(b + 40);
    r.ev = make_float4(lds32(b + 48), lds32(b + 56), lds32(b + 64), -lds32(b + 72));
    r.off = __float_as_int(lds32(b + 80));
    r.info = __float_as_int(lds32(b + 88));
    return r;
}

// one event: close a target level with run + I (NaN-skipping block sum, coarsen.py:211), restart the
// running sum of the next one.  acc_lane = &acc[0][lane]; slot 0 collects what belongs to no level.
__device__ __forceinline__ void apply_event(const float4 ev, int off, float q, float x, float y, bool keep, float &run,
                                            uint32_t acc_lane)
{
    const float I = ev.x * q + ev.y * x + ev.z * y;
    const float tot = run + I;
    const uint32_t a = acc_lane + off;
    sts32(a, lds32(a) + (tot == tot ? tot : 0.f));
    run = keep ? run : -(ev.w * I);
}

// events 1 .. n-1 of a layer cut by several coarse interfaces (rare: formed on the fly, out of line)
__device__ __noinline__ float extra_events(WG g, int c, int L, int info, float q, float x, float y, float run,
                                           uint32_t acc_lane, int nz, int mask_below, float toa)
{
    const int n = info >> 8, i0 = info & 255;
    for (int e = 1; e < n; ++e) {
        int slot;
        const float4 ev = event_record(g, c, L, i0 + e, nz, mask_below, toa, slot);
        apply_event(ev, slot * ACC_STRIDE, q, x, y, false, run, acc_lane);
    }
    return run;
}

// Limited parabola of layer L in deviation form, x = AL - q, y = AR - q, A6 = -3 (x + y)
// (ppm_limiters lmt = 0, mappm.f90:875-894), then the events of the layer and its whole-layer term.
__device__ __forceinline__ void finish_layer(const WG &g, int c, int L, float q, float x, float y, float dc, float &run,
                                             const Rec &r, uint32_t acc_lane, int nz, int mask_below, float toa)
{
    const float da1 = y - x, s = x + y;
    const float t = s * da1, da2 = da1 * (da1 * (1.f / 3.f));
    const float m2x = -2.f * x, m2y = -2.f * y;
    const bool c1 = t > da2, c2 = -t > da2, flat = dc == 0.f;      // c1 and c2 exclude each other (da2 >= 0)
    y = c1 ? m2x : y;                  // a6da < -da2: AR = AL - 3 (AL - q)
    x = c2 ? m2y : x;                  // a6da >  da2: AL = AR - 3 (AR - q)
    x = flat ? 0.f : x;
    y = flat ? 0.f : y;
    apply_event(r.ev, r.off, q, x, y, r.info < 256, run, acc_lane);
    if (r.info >= (2 << 8))            // a layer cut by several coarse interfaces: the others out of line
        run = extra_events(g, c, L, r.info, q, x, y, run, acc_lane, nz, mask_below, toa);
    run = fmaf(r.dpw, q, run);
}

// one-sided cubic at the column top (mappm.f90:689-707), evaluated at step t = 2 once level 3 is
// visible: returns (x, y, dc) of layer 0 and x of layer 1
__device__ __noinline__ float4 top_fix(uint32_t rec, int c, uint32_t edge, float qn, float dqn, float dcn, float q1, float q0,
                                       float q3)
{
    const uint32_t b = rec + (3 * (C / 2) + (c >> 1)) * REC_BLOCK + 4 * (c & 1);      // record of step 3
    const float E1 = lds32(b), E2 = lds32(b + 8), Fm = lds32(b + 16), G = lds32(b + 24), nH = lds32(b + 32);
    const float dq3 = q3 - qn;
    const float df2 = E1 * dq3 + E2 * dqn;
    const float na = -dqn;
    const float hi3 = max3(na, dq3, 0.f), nlo = min3(na, dq3, 0.f);
    const float dc2 = max3(min3(df2, hi3, -nlo), -hi3, nlo);
    const float al2 = qn + (Fm * dqn + G * dcn + nH * dc2);
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

// the per-lane description of a field: parents' boxes and q = (a - c1 b) (c2 b + c3)
struct Field {
    int fa, fb;
    float c1, c2, c3;
};

// one 128-B row of a lane's own field: its 8 columns of the warp's cell, as 4 column pairs
// (budgets.py:151 products and :80 storage terms are formed here; exact for plain fields: (a - 0)(0 + 1))
__device__ __forceinline__ void load_row(u64 (&q)[4], uint32_t stage, int kk, int w, const Field &f, int any_derived)
{
    const uint32_t ra = stage + f.fa * BOX_BYTES + kk * 128;
    lds2x64(swz(ra, 2 * w), q[0], q[1]);
    lds2x64(swz(ra, 2 * w + 1), q[2], q[3]);
    if (any_derived) {                                   // uniform across the launch
        const uint32_t rb = stage + f.fb * BOX_BYTES + kk * 128;
        u64 b[4];
        lds2x64(swz(rb, 2 * w), b[0], b[1]);
        lds2x64(swz(rb, 2 * w + 1), b[2], b[3]);
        const u64 n1 = pk(-f.c1, -f.c1), m2 = pk(f.c2, f.c2), a3 = pk(f.c3, f.c3);
#pragma unroll
        for (int p = 0; p < 4; ++p) q[p] = mul2(fma2(b[p], n1, q[p]), fma2(b[p], m2, a3));
    }
}

__device__ __forceinline__ float limited_slope(float df2, float dq_lo, float dq_hi)
{   // mappm.f90:663-665: sign(min(|df2|, max(q-,q,q+) - q, q - min(q-,q,q+)), df2)
    const float na = -dq_lo;
    const float hi3 = max3(na, dq_hi, 0.f), nlo = min3(na, dq_hi, 0.f);
    return max3(min3(df2, hi3, -nlo), -hi3, nlo);
}

// Rolling state of a lane's 8 columns as 4 column pairs: A = q(t-1), B = q(t-2), dq = delq(t-1),
// dc = dc(t-2), xp = AL(t-2) - q(t-2), run = running integral of the target level that is open
struct State {
    u64 A[4], B[4], dq[4], dc[4], xp[4], run[4];
};

// One step of the tandem walk for the 8 columns of a row: level t arrives, layer t-2 is finished.
// Interior levels where no column is cut by more than one coarse interface: straight line, two columns
// per arithmetic instruction; the min/max, compare and select parts run on the halves.
__device__ __forceinline__ void hot_step(uint32_t rec, uint32_t stage, int kk, int w, const Field &f, int any_derived,
                                         State &s, uint32_t acc_lane)
{
    u64 q[4];
    load_row(q, stage, kk, w, f, any_derived);
    const uint32_t rb = rec + kk * (C / 2) * REC_BLOCK;
    const u64 third = pk(1.f / 3.f, 1.f / 3.f), neg2 = pk(-2.f, -2.f);
    u64 E1, E2, Fm, G, NH, DPW, EX, EY, EZ, NW;
    int4 oi;
    lds2x64(rb, E1, E2);        lds2x64(rb + 16, Fm, G);    lds2x64(rb + 32, NH, DPW);
    lds2x64(rb + 48, EX, EY);   lds2x64(rb + 64, EZ, NW);   oi = lds128i(rb + 80);
#pragma unroll
    for (int p = 0; p < 4; ++p) {
        u64 E1n = E1, E2n = E2, Fmn = Fm, Gn = G, NHn = NH, DPWn = DPW, EXn = EX, EYn = EY, EZn = EZ, NWn = NW;
        int4 oin = oi;
        if (p + 1 < 4) {
            const uint32_t rn = rb + (p + 1) * REC_BLOCK;
            lds2x64(rn, E1n, E2n);        lds2x64(rn + 16, Fmn, Gn);    lds2x64(rn + 32, NHn, DPWn);
            lds2x64(rn + 48, EXn, EYn);   lds2x64(rn + 64, EZn, NWn);   oin = lds128i(rn + 80);
        }
        const u64 qn = q[p], dq1 = s.dq[p];
        const u64 dqn = sub2(qn, s.A[p]);                                   // delq(t)
        const u64 df2 = fma2(E1, dqn, mul2(E2, dq1));                       // layer t-1, mappm.f90:661
        float dfa, dfb, d1a, d1b, dna, dnb;
        upk(df2, dfa, dfb); upk(dq1, d1a, d1b); upk(dqn, dna, dnb);
        const u64 dcn = pk(limited_slope(dfa, d1a, dna), limited_slope(dfb, d1b, dnb));    // dc(t-1)
        const u64 xn = fma2(NH, dcn, fma2(G, s.dc[p], mul2(Fm, dq1)));      // AL(t-1) - q(t-1), :674-683
        // layer t-2: limiter in deviation form (ppm_limiters lmt = 0)
        const u64 x = s.xp[p], y = add2(xn, dq1);
        const u64 da1 = sub2(y, x), sm = add2(x, y);
        const u64 tt = mul2(sm, da1), da2 = mul2(da1, mul2(da1, third));
        const u64 m2x = mul2(x, neg2), m2y = mul2(y, neg2);
        float xa, xb, ya, yb, ta, tb, ga, gb, mxa, mxb, mya, myb, ca, cb;
        upk(x, xa, xb); upk(y, ya, yb); upk(tt, ta, tb); upk(da2, ga, gb);
        upk(m2x, mxa, mxb); upk(m2y, mya, myb); upk(s.dc[p], ca, cb);
        ya = ta > ga ? mxa : ya;    xa = -ta > ga ? mya : xa;
        yb = tb > gb ? mxb : yb;    xb = -tb > gb ? myb : xb;
        xa = ca == 0.f ? 0.f : xa;  ya = ca == 0.f ? 0.f : ya;
        xb = cb == 0.f ? 0.f : xb;  yb = cb == 0.f ? 0.f : yb;
        const u64 xl = pk(xa, xb), yl = pk(ya, yb), qB = s.B[p];
        // the (at most one) event of each column: close a target level, restart the next
        const u64 I = fma2(EZ, yl, fma2(EY, xl, mul2(EX, qB)));
        const u64 tot = add2(s.run[p], I);
        float tota, totb;
        upk(tot, tota, totb);
        const uint32_t aa = acc_lane + oi.x, ab = acc_lane + oi.y;
        sts32(aa, lds32(aa) + (tota == tota ? tota : 0.f));
        sts32(ab, lds32(ab) + (totb == totb ? totb : 0.f));
        float ra, rb_, na, nb;
        upk(s.run[p], ra, rb_);
        upk(mul2(NW, I), na, nb);
        ra = oi.z < 256 ? ra : na;
        rb_ = oi.w < 256 ? rb_ : nb;
        s.run[p] = fma2(DPW, qB, pk(ra, rb_));
        s.B[p] = s.A[p]; s.A[p] = qn; s.dq[p] = dqn; s.dc[p] = dcn; s.xp[p] = xn;
        E1 = E1n; E2 = E2n; Fm = Fmn; G = Gn; NH = NHn; DPW = DPWn; EX = EXn; EY = EYn; EZ = EZn; NW = NWn;
        oi = oin;
    }
}

// The same step for the ends of the column (t < 3, t >= nz: one-sided cubics, nothing to finish, below the
// surface) and for levels where some column is cut by several coarse interfaces: column by column.
__device__ __forceinline__ void general_step(const WG &g, uint32_t stage, int t, int kk, int w, const Field &f,
                                             int any_derived, State &s, uint32_t acc_lane, int nz, int mask_below,
                                             float toa)
{
    float A[C], B[C], dq1[C], dcp[C], xp[C], run[C], qv[C], q3[C];
#pragma unroll
    for (int p = 0; p < 4; ++p) {
        upk(s.A[p], A[2 * p], A[2 * p + 1]);       upk(s.B[p], B[2 * p], B[2 * p + 1]);
        upk(s.dq[p], dq1[2 * p], dq1[2 * p + 1]);  upk(s.dc[p], dcp[2 * p], dcp[2 * p + 1]);
        upk(s.xp[p], xp[2 * p], xp[2 * p + 1]);    upk(s.run[p], run[2 * p], run[2 * p + 1]);
    }
    const bool ends = (t == 2) || (t >= nz);
    if (t < nz) {
        u64 q[4];
        load_row(q, stage, kk, w, f, any_derived);
#pragma unroll
        for (int p = 0; p < 4; ++p) upk(q[p], qv[2 * p], qv[2 * p + 1]);
    } else {
#pragma unroll
        for (int c = 0; c < C; ++c) qv[c] = A[c];
    }
    if (t == 2) {                                                   // level 3 is in the same stage (KC = 4)
        u64 q[4];
        load_row(q, stage, 3, w, f, any_derived);
#pragma unroll
        for (int p = 0; p < 4; ++p) upk(q[p], q3[2 * p], q3[2 * p + 1]);
    }
#pragma unroll
    for (int c = 0; c < C; ++c) {
        const Rec r = load_rec(g.rec, kk, c);
        const float qn = qv[c];
        float dqn = qn - A[c];                                              // delq(t)
        const float df2 = r.E1 * dqn + r.E2 * dq1[c];                       // layer t-1, mappm.f90:661
        float dcn = limited_slope(df2, dq1[c], dqn);                        // dc(t-1)
        float xn = r.Fm * dq1[c] + r.G * dcp[c] - r.H * dcn;                // AL(t-1) - q(t-1), :674-683
        float x = xp[c], y = xn + dq1[c], dc = dcp[c];                      // layer t-2
        if (ends) {
            if (t == 2) {
                const float4 v = top_fix(g.rec, c, g.edge + 4 * (c * 24), qn, dqn, dcn, A[c], B[c], q3[c]);
                x = v.x; y = v.y; dc = v.z; xn = v.w;
            } else if (t == nz) {
                const float4 v = bottom_fix(g.edge + 4 * (c * 24 + 12), A[c], B[c], xp[c]);
                y = v.x; xn = v.y; dcn = v.z; dqn = v.w;                    // layer nz-1 is finished next step
            } else if (t == nz + 1) {
                y = dq1[c];
            } else {                                                        // below the surface
                x = 0.f; y = 0.f; dc = 0.f;
            }
        }
        finish_layer(g, c, t - 2, B[c], x, y, dc, run[c], r, acc_lane, nz, mask_below, toa);
        B[c] = A[c]; A[c] = qn; dq1[c] = dqn; dcp[c] = dcn; xp[c] = xn;
    }
#pragma unroll
    for (int p = 0; p < 4; ++p) {
        s.A[p] = pk(A[2 * p], A[2 * p + 1]);       s.B[p] = pk(B[2 * p], B[2 * p + 1]);
        s.dq[p] = pk(dq1[2 * p], dq1[2 * p + 1]);  s.dc[p] = pk(dcp[2 * p], dcp[2 * p + 1]);
        s.xp[p] = pk(xp[2 * p], xp[2 * p + 1]);    s.run[p] = pk(run[2 * p], run[2 * p + 1]);
    }
}

// 128 registers at launch (2 CTAs of 256 threads per SM); the consumer warpgroup then takes 184 and the
// geometry warpgroup drops to 72 (setmaxnreg)
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
    g.acc = wb + lay.acc;   g.rec = wb + lay.rec;   g.pe1 = wb + lay.pe1;
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
    const uint32_t acc_lane = g.acc + 4 * lane;                 // every lane owns a column of the accumulator
    for (int i = lane; i < (nz + 1) * 32; i += 32) sts32(g.acc + 4 * i, 0.f);
    __syncwarp();
    Field fld;
    fld.fa = active ? p.fa[lane] : 0;
    fld.fb = active ? p.fb[lane] : 0;
    fld.c1 = active ? p.fc1[lane] : 0.f;
    fld.c2 = active ? p.fc2[lane] : 0.f;
    fld.c3 = active ? p.fc3[lane] : 1.f;
    const int any_derived = p.any_derived;

    for (int row = 0; row < 8; ++row) {
        State st;
#pragma unroll
        for (int i = 0; i < 4; ++i) { st.A[i] = 0; st.B[i] = 0; st.dq[i] = 0; st.dc[i] = 0; st.xp[i] = 0; st.run[i] = 0; }
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
            gb.nmax = g.nmax + b * KC * 4;
#pragma unroll 1
            for (int kk = 0; kk < KC; ++kk) {
                const int t = t0 + kk;
                if (t > nz + 2) break;
                if (t >= 3 && t < nz && lds32(gb.nmax + 4 * kk) <= 1.f)
                    hot_step(gb.rec, stage, kk, w, fld, any_derived, st, acc_lane);
                else
                    general_step(gb, stage, t, kk, w, fld, any_derived, st, acc_lane, nz, mask_below, toa);
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
            out[(o * nz + k) * hc + (long)by * nxc + cx] = lds32(acc_lane + (k + 1) * ACC_STRIDE) / lds32(g.den + 4 * k);
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
