// Streaming pressure-level coarse-graining for sm_100a: f = 8, iv = 1, kord = 1, nx % 32 == 0.
// (regridz.py:115-147 + coarsen.py:183, budgets.py:49-57; mappm.f90:10-126, :614-851, :854-931)
//
// A CTA owns 4 x-adjacent coarse cells = 32 fine columns x 8 rows x nz levels, walked row by row.
//
//   * Field warps ("lanes over fields"): warp w owns cell w for the whole launch.  LANE v IS
//     FIELD v and every thread carries the 8 fine columns of the current row, so everything
//     that depends only on pressure (the geometry: PPM stencil coefficients of every layer,
//     where each coarse interface cuts the fine column and with which weight) is the same
//     for all lanes.  It is read from shared memory with broadcast loads: no per-field integer
//     work, no divergence.  Field data arrives by TMA (cp.async.bulk.tensor, one
//     [4 levels][32 columns] box per source array and stage, 128B-swizzled) into a 2-stage
//     ring; a lane reads its own field's 8 columns with two conflict-free LDS.128.
//   * One geometry warp ("lanes over columns"): lane x walks fine column x of the row from
//     the top with its interface pressures, the two-pointer merge against the coarse
//     interfaces and the PPM stencil in registers, and writes one record per (column, level)
//     for the field warps; it also issues the TMA loads.
//   * The vertical walk is a tandem walk: source layers in order (rolling PPM state in
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
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <vector>

namespace cg {
namespace stream {

#ifndef STREAM_UNROLL
#define STREAM_UNROLL 1
#endif
constexpr int STEP_UNROLL = STREAM_UNROLL;   // steps of the walk per loop body
constexpr int C = 8;                 // fine columns per thread (one row of a coarse cell)
constexpr int NCELL = 4;             // coarse cells per CTA = field warps
constexpr int NCOL = NCELL * C;      // fine columns per CTA row = lanes of the geometry warp
constexpr int KC = 4;                // levels per pipeline stage
constexpr int BOX_ROWS = KC + 1;     // 128-B rows per source box: odd stride -> 8 distinct swizzle phases over 8 lanes
constexpr int BOX_BYTES = BOX_ROWS * 128;
constexpr int NSTAGE = 2;            // == geometry buffers: both are released by the same barrier
constexpr int MAXF = CG_MAX_FIELDS;  // output fields per launch (lanes)
constexpr int MAXSRC = 2 * MAXF;     // distinct parent arrays
constexpr int NTHREADS = 2 * NCELL * 32;         // field warpgroup + helper warpgroup (its first warp builds the geometry)
// setmaxnreg targets (launch value 128 x 256 threads; multiples of 8, sum 256).  Measured: 184 / 72 6.50 ms,
// 176 / 80 and 168 / 88 6.37, 160 / 96 6.49, 192 / 64 6.53 - the geometry warps' code wants its 80 registers
#ifndef CG_FIELD_REGS
#define CG_FIELD_REGS 176
#endif
constexpr int CONSUMER_REGS = CG_FIELD_REGS, HELPER_REGS = 256 - CG_FIELD_REGS;   // setmaxnreg targets; launch value 128 x 256 threads
// geometry record of one step and one PAIR of adjacent columns (a, b): 6 x 16 B, pair-interleaved so that
// the field warps load every coefficient as an f32x2 operand
//   +0  E1a E1b E2a E2b | +16 Fa Fb Ga Gb | +32 -Ha -Hb dpwa dpwb | +48 exa exb eya eyb
//   +64 eza ezb -ewa -ewb | +80 offa offb infoa infob
// off = accumulator byte offset of the level the first event closes; info = number of coarse interfaces in
// the layer (0: none, the running sum is kept; n >= 2: n - 1 further events sit in the overflow list;
// INFO_SLOW: more than MAXE of them, formed on the fly from the column's pressures parked in the list)
constexpr int REC_BLOCK = 96;
constexpr int REC_BYTES = KC * (NCOL / 2) * REC_BLOCK;
constexpr int MAXE = 2;              // listed further events per (column, layer)
constexpr int INFO_SLOW = 255;
constexpr int OVA_BYTES = KC * NCOL * MAXE * 16;     // (x, y, z, -w) per event
constexpr int OVB_BYTES = KC * NCOL * MAXE * 4;      // accumulator byte offset per event
constexpr int TAILMAX = 16;          // listed further events of the layer below the surface: (x, offset) per event
constexpr int TAIL_BYTES = NCOL * TAILMAX * 8;
// context block in shared memory for the out-of-line paths (word offsets)
enum { CTX_NZ, CTX_MASK, CTX_TOA, CTX_REC, CTX_OVA, CTX_OVB, CTX_EDGE, CTX_TAIL, CTX_CELL0, CTX_CELLB, CTX_ACC, CTX_PE2, CTX_RDP2,
       CTX_WORDS };
constexpr int ACC_STRIDE = 128;      // bytes per accumulator slot: one float per lane
constexpr int DUMMY_SLOT = 0;        // accumulator slot k + 1 belongs to target level k
constexpr int EDGE_WORDS = 24;       // per column: one-sided cubic coefficients, top (0..8) and bottom (12..20)

struct Params {
    CUtensorMap maps[MAXSRC];
    float *outs[MAXF];
    int fa[MAXF], fb[MAXF];          // TMA sources of the parents (fb = fa for a plain field)
    const float *area;               // (n_outer, ny, nx)
    const float *psurf;              // (n_outer, ny, nx) sequential-cumsum surface pressure, mask_below only
    const float *delp_coarse;        // (n_outer, nz, ny/8, nx/8)
    const float *delp;               // (n_outer, nz, ny, nx), read by the geometry warp
    float fc1[MAXF], fc2[MAXF], fc3[MAXF];   // q = (a - c1 b) (c2 b + c3): plain (0,0,1), product (0,1,0), diff (1,0,1/scale)
    int any_derived;                 // some field is a product or a difference
    int nsrc;                        // TMA sources (field parents)
    int nfields, nz, ny, nx, mask_below;
    float toa;
    int *flags;                      // per CTA: the fast variant met a non-finite field value (NaN-safe variant reruns the CTA)
};

// ---- shared-memory carve-up --------------------------------------------------------------------
struct Layout {
    int ring, bars, cell0, cell_bytes;          // per cell: acc, pe2, rdp2, den
    int acc, pe2, rdp2, den;
    int rec, ova, ovb, tail, ctx, edge, colw, colps, total;
};

__host__ __device__ inline Layout make_layout(int nsrc, int nz)
{
    Layout L;
    int o = 0;
    L.ring = o;  o += NSTAGE * nsrc * BOX_BYTES;
    L.bars = o;  o += 128;
    L.cell0 = o;
    int w = 0;
    L.acc = w;   w += (nz + 1) * ACC_STRIDE;
    L.pe2 = w;   w += (nz + 2) * 4;              w = (w + 15) & ~15;
    L.rdp2 = w;  w += (nz + 1) * 4;              w = (w + 15) & ~15;
    L.den = w;   w += (nz + 1) * 4;
    L.cell_bytes = (w + 127) & ~127;
    o += NCELL * L.cell_bytes;
    L.rec = o;   o += 2 * REC_BYTES;             // double-buffered between the geometry warp and the field warps
    L.ova = o;   o += 2 * OVA_BYTES;
    L.ovb = o;   o += 2 * OVB_BYTES;
    L.tail = o;  o += TAIL_BYTES;
    L.ctx = o;   o += 64;
    L.edge = o;  o += NCOL * EDGE_WORDS * 4;
    L.colw = o;  o += 8 * NCOL * 4;              // [row][column of the CTA]
    L.colps = o; o += 8 * NCOL * 4;
    L.total = o;
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
// acc[slot][lane] += v.  (red.shared.add.f32 compiles to a compare-and-swap loop on sm_100a: 9.2 ms against 6.7 ms.)
__device__ __forceinline__ void red_add(uint32_t addr, float v)
{
    float o;
    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(o) : "r"(addr));
    asm volatile("st.shared.f32 [%0], %1;" ::"r"(addr), "f"(o + v) : "memory");
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
// a value the compiler must keep in a register instead of recomputing it where it is used
__device__ __forceinline__ uint32_t opaque(uint32_t v)
{
    uint32_t r;
    asm volatile("mov.u32 %0, %1;" : "=r"(r) : "r"(v));
    return r;
}
__device__ __forceinline__ int ldsi32(uint32_t addr)
{
    int v;
    asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(addr));
    return v;
}
__device__ __forceinline__ void stsi32(uint32_t addr, int v)
{
    asm volatile("st.shared.b32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}

// ---- geometry: what all fields share -----------------------------------------------------------
// All of it is a function of pressure only (mappm.f90:41-46, :58-124, :651-683).
struct CellG {                       // shared-window byte addresses of one coarse cell's arrays
    uint32_t pe2, rdp2;              // coarse interfaces (nz + 1), 1 / thickness of every coarse level
};
struct ColG {                        // one fine column
    float w, ps;                     // block-average weight (NaN -> 0); fine surface pressure (mask_below)
};

// weight of a fine column in target level k (regridz.py:150-161 when mask_below)
__device__ __forceinline__ float level_weight(const CellG &g, const ColG &col, int k, int mask_below)
{
    if (!mask_below) return col.w;
    return lds32(g.pe2 + 4 * (k + 1)) < col.ps ? col.w : 0.f;
}
// ... times 1 / thickness of that level
__device__ __forceinline__ float level_scale(const CellG &g, const ColG &col, int k, int mask_below)
{
    return level_weight(g, col, k, mask_below) * lds32(g.rdp2 + 4 * k);
}

// Event of target interface i met in source layer L = [pt, pb] of a column (L == nz: below the
// surface, pt = surface pressure).  I = x*q + y*x' + z*y' closes a target level: run + I goes to
// accumulator slot `slot`; the running sum of the next level restarts at -w * I.
struct Event {
    float x, y, z, w;
    int slot;
};
__device__ __forceinline__ Event make_event(const CellG &g, const ColG &col, float pt, float pb, int L, int i, int nz,
                                            int mask_below, float toa)
{
    const int kn = nz;
    Event e = {0.f, 0.f, 0.f, 0.f, DUMMY_SLOT};
    if (L < nz) {
        const float pi = lds32(g.pe2 + 4 * i);
        if (pi <= toa) {                                                // mappm.f90:62-64, only met in L == 0
            if (i < kn) { e.x = level_weight(g, col, i, mask_below); e.slot = i + 1; }
        } else {
            const float dpl = pb - pt;
            const float s = __fdividef(pi - pt, dpl);
            const int ke = i - 1;
            const bool end_ok = ke >= 0 && lds32(g.pe2 + 4 * (ke < 0 ? 0 : ke)) > toa;
            const bool start_ok = i < kn;
            const float sc_e = end_ok ? level_scale(g, col, ke, mask_below) : 0.f;
            const float sc_s = start_ok ? level_scale(g, col, i, mask_below) : 0.f;
            const float sc = end_ok ? sc_e : sc_s;
            const float a = sc * dpl * s, oms = 1.f - s;
            e.x = a;
            e.y = a * oms * oms;
            e.z = -a * s * oms;
            e.w = end_ok ? (sc_e != 0.f ? __fdividef(sc_s, sc_e) : 0.f) : 1.f;
            e.slot = end_ok ? ke + 1 : DUMMY_SLOT;
        }
    } else {                                                            // mappm.f90:65-67, :115-120
        const float psf = pt;
        const int ke = i - 1;                                           // level whose bottom is interface i
        if (ke >= 0 && ke < kn) {
            const float top = lds32(g.pe2 + 4 * ke);
            if (top >= psf) { e.x = level_weight(g, col, ke, mask_below); e.slot = ke + 1; }
            else if (top > toa) { e.x = level_scale(g, col, ke, mask_below) * (lds32(g.pe2 + 4 * i) - psf); e.slot = ke + 1; }
        }
    }
    return e;
}

// Step t of column x (geometry warp g handles t = 4 * chunk + g): level t arrives.  P(t-3) .. P(t+1) are the
// column's interface pressures, i1 the coarse interfaces met so far.  Writes the column's half of the pair record:
//   E1, E2, F-1, G, H: PPM stencil of layer t-1 (mappm.f90:651-683, delp-only factors)
//   dpw: weight of the whole layer t-2 in the target level open at its bottom
//   ex, ey, ez, ew, off: the first event of layer t-2 (a no-op when it has none); further events go to the lists
template <bool NS>
__device__ __forceinline__ void geometry_step(int &icnt, float pm3, float pm2, float pm1, float p0, float pp1,
                                              const CellG &g, const ColG &col, int t, int kk, int x, uint32_t rec,
                                              uint32_t ova, uint32_t ovb, uint32_t tail, uint32_t edge, int nz,
                                              int mask_below, float toa)
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
    // events of layer L = t - 2: the coarse interfaces in (top, bottom] of the layer, by the two-pointer merge
    // (mappm.f90:70-75 keeps k0 the same way); L == nz is everything below the surface
    const int L = t - 2, kn = nz;
    const float pt = pm2, pb = pm1;
    int i0 = icnt, i1 = icnt, open_slot = DUMMY_SLOT;
    float dpw = 0.f;
    if (L >= 0 && L <= nz) {
        if (L == 0) i0 = 0;
        else {      // this warp last stood at the bottom of layer L - 4: usually three more interfaces since
            const int lo = i0;
            i0 = min(i0 + 3, kn + 1);
            while (i0 > lo && !(lds32(g.pe2 + 4 * (i0 - 1)) <= pt)) --i0;
            while (i0 <= kn && lds32(g.pe2 + 4 * i0) <= pt) ++i0;
        }
        i1 = i0;
        if (L < nz) {
            while (i1 <= kn && lds32(g.pe2 + 4 * i1) <= pb) ++i1;
            const int ko = i1 - 1;                                      // level open at the bottom of L
            if (ko >= 0 && ko < kn && lds32(g.pe2 + 4 * ko) > toa) {
                dpw = level_scale(g, col, ko, mask_below) * (pb - pt);
                open_slot = ko + 1;
            }
        } else {
            i1 = kn + 1;
        }
        icnt = i1;
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
            stsi32(a + 4, ev.slot * ACC_STRIDE);
        } else {
            sts128(oa + 16 * (e - 1), make_float4(ev.x, ev.y, ev.z, -ev.w));
            stsi32(ob + 4 * (e - 1), ev.slot * ACC_STRIDE);
        }
    }
    if (n - 1 > cap) {                                                  // park what the field warps need to form them
        sts128(oa, make_float4(pt, pb, col.w, col.ps));
        stsi32(ob, i0);
        stsi32(ob + 4, n);
        info = INFO_SLOW;
    }
    const uint32_t r = rec + (kk * (NCOL / 2) + (x >> 1)) * REC_BLOCK + 4 * (x & 1);
    sts32(r, E1);           sts32(r + 8, E2);
    sts32(r + 16, Fm);      sts32(r + 24, G);
    sts32(r + 32, -H);      sts32(r + 40, dpw);
    sts32(r + 48, e0.x);    sts32(r + 56, e0.y);
    sts32(r + 64, e0.z);    sts32(r + 72, -e0.w);
    stsi32(r + 80, e0.slot * ACC_STRIDE);
    stsi32(r + 88, info);
}

// Integral weights times (q, x, y) of one event.  Events that use the layer mean alone - a target interface at or
// above the column top (mappm.f90:62-64), everything below the surface (:115-120), the empty record - carry
// z = +0; an event inside a layer always carries z = -(a s (1 - s)) with the sign bit set, even when it is zero
// (make_event).  The NaN-safe variant drops the edge terms of the former, so that a NaN edge value (e.g. from a NaN
// two levels down, through the one-sided cubic) does not reach a level the Fortran fills from q alone, and keeps
// them for the latter, where the Fortran evaluates the parabola and returns NaN.
template <bool NS>
__device__ __forceinline__ float event_sum(float ex, float ey, float ez, float q, float x, float y)
{
    if (NS && __float_as_int(ez) == 0) return ex * q;
    return ex * q + ey * x + ez * y;
}
template <bool NS>
__device__ __forceinline__ u64 event_sum2(u64 EX, u64 EY, u64 EZ, u64 q, u64 x, u64 y)
{
    if (NS) {
        float eza, ezb, xa, xb, ya, yb;
        upk(EZ, eza, ezb); upk(x, xa, xb); upk(y, ya, yb);
        const bool ma = __float_as_int(eza) == 0, mb = __float_as_int(ezb) == 0;
        x = pk(ma ? 0.f : xa, mb ? 0.f : xb);
        y = pk(ma ? 0.f : ya, mb ? 0.f : yb);
    }
    return fma2(EZ, y, fma2(EY, x, mul2(EX, q)));
}

// ---- per-field work -----------------------------------------------------------------------------
// close a target level with `tot` (NaN-skipping block sum, coarsen.py:211).  acc_lane = &acc[0][lane];
// slot 0 collects what belongs to no level.
template <bool NS>
__device__ __forceinline__ void flush(uint32_t acc_lane, int off, float tot)
{
    red_add(acc_lane + off, (!NS || tot == tot) ? tot : 0.f);
}
__device__ __forceinline__ uint32_t ctx_u32(uint32_t sctx, int i) { return (uint32_t)ldsi32(sctx + 4 * i); }

// Events 1 .. n-1 of column c of the warp's cell at step t (kcb = kk | c << 4 | buffer << 8); r = restart value
// left by event 0.  Out of line: a layer cut by several coarse interfaces is the less common case.
template <bool NS>
__device__ __noinline__ float apply_extras(uint32_t sctx, int t, int kcb, int info, float q, float x, float y, float r)
{
    const int kk = kcb & 15, c = (kcb >> 4) & 15, b = kcb >> 8;
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int nz = ldsi32(sctx + 4 * CTX_NZ);
    const uint32_t cb = ctx_u32(sctx, CTX_CELL0) + w * ctx_u32(sctx, CTX_CELLB);
    const uint32_t acc_lane = cb + ctx_u32(sctx, CTX_ACC) + 4 * lane;
    const int xc = 8 * w + c;
    if (info != INFO_SLOW && t == nz + 2) {                             // below the surface: I = x * q, restart at 0
        const uint32_t a = ctx_u32(sctx, CTX_TAIL) + xc * (TAILMAX * 8);
        for (int e = 0; e + 2 <= info; ++e) {
            const float I = lds32(a + 8 * e) * q;
            flush<NS>(acc_lane, ldsi32(a + 8 * e + 4), r + I);
            r = 0.f * I;
        }
        return r;
    }
    const uint32_t oa = ctx_u32(sctx, CTX_OVA) + b * OVA_BYTES + (kk * NCOL + xc) * (MAXE * 16);
    const uint32_t ob = ctx_u32(sctx, CTX_OVB) + b * OVB_BYTES + (kk * NCOL + xc) * (MAXE * 4);
    if (info != INFO_SLOW) {
        for (int e = 0; e + 2 <= info; ++e) {
            const float4 ev = lds128(oa + 16 * e);
            const float I = event_sum<NS>(ev.x, ev.y, ev.z, q, x, y);
            flush<NS>(acc_lane, ldsi32(ob + 4 * e), r + I);
            r = ev.w * I;                                               // stored negated
        }
        return r;
    }
    // more than the lists hold (rare): formed on the fly from the column's pressures parked in the list
    const float4 pp = lds128(oa);                                       // pt, pb, weight, surface pressure
    const int i0 = ldsi32(ob), n = ldsi32(ob + 4);
    const int mask_below = ldsi32(sctx + 4 * CTX_MASK);
    const float toa = lds32(sctx + 4 * CTX_TOA);
    CellG g;
    g.pe2 = cb + ctx_u32(sctx, CTX_PE2);
    g.rdp2 = cb + ctx_u32(sctx, CTX_RDP2);
    ColG col;
    col.w = pp.z;
    col.ps = pp.w;
    for (int e = 1; e < n; ++e) {
        const Event ev = make_event(g, col, pp.x, pp.y, t - 2, i0 + e, nz, mask_below, toa);
        const float I = event_sum<NS>(ev.x, ev.y, ev.z, q, x, y);
        flush<NS>(acc_lane, ev.slot * ACC_STRIDE, r + I);
        r = -(ev.w * I);
    }
    return r;
}

__device__ __forceinline__ float limited_slope(float df2, float dq_lo, float dq_hi)
{   // mappm.f90:663-665: sign(min(|df2|, max(q-,q,q+) - q, q - min(q-,q,q+)), df2)
    const float na = -dq_lo;
    const float hi3 = max3(na, dq_hi, 0.f), nlo = min3(na, dq_hi, 0.f);
    return max3(min3(df2, hi3, -nlo), -hi3, nlo);
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

// Rolling state of a lane's 8 columns as 4 column pairs: A = q(t-1), B = q(t-2), dq = delq(t-1),
// dc = dc(t-2), xp = AL(t-2) - q(t-2), run = running integral of the target level that is open
struct State {
    u64 A[4], B[4], dq[4], dc[4], xp[4], run[4];
    u64 chk;                         // fast variant: 0 * every field value; NaN once a non-finite value went by
};

// Further coarse interfaces of a layer at an interior step.  oa / ob: the column's entries in the lists.
// Out of line (one small body instead of eight inline copies): about a quarter of the layers have a second one.
template <bool NS>
__device__ __noinline__ float further_events(uint32_t oa, uint32_t ob, int info, float q, float x, float y, float r,
                                             uint32_t acc_lane, uint32_t sctx, int t, int kcb)
{
    if (info > MAXE + 1) return apply_extras<NS>(sctx, t, kcb, info & 1023, q, x, y, r);    // INFO_SLOW, or the tail
    const float4 ev = lds128(oa);
    const int off = ldsi32(ob);
    float I = event_sum<NS>(ev.x, ev.y, ev.z, q, x, y);
    flush<NS>(acc_lane, off, r + I);
    r = ev.w * I;                                                           // stored negated
    if (info > 2) {
        const float4 e2 = lds128(oa + 16);
        const int o2 = ldsi32(ob + 4);
        I = event_sum<NS>(e2.x, e2.y, e2.z, q, x, y);
        flush<NS>(acc_lane, o2, r + I);
        r = e2.w * I;
    }
    return r;
}

// Both columns of pair p at step t when some column has more than one further event, or below the surface.
// code = kk | p << 2 | buffer << 4 | t << 5 | info_a << 13 | info_b << 21.  Out of line, rare: everything the
// events need is recomputed from the context block so that the call site stays a handful of instructions.
template <bool NS>
__device__ __noinline__ u64 rare_pair(uint32_t sctx, int code, u64 q, u64 x, u64 y, u64 r)
{
    const int kk = code & 3, p = (code >> 2) & 3, b = (code >> 4) & 1, t = (code >> 5) & 255;
    const int info[2] = {(code >> 13) & 255, (code >> 21) & 255};
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int nz = ldsi32(sctx + 4 * CTX_NZ);
    const uint32_t acc_lane = ctx_u32(sctx, CTX_CELL0) + w * ctx_u32(sctx, CTX_CELLB) + ctx_u32(sctx, CTX_ACC) + 4 * lane;
    const int big = t > nz ? 1024 : 0;
    float qv[2], xv[2], yv[2], rv[2];
    upk(q, qv[0], qv[1]); upk(x, xv[0], xv[1]); upk(y, yv[0], yv[1]); upk(r, rv[0], rv[1]);
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        if (info[h] <= 1) continue;
        const int c = 2 * p + h, xc = 8 * w + c;
        const uint32_t oa = ctx_u32(sctx, CTX_OVA) + b * OVA_BYTES + (kk * NCOL + xc) * (MAXE * 16);
        const uint32_t ob = ctx_u32(sctx, CTX_OVB) + b * OVB_BYTES + (kk * NCOL + xc) * (MAXE * 4);
        rv[h] = further_events<NS>(oa, ob, info[h] + big, qv[h], xv[h], yv[h], rv[h], acc_lane, sctx, t, kk | c << 4 | b << 8);
    }
    return pk(rv[0], rv[1]);
}

// One step of the tandem walk for the 8 columns of a row: level t arrives, layer t-2 is finished.
// Interior levels: two columns per arithmetic instruction; the min/max parts run on the halves.
struct PairRec {
    u64 E1, E2, Fm, G, NH, DPW, EX, EY, EZ, NW;
    int4 oi;
};
__device__ __forceinline__ void load_pair_rec(PairRec &r, uint32_t rb)
{
    lds2x64(rb, r.E1, r.E2);        lds2x64(rb + 16, r.Fm, r.G);    lds2x64(rb + 32, r.NH, r.DPW);
    lds2x64(rb + 48, r.EX, r.EY);   lds2x64(rb + 64, r.EZ, r.NW);   r.oi = lds128i(rb + 80);
}
struct PairFront {
    u64 dqn, dcn, xn, xl, yl, tot, rs;
};

template <bool NS>
__device__ __forceinline__ void pair_front(PairFront &o, const PairRec &r, const State &s, int p, u64 qn)
{
    const u64 third = pk(1.f / 3.f, 1.f / 3.f), neg2 = pk(-2.f, -2.f);
    const u64 dq1 = s.dq[p];
    const u64 dqn = sub2(qn, s.A[p]);                                   // delq(t)
    const u64 df2 = fma2(r.E1, dqn, mul2(r.E2, dq1));                   // layer t-1, mappm.f90:661
    float dfa, dfb, d1a, d1b, dna, dnb;
    upk(df2, dfa, dfb); upk(dq1, d1a, d1b); upk(dqn, dna, dnb);
    const u64 dcn = pk(limited_slope(dfa, d1a, dna), limited_slope(dfb, d1b, dnb));    // dc(t-1)
    const u64 xn = fma2(r.NH, dcn, fma2(r.G, s.dc[p], mul2(r.Fm, dq1)));    // AL(t-1) - q(t-1), :674-683
    // layer t-2: limiter in deviation form (ppm_limiters lmt = 0)
    const u64 x = s.xp[p], y = add2(xn, dq1);
    const u64 da1 = sub2(y, x);
    float xa, xb, ya, yb, ca, cb;
    upk(s.dc[p], ca, cb);
    if (NS) {
        const u64 sm = add2(x, y);
        const u64 tt = mul2(sm, da1), da2 = mul2(da1, mul2(da1, third));
        const u64 m2x = mul2(x, neg2), m2y = mul2(y, neg2);
        float ta, tb, ga, gb, mxa, mxb, mya, myb;
        upk(x, xa, xb); upk(y, ya, yb); upk(tt, ta, tb); upk(da2, ga, gb);
        upk(m2x, mxa, mxb); upk(m2y, mya, myb);
        ya = ta > ga ? mxa : ya;    xa = -ta > ga ? mya : xa;
        yb = tb > gb ? mxb : yb;    xb = -tb > gb ? myb : xb;
        xa = ca == 0.f ? 0.f : xa;  ya = ca == 0.f ? 0.f : ya;
        xb = cb == 0.f ? 0.f : xb;  yb = cb == 0.f ? 0.f : yb;
        o.xl = pk(xa, xb); o.yl = pk(ya, yb);
    } else {
        // the same limiter without compares: with X = sg x, Y = sg y, sg = sign(y - x), the two cases
        // a6da < -da2, a6da > da2 are (Y - X)(Y + 2X) > 0 and (X - Y)(X + 2Y) > 0 with Y >= X, i.e.
        // Y' = min(Y, -2X), X' = max(X, -2Y) (they exclude each other); flat layers through a 0 / 1 factor
        float da, db;
        upk(da1, da, db);
        const float sga = __int_as_float((__float_as_int(da) & 0x80000000) | 0x3f800000);
        const float sgb = __int_as_float((__float_as_int(db) & 0x80000000) | 0x3f800000);
        const u64 sg = pk(sga, sgb), sgm = mul2(sg, pk(ca != 0.f ? 1.f : 0.f, cb != 0.f ? 1.f : 0.f));
        const u64 X = mul2(sg, x), Y = mul2(sg, y);
        const u64 m2X = mul2(X, neg2), m2Y = mul2(Y, neg2);
        float Xa, Xb, Ya, Yb, mXa, mXb, mYa, mYb;
        upk(X, Xa, Xb); upk(Y, Ya, Yb); upk(m2X, mXa, mXb); upk(m2Y, mYa, mYb);
        const u64 Xl = pk(fmaxf(Xa, mYa), fmaxf(Xb, mYb)), Yl = pk(fminf(Ya, mXa), fminf(Yb, mXb));
        o.xl = mul2(sgm, Xl);
        o.yl = mul2(sgm, Yl);
    }
    // the first event of each column: close a target level, restart the next
    const u64 I = event_sum2<NS>(r.EX, r.EY, r.EZ, s.B[p], o.xl, o.yl);
    o.tot = add2(s.run[p], I);
    o.rs = mul2(r.NW, I);                                               // restart of the next level
    if (NS) {                                                           // no event: the running sum is kept
        float ra, rb_, na, nb;
        upk(s.run[p], ra, rb_);
        upk(o.rs, na, nb);
        o.rs = pk(r.oi.z == 0 ? ra : na, r.oi.w == 0 ? rb_ : nb);
    }
    o.dqn = dqn; o.dcn = dcn; o.xn = xn;
}

// Both columns of pair p at one of the end steps of the walk (t < 3: nothing to finish yet / one-sided cubic at the
// top, mappm.f90:689-707; t >= nz: one-sided cubic at the bottom, :729-747, and the layer below the surface), packed
// like the hot step.  State of the pair: A = q(t-1), B = q(t-2), dq1 = delq(t-1), dcp = dc(t-2), xp = AL(t-2) - q(t-2),
// run.  Returns (delq(t), dc(t-1), AL(t-1) - q(t-1), run).  Out of line: 3 of nz + 3 steps.
struct EndOut { u64 dq, dc, x, run; };
__device__ __forceinline__ u64 clamp2(u64 v, u64 a, u64 b)            // v between a and b, per half
{
    float va, vb, aa, ab, ba, bb;
    upk(v, va, vb); upk(a, aa, ab); upk(b, ba, bb);
    return pk(fminf(fmaxf(va, fminf(aa, ba)), fmaxf(aa, ba)), fminf(fmaxf(vb, fminf(ab, bb)), fmaxf(ab, bb)));
}
__device__ __forceinline__ u64 slope2(u64 df2, u64 dlo, u64 dhi)
{
    float fa, fb, la, lb, ha, hb;
    upk(df2, fa, fb); upk(dlo, la, lb); upk(dhi, ha, hb);
    return pk(limited_slope(fa, la, ha), limited_slope(fb, lb, hb));
}
template <bool NS>
__device__ __noinline__ EndOut end_pair(uint32_t sctx, int t, int kpb, u64 qn, u64 q3, u64 A, u64 B, u64 dq1, u64 dcp, u64 xp,
                                        u64 run)
{
    const int kk = kpb & 15, p = (kpb >> 4) & 3, b = kpb >> 8;
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int nz = ldsi32(sctx + 4 * CTX_NZ);
    const uint32_t rec = ctx_u32(sctx, CTX_REC) + b * REC_BYTES;
    PairRec r;
    load_pair_rec(r, rec + (kk * (NCOL / 2) + 4 * w + p) * REC_BLOCK);
    const u64 half = pk(0.5f, 0.5f);
    u64 dqn = sub2(qn, A);                                                  // delq(t)
    const u64 df2 = fma2(r.E1, dqn, mul2(r.E2, dq1));                       // layer t-1, mappm.f90:661
    u64 dcn = slope2(df2, dq1, dqn);                                        // dc(t-1)
    u64 xn = fma2(r.NH, dcn, fma2(r.G, dcp, mul2(r.Fm, dq1)));              // AL(t-1) - q(t-1), :674-683
    u64 x = xp, y = add2(xn, dq1), dc = dcp;                                // layer t-2
    const uint32_t edge = ctx_u32(sctx, CTX_EDGE) + 4 * ((4 * w + p) * (2 * EDGE_WORDS));
    if (t == 2) {
        // level 3 is visible (same stage): AL of layer 2 from the record of step 3, then the cubic through
        // the top two layers and that edge
        const uint32_t r3 = rec + (3 * (NCOL / 2) + 4 * w + p) * REC_BLOCK;
        u64 e1, e2, fm, gg, nh, unused;
        lds2x64(r3, e1, e2); lds2x64(r3 + 16, fm, gg); lds2x64(r3 + 32, nh, unused);
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
        y = dq1;
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
    const uint32_t acc_lane = ctx_u32(sctx, CTX_CELL0) + w * ctx_u32(sctx, CTX_CELLB) + ctx_u32(sctx, CTX_ACC) + 4 * lane;
    const u64 I = event_sum2<NS>(r.EX, r.EY, r.EZ, q, x, y);
    float ta_, tb_, ra, rb_, na, nb, qa, qb, xa, xb, ya, yb;
    upk(add2(run, I), ta_, tb_); upk(run, ra, rb_); upk(mul2(r.NW, I), na, nb);     // -w is what the record holds
    flush<NS>(acc_lane, r.oi.x, ta_);
    flush<NS>(acc_lane, r.oi.y, tb_);
    float rsa = (NS && r.oi.z == 0) ? ra : na, rsb = (NS && r.oi.w == 0) ? rb_ : nb;
    if ((r.oi.z | r.oi.w) > 1) {
        upk(q, qa, qb); upk(x, xa, xb); upk(y, ya, yb);
        if (r.oi.z > 1) rsa = apply_extras<NS>(sctx, t, kk | (2 * p) << 4 | b << 8, r.oi.z, qa, xa, ya, rsa);
        if (r.oi.w > 1) rsb = apply_extras<NS>(sctx, t, kk | (2 * p + 1) << 4 | b << 8, r.oi.w, qb, xb, yb, rsb);
    }
    EndOut o;
    o.dq = dqn; o.dc = dcn; o.x = xn;
    o.run = fma2(r.DPW, q, pk(rsa, rsb));
    return o;
}

// The same step as one straight-line block: fronts and accumulator updates of the four pairs without a branch
// (two columns closing the same level are added one after the other), then the further events of all pairs behind
// ONE warp-uniform branch, then the running sums.  Branches end scheduling regions: this leaves the compiler one long
// region to interleave the four dependency chains in.
template <bool NS>
__device__ __forceinline__ void close_levels(const PairFront &o, const int4 &oi, uint32_t acc_lane)
{
    float tota, totb;
    upk(o.tot, tota, totb);
    if (NS) {
        tota = tota == tota ? tota : 0.f;
        totb = totb == totb ? totb : 0.f;
    }
    const uint32_t aa = acc_lane + oi.x, ab = acc_lane + oi.y;
    red_add(aa, tota);
    red_add(ab, totb);
}
template <bool NS>
__device__ __forceinline__ u64 pair_extras(const PairFront &o, const int4 &oi, u64 qB, int p, uint32_t acc_lane, uint32_t ova,
                                           uint32_t ovb, uint32_t sctx, int t, int nz, int kk, int b)
{
    u64 rs = o.rs;
    if ((oi.z | oi.w) > 1) {
        const uint32_t oa = ova + (2 * p) * (MAXE * 16), ob = ovb + (2 * p) * (MAXE * 4);
        if (max(oi.z, oi.w) == 2 && t <= nz) {
            float qa, qb, ra, rb_, xa, xb, ya, yb;
            upk(qB, qa, qb); upk(rs, ra, rb_); upk(o.xl, xa, xb); upk(o.yl, ya, yb);
            if (oi.z == 2) {
                const float4 ev = lds128(oa);
                const float I = event_sum<NS>(ev.x, ev.y, ev.z, qa, xa, ya);
                flush<NS>(acc_lane, ldsi32(ob), ra + I);
                ra = ev.w * I;                                              // stored negated
            }
            if (oi.w == 2) {
                const float4 ev = lds128(oa + MAXE * 16);
                const float I = event_sum<NS>(ev.x, ev.y, ev.z, qb, xb, yb);
                flush<NS>(acc_lane, ldsi32(ob + MAXE * 4), rb_ + I);
                rb_ = ev.w * I;
            }
            rs = pk(ra, rb_);
        } else {
            rs = rare_pair<NS>(sctx, kk | p << 2 | b << 4 | t << 5 | oi.z << 13 | oi.w << 21, qB, o.xl, o.yl, rs);
        }
    }
    return rs;
}
template <bool NS>
__device__ __forceinline__ void hot_step2(uint32_t sctx, uint32_t rec, uint32_t ova, uint32_t ovb, uint32_t stage, int t, int nz, int kk, int b,
                                          int w, const Field &f, int any_derived, State &s, uint32_t acc_lane)
{
    u64 q[4];
    load_row(q, stage, kk, w, f, any_derived);
    if (t >= nz) {
#pragma unroll
        for (int p = 0; p < 4; ++p) q[p] = s.A[p];
    }
    PairRec r0, r1, r2, r3;
    PairFront f0, f1, f2, f3;
    load_pair_rec(r0, rec);
    load_pair_rec(r1, rec + REC_BLOCK);
    pair_front<NS>(f0, r0, s, 0, q[0]);
    pair_front<NS>(f1, r1, s, 1, q[1]);
    load_pair_rec(r2, rec + 2 * REC_BLOCK);
    load_pair_rec(r3, rec + 3 * REC_BLOCK);
    close_levels<NS>(f0, r0.oi, acc_lane);
    close_levels<NS>(f1, r1.oi, acc_lane);
    pair_front<NS>(f2, r2, s, 2, q[2]);
    pair_front<NS>(f3, r3, s, 3, q[3]);
    close_levels<NS>(f2, r2.oi, acc_lane);
    close_levels<NS>(f3, r3.oi, acc_lane);
    u64 rs0 = f0.rs, rs1 = f1.rs, rs2 = f2.rs, rs3 = f3.rs;
    if ((r0.oi.z | r0.oi.w | r1.oi.z | r1.oi.w | r2.oi.z | r2.oi.w | r3.oi.z | r3.oi.w) > 1) {
        rs0 = pair_extras<NS>(f0, r0.oi, s.B[0], 0, acc_lane, ova, ovb, sctx, t, nz, kk, b);
        rs1 = pair_extras<NS>(f1, r1.oi, s.B[1], 1, acc_lane, ova, ovb, sctx, t, nz, kk, b);
        rs2 = pair_extras<NS>(f2, r2.oi, s.B[2], 2, acc_lane, ova, ovb, sctx, t, nz, kk, b);
        rs3 = pair_extras<NS>(f3, r3.oi, s.B[3], 3, acc_lane, ova, ovb, sctx, t, nz, kk, b);
    }
    s.run[0] = fma2(r0.DPW, s.B[0], rs0);   s.run[1] = fma2(r1.DPW, s.B[1], rs1);
    s.run[2] = fma2(r2.DPW, s.B[2], rs2);   s.run[3] = fma2(r3.DPW, s.B[3], rs3);
    const PairFront *fr[4] = {&f0, &f1, &f2, &f3};
#pragma unroll
    for (int p = 0; p < 4; ++p) {
        if (!NS) s.chk = fma2(q[p], 0ull, s.chk);
        s.B[p] = s.A[p]; s.A[p] = q[p]; s.dq[p] = fr[p]->dqn; s.dc[p] = fr[p]->dcn; s.xp[p] = fr[p]->xn;
    }
}

// The same step at the ends of the column: pair by pair through end_pair
template <bool NS>
__device__ __forceinline__ void end_step(uint32_t sctx, uint32_t stage, int t, int kk, int b, int w, int nz, const Field &f,
                                         int any_derived, State &s)
{
    u64 q[4], q3[4];
#pragma unroll
    for (int p = 0; p < 4; ++p) { q[p] = s.A[p]; q3[p] = 0; }
    if (t < nz) load_row(q, stage, kk, w, f, any_derived);
    if (t == 2) {
        // the walk starts here: levels 0 and 1 (same stage) only seed the rolling state - B = q(0), A = q(1),
        // delq(1), everything else zero - so steps 0 and 1 are not run; level 3 is visible too (KC = 4)
        u64 l0[4], l1[4];
        load_row(l0, stage, 0, w, f, any_derived);
        load_row(l1, stage, 1, w, f, any_derived);
        load_row(q3, stage, 3, w, f, any_derived);
#pragma unroll
        for (int p = 0; p < 4; ++p) {
            s.B[p] = l0[p]; s.A[p] = l1[p]; s.dq[p] = sub2(l1[p], l0[p]);
            if (!NS) s.chk = fma2(l1[p], 0ull, fma2(l0[p], 0ull, s.chk));
        }
    }
    if (!NS && t < nz) {
#pragma unroll
        for (int p = 0; p < 4; ++p) s.chk = fma2(q[p], 0ull, s.chk);
    }
#pragma unroll
    for (int p = 0; p < 4; ++p) {
        const EndOut v = end_pair<NS>(sctx, t, kk | p << 4 | b << 8, q[p], q3[p], s.A[p], s.B[p], s.dq[p], s.dc[p], s.xp[p], s.run[p]);
        s.B[p] = s.A[p]; s.A[p] = q[p];
        s.dq[p] = v.dq; s.dc[p] = v.dc; s.xp[p] = v.x; s.run[p] = v.run;
    }
}

// 128 registers at launch (2 CTAs of 256 threads per SM); the field warpgroup then takes 176 and the
// geometry warpgroup drops to 80 (setmaxnreg).
// Two variants.  NS = false (fast): assumes every field value is finite, which lets the block sums drop their
// NaN tests and selects; it checks that assumption (0 * q summed over everything it reads) and flags the CTA
// otherwise.  NS = true: NaN-skipping sums exactly like the reference's nansum; launched after the fast one, its
// CTAs return at once unless flagged (p.flags == nullptr: runs everywhere, CG_PLEVEL_NANSAFE=1).
template <bool NS>
__global__ void __maxnreg__(128)
plevel_stream_kernel(const __grid_constant__ Params p)
{
    const int cta = (blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
    if (NS && p.flags && p.flags[cta] == 0) return;
    extern __shared__ __align__(1024) unsigned char smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int nz = p.nz, V = p.nfields, nsrc = p.nsrc;
    const Layout lay = make_layout(nsrc, nz);
    const uint32_t sbase = smem_u32(smem);
    const uint32_t ring = sbase + lay.ring;
    // mbarriers: full[NSTAGE] (field ring), gfull[2] (geometry buffers); stage / buffer b is handed back to the
    // geometry warps through named barrier 2 + b (bar.arrive by the field warps, bar.sync by the geometry warps:
    // a blocked warp does not poll)
    const uint32_t bars = sbase + lay.bars;
    const uint32_t gfull = bars + 8 * 2 * NSTAGE;
    const uint32_t sctx = sbase + lay.ctx;
    const int stage_bytes = nsrc * BOX_BYTES;
    const int nch = (nz + 3 + KC - 1) / KC;                     // steps t = 0 .. nz + 2
    const int bx = blockIdx.x, by = blockIdx.y;
    const long o = blockIdx.z;
    const float toa = p.toa;
    const int mask_below = p.mask_below;
    const uint32_t rec0 = sbase + lay.rec, ova0 = sbase + lay.ova, ovb0 = sbase + lay.ovb, edge = sbase + lay.edge;
    const uint32_t tail = sbase + lay.tail;

    if (tid == 0) {
        if (!NS) p.flags[cta] = 0;
        for (int s = 0; s < NSTAGE; ++s) {
            mbar_init(bars + 8 * s, NCELL);                 // full: one expect_tx arrival per geometry warp
            mbar_init(gfull + 8 * s, NCELL);                // geometry buffer: one arrival per geometry warp
        }
        asm volatile("fence.mbarrier_init.release.cluster;");
        stsi32(sctx + 4 * CTX_NZ, nz);              stsi32(sctx + 4 * CTX_MASK, mask_below);
        sts32(sctx + 4 * CTX_TOA, toa);             stsi32(sctx + 4 * CTX_REC, (int)rec0);
        stsi32(sctx + 4 * CTX_OVA, (int)ova0);      stsi32(sctx + 4 * CTX_OVB, (int)ovb0);
        stsi32(sctx + 4 * CTX_EDGE, (int)edge);     stsi32(sctx + 4 * CTX_TAIL, (int)tail);
        stsi32(sctx + 4 * CTX_CELL0, (int)(sbase + lay.cell0));
        stsi32(sctx + 4 * CTX_CELLB, lay.cell_bytes);
        stsi32(sctx + 4 * CTX_ACC, lay.acc);        stsi32(sctx + 4 * CTX_PE2, lay.pe2);
        stsi32(sctx + 4 * CTX_RDP2, lay.rdp2);
    }
    __syncthreads();

    const int ny = p.ny, nx = p.nx, nyc = ny >> 3, nxc = nx >> 3;
    const long h = (long)ny * nx, hc = (long)nyc * nxc;

    if (warp >= NCELL) {
        // ---------------- geometry warp gw: lane = fine column x of the CTA's 32, level 4 * chunk + gw ----
        asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(HELPER_REGS));
        const int gw = warp - NCELL;
        const int x = lane, cell = x >> 3;
        const uint32_t colw = sbase + lay.colw, colps = sbase + lay.colps;
        // CTA prologue, warp gw for cell gw: coarse interfaces (sequential cumsum, vertically_dependent.py:61-63),
        // 1 / thickness, sum of the 64 columns' weights per target level (coarsen.py:212)
        {
            const uint32_t cbw = sbase + lay.cell0 + gw * lay.cell_bytes;
            for (int k = lane; k < nz; k += 32)
                sts32(cbw + lay.rdp2 + 4 * k, p.delp_coarse[(o * nz + k) * hc + (long)by * nxc + NCELL * bx + gw]);
            for (int r = 2 * gw; r < 2 * gw + 2; ++r) {
                const long col = o * h + (long)(8 * by + r) * nx + NCOL * bx + x;
                const float a = p.area[col];
                sts32(colw + 4 * (r * NCOL + x), (a == a) ? a : 0.f);
                sts32(colps + 4 * (r * NCOL + x), mask_below ? p.psurf[col] : 0.f);
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
            asm volatile("bar.sync 1, 128;" ::: "memory");          // the geometry warpgroup: all cells' arrays are in place
            // the cell's 64 columns: rows lane / 8 and 4 + lane / 8, column lane % 8
            const int xc = 8 * gw + (lane & 7), r0 = lane >> 3;
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
        CellG g;
        g.pe2 = sbase + lay.cell0 + cell * lay.cell_bytes + lay.pe2;
        g.rdp2 = sbase + lay.cell0 + cell * lay.cell_bytes + lay.rdp2;
        const float *dcol = p.delp + o * nz * h + (long)(8 * by) * nx + NCOL * bx + x;
        float dn[KC];
        const float *dptr = dcol + (long)(gw - 3) * h;               // level t - 3 of the chunk to load, t = 4 chunk + gw
        int dlvl = gw - 3;
        auto load_delp = [&]() {                                     // thicknesses of levels t-3 .. t, then on to the next chunk
#pragma unroll
            for (int i = 0; i < KC; ++i) {
                const int lvl = dlvl + i;
                dn[i] = (lvl >= 0 && lvl < nz) ? ldg_stream(dptr + i * h) : 0.f;
            }
            dptr += KC * h;
            dlvl += KC;
        };
        load_delp();
        const int mine = (nsrc - gw + NCELL - 1) / NCELL;            // this warp's share of the TMA boxes: gw, gw + 4, ... (<= 16)
        const CUtensorMap *mymap = &p.maps[gw + NCELL * (lane < mine ? lane : 0)];
        const uint32_t mydst = ring + (gw + NCELL * lane) * BOX_BYTES;
        bool bad_geometry = false;                                   // a thickness that is not positive and finite
        for (int row = 0; row < 8; ++row) {
            ColG col;
            col.w = lds32(colw + 4 * (row * NCOL + x));
            col.ps = lds32(colps + 4 * (row * NCOL + x));
            float plast = toa;                                       // P(t-3); interfaces above the top sit at toa
            int icnt = 0;
            for (int chunk = 0; chunk < nch; ++chunk) {
                const int it = row * nch + chunk, b = it & 1;
                // interface pressures by the reference's sequential cumsum (vertically_dependent.py:61-63):
                // every warp carries the running sum of its own column walk, same order of additions
                const float pm3 = plast, pm2 = pm3 + dn[0], pm1 = pm2 + dn[1], p0 = pm1 + dn[2], pp1 = p0 + dn[3];
                plast = pp1;
                if (!NS) {                                           // the fast variant's sums have no NaN tests: see below
                    const int t1 = chunk * KC + gw;
                    if (t1 >= 1 && t1 <= nz) bad_geometry = bad_geometry || !(p0 - pm1 > 0.f) || !(p0 - pm1 < 3.0e38f);
                }
                if (chunk + 1 == nch) {                              // on to the next row
                    dptr = dcol + (long)(gw - 3) * h + (long)(row + 1) * nx;
                    dlvl = gw - 3;
                }
                if (chunk + 1 < nch || row + 1 < 8) load_delp();
                // stage b of the field ring and geometry buffer b are released together by the field warps
                if (it >= 2) asm volatile("bar.sync %0, 256;" ::"r"(2 + b) : "memory");
                if (lane == 0) mbar_expect_tx(bars + 8 * b, mine * KC * 128);
                __syncwarp();
                if (lane < mine)
                    tma_load_3d(mydst + b * stage_bytes, mymap, NCOL * bx, 8 * by + row, (int)(o * nz) + chunk * KC, bars + 8 * b);
                if (chunk * KC + gw >= 2)                            // (the records of steps 0 and 1 are not read)
                geometry_step<NS>(icnt, pm3, pm2, pm1, p0, pp1, g, col, chunk * KC + gw, gw, x, rec0 + b * REC_BYTES,
                              ova0 + b * OVA_BYTES, ovb0 + b * OVB_BYTES, tail, edge, nz, mask_below, toa);
                __syncwarp();
                if (lane == 0) mbar_arrive(gfull + 8 * b);
            }
        }
        // a source layer without thickness (0 / 0 in mappm.f90:75) or a NaN thickness: the NaN-safe variant reruns the CTA
        if (!NS && bad_geometry) p.flags[cta] = 1;
        return;
    }

    // ---------------- field warp of cell (by, 4 bx + w): lanes = fields -------------------------------
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(CONSUMER_REGS));
    const int w = warp;
    const uint32_t cb = sbase + lay.cell0 + w * lay.cell_bytes;
    const int cx_ = NCELL * bx + w;
    const bool active = lane < V;
    // (opaque(): keep these in registers; otherwise the layout arithmetic is redone inside the step loop)
    const uint32_t acc_lane = opaque(cb + lay.acc + 4 * lane);  // every lane owns a column of the accumulator
    const uint32_t rec_w = opaque(rec0 + 4 * w * REC_BLOCK), ova_w = opaque(ova0 + 8 * w * (MAXE * 16)),
                   ovb_w = opaque(ovb0 + 8 * w * (MAXE * 4));
    const uint32_t ring_r = opaque(ring);
    const int stage_bytes_r = (int)opaque((uint32_t)stage_bytes), nz_r = (int)opaque((uint32_t)nz);
    for (int i = lane; i < (nz + 1) * 32; i += 32) sts32(cb + lay.acc + 4 * i, 0.f);
    __syncwarp();
    Field fld;
    fld.fa = active ? p.fa[lane] : 0;
    fld.fb = active ? p.fb[lane] : 0;
    fld.c1 = active ? p.fc1[lane] : 0.f;
    fld.c2 = active ? p.fc2[lane] : 0.f;
    fld.c3 = active ? p.fc3[lane] : 1.f;
    const int any_derived = p.any_derived;

    u64 chk = 0;
    for (int row = 0; row < 8; ++row) {
        State st;
#pragma unroll
        for (int i = 0; i < 4; ++i) { st.A[i] = 0; st.B[i] = 0; st.dq[i] = 0; st.dc[i] = 0; st.xp[i] = 0; st.run[i] = 0; }
        st.chk = chk;

        for (int chunk = 0; chunk < nch; ++chunk) {
            const int it = row * nch + chunk, b = it & 1, t0 = chunk * KC;
            mbar_wait(gfull + 8 * b, (it >> 1) & 1);
            mbar_wait(bars + 8 * b, (it >> 1) & 1);
            const uint32_t stage = ring_r + b * stage_bytes_r;
#pragma unroll STEP_UNROLL
            for (int kk = t0 == 0 ? 2 : 0; kk < KC; ++kk) {     // (t = 0, 1 finish nothing: the end step at t = 2 seeds the state)
                const int t = t0 + kk;
                if (t > nz_r + 2) break;
                if (t != 2 && t != nz_r)
                    hot_step2<NS>(sctx, rec_w + b * REC_BYTES + kk * ((NCOL / 2) * REC_BLOCK), ova_w + b * OVA_BYTES + kk * (NCOL * MAXE * 16),
                             ovb_w + b * OVB_BYTES + kk * (NCOL * MAXE * 4), stage, t, nz_r, kk, b, w, fld, any_derived, st, acc_lane);
                else
                    end_step<NS>(sctx, stage, t, kk, b, w, nz_r, fld, any_derived, st);
            }
            __syncwarp();
            if (it + 2 < 8 * nch) asm volatile("bar.arrive %0, 256;" ::"r"(2 + b) : "memory");
        }
        chk = st.chk;
    }
    if (!NS) {
        float c0, c1;
        upk(chk, c0, c1);
        if (__any_sync(0xffffffffu, c0 != c0 || c1 != c1) && lane == 0) p.flags[cta] = 1;
    }

    // coarse outputs: block sum / weight sum (coarsen.py:213); 0/0 -> NaN like the reference
    __syncwarp();
    if (active) {
        float *out = p.outs[lane];
        for (int k = 0; k < nz; ++k)
            out[(o * nz + k) * hc + (long)by * nxc + cx_] = lds32(acc_lane + (k + 1) * ACC_STRIDE) / lds32(cb + lay.den + 4 * k);
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
    const stream::Layout lay = stream::make_layout(2 * nfields, nz);
    return lay.total <= 227 * 1024 - 1024;
}

long stream_workspace_bytes(long n_outer, int nz, int ny, int nx)
{
    const long h = (long)ny * nx, hc = (long)(ny / 8) * (nx / 8);
    return 4 * n_outer * nz * hc + 256 + 4 * n_outer * h + 256 + 4 * n_outer * (ny / 8) * (nx / 32) + 256;
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
                         const float *area, const float *delp_coarse, const float *psurf, int *flags, long n_outer,
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
    const Layout lay = make_layout(nsrc, nz);
    if (lay.total > 227 * 1024 - 1024) return CG_ENOTSUP;
    p.area = area; p.psurf = psurf; p.delp_coarse = delp_coarse; p.delp = delp;
    p.nsrc = nsrc; p.nfields = nfields; p.nz = nz; p.ny = ny; p.nx = nx; p.mask_below = mask_below; p.toa = toa;
    // function attributes are per device and cheap to set: on every call (a process may drive several devices)
    {
        for (int v = 0; v < 2; ++v) {
            const void *fn = v ? (const void *)plevel_stream_kernel<true> : (const void *)plevel_stream_kernel<false>;
            cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - 1024);
            if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(plevel_stream)");
            e = cudaFuncSetAttribute(fn, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
            if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(carveout)");
        }
    }
    dim3 grid(nx / 32, ny / 8, (unsigned)n_outer);
    // option plevel_nansafe: the NaN-safe variant alone (tests); default: the fast variant, then the NaN-safe one
    // over the CTAs that met a non-finite field value
    const bool nansafe_only = options().plevel_nansafe;
    if (!nansafe_only) {
        p.flags = flags;
        plevel_stream_kernel<false><<<grid, NTHREADS, lay.total, st>>>(p);
        CG_LAUNCH_CHECK("cg_pressure_level_average(stream)");
    }
    p.flags = nansafe_only ? nullptr : flags;
    plevel_stream_kernel<true><<<grid, NTHREADS, lay.total, st>>>(p);
    CG_LAUNCH_CHECK("cg_pressure_level_average(stream, NaN-safe pass)");
    return CG_OK;
}

}  // namespace cg
