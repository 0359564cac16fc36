// Shared pieces of the streaming pressure-level kernels (sm_100a): PTX wrappers (mbarrier, TMA, shared-memory
// vector accesses, packed f32x2 arithmetic) and the pressure-only geometry of the tandem walk - what a coarse
// interface met inside a fine layer contributes (mappm.f90:58-124), as weights that every field reuses.
#pragma once
#include "common.cuh"
#include <cuda.h>

namespace cg {
namespace plv {

typedef unsigned long long u64;

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
__device__ __forceinline__ int ldsi32(uint32_t addr)
{
    int v;
    asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(addr));
    return v;
}
__device__ __forceinline__ int4 lds128i(uint32_t addr)
{
    int4 v;
    asm volatile("ld.shared.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
    return v;
}
__device__ __forceinline__ void lds2x64(uint32_t addr, unsigned long long &a, unsigned long long &b)
{
    asm volatile("ld.shared.v2.b64 {%0,%1}, [%2];" : "=l"(a), "=l"(b) : "r"(addr));
}
__device__ __forceinline__ unsigned long long lds64(uint32_t addr)
{
    unsigned long long a;
    asm volatile("ld.shared.b64 %0, [%1];" : "=l"(a) : "r"(addr));
    return a;
}
__device__ __forceinline__ void sts32(uint32_t addr, float v)
{
    asm volatile("st.shared.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory");
}
__device__ __forceinline__ void stsi32(uint32_t addr, int v)
{
    asm volatile("st.shared.b32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}
__device__ __forceinline__ void sts128(uint32_t addr, float4 v)
{
    asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
// acc[slot][item] += v  (no native fp32 shared-memory reduction on sm_100a: red.shared.add.f32 is a CAS loop)
__device__ __forceinline__ void red_add(uint32_t addr, float v)
{
    float o;
    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(o) : "r"(addr));
    asm volatile("st.shared.f32 [%0], %1;" ::"r"(addr), "f"(o + v) : "memory");
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
// 1.0f when a > b, else 0.0f (one FSET.BF)
__device__ __forceinline__ float gt01(float a, float b)
{
    float r;
    asm("set.gt.f32.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));
    return r;
}
// byte address of 16-B chunk `chunk` of the 128-B row starting at absolute shared address `row`
// under CU_TENSOR_MAP_SWIZZLE_128B (chunk index xor row index mod 8; verified by tools/tma_probe.cu)
__device__ __forceinline__ uint32_t swz(uint32_t row, int chunk)
{
    return row + ((((row >> 7) ^ (uint32_t)chunk) & 7u) << 4);
}
// a value the compiler must keep in a register instead of recomputing it where it is used
__device__ __forceinline__ uint32_t opaque(uint32_t v)
{
    uint32_t r;
    asm volatile("mov.u32 %0, %1;" : "=r"(r) : "r"(v));
    return r;
}

// ---- f32x2: two adjacent columns per instruction ---------------------------------------------------
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

// ---- geometry: what all fields share -----------------------------------------------------------
constexpr int DUMMY_SLOT = 0;        // accumulator slot k + 1 belongs to target level k; slot 0 collects the rest

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

// Integral weights times (q, x, y) of one event.  Events that use the layer mean alone - a target interface at or
// above the column top (mappm.f90:62-64), everything below the surface (:115-120), the empty record - carry
// z = +0; an event inside a layer always carries z = -(a s (1 - s)) with the sign bit set, even when it is zero
// (make_event).  The NaN-safe variant drops the edge terms of the former, so that a NaN edge value does not reach
// a level the Fortran fills from q alone, and keeps them for the latter, where the Fortran evaluates the parabola.
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

// mappm.f90:663-665: sign(min(|df2|, max(q-,q,q+) - q, q - min(q-,q,q+)), df2), by min / max alone
__device__ __forceinline__ float limited_slope(float df2, float dq_lo, float dq_hi)
{
    const float na = -dq_lo;
    const float hi3 = max3(na, dq_hi, 0.f), nlo = min3(na, dq_hi, 0.f);
    return max3(min3(df2, hi3, -nlo), -hi3, nlo);
}
__device__ __forceinline__ u64 slope2(u64 df2, u64 dlo, u64 dhi)
{
    float fa, fb, la, lb, ha, hb;
    upk(df2, fa, fb); upk(dlo, la, lb); upk(dhi, ha, hb);
    return pk(limited_slope(fa, la, ha), limited_slope(fb, lb, hb));
}
__device__ __forceinline__ u64 clamp2(u64 v, u64 a, u64 b)            // v between a and b, per half
{
    float va, vb, aa, ab, ba, bb;
    upk(v, va, vb); upk(a, aa, ab); upk(b, ba, bb);
    return pk(fminf(fmaxf(va, fminf(aa, ba)), fmaxf(aa, ba)), fminf(fmaxf(vb, fminf(ab, bb)), fmaxf(ab, bb)));
}

}  // namespace plv
}  // namespace cg
