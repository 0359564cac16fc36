// Fused pressure-level coarse-graining, f = 8, iv = 1, kord = 1, 65 <= nz <= 80, nx % 16 == 0.
// (regridz.py:115-147 + coarsen.py:183, budgets.py:49-57; mappm.f90:10-126, :614-851, :854-931)
//
// What bounds it: the per-cell PPM work (~60 issue slots per cell-field), not HBM; see
// DESIGN.md "fused kernel".  Every fine input element is read from HBM exactly once per
// launch (delp twice, the second time from L2); nothing fine-grid is written.
//
// Mapping.  A CTA (512 threads) owns two x-adjacent coarse cells = 8 rows x 16 columns of
// fine columns and walks them in 4 passes of 2 rows x 16 columns.  In a pass, warp w owns
// x-column w: lanes 0-15 the column in row 0, lanes 16-31 the column in row 1; lane l of a
// half-warp owns the 5 consecutive layers 5l..5l+4 of its column ("lanes over layers").
//   * everything that depends only on pressure — the 5 PPM stencil coefficients per layer,
//     and for every target layer the overlap geometry reduced to linear functionals
//     (3 coefficients on (AL,AR,A6) of the first and last source layer + a whole-layer
//     scale) — is computed once per pass and kept in REGISTERS, then reused for every field;
//   * per field the lanes build the limited parabolas of their 5 layers (neighbour values
//     by warp shuffle), publish them to a per-warp shared-memory panel, and evaluate their 5
//     target layers by gathering the source layers the geometry names (data-dependent index
//     -> shared memory, bank-conflict free by layout);
//   * area(-masked) weights are folded into the coefficients, so the gathered value IS the
//     column's contribution to the block sum; the two rows are merged by shuffle, the 8
//     x-columns of a cell by a deferred shared-memory reduction (double-buffered, one
//     __syncthreads per tile, fixed order -> bitwise reproducible).
// Tiles (2 rows x 16 columns x nz levels = 64-B global segments) are prefetched two tiles
// ahead into registers with 128-bit loads and transposed into a padded [k][33] smem layout.
// Products / storage terms (cg_field) are formed while staging.
#include "common.cuh"
#include <cmath>

namespace cg {

constexpr int T = 5;            // layers per lane
constexpr int KP = 16 * T;      // padded levels per column (80)
constexpr int TS = 33;          // tile row stride (floats): conflict-free transposed access
constexpr int NWARP = 16;
constexpr int NTHREADS = NWARP * 32;
constexpr int MAXV = CG_MAX_FIELDS;

struct FusedParams {
    cg_field fields[MAXV];
    float *outs[MAXV];
    const float *delp;
    const float *area;
    float *delp_coarse;  // may be null
    int nfields, nz, ny, nx, mask_below;
    float toa;
};

struct Smem {
    float tile[2][KP * TS];          // staged q (or delp) tiles, [k][h*16 + w]
    float panel[NWARP][4][2 * KP];   // per warp: AL', AR', A6' (or pe1) and dp, [half*80 + k]
    float part[2][NWARP][KP];        // per-warp partial block sums of the current / previous tile
    float num[2][MAXV][KP];          // block sums per cell, field, level
    float den[2][KP];                // sum of (masked) weights per cell, level
    float dsum[2][KP];               // sum of area * delp per cell, level
    float pe2[2][KP + 1];            // coarse interface pressures per cell
    float asum[2];
};

// ---- tile staging -----------------------------------------------------------------------
// A tile is nz x 2 x 16 floats; as float4 chunks (k, h, xq): nz*8 chunks, thread t takes
// chunks t and t + 512.
struct TileRegs { float4 v[2]; };

__device__ __forceinline__ float4 combine(const float4 a, const float4 b, int op, float scale)
{
    if (op == CG_FIELD_PRODUCT) return make_float4(a.x * b.x, a.y * b.y, a.z * b.z, a.w * b.w);
    return make_float4((a.x - b.x) / scale, (a.y - b.y) / scale, (a.z - b.z) / scale, (a.w - b.w) / scale);
}

__device__ __forceinline__ void tile_load(TileRegs &r, const float *__restrict__ a,
                                          const float *__restrict__ b, int op, float scale,
                                          long base /* offset of (k=0, row y0, x0) */, long lev_stride,
                                          int nx, int nz)
{
#pragma unroll
    for (int c = 0; c < 2; ++c) {
        const int chunk = threadIdx.x + c * NTHREADS;
        if (chunk < nz * 8) {
            const int k = chunk >> 3, h = (chunk >> 2) & 1, xq = chunk & 3;
            const long off = base + k * lev_stride + (long)h * nx + 4 * xq;
            float4 v = ldg_stream4(a + off);
            if (op != CG_FIELD_PLAIN) v = combine(v, ldg_stream4(b + off), op, scale);
            r.v[c] = v;
        }
    }
}

__device__ __forceinline__ void tile_store(float *__restrict__ tile, const TileRegs &r, int nz)
{
#pragma unroll
    for (int c = 0; c < 2; ++c) {
        const int chunk = threadIdx.x + c * NTHREADS;
        if (chunk < nz * 8) {
            const int k = chunk >> 3, h = (chunk >> 2) & 1, xq = chunk & 3;
            float *d = tile + k * TS + h * 16 + 4 * xq;
            d[0] = r.v[c].x; d[1] = r.v[c].y; d[2] = r.v[c].z; d[3] = r.v[c].w;
        }
    }
}

__device__ __forceinline__ float nanzero(float v) { return v == v ? v : 0.f; }
__device__ __forceinline__ float max3f(float a, float b, float c) { return fmaxf(fmaxf(a, b), c); }
__device__ __forceinline__ float min3f(float a, float b, float c) { return fminf(fminf(a, b), c); }

// ---- per-lane geometry (registers, computed once per pass, reused by every field) ----------
template <int NZ> struct Geo {
    // PPM stencil of the lane's own layers j = k0+i (mappm.f90:658-683 with the delp-only
    // factors folded):  df2 = E1*delq(j) + E2*delq(j-1)
    //                   al  = q(j-1) + F*delq(j-1) + G*dc(j-1) - H*dc(j)
    float E1[T], E2[T], F[T], G[T], H[T];
    // one-sided cubic at the column's top (lane 0) / bottom (lane LB) edge, :689-747, in the
    // common form  qm = sa*qe + sb*qi; dq = sg*(qi-qe); c1 = (ali-qm-sd2*dq)*sD;
    //              c3 = dq - c1*sP; inner = qm - c1*sR; edge = c1*sE - c3*sd1 + inner
    float sa, sb, sg, sD, sd2, sP, sR, sE, sd1, ssign;
    // overlap of the lane's target layers k = k0+i with the column (mappm.f90:58-124) as
    // linear functionals: value = ct . (AL,AR,A6)[Lt] + cb . (AL,AR,A6)[Lb]
    //                             + ww * sum_{Lt<L<Lb} dp[L]*q[L];  weights folded in.
    int pack[T];           // Lt | Lb << 8 | nwhole << 16
    float ct[T][3], cb[T][3], ww[T];
};

__host__ __device__ constexpr int lane_of(int k) { return k / T; }
__host__ __device__ constexpr int slot_of(int k) { return k % T; }

// dp[j] from the staged delp tile, 1.0 outside the column (keeps the unused stencils finite)
template <int NZ>
__device__ __forceinline__ float dp_at(const float *tile, int col, int j)
{
    return (j >= 0 && j < NZ) ? tile[j * TS + col] : 1.0f;
}

template <int NZ>
__device__ void build_geometry(Geo<NZ> &g, const float *__restrict__ tile, float *__restrict__ pe_panel,
                               float *__restrict__ dp_panel, const float *__restrict__ pe2,
                               float area_w, int mask_below, float toa, int col, int l, int half,
                               float (&den_out)[T])
{
    static_assert(slot_of(NZ - 1) >= 1, "bottom layer and its neighbour must share a lane");
    const int k0 = T * l;
    constexpr unsigned FULL = 0xffffffffu;

    // --- own interface pressures by the reference's sequential cumsum (bitwise the same
    //     order as vertically_dependent.py:61-63): lane s extends the chain of lane s-1.
    float dpo[T], pe[T + 1];
#pragma unroll
    for (int i = 0; i < T; ++i) dpo[i] = dp_at<NZ>(tile, col, k0 + i);
    float carry = toa;
    for (int s = 0; s < 16; ++s) {
        float mine = carry;
        if (l == s) {
            pe[0] = carry;
#pragma unroll
            for (int i = 0; i < T; ++i) pe[i + 1] = pe[i] + dpo[i];
            mine = pe[T];
        }
        carry = __shfl_sync(FULL, mine, (half << 4) | s);
    }
    // layer thickness as the reference forms it: dp1 = pe1(k+1) - pe1(k)  (mappm.f90:43),
    // i.e. from the rounded interface sums, not the raw delp
#pragma unroll
    for (int i = 0; i < T; ++i) {
        pe_panel[half * KP + k0 + i] = pe[i];       // pe1[0..79]; pe1[NZ] is the surface
        dp_panel[half * KP + k0 + i] = (k0 + i < NZ) ? pe[i + 1] - pe[i] : 1.0f;
    }
    __syncwarp();
    const float *P1 = pe_panel + half * KP;         // pe1[L], L = 0..NZ
    const float *DP = dp_panel + half * KP;
    const float p_surf = P1[NZ];

    // --- PPM stencil coefficients of the own layers
    float dpw[T + 3];                                // dp1[k0-2 .. k0+T]
#pragma unroll
    for (int m = 0; m < T + 3; ++m) {
        const int j = k0 - 2 + m;
        dpw[m] = (j >= 0 && j < NZ) ? DP[j] : 1.0f;
    }
#pragma unroll
    for (int i = 0; i < T; ++i) {
        const int j = k0 + i;
        const float dm2 = dpw[i], dm1 = dpw[i + 1], d0 = dpw[i + 2], dp1 = dpw[i + 3];
        const float d4m = dm2 + dm1, d4 = dm1 + d0, d4p = d0 + dp1;   // d4(j-1), d4(j), d4(j+1)
        const float c1 = (dm1 + 0.5f * d0) / d4p;
        const float c2 = (dp1 + 0.5f * d0) / d4;
        const float inv = d0 / (d4 + dp1);
        const bool dc_ok = j >= 1 && j <= NZ - 2;
        g.E1[i] = dc_ok ? c1 * inv : 0.f;
        g.E2[i] = dc_ok ? c2 * inv : 0.f;
        const float cq = dm1 / d4;
        const float a1 = d4m / (d4 + dm1);
        const float a2 = d4p / (d4 + d0);
        const float gg = 2.f / (d4m + d4p);
        const bool al_ok = j >= 2 && j <= NZ - 2;
        g.F[i] = al_ok ? cq * (1.f + gg * d0 * (a1 - a2)) : 0.f;
        g.G[i] = al_ok ? gg * d0 * a2 : 0.f;
        g.H[i] = al_ok ? gg * dm1 * a1 : 0.f;
    }

    // --- edge cubics: top for lane 0, bottom for lane LB (other lanes: harmless values)
    {
        constexpr int LB = lane_of(NZ - 1);
        const bool top = (l == 0);
        const float d1 = top ? DP[0] : DP[NZ - 1];
        const float d2 = top ? DP[1] : DP[NZ - 2];
        const float D = d2 * (2.f * d2 * d2 + d1 * (d2 + 3.f * d1));
        const float Pq = d2 * (5.f * d1 + d2) - 3.f * d1 * d1;
        const float Rr = d1 * d2 * (d2 + 3.f * d1);
        g.sa = d2 / (d1 + d2);
        g.sb = d1 / (d1 + d2);
        g.sg = 2.f / (d1 + d2);
        g.sd2 = d2;
        g.sd1 = d1;
        g.sD = (top ? 4.f : 1.f) / D;
        g.sP = (top ? 0.5f : 2.f) * Pq;
        g.sR = (top ? 0.25f : 1.f) * Rr;
        g.sE = (top ? 2.f : 8.f) * d1 * d1 * d1;
        g.ssign = top ? 0.5f : -0.5f;
        (void)LB;
    }

    // --- overlap geometry of the own target layers
    const float r3 = 1.f / 3.f, r23 = 2.f / 3.f, sixth = 1.f / 6.f;
#pragma unroll
    for (int i = 0; i < T; ++i) {
        const int k = k0 + i;
        float ct0 = 0.f, ct1 = 0.f, ct2 = 0.f, cb0 = 0.f, cb1 = 0.f, cb2 = 0.f, ww = 0.f, W = 0.f;
        int Lt = 0, Lb = 0, nwhole = 0;
        if (k < NZ) {
            const float top = pe2[k], bot = pe2[k + 1];
            W = (!mask_below || bot < p_surf) ? area_w : 0.f;          // regridz.py:150-161
            if (top <= P1[0]) {                                        // mappm.f90:62-64
                Lt = 0; ct0 = 0.5f * W; ct1 = 0.5f * W; ct2 = sixth * W;
            } else if (top >= p_surf) {                                // :65-67
                Lt = NZ - 1; ct0 = 0.5f * W; ct1 = 0.5f * W; ct2 = sixth * W;
            } else {
                int lo = 0, hi = NZ - 1;                               // first L with top <= pe1[L+1]
                while (lo < hi) {
                    const int mid = (lo + hi) >> 1;
                    if (top <= P1[mid + 1]) hi = mid; else lo = mid + 1;
                }
                Lt = lo;
                const float pl0 = P1[Lt], pl1 = P1[Lt + 1], dpl = DP[Lt];
                const float pl = (top - pl0) / dpl;
                if (bot <= pl1) {                                      // :76-83
                    const float pr = (bot - pl0) / dpl;
                    const float hs = 0.5f * (pr + pl);
                    const float tt = r3 * (pr * (pr + pl) + pl * pl);
                    ct0 = W * (1.f - hs); ct1 = W * hs; ct2 = W * (hs - tt);
                } else {                                               // :85-120
                    const float delp_t = pl1 - top;
                    const float hs = 0.5f * (1.f + pl);
                    const float tt = r3 * (1.f + pl * (1.f + pl));
                    float dpsum = delp_t;
                    int L = Lt + 1;
                    while (L < NZ && bot > P1[L + 1]) { dpsum = dpsum + DP[L]; ++L; ++nwhole; }
                    float b0 = 0.f, b1 = 0.f, b2 = 0.f;
                    if (L < NZ) {
                        Lb = L;
                        const float delp_b = bot - P1[L];
                        const float esl = delp_b / DP[L];
                        dpsum = dpsum + delp_b;
                        b0 = delp_b * (1.f - 0.5f * esl);
                        b1 = delp_b * 0.5f * esl;
                        b2 = delp_b * 0.5f * esl * (1.f - r23 * esl);
                    } else {
                        Lb = NZ - 1;
                        const float ext = bot - p_surf;
                        if (ext > 0.f) {
                            dpsum = dpsum + ext;
                            b0 = 0.5f * ext; b1 = 0.5f * ext; b2 = sixth * ext;
                        }
                    }
                    const float s = W / dpsum;
                    ct0 = s * delp_t * (1.f - hs); ct1 = s * delp_t * hs; ct2 = s * delp_t * (hs - tt);
                    cb0 = s * b0; cb1 = s * b1; cb2 = s * b2;
                    ww = s;
                }
            }
        }
        g.pack[i] = Lt | (Lb << 8) | (nwhole << 16);
        g.ct[i][0] = ct0; g.ct[i][1] = ct1; g.ct[i][2] = ct2;
        g.cb[i][0] = cb0; g.cb[i][1] = cb1; g.cb[i][2] = cb2;
        g.ww[i] = ww;
        den_out[i] = W == W ? W : 0.f;
    }
    __syncwarp();
}

// ---- one field tile: limited parabolas of the own layers, then the own target layers -------
template <int NZ>
__device__ __forceinline__ void process_field(const Geo<NZ> &g, const float *__restrict__ tile,
                                              float *__restrict__ pw /* this warp's panel */,
                                              int col, int l, int half, float (&out)[T])
{
    constexpr unsigned FULL = 0xffffffffu;
    constexpr int LB = lane_of(NZ - 1), SB = slot_of(NZ - 1);
    const int k0 = T * l;
    float *AL = pw + half * KP, *AR = pw + 2 * KP + half * KP, *A6 = pw + 4 * KP + half * KP;
    const float *DP = pw + 6 * KP + half * KP;

    float q[T + 2];                                  // q[k0-1 .. k0+T], clamped at the ends
#pragma unroll
    for (int m = 0; m < T + 2; ++m) {
        int j = k0 - 1 + m;
        j = j < 0 ? 0 : (j > NZ - 1 ? NZ - 1 : j);
        q[m] = tile[j * TS + col];
    }
    float delq[T + 1];                               // delq[m] = q(k0+m) - q(k0+m-1)
#pragma unroll
    for (int m = 0; m < T + 1; ++m) delq[m] = q[m + 1] - q[m];

    float dc[T];                                     // limited slope, mappm.f90:658-668
#pragma unroll
    for (int i = 0; i < T; ++i) {
        const float df2 = g.E1[i] * delq[i + 1] + g.E2[i] * delq[i];
        const float hi = max3f(q[i], q[i + 1], q[i + 2]) - q[i + 1];
        const float lo = q[i + 1] - min3f(q[i], q[i + 1], q[i + 2]);
        dc[i] = copysignf(fminf(fminf(fabsf(df2), hi), lo), df2);
    }
    const float dc_m = __shfl_up_sync(FULL, dc[T - 1], 1, 16);

    float al[T];                                     // 4th-order edge value, :674-683
#pragma unroll
    for (int i = 0; i < T; ++i)
        al[i] = q[i] + g.F[i] * delq[i] + g.G[i] * (i == 0 ? dc_m : dc[i - 1]) - g.H[i] * dc[i];

    float ar_bot;
    {                                                // one-sided cubics at the column ends, :689-747
        const bool top = (l == 0);
        const float qe = top ? q[1] : q[SB + 1];
        const float qi = top ? q[2] : q[SB];
        const float ali = top ? al[2] : al[SB - 1];
        const float qm = g.sa * qe + g.sb * qi;
        const float dq = g.sg * (qi - qe);
        const float c1 = (ali - qm - g.sd2 * dq) * g.sD;
        const float c3 = dq - c1 * g.sP;
        float inner = qm - c1 * g.sR;
        const float edge = c1 * g.sE - c3 * g.sd1 + inner;
        inner = fminf(fmaxf(inner, fminf(qe, qi)), fmaxf(qe, qi));
        const float dce = g.ssign * (inner - qe);
        if (top) { al[0] = edge; al[1] = inner; dc[0] = dce; }
        if (l == LB) { al[SB] = inner; dc[SB] = dce; }
        ar_bot = edge;
    }
    const float al_p = __shfl_down_sync(FULL, al[0], 1, 16);

#pragma unroll
    for (int i = 0; i < T; ++i) {                    // A6 and the standard PPM limiter, :773-778, :854-894
        const float qv = q[i + 1];
        float a_l = al[i];
        float a_r = (i < T - 1) ? al[i + 1] : al_p;
        if (i == SB && l == LB) a_r = ar_bot;
        float a6 = 3.f * (2.f * qv - (a_l + a_r));
        const float da1 = a_r - a_l;
        const float da2 = da1 * da1;
        const float a6da = a6 * da1;
        const float a6_lo = 3.f * (a_l - qv), ar_lo = a_l - a6_lo;     // a6da < -da2
        const float a6_hi = 3.f * (a_r - qv), al_hi = a_r - a6_hi;     // a6da >  da2
        const bool flat = dc[i] == 0.f, lo_c = a6da < -da2, hi_c = a6da > da2;
        const float nl = flat ? qv : (lo_c ? a_l : (hi_c ? al_hi : a_l));
        const float nr = flat ? qv : (lo_c ? ar_lo : a_r);
        const float n6 = flat ? 0.f : (lo_c ? a6_lo : (hi_c ? a6_hi : a6));
        AL[k0 + i] = nl;
        AR[k0 + i] = nr;
        A6[k0 + i] = n6;
    }
    __syncwarp();

#pragma unroll
    for (int i = 0; i < T; ++i) {                    // overlap integrals as linear functionals
        const int pk = g.pack[i];
        const int Lt = pk & 255, Lb = (pk >> 8) & 255, nw = pk >> 16;
        float t = g.ct[i][0] * AL[Lt] + g.ct[i][1] * AR[Lt] + g.ct[i][2] * A6[Lt];
        t += g.cb[i][0] * AL[Lb] + g.cb[i][1] * AR[Lb] + g.cb[i][2] * A6[Lb];
        if (nw) {
            float s = 0.f;
            for (int L = Lt + 1; L <= Lt + nw; ++L) s += DP[L] * tile[L * TS + col];
            t += g.ww[i] * s;
        }
        out[i] = nanzero(t);
    }
    __syncwarp();
}

// ---- the kernel ---------------------------------------------------------------------------
template <int NZ>
__global__ void __launch_bounds__(NTHREADS, 1)
plevel_fused_kernel(const __grid_constant__ FusedParams p)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    Smem &S = *reinterpret_cast<Smem *>(smem_raw);
    constexpr unsigned FULL = 0xffffffffu;
    const int tid = threadIdx.x, w = tid >> 5, lane = tid & 31, half = lane >> 4, l = lane & 15;
    const int col = half * 16 + w;            // column inside a tile
    const int cell = w >> 3;                  // which of the CTA's two coarse cells
    const int V = p.nfields, ny = p.ny, nx = p.nx;
    const int nyc = ny >> 3, nxc = nx >> 3;
    const long h = (long)ny * nx, hc = (long)nyc * nxc;
    const long o = blockIdx.z;
    const int by = blockIdx.y, bx = blockIdx.x;
    const long base3 = o * NZ * h + (long)(8 * by) * nx + 16 * bx;
    const long base2 = o * h + (long)(8 * by) * nx + 16 * bx;

    for (int i = tid; i < 2 * MAXV * KP; i += NTHREADS) (&S.num[0][0][0])[i] = 0.f;
    for (int i = tid; i < 2 * KP; i += NTHREADS) { (&S.den[0][0])[i] = 0.f; (&S.dsum[0][0])[i] = 0.f; }
    if (tid < 2) S.asum[tid] = 0.f;
    __syncthreads();

    // ---- phase A: coarse delp of the two cells (coarsen.py:210-213) and their interfaces
    for (int pass = 0; pass < 4; ++pass) {
        TileRegs r;
        tile_load(r, p.delp, nullptr, CG_FIELD_PLAIN, 1.f, base3 + (long)(2 * pass) * nx, h, nx, NZ);
        tile_store(S.tile[0], r, NZ);
        __syncthreads();
        const float a = p.area[base2 + (long)(2 * pass + half) * nx + w];
        float aw = nanzero(a);
        aw += __shfl_xor_sync(FULL, aw, 16);
#pragma unroll
        for (int i = 0; i < T; ++i) {
            const int k = T * l + i;
            float v = k < NZ ? nanzero(S.tile[0][k * TS + col] * a) : 0.f;
            v += __shfl_xor_sync(FULL, v, 16);
            if (half == 0) S.part[0][w][k] = v;
        }
        if (lane == 0) S.part[1][w][0] = aw;
        __syncthreads();
        if (tid < 2 * KP) {
            const int c = tid / KP, k = tid % KP;
            float s = 0.f;
#pragma unroll
            for (int ww = 0; ww < 8; ++ww) s += S.part[0][c * 8 + ww][k];
            S.dsum[c][k] += s;
        } else if (tid < 2 * KP + 2) {
            const int c = tid - 2 * KP;
            float s = 0.f;
#pragma unroll
            for (int ww = 0; ww < 8; ++ww) s += S.part[1][c * 8 + ww][0];
            S.asum[c] += s;
        }
        __syncthreads();
    }
    if (tid < 2) {                             // sequential cumsum, vertically_dependent.py:61-63
        const int c = tid;
        float pe = p.toa;
        S.pe2[c][0] = pe;
        const float asum = S.asum[c];
        for (int k = 0; k < NZ; ++k) {
            const float dcv = S.dsum[c][k] / asum;
            if (p.delp_coarse) p.delp_coarse[o * NZ * hc + (long)k * hc + (long)by * nxc + 2 * bx + c] = dcv;
            pe = pe + dcv;
            S.pe2[c][k + 1] = pe;
        }
    }
    __syncthreads();

    // ---- phase B: per pass one geometry tile (delp) then one tile per field, software-pipelined
    const int per_pass = 1 + V, n_items = 4 * per_pass;
    auto load_item = [&](int it, TileRegs &r) {
        const int pass = it / per_pass, j = it % per_pass;
        const long off = base3 + (long)(2 * pass) * nx;
        if (j == 0) tile_load(r, p.delp, nullptr, CG_FIELD_PLAIN, 1.f, off, h, nx, NZ);
        else tile_load(r, p.fields[j - 1].a, p.fields[j - 1].b, p.fields[j - 1].op, p.fields[j - 1].scale, off, h, nx, NZ);
    };
    auto reduce_item = [&](int it) {          // deferred: item `it` left its partials in part[it & 1]
        if (tid < 2 * KP) {
            const int c = tid / KP, k = tid % KP, j = it % per_pass;
            float s = 0.f;
#pragma unroll
            for (int ww = 0; ww < 8; ++ww) s += S.part[it & 1][c * 8 + ww][k];
            if (j == 0) S.den[c][k] += s;
            else S.num[c][j - 1][k] += s;
        }
    };

    TileRegs r;
    load_item(0, r);
    tile_store(S.tile[0], r, NZ);
    if (n_items > 1) load_item(1, r);
    __syncthreads();

    Geo<NZ> g;
    float *pw = &S.panel[w][0][0];
    for (int it = 0; it < n_items; ++it) {
        if (it + 1 < n_items) tile_store(S.tile[(it + 1) & 1], r, NZ);
        if (it + 2 < n_items) load_item(it + 2, r);
        if (it > 0) reduce_item(it - 1);

        const int pass = it / per_pass, j = it % per_pass;
        const float *tile = S.tile[it & 1];
        float vals[T];
        if (j == 0) {
            const float a = p.area[base2 + (long)(2 * pass + half) * nx + w];
            build_geometry<NZ>(g, tile, pw, pw + 6 * KP, S.pe2[cell], a, p.mask_below, p.toa, col, l, half, vals);
        } else {
            process_field<NZ>(g, tile, pw, col, l, half, vals);
        }
#pragma unroll
        for (int i = 0; i < T; ++i) {
            const float v = vals[i] + __shfl_xor_sync(FULL, vals[i], 16);
            if (half == 0) S.part[it & 1][w][T * l + i] = v;
        }
        __syncthreads();
    }
    reduce_item(n_items - 1);
    __syncthreads();

    // ---- coarse outputs: block sum / weight sum (coarsen.py:213); 0/0 -> NaN like the reference
    for (int idx = tid; idx < 2 * V * NZ; idx += NTHREADS) {
        const int c = idx / (V * NZ), v = (idx / NZ) % V, k = idx % NZ;
        p.outs[v][o * NZ * hc + (long)k * hc + (long)by * nxc + 2 * bx + c] = S.num[c][v][k] / S.den[c][k];
    }
}

int fused_supported(int nz, int ny, int nx, int f)
{
    return f == 8 && nz == 79 && nx % 16 == 0 && ny % 8 == 0;
}

long fused_workspace_bytes(int, long, int, int, int, int) { return 0; }

int launch_plevel_fused(const cg_field *fields, float *const *outs, int nfields, const float *delp,
                        const float *area, float *delp_coarse, long n_outer, int nz, int ny, int nx,
                        int f, int mask_below, float toa, void *, long, cudaStream_t st)
{
    if (!fused_supported(nz, ny, nx, f)) return CG_ENOTSUP;
    FusedParams p;
    bool aligned = (((uintptr_t)delp) & 15) == 0;
    for (int v = 0; v < nfields; ++v) {
        p.fields[v] = fields[v];
        p.outs[v] = outs[v];
        aligned = aligned && (((uintptr_t)fields[v].a | (uintptr_t)fields[v].b) & 15) == 0;
    }
    if (!aligned) return CG_ENOTSUP;           // caller falls back to the composed path
    p.delp = delp; p.area = area; p.delp_coarse = delp_coarse;
    p.nfields = nfields; p.nz = nz; p.ny = ny; p.nx = nx; p.mask_below = mask_below; p.toa = toa;
    static_assert(sizeof(Smem) <= 227 * 1024, "shared memory budget");
    cudaError_t e = cudaFuncSetAttribute(plevel_fused_kernel<79>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Smem));
    if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(plevel_fused)");
    dim3 grid(nx / 16, ny / 8, (unsigned)n_outer);
    plevel_fused_kernel<79><<<grid, NTHREADS, sizeof(Smem), st>>>(p);
    CG_LAUNCH_CHECK("cg_pressure_level_average(fused)");
    return CG_OK;
}

}  // namespace cg
