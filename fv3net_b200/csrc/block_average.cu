// Weighted block / edge averages (coarsen.py:183-218, :221-273 of the reference).
//
// f = 8: one thread owns one coarse cell of one level (no shuffle): it streams the 8 x 8 window of
// every field past weights it keeps in registers, so weights (and delp for mass
// weighting) are read once per block, not once per field.  A window
// row is 32 B = one DRAM sector; consecutive threads take consecutive windows,
// so a warp reads 1 KB contiguous per row with 128-bit loads.
// f = 16: a warp owns four cells, whole 256-byte rows per load, xor-shuffle reduction (wavg_f16_kernel).
// HBM-bound: algorithmic bytes = every input element once + every output once.
#include "common.cuh"

namespace cg {

struct FieldPtrs {
    const void *in[CG_MAX_FIELDS];
    void *out[CG_MAX_FIELDS];
};

template <typename T> __device__ __forceinline__ T nanzero(T v) { return v == v ? v : T(0); }

// ---------------------------------------------------------------------------------
// Generic path: any window, any dtype mix, scalar loads.
template <typename T, typename TW, int WMODE>
__global__ void __launch_bounds__(256)
wavg_generic_kernel(FieldPtrs fp, int nfields, const TW *__restrict__ w3d,
                    const TW *__restrict__ w2d, long n_outer, int nz, int ny, int nx,
                    int fy, int fx, int sy, int sx, int nyo, int nxo)
{
    using P = typename Promote<T, TW>::type;
    const long total = n_outer * nz * (long)nyo * nxo;
    for (long idx = blockIdx.x * (long)blockDim.x + threadIdx.x; idx < total;
         idx += (long)gridDim.x * blockDim.x) {
        const int I = (int)(idx % nxo);
        const int J = (int)((idx / nxo) % nyo);
        const long ok = idx / ((long)nxo * nyo);  // o*nz + k
        const long o = ok / nz;
        const long base3 = ok * (long)ny * nx + (long)(J * sy) * nx + I * sx;
        const long base2 = o * (long)ny * nx + (long)(J * sy) * nx + I * sx;

        TW den = TW(0);
        for (int y = 0; y < fy; ++y)
            for (int x = 0; x < fx; ++x) {
                TW w;
                if (WMODE == CG_W2D) w = w2d[base2 + (long)y * nx + x];
                else if (WMODE == CG_W3D) w = w3d[base3 + (long)y * nx + x];
                else w = w3d[base3 + (long)y * nx + x] * w2d[base2 + (long)y * nx + x];
                den += nanzero(w);
            }
        for (int f = 0; f < nfields; ++f) {
            const T *__restrict__ obj = (const T *)fp.in[f];
            P num = P(0);
            for (int y = 0; y < fy; ++y)
                for (int x = 0; x < fx; ++x) {
                    TW w;
                    if (WMODE == CG_W2D) w = w2d[base2 + (long)y * nx + x];
                    else if (WMODE == CG_W3D) w = w3d[base3 + (long)y * nx + x];
                    else w = w3d[base3 + (long)y * nx + x] * w2d[base2 + (long)y * nx + x];
                    num += nanzero(P(obj[base3 + (long)y * nx + x]) * P(w));
                }
            ((P *)fp.out[f])[idx] = num / P(den);
        }
    }
}

// ---------------------------------------------------------------------------------
// Fast path: fp32, cell-centred f = 8, nx % 8 == 0 (rows 32-B aligned).
template <int WMODE>
__global__ void __launch_bounds__(128)
wavg_f8_kernel(FieldPtrs fp, int nfields, const float *__restrict__ w3d,
               const float *__restrict__ w2d, long n_cells /* n_outer*nz*nyo*nxo */,
               int nz, int ny, int nx)
{
    const int nxo = nx >> 3, nyo = ny >> 3;
    const long idx = blockIdx.x * (long)blockDim.x + threadIdx.x;
    if (idx >= n_cells) return;
    const int I = (int)(idx % nxo);
    const int J = (int)((idx / nxo) % nyo);
    const long ok = idx / ((long)nxo * nyo);
    const long o = ok / nz;
    const long base3 = ok * (long)ny * nx + (long)(J * 8) * nx + I * 8;
    const long base2 = o * (long)ny * nx + (long)(J * 8) * nx + I * 8;

    float4 w[8][2];
    float den = 0.f;
#pragma unroll
    for (int y = 0; y < 8; ++y)
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            float4 a;
            if (WMODE == CG_W2D) {
                a = *reinterpret_cast<const float4 *>(w2d + base2 + (long)y * nx + 4 * h);
            } else if (WMODE == CG_W3D) {
                a = ldg_stream4(w3d + base3 + (long)y * nx + 4 * h);
            } else {
                const float4 d = ldg_stream4(w3d + base3 + (long)y * nx + 4 * h);
                const float4 b = *reinterpret_cast<const float4 *>(w2d + base2 + (long)y * nx + 4 * h);
                a = make_float4(d.x * b.x, d.y * b.y, d.z * b.z, d.w * b.w);
            }
            w[y][h] = a;
            den += nanzero(a.x);
            den += nanzero(a.y);
            den += nanzero(a.z);
            den += nanzero(a.w);
        }
    for (int f = 0; f < nfields; ++f) {
        const float *__restrict__ obj = (const float *)fp.in[f];
        float4 v[8][2];
#pragma unroll
        for (int y = 0; y < 8; ++y)
#pragma unroll
            for (int h = 0; h < 2; ++h) v[y][h] = ldg_stream4(obj + base3 + (long)y * nx + 4 * h);
        float num = 0.f;
#pragma unroll
        for (int y = 0; y < 8; ++y)
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                num += nanzero(v[y][h].x * w[y][h].x);
                num += nanzero(v[y][h].y * w[y][h].y);
                num += nanzero(v[y][h].z * w[y][h].z);
                num += nanzero(v[y][h].w * w[y][h].w);
            }
        ((float *)fp.out[f])[idx] = num / den;
    }
}

// ---------------------------------------------------------------------------------
// Fast path: fp32, cell-centred f = 16, nx % 64 == 0 (four cells per warp).  A 16 x 16 window is too large for one
// thread's registers, so a warp owns four x-adjacent cells: lanes 0-15 hold the 16-byte chunks of one 256-byte row of
// the four windows, lanes 16-31 those of the next row (every load instruction moves two whole rows, 512 B); a lane
// keeps its 8 weight chunks in registers for all fields and the 8 lanes of a cell (4 chunks x 2 rows) are summed by
// xor-shuffles - a fixed tree, so runs are bitwise reproducible.
template <int WMODE>
__global__ void __launch_bounds__(128)
wavg_f16_kernel(FieldPtrs fp, int nfields, const float *__restrict__ w3d, const float *__restrict__ w2d,
                long n_groups /* n_outer*nz*nyo*(nxo/4) */, int nz, int ny, int nx)
{
    const int nxo = nx >> 4, nyo = ny >> 4, ngx = nxo >> 2;
    const long G = (blockIdx.x * (long)blockDim.x + threadIdx.x) >> 5;
    if (G >= n_groups) return;                                            // (whole warps: no partial shuffles)
    const int lane = threadIdx.x & 31;
    const int I4 = (int)(G % ngx);
    const int J = (int)((G / ngx) % nyo);
    const long ok = G / ((long)ngx * nyo);
    const long o = ok / nz;
    const long off = (long)(J * 16 + (lane >> 4)) * nx + I4 * 64 + (lane & 15) * 4;
    const long base3 = ok * (long)ny * nx + off, base2 = o * (long)ny * nx + off;
    const long step = 2L * nx;

    float4 w[8];
    float den = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        float4 a;
        if (WMODE == CG_W2D) {
            a = *reinterpret_cast<const float4 *>(w2d + base2 + i * step);
        } else if (WMODE == CG_W3D) {
            a = ldg_stream4(w3d + base3 + i * step);
        } else {
            const float4 d = ldg_stream4(w3d + base3 + i * step);
            const float4 b = *reinterpret_cast<const float4 *>(w2d + base2 + i * step);
            a = make_float4(d.x * b.x, d.y * b.y, d.z * b.z, d.w * b.w);
        }
        w[i] = a;
        den += nanzero(a.x);
        den += nanzero(a.y);
        den += nanzero(a.z);
        den += nanzero(a.w);
    }
    den += __shfl_xor_sync(0xffffffffu, den, 1);
    den += __shfl_xor_sync(0xffffffffu, den, 2);
    den += __shfl_xor_sync(0xffffffffu, den, 16);
    const long cell = ((ok * nyo + J) * (long)nxo) + I4 * 4 + ((lane & 15) >> 2);
    for (int f = 0; f < nfields; ++f) {
        const float *__restrict__ obj = (const float *)fp.in[f];
        float4 v[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) v[i] = ldg_stream4(obj + base3 + i * step);
        float num = 0.f;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            num += nanzero(v[i].x * w[i].x);
            num += nanzero(v[i].y * w[i].y);
            num += nanzero(v[i].z * w[i].z);
            num += nanzero(v[i].w * w[i].w);
        }
        num += __shfl_xor_sync(0xffffffffu, num, 1);
        num += __shfl_xor_sync(0xffffffffu, num, 2);
        num += __shfl_xor_sync(0xffffffffu, num, 16);
        if ((lane & 19) == 0) ((float *)fp.out[f])[cell] = num / den;     // lanes 0, 4, 8, 12: one per cell
    }
}

template <typename T, typename TW>
static int launch_generic(const FieldPtrs &fp, int nfields, const void *w3d, const void *w2d,
                          int wmode, long n_outer, int nz, int ny, int nx, int fy, int fx,
                          int sy, int sx, int nyo, int nxo, cudaStream_t st)
{
    const long total = n_outer * nz * (long)nyo * nxo;
    long blocks = (total + 255) / 256;
    const long cap = (long)sm_count() * 32;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
#define CG_GO(M)                                                                              \
    wavg_generic_kernel<T, TW, M><<<(unsigned)blocks, 256, 0, st>>>(                          \
        fp, nfields, (const TW *)w3d, (const TW *)w2d, n_outer, nz, ny, nx, fy, fx, sy, sx,   \
        nyo, nxo)
    if (wmode == CG_W2D) CG_GO(CG_W2D);
    else if (wmode == CG_W3D) CG_GO(CG_W3D);
    else CG_GO(CG_WMASS);
#undef CG_GO
    CG_LAUNCH_CHECK("cg_weighted_block_average");
    return CG_OK;
}

}  // namespace cg

extern "C" int cg_weighted_block_average(const void *const *objs, void *const *outs, int nfields,
                                         const void *w3d, const void *w2d, int wmode,
                                         long n_outer, int nz, int ny, int nx,
                                         int fy, int fx, int sy, int sx, int nyo, int nxo,
                                         int dtype, int wdtype, void *stream)
{
    using namespace cg;
    CG_REQUIRE(nfields >= 1 && nfields <= CG_MAX_FIELDS, "nfields must be in [1, %d]", CG_MAX_FIELDS);
    CG_REQUIRE(n_outer >= 0 && nz >= 1 && ny >= 1 && nx >= 1, "bad shape");
    CG_REQUIRE(fy >= 1 && fx >= 1 && sy >= 1 && sx >= 1 && nyo >= 0 && nxo >= 0, "bad window");
    CG_REQUIRE(nyo == 0 || (long)(nyo - 1) * sy + fy <= ny, "windows exceed ny");
    CG_REQUIRE(nxo == 0 || (long)(nxo - 1) * sx + fx <= nx, "windows exceed nx");
    CG_REQUIRE(wmode == CG_W2D || wmode == CG_W3D || wmode == CG_WMASS, "bad weight mode");
    CG_REQUIRE((wmode == CG_W3D || w2d) && (wmode == CG_W2D || w3d), "missing weight pointer");
    CG_REQUIRE((dtype == CG_F32 || dtype == CG_F64) && (wdtype == CG_F32 || wdtype == CG_F64), "bad dtype");
    cudaStream_t st = (cudaStream_t)stream;
    if (n_outer * (long)nz * nyo * nxo == 0) return CG_OK;

    FieldPtrs fp;
    for (int i = 0; i < nfields; ++i) {
        CG_REQUIRE(objs[i] && outs[i], "null field pointer");
        fp.in[i] = objs[i];
        fp.out[i] = outs[i];
    }
    const bool fast = dtype == CG_F32 && wdtype == CG_F32 && fy == 8 && fx == 8 && sy == 8 &&
                      sx == 8 && nx % 8 == 0 && nyo == ny / 8 && nxo == nx / 8 && ny % 8 == 0;
    if (fast) {
        bool aligned = (((uintptr_t)w3d | (uintptr_t)w2d) & 15) == 0;
        for (int i = 0; i < nfields; ++i) aligned = aligned && (((uintptr_t)objs[i]) & 15) == 0;
        if (aligned) {
            const long n_cells = n_outer * nz * (long)nyo * nxo;
            const unsigned blocks = cdiv(n_cells, 128);
            if (wmode == CG_W2D)
                wavg_f8_kernel<CG_W2D><<<blocks, 128, 0, st>>>(fp, nfields, (const float *)w3d, (const float *)w2d, n_cells, nz, ny, nx);
            else if (wmode == CG_W3D)
                wavg_f8_kernel<CG_W3D><<<blocks, 128, 0, st>>>(fp, nfields, (const float *)w3d, (const float *)w2d, n_cells, nz, ny, nx);
            else
                wavg_f8_kernel<CG_WMASS><<<blocks, 128, 0, st>>>(fp, nfields, (const float *)w3d, (const float *)w2d, n_cells, nz, ny, nx);
            CG_LAUNCH_CHECK("cg_weighted_block_average(f8)");
            return CG_OK;
        }
    }
    const bool fast16 = dtype == CG_F32 && wdtype == CG_F32 && fy == 16 && fx == 16 && sy == 16 && sx == 16 &&
                        nx % 64 == 0 && ny % 16 == 0 && nyo == ny / 16 && nxo == nx / 16;
    if (fast16) {
        bool aligned = (((uintptr_t)w3d | (uintptr_t)w2d) & 15) == 0;
        for (int i = 0; i < nfields; ++i) aligned = aligned && (((uintptr_t)objs[i]) & 15) == 0;
        if (aligned) {
            const long n_groups = n_outer * nz * (long)nyo * (nxo / 4);
            const unsigned blocks = cdiv(n_groups, 4);                    // 4 warps per block
            if (wmode == CG_W2D)
                wavg_f16_kernel<CG_W2D><<<blocks, 128, 0, st>>>(fp, nfields, (const float *)w3d, (const float *)w2d, n_groups, nz, ny, nx);
            else if (wmode == CG_W3D)
                wavg_f16_kernel<CG_W3D><<<blocks, 128, 0, st>>>(fp, nfields, (const float *)w3d, (const float *)w2d, n_groups, nz, ny, nx);
            else
                wavg_f16_kernel<CG_WMASS><<<blocks, 128, 0, st>>>(fp, nfields, (const float *)w3d, (const float *)w2d, n_groups, nz, ny, nx);
            CG_LAUNCH_CHECK("cg_weighted_block_average(f16)");
            return CG_OK;
        }
    }
#define CG_DISPATCH(T, TW) \
    return launch_generic<T, TW>(fp, nfields, w3d, w2d, wmode, n_outer, nz, ny, nx, fy, fx, sy, sx, nyo, nxo, st)
    if (dtype == CG_F32 && wdtype == CG_F32) CG_DISPATCH(float, float);
    if (dtype == CG_F64 && wdtype == CG_F32) CG_DISPATCH(double, float);
    if (dtype == CG_F32 && wdtype == CG_F64) CG_DISPATCH(float, double);
    CG_DISPATCH(double, double);
#undef CG_DISPATCH
}
