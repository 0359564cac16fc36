// NaN-skipping window reductions, block upsample, interface pressures, weight
// masking and tile-edge interpolation.  All are pure streaming kernels (HBM-bound,
// every element read/written once); grid-stride loops sized from the SM count.
#include "common.cuh"
#include <cmath>

namespace cg {

static inline unsigned grid_for(long total, int threads)
{
    long blocks = (total + threads - 1) / threads;
    const long cap = (long)sm_count() * 16;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return (unsigned)blocks;
}

// coarsen.py:629-683 / :795-840 and budgets.py:59 (xarray coarsen sum/min/max/mean).
// PROPAGATE: numpy's plain sum/mean/min/max (a NaN in the window gives NaN) instead of
// xarray's NaN-skipping reductions.
template <typename T, int METHOD, bool PROPAGATE>
__global__ void __launch_bounds__(256)
block_reduce_kernel(const T *__restrict__ in, T *__restrict__ out, long n_outer, int ny, int nx,
                    int fy, int fx, int sy, int sx, int nyo, int nxo)
{
    const long total = n_outer * (long)nyo * nxo;
    for (long idx = blockIdx.x * (long)blockDim.x + threadIdx.x; idx < total;
         idx += (long)gridDim.x * blockDim.x) {
        const int I = (int)(idx % nxo);
        const int J = (int)((idx / nxo) % nyo);
        const long o = idx / ((long)nxo * nyo);
        const T *p = in + o * (long)ny * nx + (long)(J * sy) * nx + I * sx;
        T acc = T(0);
        bool any = false;
        int cnt = 0;
        bool saw_nan = false;
        for (int y = 0; y < fy; ++y)
            for (int x = 0; x < fx; ++x) {
                const T v = p[(long)y * nx + x];
                if (v != v) { saw_nan = true; continue; }
                if (METHOD == CG_SUM || METHOD == CG_MEAN) acc += v;
                else if (METHOD == CG_MIN) acc = any ? (v < acc ? v : acc) : v;
                else acc = any ? (v > acc ? v : acc) : v;
                any = true;
                ++cnt;
            }
        T r;
        if (METHOD == CG_SUM) r = acc;                       // all-NaN window -> 0 (nansum)
        else if (METHOD == CG_MEAN) r = acc / T(cnt);        // all-NaN -> 0/0 = NaN (nanmean)
        else r = any ? acc : T(NAN);                         // nanmin/nanmax of all-NaN -> NaN
        if (PROPAGATE && saw_nan) r = T(NAN);
        out[idx] = r;
    }
}

// coarsen.py:843-897: out[y,x] = in[min(y/fy, nyc-1), min(x/fx, nxc-1)]
// Write-bound: a warp owns an output row (8 rows per block), a lane four consecutive x written as one 16-B
// store where the row base is aligned; the coarse row it repeats stays in L1.  No 64-bit index arithmetic
// per element.
template <typename T>
__global__ void __launch_bounds__(256)
upsample_kernel(const T *__restrict__ in, T *__restrict__ out, long n_rows, int nyc, int nxc,
                int fy, int fx, int ny, int nx, int vec_ok)
{
    constexpr int V = 16 / sizeof(T);
    struct alignas(16) Vec { T v[V]; };
    const int lane = threadIdx.x & 31;
    for (long row = (long)blockIdx.x * 8 + (threadIdx.x >> 5); row < n_rows; row += (long)gridDim.x * 8) {
        const long o = row / ny;
        const int y = (int)(row - o * ny);
        const T *src = in + (o * nyc + min(y / fy, nyc - 1)) * (long)nxc;
        T *dst = out + row * nx;
        if (vec_ok == 2) {                           // fx a multiple of the vector width: one coarse value per 16-B store
            for (int x0 = V * lane; x0 < nx; x0 += V * 32) {
                const T c = src[min(x0 / fx, nxc - 1)];
                Vec r;
#pragma unroll
                for (int j = 0; j < V; ++j) r.v[j] = c;
                *reinterpret_cast<Vec *>(dst + x0) = r;
            }
        } else if (vec_ok) {
            for (int x0 = V * lane; x0 < nx; x0 += V * 32) {
                Vec r;
#pragma unroll
                for (int j = 0; j < V; ++j) r.v[j] = src[min((x0 + j) / fx, nxc - 1)];
                *reinterpret_cast<Vec *>(dst + x0) = r;
            }
        } else {
            for (int x = lane; x < nx; x += 32) dst[x] = src[min(x / fx, nxc - 1)];
        }
    }
}

// vertically_dependent.py:61-63: p[0] = toa; p[k+1] = p[k] + delp[k], in T, top down.
template <typename T>
__global__ void __launch_bounds__(256)
interface_kernel(const T *__restrict__ delp, T *__restrict__ pe, T toa, long n_outer, int nz, long h)
{
    const long total = n_outer * h;
    for (long idx = blockIdx.x * (long)blockDim.x + threadIdx.x; idx < total;
         idx += (long)gridDim.x * blockDim.x) {
        const long o = idx / h, i = idx % h;
        const T *d = delp + o * nz * h + i;
        T *p = pe + o * (nz + 1) * h + i;
        T acc = toa;
        p[0] = acc;
        for (int k = 0; k < nz; ++k) {
            acc = acc + d[(long)k * h];
            p[(long)(k + 1) * h] = acc;
        }
    }
}

// regridz.py:150-161
template <typename T, typename TW>
__global__ void __launch_bounds__(256)
mask_weights_kernel(const TW *__restrict__ w, const T *__restrict__ pc, const T *__restrict__ pf,
                    TW *__restrict__ out, long n_outer, int nz, int nzf, long h)
{
    const long total = n_outer * nz * h;
    for (long idx = blockIdx.x * (long)blockDim.x + threadIdx.x; idx < total;
         idx += (long)gridDim.x * blockDim.x) {
        const long i = idx % h;
        const int k = (int)((idx / h) % nz);
        const long o = idx / (h * nz);
        const T lower = pc[o * (nz + 1) * h + (long)(k + 1) * h + i];
        const T surf = pf[o * (nzf + 1) * h + (long)nzf * h + i];
        out[idx] = lower < surf ? w[o * h + i] : TW(0);
    }
}

// regridz.py:86-100 with xgcm.py:7-34: neighbour tile and axis of each face.
// axis 0 = y-edges, 1 = x-edges.  nb[t][axis][side] = {tile, neighbour axis (0=y,1=x)}.
struct FaceTable { int tile[6][2][2]; int ax[6][2][2]; };
static FaceTable face_table()
{
    // {left, right} neighbours along "x" and "y" per tile, as (tile, 'x'|'y')
    static const int X[6][2][2] = {{{4, 0}, {1, 1}}, {{0, 1}, {3, 0}}, {{0, 0}, {3, 1}},
                                   {{2, 1}, {5, 0}}, {{2, 0}, {5, 1}}, {{4, 1}, {1, 0}}};
    static const int Y[6][2][2] = {{{5, 0}, {2, 1}}, {{5, 1}, {2, 0}}, {{1, 0}, {4, 1}},
                                   {{1, 1}, {4, 0}}, {{3, 0}, {0, 1}}, {{3, 1}, {0, 0}}};
    FaceTable f;
    for (int t = 0; t < 6; ++t)
        for (int s = 0; s < 2; ++s) {
            f.tile[t][0][s] = Y[t][s][0]; f.ax[t][0][s] = Y[t][s][1];
            f.tile[t][1][s] = X[t][s][0]; f.ax[t][1][s] = X[t][s][1];
        }
    return f;
}

template <typename T>
__device__ __forceinline__ T halo_value(const T *c, const FaceTable &ft, int t, int axis, int side,
                                        int k, int nz, int n, int m /* transverse index */)
{
    const int nt = ft.tile[t][axis][side], na = ft.ax[t][axis][side];
    const int line = side == 0 ? n - 1 : 0;      // left pad: neighbour's last line; right: first
    const int mm = (na != axis) ? n - 1 - m : m; // reversed when the neighbour axis differs
    const T *base = c + ((long)nt * nz + k) * n * n;
    return na == 0 ? base[(long)line * n + mm] : base[(long)mm * n + line];
}

template <typename T>
__global__ void __launch_bounds__(256)
interp_edges_kernel(const T *__restrict__ c, T *__restrict__ out, FaceTable ft, int nz, int n, int axis)
{
    const int nyo = axis == 0 ? n + 1 : n, nxo = axis == 1 ? n + 1 : n;
    const long total = 6L * nz * nyo * nxo;
    for (long idx = blockIdx.x * (long)blockDim.x + threadIdx.x; idx < total;
         idx += (long)gridDim.x * blockDim.x) {
        const int x = (int)(idx % nxo);
        const int y = (int)((idx / nxo) % nyo);
        const int k = (int)((idx / ((long)nxo * nyo)) % nz);
        const int t = (int)(idx / ((long)nxo * nyo * nz));
        const T *base = c + ((long)t * nz + k) * n * n;
        const int e = axis == 0 ? y : x;   // edge index along the interpolated axis
        const int m = axis == 0 ? x : y;   // transverse index
        T a, b;
        if (e == 0) a = halo_value(c, ft, t, axis, 0, k, nz, n, m);
        else a = axis == 0 ? base[(long)(e - 1) * n + m] : base[(long)m * n + e - 1];
        if (e == n) b = halo_value(c, ft, t, axis, 1, k, nz, n, m);
        else b = axis == 0 ? base[(long)e * n + m] : base[(long)m * n + e];
        out[idx] = T(0.5) * (a + b);
    }
}

}  // namespace cg

using namespace cg;

extern "C" int cg_block_reduce(const void *in, void *out, int method, long n_outer, int ny, int nx,
                               int fy, int fx, int sy, int sx, int nyo, int nxo, int dtype,
                               void *stream)
{
    CG_REQUIRE(in && out, "null pointer");
    CG_REQUIRE(n_outer >= 0 && ny >= 1 && nx >= 1 && fy >= 1 && fx >= 1 && sy >= 1 && sx >= 1, "bad shape");
    CG_REQUIRE(nyo == 0 || (long)(nyo - 1) * sy + fy <= ny, "windows exceed ny");
    CG_REQUIRE(nxo == 0 || (long)(nxo - 1) * sx + fx <= nx, "windows exceed nx");
    const bool propagate = (method & CG_NAN_PROPAGATE) != 0;
    method &= ~CG_NAN_PROPAGATE;
    CG_REQUIRE(method >= CG_SUM && method <= CG_MEAN, "bad method");
    CG_REQUIRE(dtype == CG_F32 || dtype == CG_F64, "bad dtype");
    const long total = n_outer * (long)nyo * nxo;
    if (total == 0) return CG_OK;
    cudaStream_t st = (cudaStream_t)stream;
    const unsigned g = grid_for(total, 256);
#define GO(T, M) do { if (propagate) block_reduce_kernel<T, M, true><<<g, 256, 0, st>>>((const T *)in, (T *)out, n_outer, ny, nx, fy, fx, sy, sx, nyo, nxo); \
                      else block_reduce_kernel<T, M, false><<<g, 256, 0, st>>>((const T *)in, (T *)out, n_outer, ny, nx, fy, fx, sy, sx, nyo, nxo); } while (0)
#define GOM(T) do { if (method == CG_SUM) GO(T, CG_SUM); else if (method == CG_MIN) GO(T, CG_MIN); \
                    else if (method == CG_MAX) GO(T, CG_MAX); else GO(T, CG_MEAN); } while (0)
    if (dtype == CG_F32) GOM(float); else GOM(double);
#undef GOM
#undef GO
    CG_LAUNCH_CHECK("cg_block_reduce");
    return CG_OK;
}

extern "C" int cg_block_upsample(const void *in, void *out, long n_outer, int nyc, int nxc,
                                 int fy, int fx, int stag_y, int stag_x, int dtype, void *stream)
{
    CG_REQUIRE(in && out, "null pointer");
    CG_REQUIRE(n_outer >= 0 && nyc >= 1 && nxc >= 1 && fy >= 1 && fx >= 1, "bad shape");
    CG_REQUIRE(dtype == CG_F32 || dtype == CG_F64, "bad dtype");
    const int ny = stag_y ? (nyc - 1) * fy + 1 : nyc * fy;
    const int nx = stag_x ? (nxc - 1) * fx + 1 : nxc * fx;
    const long total = n_outer * (long)ny * nx;
    if (total == 0) return CG_OK;
    cudaStream_t st = (cudaStream_t)stream;
    const long n_rows = n_outer * ny;
    const long blocks = (n_rows + 7) / 8;
    const unsigned g = (unsigned)(blocks < 0x7fffffffL ? blocks : 0x7fffffffL);
    const int per16 = dtype == CG_F32 ? 4 : 2;
    int vec_ok = nx % per16 == 0 && ((uintptr_t)out & 15) == 0;
    if (vec_ok && fx % per16 == 0) vec_ok = 2;
    if (dtype == CG_F32) upsample_kernel<float><<<g, 256, 0, st>>>((const float *)in, (float *)out, n_rows, nyc, nxc, fy, fx, ny, nx, vec_ok);
    else upsample_kernel<double><<<g, 256, 0, st>>>((const double *)in, (double *)out, n_rows, nyc, nxc, fy, fx, ny, nx, vec_ok);
    CG_LAUNCH_CHECK("cg_block_upsample");
    return CG_OK;
}

extern "C" int cg_pressure_at_interface(const void *delp, void *pe, double toa, long n_outer, int nz,
                                        long h, int dtype, void *stream)
{
    CG_REQUIRE(delp && pe, "null pointer");
    CG_REQUIRE(n_outer >= 0 && nz >= 1 && h >= 0, "bad shape");
    CG_REQUIRE(dtype == CG_F32 || dtype == CG_F64, "bad dtype");
    if (n_outer * h == 0) return CG_OK;
    cudaStream_t st = (cudaStream_t)stream;
    const unsigned g = grid_for(n_outer * h, 256);
    if (dtype == CG_F32) interface_kernel<float><<<g, 256, 0, st>>>((const float *)delp, (float *)pe, (float)toa, n_outer, nz, h);
    else interface_kernel<double><<<g, 256, 0, st>>>((const double *)delp, (double *)pe, toa, n_outer, nz, h);
    CG_LAUNCH_CHECK("cg_pressure_at_interface");
    return CG_OK;
}

extern "C" int cg_mask_weights(const void *w2d, const void *pc, const void *pf, void *out, long n_outer,
                               int nz, int nzf, long h, int dtype, int wdtype, void *stream)
{
    CG_REQUIRE(w2d && pc && pf && out, "null pointer");
    CG_REQUIRE(n_outer >= 0 && nz >= 1 && nzf >= 1 && h >= 0, "bad shape");
    CG_REQUIRE((dtype == CG_F32 || dtype == CG_F64) && (wdtype == CG_F32 || wdtype == CG_F64), "bad dtype");
    const long total = n_outer * nz * h;
    if (total == 0) return CG_OK;
    cudaStream_t st = (cudaStream_t)stream;
    const unsigned g = grid_for(total, 256);
#define GO(T, TW) mask_weights_kernel<T, TW><<<g, 256, 0, st>>>((const TW *)w2d, (const T *)pc, (const T *)pf, (TW *)out, n_outer, nz, nzf, h)
    if (dtype == CG_F32 && wdtype == CG_F32) GO(float, float);
    else if (dtype == CG_F64 && wdtype == CG_F32) GO(double, float);
    else if (dtype == CG_F32 && wdtype == CG_F64) GO(float, double);
    else GO(double, double);
#undef GO
    CG_LAUNCH_CHECK("cg_mask_weights");
    return CG_OK;
}

extern "C" int cg_interp_to_edges(const void *c, void *out, int nz, int n, int axis, int dtype,
                                  void *stream)
{
    CG_REQUIRE(c && out, "null pointer");
    CG_REQUIRE(nz >= 1 && n >= 1 && (axis == 0 || axis == 1), "bad shape/axis");
    CG_REQUIRE(dtype == CG_F32 || dtype == CG_F64, "bad dtype");
    const long total = 6L * nz * (long)(n + 1) * n;
    cudaStream_t st = (cudaStream_t)stream;
    const unsigned g = grid_for(total, 256);
    const FaceTable ft = face_table();
    if (dtype == CG_F32) interp_edges_kernel<float><<<g, 256, 0, st>>>((const float *)c, (float *)out, ft, nz, n, axis);
    else interp_edges_kernel<double><<<g, 256, 0, st>>>((const double *)c, (double *)out, ft, nz, n, axis);
    CG_LAUNCH_CHECK("cg_interp_to_edges");
    return CG_OK;
}
