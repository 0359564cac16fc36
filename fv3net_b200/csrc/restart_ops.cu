// Small kernels of the restart-coarsening recipes (coarsen_restarts.py): the element-wise
// selections / products that sit between the block kernels, and the per-column thermodynamics
// (hydrostatic balance, blending weights).  All of it is one pass over HBM: each thread owns
// one element or one column, neighbouring threads neighbouring addresses.
#include "common.cuh"
#include <cmath>

namespace cg {

template <typename T> __device__ __forceinline__ T nan_of();
template <> __device__ __forceinline__ float nan_of<float>() { return __int_as_float(0x7fc00000); }
template <> __device__ __forceinline__ double nan_of<double>() { return __longlong_as_double(0x7ff8000000000000ll); }

// np.isclose(a, b) with its defaults (xarray_utils.py:7-30: rtol 1e-5, atol 1e-8, equal_nan False)
template <typename T> __device__ __forceinline__ bool is_close(T a, T b)
{
    if (a != a || b != b) return false;
    if (isinf(a) || isinf(b)) return a == b;
    return fabs((double)a - (double)b) <= 1e-8 + 1e-5 * fabs((double)b);
}

template <typename T>
__global__ void elementwise_kernel(int op, const T *__restrict__ a, const T *__restrict__ b, const T *__restrict__ c, T s,
                                   T *__restrict__ out, long n)
{
    const long stride = (long)gridDim.x * blockDim.x;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const T x = a[i];
        const T y = b ? b[i] : s;
        T r;
        switch (op) {
        case CG_EW_MUL: r = x * y; break;
        case CG_EW_ISCLOSE: r = is_close(x, y) ? T(1) : T(0); break;
        case CG_EW_AND: r = (x != T(0) && y != T(0)) ? T(1) : T(0); break;
        case CG_EW_WHERE: r = c[i] != T(0) ? x : y; break;
        case CG_EW_GT: r = x > y ? T(1) : T(0); break;
        case CG_EW_LT: r = x < y ? T(1) : T(0); break;
        case CG_EW_BLEND: r = x * y + (T(1) - x) * c[i]; break;
        case CG_EW_FILLNA: r = x != x ? y : x; break;
        case CG_EW_SUB: r = x - y; break;
        case CG_EW_DIV: r = x / y; break;
        default: r = nan_of<T>();
        }
        out[i] = r;
    }
}

// vertically_dependent.py:189-208 surface_pressure_from_delp: sum over z (top-down) + toa
template <typename T>
__global__ void surface_pressure_sum_kernel(const T *__restrict__ delp, T *__restrict__ out, T toa, long n_outer, int nz, long h)
{
    const long n = n_outer * h;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
        const long o = i / h, c = i - o * h;
        const T *p = delp + o * nz * h + c;
        T s = p[0];
        for (int k = 1; k < nz; ++k) s += p[(long)k * h];
        out[i] = s + toa;
    }
}

// the same with xarray's default skipna: NaN thicknesses count as zero (DataArray.sum -> nansum)
template <typename T>
__global__ void surface_pressure_nansum_kernel(const T *__restrict__ delp, T *__restrict__ out, T toa, long n_outer, int nz, long h)
{
    const long n = n_outer * h;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
        const long o = i / h, c = i - o * h;
        const T *p = delp + o * nz * h + c;
        T s = 0;
        for (int k = 0; k < nz; ++k) {
            const T d = p[(long)k * h];
            s += d == d ? d : T(0);
        }
        out[i] = s + toa;
    }
}

// vertically_dependent.py:153-179 pressure_at_midpoint_log: delp / diff(log(p_interface)), interfaces by the
// sequential cumsum from toa (:61-63)
template <typename T>
__global__ void midpoint_log_kernel(const T *__restrict__ delp, T *__restrict__ out, T toa, long n_outer, int nz, long h)
{
    const long n = n_outer * h;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
        const long o = i / h, c = i - o * h;
        const T *p = delp + o * nz * h + c;
        T *w = out + o * nz * h + c;
        T pi = toa, lp = log(pi);
        for (int k = 0; k < nz; ++k) {
            const T d = p[(long)k * h];
            pi = pi + d;
            const T ln = log(pi);
            w[(long)k * h] = d / (ln - lp);
            lp = ln;
        }
    }
}

// coarsen.py:54-80 shift_edge_var_to_center: 0.5 * (a[i] + a[i-1]) along the staggered dim, (outer, n, inner) ->
// (outer, n - 1, inner)
template <typename T>
__global__ void shift_to_center_kernel(const T *__restrict__ a, T *__restrict__ out, long n_outer, int n, long inner)
{
    const long total = n_outer * (n - 1) * inner;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        const long o = i / ((n - 1) * inner), r = i - o * (n - 1) * inner;
        const long src = o * n * inner + r;                             // element (o, j, x) with j = r / inner
        out[i] = T(0.5) * (a[src + inner] + a[src]);
    }
}

// coarsen_restarts.py:519-581 (compute_blending_weights on one coarse column):
//   ps = sum(delp) + toa;  pfull_k = delp_k / (log pi_{k+1} - log pi_k)  (pressure_at_midpoint_log)
//   pb = sigma * min over the block of the fine surface pressure (given)
//   w_k = pfull_k > pb ? (ps - pfull_k) / (ps - pb) : 1
template <typename T>
__global__ void blending_weights_kernel(const T *__restrict__ delp, const T *__restrict__ ps_min, T *__restrict__ out, T toa,
                                        T sigma, long n_outer, int nz, long h)
{
    const long n = n_outer * h;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
        const long o = i / h, c = i - o * h;
        const T *p = delp + o * nz * h + c;
        T *w = out + o * nz * h + c;
        T s = p[0];
        for (int k = 1; k < nz; ++k) s += p[(long)k * h];
        const T ps = s + toa, pb = sigma * ps_min[i];
        T pi = toa, lp = log(pi);
        for (int k = 0; k < nz; ++k) {
            const T d = p[(long)k * h];
            pi += d;
            const T ln = log(pi);
            const T pfull = d / (ln - lp);
            lp = ln;
            w[(long)k * h] = pfull > pb ? (ps - pfull) / (ps - pb) : T(1);
        }
    }
}

// coarsen_restarts.py:916-938 _impose_hydrostatic_balance on one coarse column:
//   z_top = phis / g + sum_{k = nz-1 .. 0} (-DZ_k)          (height_at_interface, reversed cumsum)
//   dz_k  = -(log pi_{k+1} - log pi_k) * Rd * T_k (1 + (Rv/Rd - 1) q_k) / g      (hydrostatic_dz)
//   phis' = g * (z_top + sum_k dz_k)                         (dz_and_top_to_phis)
template <typename T>
__global__ void hydrostatic_balance_kernel(const T *__restrict__ dz, const T *__restrict__ phis, const T *__restrict__ temp,
                                           const T *__restrict__ q, const T *__restrict__ delp, T *__restrict__ dz_out,
                                           T *__restrict__ phis_out, T toa, long n_outer, int nz, long h)
{
    const T g = T(9.80665), rd = T(287.05), rv = T(461.5);
    const long n = n_outer * h;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
        const long o = i / h, c = i - o * h;
        const long base = o * nz * h + c;
        T top = phis[i] / g;
        for (int k = nz - 1; k >= 0; --k) top += -dz[base + (long)k * h];
        T pi = toa, lp = log(pi), sum = T(0);
        for (int k = 0; k < nz; ++k) {
            const long a = base + (long)k * h;
            pi += delp[a];
            const T ln = log(pi);
            const T tv = temp[a] * (T(1) + (rv / rd - T(1)) * q[a]);
            const T d = -(ln - lp) * rd * tv / g;
            lp = ln;
            dz_out[a] = d;
            sum = k == 0 ? d : sum + d;
        }
        phis_out[i] = g * (top + sum);
    }
}

static inline unsigned grid_for(long n, int threads)
{
    const long want = (n + threads - 1) / threads;
    const long cap = (long)sm_count() * 16;
    return (unsigned)(want < 1 ? 1 : (want < cap ? want : cap));
}

}  // namespace cg

extern "C" int cg_elementwise(int op, const void *a, const void *b, const void *c, double scalar, void *out, long n,
                              int dtype, void *stream)
{
    using namespace cg;
    CG_REQUIRE(op >= CG_EW_MUL && op <= CG_EW_DIV, "cg_elementwise: unknown op %d", op);
    CG_REQUIRE(a && out && n >= 0, "cg_elementwise: null array");
    CG_REQUIRE((op != CG_EW_WHERE && op != CG_EW_BLEND) || c, "cg_elementwise: op %d needs a third array", op);
    CG_REQUIRE(op != CG_EW_BLEND || b, "cg_elementwise: blend needs three arrays");
    CG_REQUIRE(dtype == CG_F32 || dtype == CG_F64, "cg_elementwise: bad dtype %d", dtype);
    if (n == 0) return CG_OK;
    cudaStream_t st = (cudaStream_t)stream;
    const int threads = 256;
    if (dtype == CG_F32)
        elementwise_kernel<float><<<grid_for(n, threads), threads, 0, st>>>(op, (const float *)a, (const float *)b,
                                                                            (const float *)c, (float)scalar, (float *)out, n);
    else
        elementwise_kernel<double><<<grid_for(n, threads), threads, 0, st>>>(op, (const double *)a, (const double *)b,
                                                                             (const double *)c, scalar, (double *)out, n);
    CG_LAUNCH_CHECK("cg_elementwise");
    return CG_OK;
}

extern "C" int cg_surface_pressure(const void *delp, void *out, double toa, long n_outer, int nz, long h, int dtype,
                                   void *stream)
{
    using namespace cg;
    CG_REQUIRE(delp && out, "cg_surface_pressure: null array");
    CG_REQUIRE(n_outer >= 0 && nz >= 1 && h >= 0, "cg_surface_pressure: bad shape");
    CG_REQUIRE(dtype == CG_F32 || dtype == CG_F64, "cg_surface_pressure: bad dtype %d", dtype);
    const long n = n_outer * h;
    if (n == 0) return CG_OK;
    cudaStream_t st = (cudaStream_t)stream;
    if (dtype == CG_F32)
        surface_pressure_sum_kernel<float><<<grid_for(n, 128), 128, 0, st>>>((const float *)delp, (float *)out, (float)toa,
                                                                             n_outer, nz, h);
    else
        surface_pressure_sum_kernel<double><<<grid_for(n, 128), 128, 0, st>>>((const double *)delp, (double *)out, toa,
                                                                              n_outer, nz, h);
    CG_LAUNCH_CHECK("cg_surface_pressure");
    return CG_OK;
}

extern "C" int cg_surface_pressure_nansum(const void *delp, void *out, double toa, long n_outer, int nz, long h, int dtype,
                                          void *stream)
{
    using namespace cg;
    CG_REQUIRE(delp && out, "cg_surface_pressure_nansum: null array");
    CG_REQUIRE(n_outer >= 0 && nz >= 1 && h >= 0, "cg_surface_pressure_nansum: bad shape");
    CG_REQUIRE(dtype == CG_F32 || dtype == CG_F64, "cg_surface_pressure_nansum: bad dtype %d", dtype);
    const long n = n_outer * h;
    if (n == 0) return CG_OK;
    cudaStream_t st = (cudaStream_t)stream;
    if (dtype == CG_F32)
        surface_pressure_nansum_kernel<float><<<grid_for(n, 128), 128, 0, st>>>((const float *)delp, (float *)out, (float)toa,
                                                                                n_outer, nz, h);
    else
        surface_pressure_nansum_kernel<double><<<grid_for(n, 128), 128, 0, st>>>((const double *)delp, (double *)out, toa,
                                                                                 n_outer, nz, h);
    CG_LAUNCH_CHECK("cg_surface_pressure_nansum");
    return CG_OK;
}

extern "C" int cg_pressure_at_midpoint_log(const void *delp, void *out, double toa, long n_outer, int nz, long h, int dtype,
                                           void *stream)
{
    using namespace cg;
    CG_REQUIRE(delp && out, "cg_pressure_at_midpoint_log: null array");
    CG_REQUIRE(n_outer >= 0 && nz >= 1 && h >= 0, "cg_pressure_at_midpoint_log: bad shape");
    CG_REQUIRE(dtype == CG_F32 || dtype == CG_F64, "cg_pressure_at_midpoint_log: bad dtype %d", dtype);
    const long n = n_outer * h;
    if (n == 0) return CG_OK;
    cudaStream_t st = (cudaStream_t)stream;
    if (dtype == CG_F32)
        midpoint_log_kernel<float><<<grid_for(n, 128), 128, 0, st>>>((const float *)delp, (float *)out, (float)toa, n_outer, nz, h);
    else
        midpoint_log_kernel<double><<<grid_for(n, 128), 128, 0, st>>>((const double *)delp, (double *)out, toa, n_outer, nz, h);
    CG_LAUNCH_CHECK("cg_pressure_at_midpoint_log");
    return CG_OK;
}

extern "C" int cg_shift_to_center(const void *in, void *out, long n_outer, int n, long inner, int dtype, void *stream)
{
    using namespace cg;
    CG_REQUIRE(in && out, "cg_shift_to_center: null array");
    CG_REQUIRE(n_outer >= 0 && n >= 1 && inner >= 0, "cg_shift_to_center: bad shape");
    CG_REQUIRE(dtype == CG_F32 || dtype == CG_F64, "cg_shift_to_center: bad dtype %d", dtype);
    const long total = n_outer * (n - 1) * inner;
    if (total == 0) return CG_OK;
    cudaStream_t st = (cudaStream_t)stream;
    if (dtype == CG_F32)
        shift_to_center_kernel<float><<<grid_for(total, 256), 256, 0, st>>>((const float *)in, (float *)out, n_outer, n, inner);
    else
        shift_to_center_kernel<double><<<grid_for(total, 256), 256, 0, st>>>((const double *)in, (double *)out, n_outer, n, inner);
    CG_LAUNCH_CHECK("cg_shift_to_center");
    return CG_OK;
}

extern "C" int cg_blending_weights(const void *delp_coarse, const void *ps_min, void *out, double toa, double sigma,
                                   long n_outer, int nz, long h, int dtype, void *stream)
{
    using namespace cg;
    CG_REQUIRE(delp_coarse && ps_min && out, "cg_blending_weights: null array");
    CG_REQUIRE(n_outer >= 0 && nz >= 1 && h >= 0, "cg_blending_weights: bad shape");
    CG_REQUIRE(dtype == CG_F32 || dtype == CG_F64, "cg_blending_weights: bad dtype %d", dtype);
    const long n = n_outer * h;
    if (n == 0) return CG_OK;
    cudaStream_t st = (cudaStream_t)stream;
    if (dtype == CG_F32)
        blending_weights_kernel<float><<<grid_for(n, 128), 128, 0, st>>>((const float *)delp_coarse, (const float *)ps_min,
                                                                         (float *)out, (float)toa, (float)sigma, n_outer, nz, h);
    else
        blending_weights_kernel<double><<<grid_for(n, 128), 128, 0, st>>>((const double *)delp_coarse, (const double *)ps_min,
                                                                          (double *)out, toa, sigma, n_outer, nz, h);
    CG_LAUNCH_CHECK("cg_blending_weights");
    return CG_OK;
}

extern "C" int cg_hydrostatic_balance(const void *dz, const void *phis, const void *temp, const void *sphum,
                                      const void *delp, void *dz_out, void *phis_out, double toa, long n_outer, int nz,
                                      long h, int dtype, void *stream)
{
    using namespace cg;
    CG_REQUIRE(dz && phis && temp && sphum && delp && dz_out && phis_out, "cg_hydrostatic_balance: null array");
    CG_REQUIRE(n_outer >= 0 && nz >= 1 && h >= 0, "cg_hydrostatic_balance: bad shape");
    CG_REQUIRE(dtype == CG_F32 || dtype == CG_F64, "cg_hydrostatic_balance: bad dtype %d", dtype);
    const long n = n_outer * h;
    if (n == 0) return CG_OK;
    cudaStream_t st = (cudaStream_t)stream;
    if (dtype == CG_F32)
        hydrostatic_balance_kernel<float><<<grid_for(n, 128), 128, 0, st>>>(
            (const float *)dz, (const float *)phis, (const float *)temp, (const float *)sphum, (const float *)delp,
            (float *)dz_out, (float *)phis_out, (float)toa, n_outer, nz, h);
    else
        hydrostatic_balance_kernel<double><<<grid_for(n, 128), 128, 0, st>>>(
            (const double *)dz, (const double *)phis, (const double *)temp, (const double *)sphum, (const double *)delp,
            (double *)dz_out, (double *)phis_out, toa, n_outer, nz, h);
    CG_LAUNCH_CHECK("cg_hydrostatic_balance");
    return CG_OK;
}
