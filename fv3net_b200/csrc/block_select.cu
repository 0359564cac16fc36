// block_median / _block_mode (coarsen.py:557-588, :750-786 of the reference).
//
// Median: one thread selects the two middle order statistics of its f x f block in
// registers with a Batcher odd-even merge network (min/max only, no data-dependent
// branches, no shared memory).  For 64 values the two 32-value halves are sorted
// (2 x 191 comparators) and the middle pair is read off one cross-compare layer:
//   L[i] = min(a[i], b[31-i]) are the 32 smallest, U[i] = max(...) the 32 largest,
//   so lo = max(L), hi = min(U)  ->  median = (lo + hi) / 2 in the input dtype,
// which is np.median's mean of the two middle elements.  Any NaN in the block
// gives NaN (numpy propagates).  Results are bit-exact w.r.t. numpy.
// Larger / odd blocks use one warp per block with a shared-memory bitonic sort.
// HBM-bound in intent: 4 B read per fine cell; ~14 issue slots per cell for f = 8.
#include "common.cuh"
#include <cmath>
#include <limits>

namespace cg {

template <typename T> __device__ __forceinline__ void cswap(T &a, T &b)
{
    const T lo = a < b ? a : b;  // inputs are NaN-free here (NaN blocks are flagged apart)
    const T hi = a < b ? b : a;
    a = lo;
    b = hi;
}
template <> __device__ __forceinline__ void cswap<float>(float &a, float &b)
{
    const float lo = fminf(a, b), hi = fmaxf(a, b);
    a = lo;
    b = hi;
}
// min / max that return NaN when either operand is NaN (FMNMX.NAN): a comparator network built from them carries a
// NaN input to every output it can reach, so the block's "any NaN -> NaN" rule needs no separate test
__device__ __forceinline__ float min_nan(float a, float b)
{
    float r;
    asm("min.NaN.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));
    return r;
}
__device__ __forceinline__ float max_nan(float a, float b)
{
    float r;
    asm("max.NaN.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));
    return r;
}
struct NanF { float v; };                                   // float whose compare-exchange propagates NaN
__device__ __forceinline__ void cswap(NanF &a, NanF &b)
{
    const float lo = min_nan(a.v, b.v), hi = max_nan(a.v, b.v);
    a.v = lo;
    b.v = hi;
}

// Batcher's odd-even merge sort on a register array; N a power of two.  Written as
// template recursion so every index is a compile-time constant (the array must
// stay in registers).
template <typename T, int N, int LO, int CNT, int R> struct OddEvenMerge {
    __device__ __forceinline__ static void run(T (&a)[N])
    {
        constexpr int M = R * 2;
        if constexpr (M < CNT) {
            OddEvenMerge<T, N, LO, CNT, M>::run(a);
            OddEvenMerge<T, N, LO + R, CNT, M>::run(a);
#pragma unroll
            for (int i = LO + R; i + R < LO + CNT; i += M) cswap(a[i], a[i + R]);
        } else {
            cswap(a[LO], a[LO + R]);
        }
    }
};
template <typename T, int N, int LO, int CNT> struct OddEvenSort {
    __device__ __forceinline__ static void run(T (&a)[N])
    {
        if constexpr (CNT > 1) {
            constexpr int M = CNT / 2;
            OddEvenSort<T, N, LO, M>::run(a);
            OddEvenSort<T, N, LO + M, M>::run(a);
            OddEvenMerge<T, N, LO, CNT, 1>::run(a);
        }
    }
};
template <typename T, int N> __device__ __forceinline__ void sort_network(T (&a)[N])
{
    OddEvenSort<T, N, 0, N>::run(a);
}

// f = 8, fp32: the config-4 shape.  Thread per coarse cell, 16 x 128-bit loads.
__global__ void __launch_bounds__(128)
median_f8_f32_kernel(const float *__restrict__ in, float *__restrict__ out, long n_cells, int ny, int nx)
{
    const int nxo = nx >> 3, nyo = ny >> 3;
    const long idx = blockIdx.x * (long)blockDim.x + threadIdx.x;
    if (idx >= n_cells) return;
    const int I = (int)(idx % nxo);
    const int J = (int)((idx / nxo) % nyo);
    const long o = idx / ((long)nxo * nyo);
    const float *p = in + o * (long)ny * nx + (long)(J * 8) * nx + I * 8;

    NanF a[32], b[32];
#pragma unroll
    for (int y = 0; y < 4; ++y)
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const float4 v = ldg_stream4(p + (long)y * nx + 4 * h);
            const float4 u = ldg_stream4(p + (long)(y + 4) * nx + 4 * h);
            a[y * 8 + h * 4 + 0].v = v.x; a[y * 8 + h * 4 + 1].v = v.y;
            a[y * 8 + h * 4 + 2].v = v.z; a[y * 8 + h * 4 + 3].v = v.w;
            b[y * 8 + h * 4 + 0].v = u.x; b[y * 8 + h * 4 + 1].v = u.y;
            b[y * 8 + h * 4 + 2].v = u.z; b[y * 8 + h * 4 + 3].v = u.w;
        }
    // every wire of both halves enters lo or hi below, and a NaN stays on a wire through every comparator: any NaN
    // in the block makes lo and hi NaN (numpy's median propagates NaN the same way)
    sort_network<NanF, 32>(a);
    sort_network<NanF, 32>(b);
    float lo = -INFINITY, hi = INFINITY;
#pragma unroll
    for (int i = 0; i < 32; ++i) {
        lo = max_nan(lo, min_nan(a[i].v, b[31 - i].v));
        hi = min_nan(hi, max_nan(a[i].v, b[31 - i].v));
    }
    out[idx] = (lo + hi) / 2.0f;
}

// Small power-of-two blocks (N = f*f in {4, 16, 64}), any dtype: full sort in registers.
template <typename T, int F>
__global__ void __launch_bounds__(128)
median_small_kernel(const T *__restrict__ in, T *__restrict__ out, long n_cells, int ny, int nx)
{
    constexpr int N = F * F;
    const int nxo = nx / F, nyo = ny / F;
    const long idx = blockIdx.x * (long)blockDim.x + threadIdx.x;
    if (idx >= n_cells) return;
    const int I = (int)(idx % nxo);
    const int J = (int)((idx / nxo) % nyo);
    const long o = idx / ((long)nxo * nyo);
    const T *p = in + o * (long)ny * nx + (long)(J * F) * nx + I * F;
    T a[N];
    bool has_nan = false;
#pragma unroll
    for (int y = 0; y < F; ++y)
#pragma unroll
        for (int x = 0; x < F; ++x) {
            a[y * F + x] = p[(long)y * nx + x];
            has_nan = has_nan || (a[y * F + x] != a[y * F + x]);
        }
    sort_network<T, N>(a);
    out[idx] = has_nan ? T(NAN) : (a[N / 2 - 1] + a[N / 2]) / T(2);
}

// Mode of small power-of-two blocks: the same register sort, then the longest run of the sorted keys (the
// smallest value among equally long runs, like scipy.stats.mode).  NaNs sort last as +inf and are counted apart.
// MODE: 1 = propagate NaN, 2 = omit NaN (see select_generic_kernel for the rules).
template <typename T, int F, int MODE>
__global__ void __launch_bounds__(128)
mode_small_kernel(const T *__restrict__ in, T *__restrict__ out, long n_cells, int ny, int nx)
{
    constexpr int N = F * F;
    const int nxo = nx / F, nyo = ny / F;
    const long idx = blockIdx.x * (long)blockDim.x + threadIdx.x;
    if (idx >= n_cells) return;
    const int I = (int)(idx % nxo);
    const int J = (int)((idx / nxo) % nyo);
    const long o = idx / ((long)nxo * nyo);
    const T *p = in + o * (long)ny * nx + (long)(J * F) * nx + I * F;
    T a[N];
    int nan_cnt = 0;
#pragma unroll
    for (int y = 0; y < F; ++y)
#pragma unroll
        for (int x = 0; x < F; ++x) {
            T v = p[(long)y * nx + x];
            if (v != v) { ++nan_cnt; v = T(INFINITY); }
            a[y * F + x] = v;
        }
    sort_network<T, N>(a);
    const int m = N - nan_cnt;
    T best = a[0];
    int best_len = 0, run = 0;
#pragma unroll
    for (int i = 0; i < N; ++i) {
        if (i < m) {
            run = (i > 0 && a[i] == a[i > 0 ? i - 1 : 0]) ? run + 1 : 1;
            if (run > best_len) { best_len = run; best = a[i]; }
        }
    }
    T r = best;
    if (m == 0 || (MODE == 1 && nan_cnt > best_len)) r = T(NAN);
    out[idx] = r;
}

// ---------------------------------------------------------------------------------
// Generic: one warp per block, n = f*f <= 1024 values sorted in shared memory
// (bitonic, padded with +inf; NaNs are counted apart and sort as +inf).
template <typename T>
__device__ void warp_bitonic_sort(T *s, int npad, int lane)
{
    for (int k = 2; k <= npad; k <<= 1)
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = lane; i < npad; i += 32) {
                const int l = i ^ j;
                if (l > i) {
                    const T x = s[i], y = s[l];
                    const bool up = (i & k) == 0;
                    if ((x > y) == up) { s[i] = y; s[l] = x; }
                }
            }
            __syncwarp();
        }
}

// MODE: 0 = median, 1 = mode (propagate NaN), 2 = mode (omit NaN)
template <typename T, int MODE>
__global__ void __launch_bounds__(128)
select_generic_kernel(const T *__restrict__ in, T *__restrict__ out, long n_cells, int ny, int nx,
                      int f, int npad)
{
    extern __shared__ unsigned char smem_raw[];
    T *smem = reinterpret_cast<T *>(smem_raw);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const long idx = blockIdx.x * (long)(blockDim.x >> 5) + warp;
    if (idx >= n_cells) return;
    T *s = smem + (long)warp * npad;
    const int nxo = nx / f, nyo = ny / f, n = f * f;
    const int I = (int)(idx % nxo);
    const int J = (int)((idx / nxo) % nyo);
    const long o = idx / ((long)nxo * nyo);
    const T *p = in + o * (long)ny * nx + (long)(J * f) * nx + I * f;
    int nan_cnt = 0;
    for (int i = lane; i < npad; i += 32) {
        T v = T(INFINITY);
        if (i < n) {
            v = p[(long)(i / f) * nx + (i % f)];
            if (v != v) { ++nan_cnt; v = T(INFINITY); }
        }
        s[i] = v;
    }
    for (int off = 16; off > 0; off >>= 1) nan_cnt += __shfl_xor_sync(0xffffffffu, nan_cnt, off);
    __syncwarp();
    warp_bitonic_sort(s, npad, lane);
    if (lane != 0) return;
    T r;
    if (MODE == 0) {
        if (nan_cnt) r = T(NAN);
        else r = (n & 1) ? s[n / 2] : (s[n / 2 - 1] + s[n / 2]) / T(2);
    } else {
        const int m = n - nan_cnt;  // valid values are s[0..m) (NaNs sorted behind as +inf)
        if (m == 0) r = T(NAN);
        else {
            // real +inf values sit before the NaN stand-ins only by convention; both are
            // +inf, so runs of +inf are counted up to m values in total.
            T best = s[0];
            int best_len = 0, run = 0;
            for (int i = 0; i < m; ++i) {
                run = (i > 0 && s[i] == s[i - 1]) ? run + 1 : 1;
                if (run > best_len) { best_len = run; best = s[i]; }
            }
            r = best;
            // 'propagate' (scipy 1.7.3 / numpy 1.21 unique): NaN is an ordinary candidate that
            // sorts last, so it wins only when strictly more frequent than every other value.
            if (MODE == 1 && nan_cnt > best_len) r = T(NAN);
        }
    }
    out[idx] = r;
}

template <typename T, int MODE>
static int launch_generic_select(const void *in, void *out, long n_cells, int ny, int nx, int f,
                                 cudaStream_t st, const char *what)
{
    int npad = 1;
    while (npad < f * f) npad <<= 1;
    CG_REQUIRE(npad <= 4096, "block of %d values is too large", f * f);
    const int warps = 4;
    const size_t smem = (size_t)warps * npad * sizeof(T);
    if (smem > 48 * 1024)
        cudaFuncSetAttribute(select_generic_kernel<T, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    select_generic_kernel<T, MODE><<<cdiv(n_cells, warps), warps * 32, smem, st>>>(
        (const T *)in, (T *)out, n_cells, ny, nx, f, npad);
    CG_LAUNCH_CHECK(what);
    return CG_OK;
}

}  // namespace cg

using namespace cg;

extern "C" int cg_block_median(const void *in, void *out, long n_outer, int ny, int nx, int f,
                               int dtype, void *stream)
{
    CG_REQUIRE(in && out, "null pointer");
    CG_REQUIRE(n_outer >= 0 && ny >= 1 && nx >= 1 && f >= 1, "bad shape");
    CG_REQUIRE(ny % f == 0 && nx % f == 0, "dimension sizes (%d, %d) are not multiples of the block size %d", ny, nx, f);
    CG_REQUIRE(dtype == CG_F32 || dtype == CG_F64, "bad dtype");
    const long n_cells = n_outer * (long)(ny / f) * (nx / f);
    if (n_cells == 0) return CG_OK;
    cudaStream_t st = (cudaStream_t)stream;
    const unsigned g = cdiv(n_cells, 128);
    if (dtype == CG_F32) {
        if (f == 8 && nx % 8 == 0 && (((uintptr_t)in) & 15) == 0) {
            median_f8_f32_kernel<<<g, 128, 0, st>>>((const float *)in, (float *)out, n_cells, ny, nx);
        } else if (f == 8) {
            median_small_kernel<float, 8><<<g, 128, 0, st>>>((const float *)in, (float *)out, n_cells, ny, nx);
        } else if (f == 4) {
            median_small_kernel<float, 4><<<g, 128, 0, st>>>((const float *)in, (float *)out, n_cells, ny, nx);
        } else if (f == 2) {
            median_small_kernel<float, 2><<<g, 128, 0, st>>>((const float *)in, (float *)out, n_cells, ny, nx);
        } else {
            return launch_generic_select<float, 0>(in, out, n_cells, ny, nx, f, st, "cg_block_median");
        }
    } else {
        if (f == 4) median_small_kernel<double, 4><<<g, 128, 0, st>>>((const double *)in, (double *)out, n_cells, ny, nx);
        else if (f == 2) median_small_kernel<double, 2><<<g, 128, 0, st>>>((const double *)in, (double *)out, n_cells, ny, nx);
        else return launch_generic_select<double, 0>(in, out, n_cells, ny, nx, f, st, "cg_block_median");
    }
    CG_LAUNCH_CHECK("cg_block_median");
    return CG_OK;
}

extern "C" int cg_block_mode(const void *in, void *out, long n_outer, int ny, int nx, int f,
                             int nan_omit, int dtype, void *stream)
{
    CG_REQUIRE(in && out, "null pointer");
    CG_REQUIRE(n_outer >= 0 && ny >= 1 && nx >= 1 && f >= 1, "bad shape");
    CG_REQUIRE(ny % f == 0 && nx % f == 0, "dimension sizes (%d, %d) are not multiples of the block size %d", ny, nx, f);
    CG_REQUIRE(dtype == CG_F32 || dtype == CG_F64, "bad dtype");
    const long n_cells = n_outer * (long)(ny / f) * (nx / f);
    if (n_cells == 0) return CG_OK;
    cudaStream_t st = (cudaStream_t)stream;
    // small power-of-two blocks: thread per block, register sort (f = 8 fp32 is the sfc_data shape)
    const unsigned g = cdiv(n_cells, 128);
#define CG_MODE_SMALL(T, F)                                                                                          \
    do {                                                                                                             \
        if (nan_omit) mode_small_kernel<T, F, 2><<<g, 128, 0, st>>>((const T *)in, (T *)out, n_cells, ny, nx);       \
        else mode_small_kernel<T, F, 1><<<g, 128, 0, st>>>((const T *)in, (T *)out, n_cells, ny, nx);                \
        CG_LAUNCH_CHECK("cg_block_mode");                                                                            \
        return CG_OK;                                                                                                \
    } while (0)
    if (dtype == CG_F32 && f == 8) CG_MODE_SMALL(float, 8);
    if (dtype == CG_F32 && f == 4) CG_MODE_SMALL(float, 4);
    if (dtype == CG_F32 && f == 2) CG_MODE_SMALL(float, 2);
    if (dtype == CG_F64 && f == 4) CG_MODE_SMALL(double, 4);
    if (dtype == CG_F64 && f == 2) CG_MODE_SMALL(double, 2);
#undef CG_MODE_SMALL
    if (dtype == CG_F32)
        return nan_omit ? launch_generic_select<float, 2>(in, out, n_cells, ny, nx, f, st, "cg_block_mode")
                        : launch_generic_select<float, 1>(in, out, n_cells, ny, nx, f, st, "cg_block_mode");
    return nan_omit ? launch_generic_select<double, 2>(in, out, n_cells, ny, nx, f, st, "cg_block_mode")
                    : launch_generic_select<double, 1>(in, out, n_cells, ny, nx, f, st, "cg_block_mode");
}
