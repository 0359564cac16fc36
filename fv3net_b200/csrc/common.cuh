// Shared helpers for libcoarsen_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdarg>
#include <cstdint>
#include "../../include/coarsen_b200.h"

namespace cg {

// thread-local last-error text and option block: the only state in the library, none of it process-wide
void set_error(const char *fmt, ...);

enum { PLEVEL_AUTO, PLEVEL_PAIR, PLEVEL_STREAM, PLEVEL_FUSED, PLEVEL_COMPOSED };
struct Options {
    int plevel_impl = PLEVEL_AUTO;   // which cg_pressure_level_average implementation runs where its shape constraints hold
    bool plevel_nansafe = false;     // the NaN-skipping variant alone, without the fast first pass
    int mappm_arrays = 0;            // 1: the generic remapper keeps its column arrays in shared memory
    int mappm_impl = 0;              // 1: the generic remapper even where the streaming one applies
    int exposed_impl = 0;            // cg_area_above_fine_surface: 0 auto, 1 strip, 2 warp per cell, 3 three kernels
};
const Options &options();            // of the calling thread (cg_set_option)

inline int cuda_fail(cudaError_t e, const char *what)
{
    set_error("%s: %s", what, cudaGetErrorString(e));
    return CG_ECUDA;
}

#define CG_REQUIRE(cond, ...)                 \
    do {                                      \
        if (!(cond)) {                        \
            cg::set_error(__VA_ARGS__);       \
            return CG_EINVAL;                 \
        }                                     \
    } while (0)

// every kernel launch of the library passes through here once (cg_launch_count())
void count_launch();

#define CG_LAUNCH_CHECK(what)                                  \
    do {                                                       \
        cg::count_launch();                                    \
        cudaError_t e__ = cudaGetLastError();                  \
        if (e__ != cudaSuccess) return cg::cuda_fail(e__, what); \
    } while (0)

// number of SMs of the current device (cached)
int sm_count();

template <typename T> struct Vec4;
template <> struct Vec4<float> { using type = float4; };
template <> struct Vec4<double> { using type = double4; };

template <typename T> __device__ __forceinline__ bool is_nan(T v) { return v != v; }

// streaming 128-bit load: read-once data should not displace L1 lines
__device__ __forceinline__ float4 ldg_stream4(const float *p)
{
    float4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ float ldg_stream(const float *p)
{
    float r;
    asm volatile("ld.global.nc.L1::no_allocate.f32 %0, [%1];" : "=f"(r) : "l"(p));
    return r;
}

template <typename A, typename B> struct Promote { using type = double; };
template <> struct Promote<float, float> { using type = float; };

inline unsigned cdiv(long a, long b) { return (unsigned)((a + b - 1) / b); }

}  // namespace cg
