// Issue-rate microbenchmark for sm_100a: which pipes bound the PPM kernel's instruction mix.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/ubench_pipes tools/ubench_pipes.cu
// Prints warp-instructions per clock per SM for each instruction (mix).
#include <cstdio>
#include <cuda_runtime.h>
#define ITERS 2048
#define U 8
typedef unsigned long long u64;

#define KERNEL(name, DECL, BODY, SINK)                                                  \
__global__ void __launch_bounds__(1024, 1) name(float *out, float seed) {               \
    DECL;                                                                               \
    for (int it = 0; it < ITERS; ++it) {                                                \
        _Pragma("unroll") for (int u = 0; u < U; ++u) { BODY; }                         \
    }                                                                                   \
    SINK;                                                                               \
}

#define DECLF float a[U], b = seed, c = seed * 0.5f; \
    _Pragma("unroll") for (int u = 0; u < U; ++u) a[u] = seed + u + threadIdx.x;
#define SINKF { float s = 0; _Pragma("unroll") for (int u = 0; u < U; ++u) s += a[u]; if (s == 12345.f) out[0] = s; }

KERNEL(k_ffma, DECLF, asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a[u]) : "f"(b), "f"(c)), SINKF)
KERNEL(k_fadd, DECLF, asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(a[u]) : "f"(b)), SINKF)
KERNEL(k_fmul, DECLF, asm volatile("mul.rn.f32 %0, %0, %1;" : "+f"(a[u]) : "f"(b)), SINKF)
KERNEL(k_fmnmx, DECLF, asm volatile("min.f32 %0, %0, %1;" : "+f"(a[u]) : "f"(b)), SINKF)
KERNEL(k_fmnmx3, DECLF, asm volatile("min.f32 %0, %0, %1, %2;" : "+f"(a[u]) : "f"(b), "f"(c)), SINKF)
KERNEL(k_lop3, DECLF, asm volatile("lop3.b32 %0, %0, %1, %2, 0xE4;" : "+r"(*(unsigned*)&a[u]) : "r"(__float_as_uint(b)), "r"(__float_as_uint(c))), SINKF)
KERNEL(k_setp_selp, DECLF, asm volatile("{.reg .pred p; setp.lt.f32 p, %0, %1; selp.f32 %0, %2, %0, p;}" : "+f"(a[u]) : "f"(b), "f"(c)), SINKF)
KERNEL(k_ffma_fmnmx, DECLF, asm volatile("fma.rn.f32 %0, %0, %1, %2; min.f32 %0, %0, %2;" : "+f"(a[u]) : "f"(b), "f"(c)), SINKF)
KERNEL(k_ffma3_fmnmx, DECLF, asm volatile("fma.rn.f32 %0, %0, %1, %2; fma.rn.f32 %0, %0, %1, %2; fma.rn.f32 %0, %0, %1, %2; min.f32 %0, %0, %2;" : "+f"(a[u]) : "f"(b), "f"(c)), SINKF)
KERNEL(k_shfl, DECLF, asm volatile("shfl.sync.down.b32 %0, %0, 1, 0x1f, 0xffffffff;" : "+f"(a[u])), SINKF)
KERNEL(k_iadd3, DECLF, asm volatile("add.s32 %0, %0, %1;" : "+r"(*(unsigned*)&a[u]) : "r"(__float_as_uint(b))), SINKF)
KERNEL(k_imad, DECLF, asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(*(unsigned*)&a[u]) : "r"(__float_as_uint(b)), "r"(__float_as_uint(c))), SINKF)

#define DECL2 u64 a[U], b, c; \
    asm("mov.b64 %0, {%1,%1};" : "=l"(b) : "f"(seed)); asm("mov.b64 %0, {%1,%1};" : "=l"(c) : "f"(seed * 0.5f)); \
    _Pragma("unroll") for (int u = 0; u < U; ++u) { float t = seed + u + threadIdx.x; asm("mov.b64 %0, {%1,%1};" : "=l"(a[u]) : "f"(t)); }
#define SINK2 { u64 s = 0; _Pragma("unroll") for (int u = 0; u < U; ++u) s ^= a[u]; if (s == 12345ull) out[0] = 1.f; }
KERNEL(k_ffma2, DECL2, asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(a[u]) : "l"(b), "l"(c)), SINK2)
KERNEL(k_fadd2, DECL2, asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(a[u]) : "l"(b)), SINK2)
KERNEL(k_fmul2, DECL2, asm volatile("mul.rn.f32x2 %0, %0, %1;" : "+l"(a[u]) : "l"(b)), SINK2)
// FFMA2 + 2 x FMNMX on the halves: does the packed op free issue slots for the alu pipe?
KERNEL(k_ffma2_fmnmx2, DECL2, asm volatile("{.reg .f32 lo, hi; fma.rn.f32x2 %0, %0, %1, %2; mov.b64 {lo,hi}, %0; min.f32 lo, lo, 3.0; min.f32 hi, hi, 3.0; mov.b64 %0, {lo,hi};}" : "+l"(a[u]) : "l"(b), "l"(c)), SINK2)
KERNEL(k_ffma2x2_fmnmx2, DECL2, asm volatile("{.reg .f32 lo, hi; fma.rn.f32x2 %0, %0, %1, %2; fma.rn.f32x2 %0, %0, %1, %2; mov.b64 {lo,hi}, %0; min.f32 lo, lo, 3.0; min.f32 hi, hi, 3.0; mov.b64 %0, {lo,hi};}" : "+l"(a[u]) : "l"(b), "l"(c)), SINK2)

// shared memory loads: conflict-free 32-bit, 64-bit, 128-bit
__global__ void __launch_bounds__(1024, 1) k_lds32(float *out, float seed) {
    __shared__ float s[4096];
    for (int i = threadIdx.x; i < 4096; i += 1024) s[i] = seed + i;
    __syncthreads();
    float acc = 0; int idx = threadIdx.x;
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int u = 0; u < U; ++u) { float v; asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"((unsigned)__cvta_generic_to_shared(&s[(idx + u * 32) & 4095]))); acc += v; }
    }
    if (acc == 12345.f) out[0] = acc;
}
__global__ void __launch_bounds__(1024, 1) k_lds128(float *out, float seed) {
    __shared__ float4 s[1024];
    for (int i = threadIdx.x; i < 1024; i += 1024) s[i] = make_float4(seed, seed, seed, seed);
    __syncthreads();
    float acc = 0; int idx = threadIdx.x;
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int u = 0; u < U; ++u) { float4 v; asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"((unsigned)__cvta_generic_to_shared(&s[(idx + u * 32) & 1023]))); acc += v.x + v.w; }
    }
    if (acc == 12345.f) out[0] = acc;
}

template <typename K> void run(const char *name, K kern, double inst_per_body, float *d) {
    int dev; cudaGetDevice(&dev); cudaDeviceProp p; cudaGetDeviceProperties(&p, dev);
    int clk_khz; cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, dev);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    kern<<<p.multiProcessorCount, 1024>>>(d, 1.0f);
    cudaDeviceSynchronize();
    float best = 1e9;
    for (int r = 0; r < 3; ++r) {
        cudaEventRecord(e0); kern<<<p.multiProcessorCount, 1024>>>(d, 1.0f); cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
    }
    double warp_inst = 32.0 * ITERS * U * inst_per_body;          // per SM (32 warps)
    double cycles = best * 1e-3 * clk_khz * 1e3;                 // at the max clock (upper bound on cycles)
    printf("%-18s %8.3f ms  %6.3f warp-inst/clk/SM (at %d MHz nominal)  err=%s\n", name, best, warp_inst / cycles, clk_khz / 1000, cudaGetErrorString(cudaGetLastError()));
}

int main() {
    float *d; cudaMalloc(&d, 1024);
    run("FFMA", k_ffma, 1, d); run("FADD", k_fadd, 1, d); run("FMUL", k_fmul, 1, d);
    run("FMNMX", k_fmnmx, 1, d); run("FMNMX3", k_fmnmx3, 1, d); run("LOP3", k_lop3, 1, d);
    run("FSETP+FSEL", k_setp_selp, 2, d); run("IADD", k_iadd3, 1, d); run("IMAD", k_imad, 1, d);
    run("FFMA+FMNMX", k_ffma_fmnmx, 2, d); run("3FFMA+FMNMX", k_ffma3_fmnmx, 4, d);
    run("SHFL", k_shfl, 1, d);
    run("FFMA2", k_ffma2, 1, d); run("FADD2", k_fadd2, 1, d); run("FMUL2", k_fmul2, 1, d);
    run("FFMA2+2FMNMX", k_ffma2_fmnmx2, 3, d); run("2FFMA2+2FMNMX", k_ffma2x2_fmnmx2, 4, d);
    run("LDS.32", k_lds32, 1, d); run("LDS.128", k_lds128, 1, d);
    return 0;
}
