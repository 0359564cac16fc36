// Micro-benchmark of the field warps' hot step (design probe, never part of the product path).
// Variants: pairs per thread (NP), steps unrolled per loop body (U), warps per CTA, CTAs per SM.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o tools/proto_step.bin tools/proto_step.cu
// The shared-memory contents are synthetic but shaped like the real records (6 x 16 B per column pair and level,
// broadcast across the lanes of a warp), field rows (one 128-B row per lane) and accumulators ([slot][lane]).
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>
typedef unsigned long long u64;

__device__ __forceinline__ u64 pk(float lo, float hi) { u64 r; asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void upk(u64 v, float &lo, float &hi) { asm("mov.b64 {%0,%1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) { u64 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ u64 add2(u64 a, u64 b) { u64 r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ u64 sub2(u64 a, u64 b) { u64 r; asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ u64 mul2(u64 a, u64 b) { u64 r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ void lds2x64(uint32_t addr, u64 &a, u64 &b) { asm volatile("ld.shared.v2.b64 {%0,%1}, [%2];" : "=l"(a), "=l"(b) : "r"(addr)); }
__device__ __forceinline__ u64 lds64(uint32_t addr) { u64 a; asm volatile("ld.shared.b64 %0, [%1];" : "=l"(a) : "r"(addr)); return a; }
__device__ __forceinline__ int4 lds128i(uint32_t addr) { int4 v; asm volatile("ld.shared.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr)); return v; }
__device__ __forceinline__ void red_add(uint32_t addr, float v)
{
    float o;
    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(o) : "r"(addr));
    asm volatile("st.shared.f32 [%0], %1;" ::"r"(addr), "f"(o + v) : "memory");
}
__device__ __forceinline__ float absmin3_sign(float a, float b, float c, float s) { return copysignf(fminf(fminf(fabsf(a), fabsf(b)), fabsf(c)), s); }
__device__ __forceinline__ float absmin2_sign(float a, float b, float s) { return copysignf(fminf(fabsf(a), fabsf(b)), s); }
__device__ __forceinline__ float gt01(float a, float b) { float r; asm("set.gt.f32.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b)); return r; }

struct PairRec { u64 E1, E2, Fm, G, NH, DPW, EX, EY, EZ, NW; int4 oi; };
__device__ __forceinline__ void load_pair_rec(PairRec &r, uint32_t rb)
{
    lds2x64(rb, r.E1, r.E2);        lds2x64(rb + 16, r.Fm, r.G);    lds2x64(rb + 32, r.NH, r.DPW);
    lds2x64(rb + 48, r.EX, r.EY);   lds2x64(rb + 64, r.EZ, r.NW);   r.oi = lds128i(rb + 80);
}
// rolling state of one column pair; the q history is a 3-slot ring so that unrolled steps need no register moves
struct Pair { u64 q[3], dq[3], dc[3], xp[3], mk[3], run; };

template <int S, int MODE>   // S = step index mod 3 (compile time in unrolled code)
__device__ __forceinline__ void pair_step(const PairRec &r, Pair &s, u64 qn, uint32_t acc_lane, int &any, float &regacc)
{
    constexpr int c = S % 3, a = (S + 2) % 3, b = (S + 1) % 3;   // new, t-1, t-2
    const u64 A = s.q[a], B = s.q[b];
    const u64 dq1 = s.dq[a];
    const u64 dqn = sub2(qn, A);
    const u64 df2 = fma2(r.E1, dqn, mul2(r.E2, dq1));
    float dfa, dfb, d1a, d1b, dna, dnb, saa, sab, sda, sdb;
    upk(df2, dfa, dfb); upk(dq1, d1a, d1b); upk(dqn, dna, dnb);
    upk(add2(dq1, dqn), saa, sab); upk(sub2(dq1, dqn), sda, sdb);
    const u64 mkn = pk(gt01(fabsf(saa), fabsf(sda)), gt01(fabsf(sab), fabsf(sdb)));
    const u64 dcn = mul2(pk(absmin3_sign(dfa, d1a, dna, d1a), absmin3_sign(dfb, d1b, dnb, d1b)), mkn);
    const u64 xn = fma2(r.NH, dcn, fma2(r.G, s.dc[a], mul2(r.Fm, dq1)));
    const u64 x = s.xp[a], y = mul2(add2(xn, dq1), s.mk[a]);
    float xa, xb, ya, yb, x2a, x2b, y2a, y2b;
    upk(x, xa, xb); upk(y, ya, yb); upk(add2(x, x), x2a, x2b); upk(add2(y, y), y2a, y2b);
    const u64 xl = pk(absmin2_sign(xa, y2a, xa), absmin2_sign(xb, y2b, xb));
    const u64 yl = pk(absmin2_sign(ya, x2a, ya), absmin2_sign(yb, x2b, yb));
    const u64 I = fma2(r.EZ, yl, fma2(r.EY, xl, mul2(r.EX, B)));
    float ta, tb;
    upk(add2(s.run, I), ta, tb);
    if (MODE & 1) regacc += ta + tb;
    else {
        red_add(acc_lane + r.oi.x, ta);
        red_add(acc_lane + r.oi.y, tb);
    }
    any |= r.oi.z | r.oi.w;
    s.run = fma2(r.DPW, B, mul2(r.NW, I));
    s.q[c] = qn; s.dq[c] = dqn; s.dc[c] = dcn; s.xp[c] = xn; s.mk[c] = mkn;
}

constexpr int REC_LEVEL = 16 * 96, NLEV = 6, QROW = 128, ACC_SLOTS = 80;
// row stride between lanes: conflict-free like the swizzled TMA boxes (8 lanes x 16 B or 16 lanes x 8 B per wavefront)
template <int NP> __host__ __device__ constexpr int qbox() { return NLEV * QROW + (NP == 1 ? 8 : 16); }

template <int NP, int NW>
struct Smem {
    static constexpr int rec = 0, q = rec + NLEV * REC_LEVEL, acc = q + 32 * (NLEV * QROW + 128), total = acc + ACC_SLOTS * 128 * NW;
};

template <int NP, int U, int NW, int MINB, int MODE>
__global__ void __launch_bounds__(NW * 32, MINB) proto(const float *init, float *out, int nsteps, long long *cycles)
{
    extern __shared__ __align__(16) unsigned char smem[];
    typedef Smem<NP, NW> L;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const uint32_t sb = (uint32_t)__cvta_generic_to_shared(smem);
    for (int i = threadIdx.x; i < L::acc / 4; i += blockDim.x) ((float *)smem)[i] = init[i];
    for (int i = threadIdx.x; i < (L::total - L::acc) / 4; i += blockDim.x) ((float *)(smem + L::acc))[i] = 0.f;
    // offsets / infos of the records: valid accumulator slots
    for (int i = threadIdx.x; i < NLEV * 16; i += blockDim.x) {
        int *oi = (int *)(smem + L::rec + i * 96 + 80);
        oi[0] = ((i * 7) % ACC_SLOTS) * 128; oi[1] = ((i * 7 + 1) % ACC_SLOTS) * 128; oi[2] = 1; oi[3] = 1;
    }
    __syncthreads();
    const uint32_t acc_lane = sb + L::acc + w * ACC_SLOTS * 128 + 4 * lane;
    // which pairs this thread owns: NP consecutive pairs starting at (w * NP) % 16 (warp-uniform, broadcast loads)
    const int p0 = (w * NP) % 16;
    Pair s[NP];
#pragma unroll
    for (int p = 0; p < NP; ++p) {
#pragma unroll
        for (int i = 0; i < 3; ++i) { s[p].q[i] = 0; s[p].dq[i] = 0; s[p].dc[i] = 0; s[p].xp[i] = 0; s[p].mk[i] = 0; }
        s[p].run = 0;
    }
    int any = 0;
    float regacc = 0.f;
    PairRec fixed[NP];
    if (MODE & 2) {
#pragma unroll
        for (int p = 0; p < NP; ++p) load_pair_rec(fixed[p], sb + L::rec + (p0 + p) * 96);
    }
    const long long c0 = clock64();
    const uint32_t qlane = sb + L::q + lane * qbox<NP>() + p0 * 8;
    const uint32_t recw = sb + L::rec + p0 * 96;
#pragma unroll 1
    for (int t = 0; t < nsteps; t += U) {
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int lv = u % NLEV;                  // (U divides NLEV or is a multiple of it)
            u64 q[NP];
            if (NP == 1) q[0] = lds64(qlane + lv * QROW);
            else {
#pragma unroll
                for (int p = 0; p < NP; p += 2) lds2x64(qlane + lv * QROW + 8 * p, q[p], q[p + 1]);
            }
#pragma unroll
            for (int p = 0; p < NP; ++p) {
                PairRec r;
                if (MODE & 2) r = fixed[p];
                else load_pair_rec(r, recw + lv * REC_LEVEL + p * 96);
                if (u % 3 == 0) pair_step<0, MODE>(r, s[p], q[p], acc_lane, any, regacc);
                else if (u % 3 == 1) pair_step<1, MODE>(r, s[p], q[p], acc_lane, any, regacc);
                else pair_step<2, MODE>(r, s[p], q[p], acc_lane, any, regacc);
            }
            if (any > 1) { out[t] = 1.f; any = 0; }
        }
    }
    const long long c1 = clock64();
    if (threadIdx.x == 0 && blockIdx.x == 0) cycles[0] = c1 - c0;
    float acc = regacc;
#pragma unroll
    for (int p = 0; p < NP; ++p) { float a, b; upk(s[p].run, a, b); acc += a + b; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

template <int NP, int U, int NW, int MINB, int MODE = 0>
static void run(const char *name, const float *dinit, float *dout, int nsm, double mhz)
{
    typedef Smem<NP, NW> L;
    auto k = proto<NP, U, NW, MINB, MODE>;
    static long long *dcyc = nullptr;
    if (!dcyc) cudaMalloc(&dcyc, 8);
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, L::total);
    cudaFuncSetAttribute(k, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
    int occ = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k, NW * 32, L::total);
    cudaFuncAttributes fa;
    cudaFuncGetAttributes(&fa, k);
    const int nsteps = 6 * 600;
    const int grid = nsm * occ;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<<<grid, NW * 32, L::total>>>(dinit, dout, nsteps, dcyc);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    k<<<grid, NW * 32, L::total>>>(dinit, dout, nsteps, dcyc);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    // warp-pair-steps per SM, and SM cycles spent per warp-pair-step per scheduler (4 schedulers)
    const double wps = (double)occ * NW * NP * nsteps;
    long long hc = 0;
    cudaMemcpy(&hc, dcyc, 8, cudaMemcpyDeviceToHost);
    const double cyc = (double)hc;   // SM cycles of CTA 0's loop (clock64), independent of the clock frequency
    (void)mhz;
    printf("%-30s regs %3d occ %d warps/SM %2d  %.3f ms (%.0f MHz eff)  %.1f sched-cycles per warp-pair-step  err=%s\n", name, fa.numRegs, occ,
           occ * NW, ms, cyc / (ms * 1e3), 4.0 * cyc / wps, cudaGetErrorString(cudaGetLastError()));
}

int main()
{
    cudaDeviceProp prop;
    cudaGetDeviceProperties(&prop, 0);
    int khz = 0;
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    const double mhz = khz / 1000.0;
    const int n = 64 * 1024;
    std::vector<float> h(n);
    srand(1);
    for (auto &v : h) v = (rand() % 2001 - 1000) / 1000.0f;
    float *dinit, *dout;
    cudaMalloc(&dinit, n * 4);
    cudaMalloc(&dout, 4 << 20);
    cudaMemcpy(dinit, h.data(), n * 4, cudaMemcpyHostToDevice);
    printf("%s, %d SMs, %.0f MHz nominal\n", prop.name, prop.multiProcessorCount, mhz);
    const int nsm = prop.multiProcessorCount;
    run<4, 3, 4, 2>("NP4 U3 4w x2", dinit, dout, nsm, mhz);
    run<4, 3, 4, 2, 1>("NP4 U3 4w x2 noacc", dinit, dout, nsm, mhz);
    run<4, 3, 4, 2, 2>("NP4 U3 4w x2 norec", dinit, dout, nsm, mhz);
    run<4, 3, 4, 2, 3>("NP4 U3 4w x2 noacc norec", dinit, dout, nsm, mhz);
    run<2, 6, 8, 1>("NP2 U6 8w x1", dinit, dout, nsm, mhz);
    run<2, 6, 8, 1, 1>("NP2 U6 8w x1 noacc", dinit, dout, nsm, mhz);
    run<2, 6, 8, 1, 3>("NP2 U6 8w x1 noacc norec", dinit, dout, nsm, mhz);
    run<2, 6, 8, 2>("NP2 U6 8w x2", dinit, dout, nsm, mhz);
    run<1, 6, 13, 1>("NP1 U6 13w x1", dinit, dout, nsm, mhz);
    run<1, 6, 13, 1, 1>("NP1 U6 13w x1 noacc", dinit, dout, nsm, mhz);
    run<1, 6, 13, 1, 3>("NP1 U6 13w x1 noacc norec", dinit, dout, nsm, mhz);
    run<1, 6, 16, 2>("NP1 U6 16w x2", dinit, dout, nsm, mhz);
    run<1, 6, 16, 2, 3>("NP1 U6 16w x2 noacc norec", dinit, dout, nsm, mhz);
    return 0;
}
