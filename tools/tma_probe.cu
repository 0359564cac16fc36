// Probe: how a [KC][1][32 floats] box lands in shared memory under CU_TENSOR_MAP_SWIZZLE_128B
// when the destination is 128-byte aligned but not 1024-byte aligned (box stride 640 B).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o build/tma_probe tools/tma_probe.cu
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <vector>

typedef CUresult (*EncodeFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                             const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                             CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__global__ void probe(const __grid_constant__ CUtensorMap tm, float *out, int x0, int y0, int z0, int dst_off)
{
    extern __shared__ __align__(1024) unsigned char smem[];
    __shared__ __align__(8) unsigned long long bar;
    float *fs = (float *)smem;
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) fs[i] = -1.f;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar)), "r"(4 * 128));
        asm volatile(
            "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
            ::"r"(smem_u32(smem + dst_off)), "l"(&tm), "r"(x0), "r"(y0), "r"(z0), "r"(smem_u32(&bar))
            : "memory");
    }
    uint32_t done = 0;
    while (!done) {
        asm volatile("{.reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p;}"
                     : "=r"(done) : "r"(smem_u32(&bar)), "r"(0));
    }
    __syncthreads();
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) out[i] = fs[i];
}

int main()
{
    const int NX = 64, NY = 4, NZ = 8;
    std::vector<float> h(NX * NY * NZ);
    for (int z = 0; z < NZ; ++z)
        for (int y = 0; y < NY; ++y)
            for (int x = 0; x < NX; ++x) h[(z * NY + y) * NX + x] = z * 10000 + y * 100 + x;
    float *d, *o;
    cudaMalloc(&d, h.size() * 4);
    cudaMalloc(&o, 4096);
    cudaMemcpy(d, h.data(), h.size() * 4, cudaMemcpyHostToDevice);
    EncodeFn enc = nullptr;
    cudaDriverEntryPointQueryResult q;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", (void **)&enc, cudaEnableDefault, &q);
    if (!enc) { printf("no cuTensorMapEncodeTiled\n"); return 1; }
    CUtensorMap tm;
    cuuint64_t dims[3] = {NX, NY, NZ};
    cuuint64_t strides[2] = {NX * 4, NX * NY * 4};
    cuuint32_t box[3] = {32, 1, 4};
    cuuint32_t es[3] = {1, 1, 1};
    CUresult r = enc(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("encode rc=%d\n", (int)r);
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 8192);
    for (int dst_off : {0, 640, 128}) {
        probe<<<1, 128, 4096>>>(tm, o, 32, 2, 1, dst_off);
        cudaError_t e = cudaDeviceSynchronize();
        printf("dst_off=%d err=%s\n", dst_off, cudaGetErrorString(e));
        std::vector<float> res(1024);
        cudaMemcpy(res.data(), o, 4096, cudaMemcpyDeviceToHost);
        for (int row = 0; row < 12; ++row) {
            printf("row %2d (byte %4d):", row, row * 128);
            for (int c = 0; c < 8; ++c) printf(" %6.0f", res[row * 32 + c * 4]);
            printf("\n");
        }
    }
    // z beyond the tensor: zero fill and full tx count?
    probe<<<1, 128, 4096>>>(tm, o, 32, 2, 6, 0);
    cudaError_t e = cudaDeviceSynchronize();
    printf("oob z: err=%s\n", cudaGetErrorString(e));
    std::vector<float> res(1024);
    cudaMemcpy(res.data(), o, 4096, cudaMemcpyDeviceToHost);
    for (int row = 0; row < 4; ++row) {
        printf("row %2d:", row);
        for (int c = 0; c < 8; ++c) printf(" %6.0f", res[row * 32 + c * 4]);
        printf("\n");
    }
    return 0;
}
