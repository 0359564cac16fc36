// Library-wide helpers: thread-local error text, device queries.
#include "common.cuh"
#include <cstring>
#include <atomic>

namespace cg {

static thread_local char g_err[512] = "";

void set_error(const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

// Behaviour switches of the calling thread (cg_set_option): tests and A/B timing pin an implementation with them.
static thread_local Options g_opt;
const Options &options() { return g_opt; }

static std::atomic<long> g_launches{0};
void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }

int sm_count()
{
    static thread_local int cached = 0, cached_dev = -1;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 148;
    if (dev != cached_dev) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        cached = n;
        cached_dev = dev;
    }
    return cached;
}

}  // namespace cg

extern "C" const char *cg_last_error(void) { return cg::g_err; }

extern "C" int cg_version(void) { return 101; }

extern "C" int cg_set_option(const char *name, const char *value)
{
    CG_REQUIRE(name && value, "null option");
    cg::Options &o = cg::g_opt;
    if (!strcmp(name, "plevel_impl")) {
        static const char *const names[] = {"auto", "pair", "stream", "fused", "composed"};
        for (int i = 0; i < 5; ++i)
            if (!strcmp(value, names[i])) { o.plevel_impl = i; return CG_OK; }
    } else if (!strcmp(name, "plevel_nansafe")) {
        if (!strcmp(value, "0") || !strcmp(value, "1")) { o.plevel_nansafe = value[0] == '1'; return CG_OK; }
    } else if (!strcmp(name, "mappm_arrays")) {
        static const char *const names[] = {"auto", "shared"};
        for (int i = 0; i < 2; ++i)
            if (!strcmp(value, names[i])) { o.mappm_arrays = i; return CG_OK; }
    } else if (!strcmp(name, "mappm_impl")) {
        static const char *const names[] = {"auto", "generic"};
        for (int i = 0; i < 2; ++i)
            if (!strcmp(value, names[i])) { o.mappm_impl = i; return CG_OK; }
    } else if (!strcmp(name, "exposed_area_impl")) {
        static const char *const names[] = {"auto", "strip", "cell", "three"};
        for (int i = 0; i < 4; ++i)
            if (!strcmp(value, names[i])) { o.exposed_impl = i; return CG_OK; }
    } else {
        cg::set_error("cg_set_option: unknown option '%s'", name);
        return CG_EINVAL;
    }
    cg::set_error("cg_set_option: '%s' is not a value of option '%s'", value, name);
    return CG_EINVAL;
}

extern "C" long cg_launch_count(void) { return cg::g_launches.load(std::memory_order_relaxed); }

extern "C" int cg_device_info(int *sm, int *cc, long *total_mem_bytes)
{
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return cg::cuda_fail(e, "cudaGetDevice");
    cudaDeviceProp p;
    e = cudaGetDeviceProperties(&p, dev);
    if (e != cudaSuccess) return cg::cuda_fail(e, "cudaGetDeviceProperties");
    if (sm) *sm = p.multiProcessorCount;
    if (cc) *cc = p.major * 10 + p.minor;
    if (total_mem_bytes) *total_mem_bytes = (long)p.totalGlobalMem;
    return CG_OK;
}

// ---- NVLS broadcast of a buffer region (the exchange step of the sharded pressure-level path) -------------------
// Every 16 bytes of [src, src + nbytes) are stored once to src + mc_offset, the multicast mapping of the symmetric
// buffer src lives in: the NVSwitch writes them into every GPU's copy (this GPU's included).  Warps store 512
// contiguous bytes per instruction, so the fabric sees full-size packets - unlike stores issued from inside the
// compute kernel, which can only make 16 bytes contiguous.
namespace cg {
__global__ void __launch_bounds__(256) multicast_store_kernel(const uint4 *__restrict__ src, long long mc_offset, long n16)
{
    const long stride = (long)gridDim.x * blockDim.x;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < n16; i += stride) {
        const uint4 v = src[i];
        asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"((const char *)(src + i) + mc_offset),
                     "f"(__uint_as_float(v.x)), "f"(__uint_as_float(v.y)), "f"(__uint_as_float(v.z)), "f"(__uint_as_float(v.w))
                     : "memory");
    }
}
}  // namespace cg

extern "C" int cg_multicast_store(const void *src, long long mc_offset, long nbytes, void *stream)
{
    CG_REQUIRE(src, "null pointer");
    CG_REQUIRE(nbytes >= 0 && nbytes % 16 == 0 && ((uintptr_t)src & 15) == 0 && mc_offset % 16 == 0,
               "cg_multicast_store moves 16-byte pieces: pointer, offset and size must be multiples of 16");
    if (nbytes == 0) return CG_OK;
    const long n16 = nbytes / 16;
    long blocks = (n16 + 255) / 256;
    const long cap = 8L * cg::sm_count();
    if (blocks > cap) blocks = cap;
    cg::multicast_store_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>((const uint4 *)src, mc_offset, n16);
    CG_LAUNCH_CHECK("cg_multicast_store");
    return CG_OK;
}
