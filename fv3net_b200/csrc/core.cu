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

extern "C" int cg_version(void) { return 100; }

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
