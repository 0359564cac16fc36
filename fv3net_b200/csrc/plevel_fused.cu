// Fused pressure-level coarse-graining kernel (placeholder until the fused kernel lands:
// fused_supported() == 0 routes every call to the composed path in plevel.cu).
#include "common.cuh"
namespace cg {
int fused_supported(int, int, int, int) { return 0; }
long fused_workspace_bytes(int, long, int, int, int, int) { return 0; }
int launch_plevel_fused(const cg_field *, float *const *, int, const float *, const float *, float *,
                        long, int, int, int, int, int, float, void *, long, cudaStream_t)
{
    set_error("fused pressure-level kernel not built");
    return CG_ENOTSUP;
}
}  // namespace cg
