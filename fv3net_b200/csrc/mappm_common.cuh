// Shared by the two PPM remappers (mappm_generic.cu, mappm_stream.cu; both compiled with --fmad=false so that every
// statement keeps the Fortran's operation order and rounding): limiter of one layer, strides, target-pressure blocks.
#pragma once
#include "common.cuh"

namespace cg {

constexpr int STREAM_TPB = 128;    // columns per block of the streaming remapper = columns per flag

__device__ __forceinline__ float max3f(float a, float b, float c) { return fmaxf(fmaxf(a, b), c); }
__device__ __forceinline__ float min3f(float a, float b, float c) { return fminf(fminf(a, b), c); }
__device__ __forceinline__ float curv(float q, float al, float ar) { return 3.f * (2.f * q - (al + ar)); }

// ppm_limiters, one layer (mappm.f90:854-931)
__device__ __forceinline__ void limit_layer(float dm, float q, float &al, float &ar, float &a6, int lmt)
{
    if (lmt == 3) return;
    if (lmt == 0) {
        if (dm == 0.f) {
            al = q; ar = q; a6 = 0.f;
        } else {
            const float da1 = ar - al;
            const float da2 = da1 * da1;
            const float a6da = a6 * da1;
            if (a6da < -da2) {
                a6 = 3.f * (al - q);
                ar = al - a6;
            } else if (a6da > da2) {
                a6 = 3.f * (ar - q);
                al = ar - a6;
            }
        }
    } else if (lmt == 1) {
        const float qmp = 2.f * dm;
        al = q - copysignf(fminf(fabsf(qmp), fabsf(al - q)), qmp);
        ar = q + copysignf(fminf(fabsf(qmp), fabsf(ar - q)), qmp);
        a6 = 3.f * (2.f * q - (al + ar));
    } else {
        const float r12 = 1.f / 12.f;
        if (fabsf(ar - al) < -a6) {
            const float d = ar - al;
            const float fmin = q + 0.25f * (d * d) / a6 + a6 * r12;
            if (fmin < 0.f) {
                if (q < ar && q < al) {
                    ar = q; al = q; a6 = 0.f;
                } else if (ar > al) {
                    a6 = 3.f * (al - q);
                    ar = al - a6;
                } else {
                    a6 = 3.f * (ar - q);
                    al = ar - a6;
                }
            }
        }
    }
}


struct Strides { long v[12]; };
// f > 0: target pressures live on the coarse grid; fine column i = (y, x) of a
// (ny, nx) tile reads coarse column (y/f, x/f) of a (ny/f, nx/f) tile.
struct Pe2Block { int f, nx, nxc; };


}  // namespace cg
