// cg_interpolate_2d — row-wise linear interpolation, the second entry point of the reference's f2py
// module (external/mappm/mappm/interpolate_2d.f90:1-29, called from vcm/interpolate.py:176-180).
// For every row i and target xp(i, j): the LAST source interval k (the Fortran's search does not stop at
// the first hit) with x(i,k) <= xp < x(i,k+1) gives y(i,k) (1 - w) + y(i,k+1) w; a target equal to a
// source point takes that point's value; anything else keeps fill_value.  float64 like the Fortran's
// real(8); compiled with --fmad=false so the two products and the sum round as the Fortran's do.
// One thread per target value; the lanes of a warp share a row wherever n_out >= 32, so the source row
// is read through broadcast loads.  Rows are addressed by (row, level) strides: the z-last layout of the
// reference and a (z, column) layout both work without a transpose.
#include "common.cuh"

namespace cg {

__global__ void interpolate_2d_kernel(const double *__restrict__ xp, const double *__restrict__ x,
                                      const double *__restrict__ y, double *__restrict__ out, double fill,
                                      long m, int n_in, int n_out, long rs_p, long ls_p, long rs_x, long ls_x,
                                      long rs_y, long ls_y, long rs_o, long ls_o)
{
    const long n = m * n_out;
    for (long t = (long)blockIdx.x * blockDim.x + threadIdx.x; t < n; t += (long)gridDim.x * blockDim.x) {
        const long i = t / n_out;
        const int j = (int)(t - i * n_out);
        const double p = xp[i * rs_p + j * ls_p];
        const double *xr = x + i * rs_x, *yr = y + i * rs_y;
        double r = fill;
        double x0 = xr[0], y0 = yr[0];
        for (int k = 0; k + 1 < n_in; ++k) {
            const double x1 = xr[(long)(k + 1) * ls_x], y1 = yr[(long)(k + 1) * ls_y];
            if (x0 <= p && p < x1) {
                const double w = (p - x0) / (x1 - x0);
                r = y0 * (1.0 - w) + y1 * w;
            } else if (x0 == p) {
                r = y0;
            } else if (x1 == p) {
                r = y1;
            }
            x0 = x1;
            y0 = y1;
        }
        out[i * rs_o + j * ls_o] = r;
    }
}

}  // namespace cg

extern "C" int cg_interpolate_2d(const double *xp, const double *x, const double *y, double *y_out, double fill_value,
                                 long m, int n_in, int n_out, const long *strides, void *stream)
{
    using namespace cg;
    CG_REQUIRE(xp && x && y && y_out && strides, "cg_interpolate_2d: null pointer");
    CG_REQUIRE(m >= 0 && n_in >= 1 && n_out >= 0, "cg_interpolate_2d: bad shape (m=%ld, n_in=%d, n_out=%d)", m, n_in, n_out);
    const long n = m * n_out;
    if (n == 0) return CG_OK;
    const int threads = 128;
    long blocks = (n + threads - 1) / threads;
    const long cap = (long)sm_count() * 16;
    if (blocks > cap) blocks = cap;
    interpolate_2d_kernel<<<(unsigned)blocks, threads, 0, (cudaStream_t)stream>>>(
        xp, x, y, y_out, fill_value, m, n_in, n_out, strides[0], strides[1], strides[2], strides[3], strides[4],
        strides[5], strides[6], strides[7]);
    CG_LAUNCH_CHECK("cg_interpolate_2d");
    return CG_OK;
}
