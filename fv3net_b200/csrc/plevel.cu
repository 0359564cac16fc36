// cg_pressure_level_average / cg_area_above_fine_surface — pressure-level
// coarse-graining (regridz.py:115-147 + coarsen.py:183; budgets.py:49-77).
//
// Two implementations sit behind one C entry point:
//   * fused (plevel_fused.cu): one kernel per batch of fields, nothing fine-grid
//     is written; used when the shape fits its specialisation (fp32, f = 8,
//     nz <= 96, nx % 8 == 0).
//   * composed (this file): coarse delp -> interface pressures -> cg_mappm with
//     block-shared target pressures -> (masked) weighted block average, through a
//     caller-provided workspace.  Any f / nz; also the cross-check for the fused
//     kernel in tests (CG_PLEVEL_IMPL=composed).
#include "common.cuh"
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <mutex>

namespace cg {

int launch_mappm(const float *pe1, const float *q1, const float *pe2, float *q2,
                 long n_outer, long n_inner, int km, int kn, const long *strides,
                 int pe2_f, int pe2_nx, int pe2_nxc, int iv, int kord, cudaStream_t st);
int fused_supported(int nz, int ny, int nx, int f);
int launch_plevel_fused(const cg_field *fields, float *const *outs, int nfields,
                        const float *delp, const float *area, float *delp_coarse,
                        long n_outer, int nz, int ny, int nx, int f, int mask_below, float toa,
                        void *workspace, long workspace_bytes, cudaStream_t st);
long fused_workspace_bytes(int nfields, long n_outer, int nz, int ny, int nx, int f);
// plevel_stream.cu: lanes-over-fields streaming kernel (TMA ring, shared-memory geometry)
int stream_supported(int nfields, int nz, int ny, int nx, int f);
long stream_workspace_bytes(long n_outer, int nz, int ny, int nx);
int launch_surface_pressure_seq(const float *delp, float *ps, float toa, long n_outer, int nz, long h, cudaStream_t st);
int pair_supported(int nfields, int nz, int ny, int nx, int f);
int launch_plevel_pair(const cg_field *fields, float *const *outs, int nfields, const float *delp,
                       const float *area, const float *delp_coarse, const float *psurf, int *flags, long n_outer,
                       int nz, int ny, int nx, int mask_below, float toa, int nansafe_only, int nrep, const long long *rep,
                       int rep_dc, int rep_kind, cudaStream_t st);
int launch_plevel_stream(const cg_field *fields, float *const *outs, int nfields, const float *delp,
                         const float *area, const float *delp_coarse, const float *psurf, int *flags, long n_outer,
                         int nz, int ny, int nx, int mask_below, float toa, cudaStream_t st);

// masked[o,k,y,x] = pe_c[o,k+1,y/f,x/f] < pe_f[o,nz,y,x] ? area[o,y,x] : 0   (regridz.py:150-161)
__global__ void __launch_bounds__(256)
mask_from_coarse_kernel(const float *__restrict__ area, const float *__restrict__ pe_c,
                        const float *__restrict__ pe_f, float *__restrict__ out, long n_outer,
                        int nz, int ny, int nx, int f)
{
    const int nyc = ny / f, nxc = nx / f;
    const long h = (long)ny * nx, hc = (long)nyc * nxc;
    const long total = n_outer * nz * h;
    for (long idx = blockIdx.x * (long)blockDim.x + threadIdx.x; idx < total;
         idx += (long)gridDim.x * blockDim.x) {
        const long i = idx % h;
        const int k = (int)((idx / h) % nz);
        const long o = idx / (h * nz);
        const int y = (int)(i / nx), x = (int)(i % nx);
        const float lower = pe_c[o * (nz + 1) * hc + (long)(k + 1) * hc + (long)(y / f) * nxc + x / f];
        const float surf = pe_f[o * (nz + 1) * h + (long)nz * h + i];
        out[idx] = lower < surf ? area[o * h + i] : 0.f;
    }
}

// budgets.py:62-77.  Thread per coarse cell and level: p_mid = delp_c / (log p[k+1] - log p[k])
// (vertically_dependent.py:153-179), compared with the fine surface pressure
// ps = sum_k delp + toa (:189-208) of every fine column of the block.
__global__ void __launch_bounds__(256)
exposed_area_kernel(const float *__restrict__ pe_c, const float *__restrict__ delp_c,
                    const float *__restrict__ ps, const float *__restrict__ area,
                    float *__restrict__ out, long n_outer, int nz, int ny, int nx, int f)
{
    const int nyc = ny / f, nxc = nx / f;
    const long hc = (long)nyc * nxc, h = (long)ny * nx;
    const long total = n_outer * nz * hc;
    for (long idx = blockIdx.x * (long)blockDim.x + threadIdx.x; idx < total;
         idx += (long)gridDim.x * blockDim.x) {
        const long ic = idx % hc;
        const int k = (int)((idx / hc) % nz);
        const long o = idx / (hc * nz);
        const int J = (int)(ic / nxc), I = (int)(ic % nxc);
        const float p0 = pe_c[o * (nz + 1) * hc + (long)k * hc + ic];
        const float p1 = pe_c[o * (nz + 1) * hc + (long)(k + 1) * hc + ic];
        const float pmid = delp_c[idx] / (logf(p1) - logf(p0));
        float acc = 0.f;
        for (int y = 0; y < f; ++y)
            for (int x = 0; x < f; ++x) {
                const long i = o * h + (long)(J * f + y) * nx + I * f + x;
                const float a = (pmid <= ps[i]) ? area[i] : 0.f;
                acc += (a == a) ? a : 0.f;
            }
        out[idx] = acc;
    }
}

// The same in one pass without temporaries: a warp owns a coarse cell.  Lanes over the cell's f x f fine
// columns sum delp top-down into the fine surface pressures (every fine value read once, whole 32-B sectors);
// lanes over levels then form the coarse log-midpoint pressures from the sequential cumsum of delp_coarse and
// add the areas of the exposed columns in the order of the loop above (y, then x), so both kernels agree bitwise.
constexpr int EXPOSED_WARPS = 8;
template <bool F8>
__global__ void __launch_bounds__(32 * EXPOSED_WARPS)
exposed_area_fused_kernel(const float *__restrict__ delp, const float *__restrict__ delp_c,
                          const float *__restrict__ area, float *__restrict__ out, long n_outer, int nz,
                          int ny, int nx, int f, float toa)
{
    extern __shared__ float sm_exposed[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int nyc = ny / f, nxc = nx / f, ncol = f * f;
    const long hc = (long)nyc * nxc, h = (long)ny * nx;
    float *ps = sm_exposed + warp * (2 * ncol + nz + 1), *ar = ps + ncol, *pe = ar + ncol;
    const long cells = n_outer * hc;
    for (long cell = (long)blockIdx.x * EXPOSED_WARPS + warp; cell < cells; cell += (long)gridDim.x * EXPOSED_WARPS) {
        const long o = cell / hc, ic = cell - o * hc;
        const int J = (int)(ic / nxc), I = (int)(ic - (long)J * nxc);
        const float *d0 = delp + o * nz * h + (long)(J * f) * nx + I * f;
        if (F8) {
            // f = 8: lane = (row lane / 8 and that row + 4, column lane % 8); eight levels = 16 loads in flight per lane,
            // added in level order per column (the order of surface_pressure_kernel)
            const float *pa = d0 + (long)(lane >> 3) * nx + (lane & 7), *pb = pa + 4L * nx;
            float sa = 0.f, sb = 0.f;
            int k = 0;
            for (; k + 8 <= nz; k += 8) {
                float a[8], b[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) { a[j] = ldg_stream(pa + (long)(k + j) * h); b[j] = ldg_stream(pb + (long)(k + j) * h); }
#pragma unroll
                for (int j = 0; j < 8; ++j) { sa += (a[j] == a[j]) ? a[j] : 0.f; sb += (b[j] == b[j]) ? b[j] : 0.f; }
            }
            for (; k < nz; ++k) {
                const float a = ldg_stream(pa + (long)k * h), b = ldg_stream(pb + (long)k * h);
                sa += (a == a) ? a : 0.f;
                sb += (b == b) ? b : 0.f;
            }
            const long ia = (long)(lane >> 3) * nx + (lane & 7), ib = ia + 4L * nx;
            ps[lane] = sa + toa;        ar[lane] = area[o * h + (long)(J * 8) * nx + I * 8 + ia];
            ps[lane + 32] = sb + toa;   ar[lane + 32] = area[o * h + (long)(J * 8) * nx + I * 8 + ib];
        } else
        for (int c0 = 0; c0 < ncol; c0 += 64) {                       // two columns per lane and sweep
            const int ca = c0 + lane, cb = c0 + 32 + lane;
            const bool va = ca < ncol, vb = cb < ncol;
            const long ia = va ? (long)(ca / f) * nx + ca % f : 0, ib = vb ? (long)(cb / f) * nx + cb % f : 0;
            float sa = 0.f, sb = 0.f;
#pragma unroll 4
            for (int k = 0; k < nz; ++k) {
                const float a = ldg_stream(d0 + (long)k * h + ia), b = ldg_stream(d0 + (long)k * h + ib);
                sa += (a == a) ? a : 0.f;
                sb += (b == b) ? b : 0.f;
            }
            if (va) { ps[ca] = sa + toa; ar[ca] = area[o * h + (long)(J * f) * nx + I * f + ia]; }
            if (vb) { ps[cb] = sb + toa; ar[cb] = area[o * h + (long)(J * f) * nx + I * f + ib]; }
        }
        for (int k = lane; k < nz; k += 32) pe[k + 1] = delp_c[(o * nz + k) * hc + ic];
        __syncwarp();
        if (lane == 0) {                                              // vertically_dependent.py:61-63, same order of additions
            float acc = toa;
            pe[0] = acc;
            for (int k = 0; k < nz; ++k) { acc = acc + pe[k + 1]; pe[k + 1] = acc; }
        }
        __syncwarp();
        for (int k = lane; k < nz; k += 32) {
            const float p0 = pe[k], p1 = pe[k + 1];
            const float pmid = delp_c[(o * nz + k) * hc + ic] / (logf(p1) - logf(p0));
            float acc = 0.f;
            for (int c = 0; c < ncol; ++c) {
                const float a = (pmid <= ps[c]) ? ar[c] : 0.f;
                acc += (a == a) ? a : 0.f;
            }
            out[(o * nz + k) * hc + ic] = acc;
        }
        __syncwarp();
    }
}

// f = 8 and a coarse row that is a multiple of 4 cells: a warp owns a strip of 4 adjacent cells, so that every
// load is one whole 128-B line (lane = fine column of the strip; rows and levels in registers).  The warp-per-cell
// form above asks DRAM for 32-B pieces of four lines per instruction: 125 us against this form's time in DESIGN.md.
__global__ void __launch_bounds__(32 * EXPOSED_WARPS, 3)         // 24 warps per SM: the C384 strips fit one wave
exposed_area_strip_kernel(const float *__restrict__ delp, const float *__restrict__ delp_c,
                          const float *__restrict__ area, float *__restrict__ out, long n_outer, int nz,
                          int ny, int nx, float toa)
{
    extern __shared__ float sm_exposed[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int nyc = ny >> 3, nxc = nx >> 3, nstrip = nxc >> 2;
    const long hc = (long)nyc * nxc, h = (long)ny * nx;
    float *ps = sm_exposed + warp * (512 + nz + 1), *ar = ps + 256, *pe = ar + 256;
    const long strips = n_outer * nyc * nstrip;
    for (long sidx = (long)blockIdx.x * EXPOSED_WARPS + warp; sidx < strips; sidx += (long)gridDim.x * EXPOSED_WARPS) {
        const long o = sidx / ((long)nyc * nstrip), rem = sidx - o * ((long)nyc * nstrip);
        const int J = (int)(rem / nstrip), I4 = (int)(rem - (long)J * nstrip);
        const float *base = delp + o * nz * h + (long)(J * 8) * nx + I4 * 32 + lane;
        float s[8];
#pragma unroll
        for (int r = 0; r < 8; ++r) s[r] = 0.f;
        int k = 0;
        for (; k + 4 <= nz; k += 4) {                                   // 32 whole lines in flight per warp
            float v[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = ldg_stream(base + (long)(k + (j >> 3)) * h + (long)(j & 7) * nx);
#pragma unroll
            for (int j = 0; j < 32; ++j) s[j & 7] += (v[j] == v[j]) ? v[j] : 0.f;      // level order per column
        }
        for (; k < nz; ++k) {
#pragma unroll
            for (int r = 0; r < 8; ++r) {
                const float v = ldg_stream(base + (long)k * h + (long)r * nx);
                s[r] += (v == v) ? v : 0.f;
            }
        }
        const int cell = lane >> 3, x = lane & 7;
#pragma unroll
        for (int r = 0; r < 8; ++r) {
            ps[cell * 64 + r * 8 + x] = s[r] + toa;
            ar[cell * 64 + r * 8 + x] = area[o * h + (long)(J * 8 + r) * nx + I4 * 32 + lane];
        }
        for (int c4 = 0; c4 < 4; ++c4) {
            const long ic = (long)J * nxc + I4 * 4 + c4;
            __syncwarp();
            for (int kk = lane; kk < nz; kk += 32) pe[kk] = delp_c[(o * nz + kk) * hc + ic];
            __syncwarp();
            // vertically_dependent.py:61-63: the sequential sum from toa, carried by every lane in a register (broadcast
            // reads, no store-to-load chain); a lane keeps the interfaces of its own levels lane, lane + 32, ...
            float pl[4][2];
            {
                float acc = toa;
#pragma unroll
                for (int q = 0; q < 4; ++q) {                           // (nz <= 128: checked by the host)
                    pl[q][0] = 0.f; pl[q][1] = 0.f;
                    const int n = min(32, nz - 32 * q);
                    for (int j = 0; j < n; ++j) {
                        const float nxt = acc + pe[32 * q + j];
                        if (j == lane) { pl[q][0] = acc; pl[q][1] = nxt; }
                        acc = nxt;
                    }
                }
            }
            const float *psc = ps + c4 * 64, *arc = ar + c4 * 64;
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int kk = 32 * q + lane;
                if (kk >= nz) break;
                const float p0 = pl[q][0], p1 = pl[q][1];
                const float pmid = pe[kk] / (logf(p1) - logf(p0));
                float acc = 0.f;
#pragma unroll 8
                for (int c = 0; c < 64; ++c) {
                    const float a = (pmid <= psc[c]) ? arc[c] : 0.f;
                    acc += (a == a) ? a : 0.f;
                }
                out[(o * nz + kk) * hc + ic] = acc;
            }
        }
        __syncwarp();
    }
}

__global__ void __launch_bounds__(256)
surface_pressure_kernel(const float *__restrict__ delp, float *__restrict__ ps, float toa,
                        long n_outer, int nz, long h)
{
    const long total = n_outer * h;
    for (long idx = blockIdx.x * (long)blockDim.x + threadIdx.x; idx < total;
         idx += (long)gridDim.x * blockDim.x) {
        const long o = idx / h, i = idx % h;
        float acc = 0.f;
        for (int k = 0; k < nz; ++k) {
            const float d = delp[o * nz * h + (long)k * h + i];
            acc += (d == d) ? d : 0.f;
        }
        ps[idx] = acc + toa;
    }
}

// derived budget inputs (budgets.py:80,151) materialised for the composed path only
__global__ void __launch_bounds__(256)
derive_field_kernel(const float *__restrict__ a, const float *__restrict__ b, float *__restrict__ out,
                    int op, float scale, long total)
{
    for (long idx = blockIdx.x * (long)blockDim.x + threadIdx.x; idx < total;
         idx += (long)gridDim.x * blockDim.x)
        out[idx] = op == CG_FIELD_PRODUCT ? a[idx] * b[idx] : (a[idx] - b[idx]) / scale;
}

static inline long align256(long n) { return (n + 255) / 256 * 256; }

static long composed_workspace_bytes(long n_outer, int nz, int ny, int nx, int f)
{
    const long h = (long)ny * nx, hc = (long)(ny / f) * (nx / f);
    return align256(4 * n_outer * (nz + 1) * h)     // fine interfaces
           + align256(4 * n_outer * (nz + 1) * hc)  // coarse interfaces
           + align256(4 * n_outer * nz * hc)        // coarse delp (when the caller passes NULL)
           + align256(4 * n_outer * nz * h)         // one remapped fine field
           + align256(4 * n_outer * nz * h)         // one derived fine field
           + align256(4 * n_outer * nz * h);        // masked weights
}

static unsigned grid_for(long total)
{
    long b = (total + 255) / 256;
    const long cap = (long)sm_count() * 16;
    return (unsigned)(b > cap ? cap : (b < 1 ? 1 : b));
}

static int composed(const cg_field *fields, float *const *outs, int nfields, const float *delp,
                    const float *area, float *delp_coarse, long n_outer, int nz, int ny, int nx,
                    int f, int mask_below, float toa, void *workspace, long workspace_bytes,
                    cudaStream_t st)
{
    const long need = composed_workspace_bytes(n_outer, nz, ny, nx, f);
    CG_REQUIRE(workspace && workspace_bytes >= need, "workspace too small: need %ld bytes, got %ld", need, workspace_bytes);
    const int nyc = ny / f, nxc = nx / f;
    const long h = (long)ny * nx, hc = (long)nyc * nxc;
    char *w = (char *)workspace;
    float *pe_f = (float *)w;  w += align256(4 * n_outer * (nz + 1) * h);
    float *pe_c = (float *)w;  w += align256(4 * n_outer * (nz + 1) * hc);
    float *dc_ws = (float *)w; w += align256(4 * n_outer * nz * hc);
    float *remap = (float *)w; w += align256(4 * n_outer * nz * h);
    float *derived = (float *)w; w += align256(4 * n_outer * nz * h);
    float *mask = (float *)w;
    float *dc = delp_coarse ? delp_coarse : dc_ws;

    int rc;
    {
        const void *in[1] = {delp};
        void *out[1] = {dc};
        rc = cg_weighted_block_average(in, out, 1, nullptr, area, CG_W2D, n_outer, nz, ny, nx, f, f,
                                       f, f, nyc, nxc, CG_F32, CG_F32, st);
        if (rc) return rc;
    }
    if ((rc = cg_pressure_at_interface(delp, pe_f, toa, n_outer, nz, h, CG_F32, st))) return rc;
    if ((rc = cg_pressure_at_interface(dc, pe_c, toa, n_outer, nz, hc, CG_F32, st))) return rc;
    if (mask_below) {
        mask_from_coarse_kernel<<<grid_for(n_outer * nz * h), 256, 0, st>>>(area, pe_c, pe_f, mask, n_outer, nz, ny, nx, f);
        CG_LAUNCH_CHECK("mask_from_coarse");
    }
    const long strides[12] = {(nz + 1) * h, 1, h, nz * h, 1, h, (nz + 1) * hc, 1, hc, nz * h, 1, h};
    for (int v = 0; v < nfields; ++v) {
        const float *src = fields[v].a;
        if (fields[v].op != CG_FIELD_PLAIN) {
            derive_field_kernel<<<grid_for(n_outer * nz * h), 256, 0, st>>>(fields[v].a, fields[v].b, derived, fields[v].op, fields[v].scale, n_outer * nz * h);
            CG_LAUNCH_CHECK("derive_field");
            src = derived;
        }
        rc = launch_mappm(pe_f, src, pe_c, remap, n_outer, h, nz, nz, strides, f, nx, nxc, 1, 1, st);
        if (rc) return rc;
        const void *in[1] = {remap};
        void *out[1] = {outs[v]};
        rc = cg_weighted_block_average(in, out, 1, mask_below ? mask : nullptr, mask_below ? nullptr : area,
                                       mask_below ? CG_W3D : CG_W2D, n_outer, nz, ny, nx, f, f, f, f,
                                       nyc, nxc, CG_F32, CG_F32, st);
        if (rc) return rc;
    }
    return CG_OK;
}

// cg_set_option("plevel_impl", composed | fused | stream | pair) pins one implementation (tests, A/B timing);
// default: pair where its shape constraints hold, else stream, else fused, else composed.
static bool impl_is(int which) { return options().plevel_impl == which; }
static bool want_composed() { return impl_is(PLEVEL_COMPOSED); }
static thread_local const char *g_last_impl = "none";
// replicas of the coarse outputs requested by cg_pressure_level_average_replicated for the call in progress
struct Replicas { int n; int dc; int kind; long long off[CG_MAX_REPLICAS]; };
static thread_local Replicas g_rep = {0, 0, 0, {0}};

static int stream_path(const cg_field *fields, float *const *outs, int nfields, const float *delp,
                       const float *area, float *delp_coarse, long n_outer, int nz, int ny, int nx,
                       int mask_below, float toa, void *workspace, long workspace_bytes, cudaStream_t st)
{
    if (!workspace || workspace_bytes < stream_workspace_bytes(n_outer, nz, ny, nx)) return CG_ENOTSUP;
    const int nyc = ny / 8, nxc = nx / 8;
    const long h = (long)ny * nx, hc = (long)nyc * nxc;
    char *w = (char *)workspace;
    float *dc = delp_coarse ? delp_coarse : (float *)w;
    float *ps = (float *)(w + align256(4 * n_outer * nz * hc));
    int *flags = (int *)(w + align256(4 * n_outer * nz * hc) + align256(4 * n_outer * h));
    const void *in[1] = {delp};
    void *out[1] = {dc};
    int rc = cg_weighted_block_average(in, out, 1, nullptr, area, CG_W2D, n_outer, nz, ny, nx, 8, 8, 8, 8,
                                       nyc, nxc, CG_F32, CG_F32, st);
    if (rc) return rc;
    if (mask_below && (rc = launch_surface_pressure_seq(delp, ps, toa, n_outer, nz, h, st))) return rc;
    // the pair-thread kernel where its shared-memory layout fits (option plevel_impl = stream pins the round-1 kernel)
    if ((!impl_is(PLEVEL_STREAM) || g_rep.n > 0) && pair_supported(nfields, nz, ny, nx, 8)) {
        rc = launch_plevel_pair(fields, outs, nfields, delp, area, dc, mask_below ? ps : nullptr, flags, n_outer, nz,
                                ny, nx, mask_below, toa, options().plevel_nansafe, g_rep.n, g_rep.off,
                                g_rep.dc && delp_coarse != nullptr, g_rep.kind, st);
        if (rc != CG_ENOTSUP) { g_last_impl = "pair"; return rc; }
    }
    if (g_rep.n > 0) return CG_ENOTSUP;         // only the pair kernel stores replicas
    if (!stream_supported(nfields, nz, ny, nx, 8)) return CG_ENOTSUP;
    g_last_impl = "stream";
    return launch_plevel_stream(fields, outs, nfields, delp, area, dc, mask_below ? ps : nullptr, flags, n_outer, nz,
                                ny, nx, mask_below, toa, st);
}

}  // namespace cg

using namespace cg;

extern "C" const char *cg_pressure_level_last_impl(void) { return g_last_impl; }

extern "C" long cg_pressure_level_workspace_bytes(int nfields, long n_outer, int nz, int ny, int nx, int f)
{
    if (f < 1 || ny < 1 || nx < 1 || nz < 1 || n_outer < 0) return 0;
    // what the implementation that will run needs: the streaming kernels keep delp_coarse, one surface pressure per
    // column and one flag per CTA (a few MB at C384); only the composed recipe needs fine-grid temporaries
    const int impl = options().plevel_impl;
    if (impl != PLEVEL_COMPOSED && impl != PLEVEL_FUSED && nfields > 0 &&
        (stream_supported(nfields, nz, ny, nx, f) || pair_supported(nfields, nz, ny, nx, f)))
        return stream_workspace_bytes(n_outer, nz, ny, nx);
    if (impl != PLEVEL_COMPOSED && impl != PLEVEL_STREAM && fused_supported(nz, ny, nx, f))
        return fused_workspace_bytes(nfields, n_outer, nz, ny, nx, f);
    return composed_workspace_bytes(n_outer, nz, ny, nx, f);
}

extern "C" long cg_pressure_level_fallback_workspace_bytes(int nfields, long n_outer, int nz, int ny, int nx, int f)
{
    (void)nfields;
    if (f < 1 || ny < 1 || nx < 1 || nz < 1 || n_outer < 0) return 0;
    return composed_workspace_bytes(n_outer, nz, ny, nx, f);
}

extern "C" int cg_pressure_level_average(const cg_field *fields, float *const *outs, int nfields,
                                         const float *delp, const float *area, float *delp_coarse,
                                         long n_outer, int nz, int ny, int nx, int f, int mask_below,
                                         double toa, void *workspace, long workspace_bytes, void *stream)
{
    CG_REQUIRE(fields && outs && delp && area, "null pointer");
    CG_REQUIRE(nfields >= 0 && nfields <= CG_MAX_FIELDS, "nfields must be in [0, %d]", CG_MAX_FIELDS);
    CG_REQUIRE(n_outer >= 0 && nz >= 4 && ny >= 1 && nx >= 1 && f >= 1, "bad shape (nz >= 4 required)");
    CG_REQUIRE(ny % f == 0 && nx % f == 0, "dimension sizes (%d, %d) are not multiples of the factor %d", ny, nx, f);
    for (int v = 0; v < nfields; ++v) {
        CG_REQUIRE(fields[v].a && outs[v], "null field pointer");
        CG_REQUIRE(fields[v].op >= CG_FIELD_PLAIN && fields[v].op <= CG_FIELD_DIFF, "bad field op");
        CG_REQUIRE(fields[v].op == CG_FIELD_PLAIN || fields[v].b, "derived field needs a second parent");
        CG_REQUIRE(fields[v].op != CG_FIELD_DIFF || fields[v].scale != 0.f, "storage term needs a non-zero dt");
    }
    if (n_outer == 0) return CG_OK;
    cudaStream_t st = (cudaStream_t)stream;
    if (!want_composed() && !impl_is(PLEVEL_FUSED) && nfields > 0 &&
        (stream_supported(nfields, nz, ny, nx, f) || pair_supported(nfields, nz, ny, nx, f))) {
        const int rc = stream_path(fields, outs, nfields, delp, area, delp_coarse, n_outer, nz, ny, nx,
                                   mask_below, (float)toa, workspace, workspace_bytes, st);
        if (rc != CG_ENOTSUP || g_rep.n > 0) return rc;     // misaligned pointers / no workspace: next implementation
    }
    if (g_rep.n > 0) return CG_ENOTSUP;
    if (!want_composed() && !impl_is(PLEVEL_STREAM) && fused_supported(nz, ny, nx, f)) {
        const int rc = launch_plevel_fused(fields, outs, nfields, delp, area, delp_coarse, n_outer, nz,
                                           ny, nx, f, mask_below, (float)toa, workspace, workspace_bytes, st);
        if (rc != CG_ENOTSUP) { g_last_impl = "fused"; return rc; }     // misaligned pointers: use the composed path
    }
    g_last_impl = "composed";
    if (!workspace || workspace_bytes < composed_workspace_bytes(n_outer, nz, ny, nx, f)) {
        // e.g. pointers the streaming kernels cannot take (16-byte alignment): the caller comes back with the
        // composed recipe's workspace (cg_pressure_level_fallback_workspace_bytes)
        set_error("cg_pressure_level_average: the streaming kernels do not take these arguments and the workspace "
                  "(%ld bytes) is too small for the composed recipe (%ld bytes)", workspace_bytes,
                  composed_workspace_bytes(n_outer, nz, ny, nx, f));
        return CG_ENOTSUP;
    }
    return composed(fields, outs, nfields, delp, area, delp_coarse, n_outer, nz, ny, nx, f, mask_below,
                    (float)toa, workspace, workspace_bytes, st);
}

extern "C" int cg_pressure_level_average_replicated(const cg_field *fields, float *const *outs, int nfields,
                                                    const float *delp, const float *area, float *delp_coarse,
                                                    long n_outer, int nz, int ny, int nx, int f, int mask_below,
                                                    double toa, int replica_kind, int n_replicas,
                                                    const long long *replica_offsets,
                                                    void *workspace, long workspace_bytes, void *stream)
{
    CG_REQUIRE(replica_kind == CG_REPLICA_PEERS || replica_kind == CG_REPLICA_MULTICAST, "bad replica_kind");
    CG_REQUIRE(n_replicas >= 0 && n_replicas <= CG_MAX_REPLICAS, "n_replicas must be in [0, %d]", CG_MAX_REPLICAS);
    CG_REQUIRE(replica_kind != CG_REPLICA_MULTICAST || n_replicas <= 1, "a multicast mapping is ONE offset");
    CG_REQUIRE(n_replicas == 0 || replica_offsets, "null replica_offsets");
    for (int r = 0; r < n_replicas; ++r)
        CG_REQUIRE(replica_offsets[r] % 16 == 0, "replica offset %d (%lld bytes) is not a multiple of 16", r, replica_offsets[r]);
    for (int v = 0; n_replicas > 0 && outs && v < nfields; ++v)
        CG_REQUIRE(((uintptr_t)outs[v] & 15) == 0, "replicated outputs are stored 16 bytes at a time: outs[%d] is not 16-byte aligned", v);
    CG_REQUIRE(n_replicas == 0 || ((uintptr_t)delp_coarse & 15) == 0, "delp_coarse is not 16-byte aligned");
    if (n_replicas == 0)
        return cg_pressure_level_average(fields, outs, nfields, delp, area, delp_coarse, n_outer, nz, ny, nx, f,
                                         mask_below, toa, workspace, workspace_bytes, stream);
    CG_REQUIRE(fields && outs && delp && area, "null pointer");
    CG_REQUIRE(nfields >= 1 && nfields <= CG_MAX_FIELDS, "nfields must be in [1, %d]", CG_MAX_FIELDS);
    CG_REQUIRE(n_outer >= 0 && nz >= 4 && ny >= 1 && nx >= 1 && f >= 1, "bad shape (nz >= 4 required)");
    if (!(f == 8 && ny % 8 == 0 && nx % 8 == 0 && pair_supported(nfields, nz, ny, nx, 8))) {
        set_error("cg_pressure_level_average_replicated: only the pair-thread kernel stores replicas (f = 8, "
                  "nx %% 32 == 0, shared-memory layout for %d fields x %d levels)", nfields, nz);
        return CG_ENOTSUP;
    }
    g_rep.n = n_replicas;
    g_rep.dc = 1;
    g_rep.kind = replica_kind;
    for (int r = 0; r < n_replicas; ++r) g_rep.off[r] = replica_offsets[r];
    int rc = cg_pressure_level_average(fields, outs, nfields, delp, area, delp_coarse, n_outer, nz, ny, nx, f,
                                       mask_below, toa, workspace, workspace_bytes, stream);
    g_rep.n = 0;
    if (rc == CG_ENOTSUP)
        set_error("cg_pressure_level_average_replicated: the pair-thread kernel does not take these arguments "
                  "(16-byte aligned field arrays, a workspace of cg_pressure_level_workspace_bytes)");
    return rc;
}

extern "C" int cg_area_above_fine_surface(const float *delp, const float *delp_coarse, const float *area,
                                          float *out, long n_outer, int nz, int ny, int nx, int f,
                                          double toa, void *stream)
{
    CG_REQUIRE(delp && delp_coarse && area && out, "null pointer");
    CG_REQUIRE(n_outer >= 0 && nz >= 1 && ny >= 1 && nx >= 1 && f >= 1, "bad shape");
    CG_REQUIRE(ny % f == 0 && nx % f == 0, "dimension sizes (%d, %d) are not multiples of the factor %d", ny, nx, f);
    if (n_outer == 0) return CG_OK;
    cudaStream_t st = (cudaStream_t)stream;
    const long h = (long)ny * nx, hc = (long)(ny / f) * (nx / f);
    const int pin = options().exposed_impl;                               // (tests: the formulations agree bit for bit)
    if (pin <= 1 && f == 8 && nz <= 128 && (nx / 8) % 4 == 0 && (size_t)EXPOSED_WARPS * (512 + nz + 1) * sizeof(float) <= 48 * 1024) {
        const long blocks = (n_outer * (ny / 8) * (nx / 32) + EXPOSED_WARPS - 1) / EXPOSED_WARPS;
        exposed_area_strip_kernel<<<(unsigned)(blocks < 0x7fffffffL ? blocks : 0x7fffffffL), 32 * EXPOSED_WARPS,
                                    (size_t)EXPOSED_WARPS * (512 + nz + 1) * sizeof(float), st>>>(
            delp, delp_coarse, area, out, n_outer, nz, ny, nx, (float)toa);
        CG_LAUNCH_CHECK("cg_area_above_fine_surface");
        return CG_OK;
    }
    const size_t smem = (size_t)EXPOSED_WARPS * (2 * f * f + nz + 1) * sizeof(float);
    if (pin <= 2 && smem <= 48 * 1024) {
        const long blocks = (n_outer * hc + EXPOSED_WARPS - 1) / EXPOSED_WARPS;
        const unsigned g = (unsigned)(blocks < 0x7fffffffL ? blocks : 0x7fffffffL);
        if (f == 8) exposed_area_fused_kernel<true><<<g, 32 * EXPOSED_WARPS, smem, st>>>(delp, delp_coarse, area, out, n_outer, nz, ny, nx, f, (float)toa);
        else exposed_area_fused_kernel<false><<<g, 32 * EXPOSED_WARPS, smem, st>>>(delp, delp_coarse, area, out, n_outer, nz, ny, nx, f, (float)toa);
        CG_LAUNCH_CHECK("cg_area_above_fine_surface");
        return CG_OK;
    }
    // very large blocks or columns: the three-kernel form with temporaries
    float *ps = nullptr, *pe_c = nullptr;
    cudaError_t e = cudaMallocAsync((void **)&ps, 4 * n_outer * h, st);
    if (e != cudaSuccess) return cuda_fail(e, "cudaMallocAsync(ps)");
    e = cudaMallocAsync((void **)&pe_c, 4 * n_outer * (nz + 1) * hc, st);
    if (e != cudaSuccess) { cudaFreeAsync(ps, st); return cuda_fail(e, "cudaMallocAsync(pe_c)"); }
    surface_pressure_kernel<<<grid_for(n_outer * h), 256, 0, st>>>(delp, ps, (float)toa, n_outer, nz, h);
    count_launch();
    int rc = cg_pressure_at_interface(delp_coarse, pe_c, toa, n_outer, nz, hc, CG_F32, st);
    if (rc == CG_OK)
    {
        exposed_area_kernel<<<grid_for(n_outer * nz * hc), 256, 0, st>>>(pe_c, delp_coarse, ps, area, out, n_outer, nz, ny, nx, f);
        count_launch();
    }
    cudaError_t le = cudaGetLastError();
    cudaFreeAsync(ps, st);
    cudaFreeAsync(pe_c, st);
    if (rc) return rc;
    if (le != cudaSuccess) return cuda_fail(le, "cg_area_above_fine_surface");
    return CG_OK;
}

// ---- allocation-free variant (no counterpart in the reference) ----
// cg_area_above_fine_surface's general form needs two temporaries (fine surface pressure, coarse interface
// pressures) which it takes from cudaMallocAsync.  Callers that may not allocate (CUDA graph capture, serving
// loops, harnesses that forbid it) get the same three kernels over a file-scope device array instead.
// Every GPU that loads the library holds the array until exit.  All calls share it and it keeps the last call's
// contents: both temporaries are fully overwritten before they are read, so nothing has to be cleared, but
// calls that overlap on the GPU (different streams, no ordering) race on it.
namespace cg {
constexpr long STATIC_SCRATCH_BYTES = 16L << 20;       // 16 MiB: a 6-tile C384 timestep needs 8.0 MiB
static __device__ __align__(256) unsigned char g_static_scratch[STATIC_SCRATCH_BYTES];

// Device address of the array on the current device.  In host code the array's name is only a stand-in for
// the symbol, not a device pointer: the address comes from cudaGetSymbolAddress, once per device.
static int static_scratch_ptr(unsigned char **out)
{
    constexpr int MAXDEV = 64;
    static std::mutex mu;
    static unsigned char *ptrs[MAXDEV] = {nullptr};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return cuda_fail(e, "cudaGetDevice");
    if (dev < 0 || dev >= MAXDEV) { set_error("device index %d out of range", dev); return CG_ENOTSUP; }
    std::lock_guard<std::mutex> lock(mu);
    if (!ptrs[dev]) {
        void *p = nullptr;
        e = cudaGetSymbolAddress(&p, g_static_scratch);
        if (e != cudaSuccess) return cuda_fail(e, "cudaGetSymbolAddress(g_static_scratch)");
        ptrs[dev] = (unsigned char *)p;
    }
    *out = ptrs[dev];
    return CG_OK;
}
}  // namespace cg

extern "C" long cg_static_scratch_bytes(void) { return cg::STATIC_SCRATCH_BYTES; }

extern "C" int cg_area_above_fine_surface_static(const float *delp, const float *delp_coarse, const float *area,
                                                 float *out, long n_outer, int nz, int ny, int nx, int f,
                                                 double toa, void *stream)
{
    CG_REQUIRE(delp && delp_coarse && area && out, "null pointer");
    CG_REQUIRE(n_outer >= 0 && nz >= 1 && ny >= 1 && nx >= 1 && f >= 1, "bad shape");
    CG_REQUIRE(ny % f == 0 && nx % f == 0, "dimension sizes (%d, %d) are not multiples of the factor %d", ny, nx, f);
    if (n_outer == 0) return CG_OK;
    const long h = (long)ny * nx, hc = (long)(ny / f) * (nx / f);
    const long ps_bytes = align256(4 * n_outer * h), pe_bytes = align256(4 * n_outer * (nz + 1) * hc);
    CG_REQUIRE(ps_bytes + pe_bytes <= STATIC_SCRATCH_BYTES,
               "cg_area_above_fine_surface_static: needs %ld bytes of scratch, the static array holds %ld; "
               "use cg_area_above_fine_surface for this size", ps_bytes + pe_bytes, STATIC_SCRATCH_BYTES);
    unsigned char *scratch = nullptr;
    int rc = static_scratch_ptr(&scratch);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    float *ps = (float *)scratch, *pe_c = (float *)(scratch + ps_bytes);
    surface_pressure_kernel<<<grid_for(n_outer * h), 256, 0, st>>>(delp, ps, (float)toa, n_outer, nz, h);
    CG_LAUNCH_CHECK("cg_area_above_fine_surface_static(surface pressure)");
    if ((rc = cg_pressure_at_interface(delp_coarse, pe_c, toa, n_outer, nz, hc, CG_F32, st))) return rc;
    exposed_area_kernel<<<grid_for(n_outer * nz * hc), 256, 0, st>>>(pe_c, delp_coarse, ps, area, out, n_outer, nz, ny, nx, f);
    CG_LAUNCH_CHECK("cg_area_above_fine_surface_static");
    return CG_OK;
}
