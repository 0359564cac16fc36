/* coarsen_b200.h — C ABI of libcoarsen_b200.so (CUDA, sm_100a).
 *
 * Drop-in boundary for the coarse-graining hot path of fv3net's `vcm` package
 * (paths below are relative to the reference checkout):
 *   external/vcm/vcm/cubedsphere/coarsen.py   block reductions (numpy/xarray there)
 *   external/vcm/vcm/cubedsphere/regridz.py   pressure-level glue
 *   external/mappm/mappm/mappm.f90            Fortran PPM remapper (f2py ABI:
 *       mappm.mappm(pe1, q1, pe2, i1, i2, iv, kord, ptop) -> q2, regridz.py:273)
 *
 * Conventions
 *   - every pointer is a DEVICE pointer owned by the caller (scratch is either a
 *     caller-provided workspace or stream-ordered cudaMallocAsync/cudaFreeAsync
 *     inside the call); `stream` is a cudaStream_t passed as
 *     void* (NULL = legacy default stream); calls are asynchronous on it;
 *   - arrays are dense, row-major, horizontal dims last: (n_outer, nz, ny, nx),
 *     2-D weights (n_outer, ny, nx); n_outer folds tile/time/anything else;
 *   - dtype codes: CG_F32 / CG_F64; "wdtype" is the dtype of the weights
 *     (the reference mixes f64 restart fields with f32 grid weights);
 *   - return value: CG_OK or a negative CG_E* code; cg_last_error() gives the
 *     text for the calling thread.  No exceptions.  The only state is per calling
 *     thread: the error string and the option block of cg_set_option; nothing is
 *     read from the environment.  There is NO CPU fallback anywhere.
 */
#ifndef COARSEN_B200_H
#define COARSEN_B200_H

#ifdef __cplusplus
extern "C" {
#endif

#define CG_OK 0
#define CG_EINVAL (-1)   /* bad shape / factor / enum */
#define CG_ENOTSUP (-2)  /* outside what the kernels take (see each entry point) */
#define CG_ECUDA (-3)    /* CUDA runtime error; see cg_last_error() */

#define CG_F32 0
#define CG_F64 1

/* reduction methods of cg_block_reduce (xarray coarsen names; NaN-skipping) */
#define CG_SUM 0
#define CG_MIN 1
#define CG_MAX 2
#define CG_MEAN 3
#define CG_NAN_PROPAGATE 16 /* OR into the method: numpy's plain sum/mean/min/max */

/* weight modes of cg_weighted_block_average */
#define CG_W2D 0   /* w: (n_outer, ny, nx), broadcast over z (area)            */
#define CG_W3D 1   /* w: (n_outer, nz, ny, nx) (masked area, precomputed mass) */
#define CG_WMASS 2 /* w = w3d * w2d formed on the fly (delp * area)            */

#define CG_MAX_FIELDS 32

const char *cg_last_error(void);
int cg_version(void);
/* Behaviour switches of the CALLING THREAD, for tests and A/B timing (defaults in brackets):
 *   "plevel_impl"    [auto] | pair | stream | fused | composed   implementation of cg_pressure_level_average
 *                    that runs wherever its shape constraints hold (otherwise the next one in that order)
 *   "plevel_nansafe" [0] | 1   the NaN-skipping variant of the streaming kernels alone, no fast first pass
 *   "mappm_arrays"   [auto] | shared   column work arrays of the generic remapper in shared memory
 *   "mappm_impl"     [auto] | generic  the generic remapper even where the streaming one applies
 *   "exposed_area_impl" [auto] | strip | cell | three   formulation of cg_area_above_fine_surface (same bits)
 * CG_EINVAL for an unknown name or value. */
int cg_set_option(const char *name, const char *value);
/* number of CUDA kernels this library has launched in this process (monotonic) */
long cg_launch_count(void);
/* sm count, major*10+minor compute capability; any may be NULL */
int cg_device_info(int *sm_count, int *cc, long *total_mem_bytes);

/* coarsen.py:183-218 weighted_block_average and :221-273 edge_weighted_block_average.
 *   out[o,k,J,I] = nansum_{win}(obj*w) / nansum_{win}(w)
 * window (fy, fx) anchored at (J*sy, I*sx); cell-centred: fy=fx=sy=sx=f;
 * edge='x': fy=1, fx=f, sy=sx=f on (ny = nyc*f+1) rows; edge='y' symmetric.
 * nfields objects share one weight array (read once per block, not per field).
 * objs/outs: HOST arrays of nfields device pointers.  Output (n_outer,nz,nyo,nxo)
 * in the promoted dtype of (dtype, wdtype); the denominator is summed in wdtype. */
int cg_weighted_block_average(const void *const *objs, void *const *outs, int nfields,
                              const void *w3d, const void *w2d, int wmode,
                              long n_outer, int nz, int ny, int nx,
                              int fy, int fx, int sy, int sx, int nyo, int nxo,
                              int dtype, int wdtype, void *stream);

/* coarsen.py:629-683 block_edge_coarsen / :795-840 block_coarsen(sum|min|max|mean),
 * budgets.py:59 block_sum: NaN-skipping window reduction, same window rule. */
int cg_block_reduce(const void *in, void *out, int method,
                    long n_outer, int ny, int nx,
                    int fy, int fx, int sy, int sx, int nyo, int nxo,
                    int dtype, void *stream);

/* coarsen.py:557-588 block_median (np.median over f x f blocks; any NaN -> NaN;
 * even count -> (lo+hi)/2 in the input dtype).  Bit-exact selection. */
int cg_block_median(const void *in, void *out, long n_outer, int ny, int nx, int f,
                    int dtype, void *stream);

/* coarsen.py:750-786 _block_mode (scipy.stats.mode: most frequent value, ties ->
 * smallest).  nan_omit=0: 'propagate' as in the reference's pinned scipy 1.7.3 (NaN is an
 * ordinary candidate sorted last), 1: 'omit' (NaNs ignored; all-NaN block -> NaN). */
int cg_block_mode(const void *in, void *out, long n_outer, int ny, int nx, int f,
                  int nan_omit, int dtype, void *stream);

/* coarsen.py:843-897 block_upsample: repeat f times along y and/or x; a
 * staggered (odd-sized) dim repeats all but its last line.  fy or fx = 1 leaves
 * that dim alone.  Output dims: stag ? (n-1)*f+1 : n*f. */
int cg_block_upsample(const void *in, void *out, long n_outer, int nyc, int nxc,
                      int fy, int fx, int stag_y, int stag_x, int dtype, void *stream);

/* calc/thermo/vertically_dependent.py:41-66 pressure_at_interface:
 * (n_outer, nz, h) -> (n_outer, nz+1, h), sequential cumsum from toa in dtype. */
int cg_pressure_at_interface(const void *delp, void *pe, double toa,
                             long n_outer, int nz, long h, int dtype, void *stream);

/* regridz.py:150-161 _mask_weights: out[o,k,i] = pc[o,k+1,i] < pf[o,nzf,i] ? w[o,i] : 0
 * pc: (n_outer, nz+1, h) coarse interfaces on the fine grid, pf: (n_outer, nzf+1, h). */
int cg_mask_weights(const void *w2d, const void *pc, const void *pf, void *out,
                    long n_outer, int nz, int nzf, long h, int dtype, int wdtype, void *stream);

/* regridz.py:86-100 (xgcm.Grid.interp with xgcm.py:7-34 face connections):
 * cell-centred (6, nz, n, n) -> edges; axis 0: y-edges (6,nz,n+1,n), 1: x-edges
 * (6,nz,n,n+1); tile-face lines come from the neighbour tile. */
int cg_interp_to_edges(const void *c, void *out, int nz, int n, int axis,
                       int dtype, void *stream);

/* mappm.f90:10-126 (ppm_profile :614, ppm_limiters :854, cs_profile :132): conservative PPM remap,
 * fp32 arithmetic like the Fortran `real`.  Strided columns: column (o,i) level k
 * of array A is A[o*os_A + i*cs_A + k*ls_A]; (ncol,nz) z-last and (tile,z,y,x)
 * layouts both work without a transpose; cs_pe2 = 0 shares target pressures.
 * strides: {os,cs,ls} for pe1, q1, pe2, q2 in that order (12 longs, elements).
 * kord <= 7: ppm_profile; kord > 7: cs_profile (:132-532) / cs_limiters (:535-611); km >= 4.
 * (iv = -2 with kord > 7 -> CG_ENOTSUP: cs_profile reads mappm's never-set `qs` there.)
 * A target layer no source layer brackets (NaN / non-monotone pressure: undefined
 * behaviour in the Fortran) yields NaN. */
int cg_mappm(const float *pe1, const float *q1, const float *pe2, float *q2,
             long n_outer, long n_inner, int km, int kn, const long *strides,
             int iv, int kord, void *stream);

/* external/mappm/mappm/interpolate_2d.f90:1-29, the f2py module's second entry point
 * (mappm.interpolate_2d(xp, x, y, fill_value) -> y_out, vcm/interpolate.py:176-180): for every
 * row i, y(i, :) given at x(i, :) linearly interpolated to xp(i, :); float64; targets outside
 * every source interval keep fill_value; the last matching interval wins, as in the Fortran.
 * Element (i, k) of an array lives at A[i*rs + k*ls]; strides = {rs, ls} for xp, x, y, y_out
 * (8 longs, elements). */
int cg_interpolate_2d(const double *xp, const double *x, const double *y, double *y_out,
                      double fill_value, long m, int n_in, int n_out, const long *strides,
                      void *stream);

/* Fused pressure-level coarse-graining (iv=1, kord=1, the only combination the
 * reference's callers use: regridz.py:168-169, budgets.py:45):
 *   regridz.py:115-147 (+ coarsen.py:183 with the masked weights)  mask_below=1
 *   workflows/fine_res_budget/budget/budgets.py:49-57               mask_below=0
 * For every f x f block: coarse delp = area-weighted block mean of delp; both
 * interface pressures by cumsum from `toa`; every field remapped column by
 * column onto the block's coarse pressures; area(-masked)-weighted block mean.
 * Nothing fine-grid is written (fused path).  A field is described by a cg_field so the
 * budget workflow's derived inputs are formed on the fly from their parents instead of
 * being materialised (budgets.py:151 second moments, :80 storage terms):
 *   CG_FIELD_PLAIN   q = a
 *   CG_FIELD_PRODUCT q = a * b                      (T * omega ...)
 *   CG_FIELD_DIFF    q = (a - b) / scale            (storage: (end - begin) / dt)
 * fields: HOST array of nfields descriptors (device pointers inside), each
 * (n_outer,nz,ny,nx) f32; outs: HOST array of device pointers (n_outer,nz,ny/f,nx/f) f32.
 * delp_coarse: optional (n_outer,nz,ny/f,nx/f) output, may be NULL.
 * workspace: see cg_pressure_level_workspace_bytes (may be 0). */
#define CG_FIELD_PLAIN 0
#define CG_FIELD_PRODUCT 1
#define CG_FIELD_DIFF 2
typedef struct cg_field {
    const float *a;
    const float *b; /* NULL for CG_FIELD_PLAIN */
    int op;
    float scale;
} cg_field;

long cg_pressure_level_workspace_bytes(int nfields, long n_outer, int nz, int ny, int nx, int f);
/* workspace of the composed recipe (fine-grid temporaries), which cg_pressure_level_average falls back to when
 * the streaming kernels cannot take the arguments (it then returns CG_ENOTSUP if the workspace is smaller) */
long cg_pressure_level_fallback_workspace_bytes(int nfields, long n_outer, int nz, int ny, int nx, int f);
/* which implementation the calling thread's last cg_pressure_level_average ran: "pair", "stream",
 * "fused", "composed" ("none" before the first call) - for tests and bench records */
const char *cg_pressure_level_last_impl(void);
int cg_pressure_level_average(const cg_field *fields, float *const *outs, int nfields,
                              const float *delp, const float *area, float *delp_coarse,
                              long n_outer, int nz, int ny, int nx, int f,
                              int mask_below, double toa,
                              void *workspace, long workspace_bytes, void *stream);

/* Multi-GPU form of cg_pressure_level_average, the compute step fused with the exchange that follows it (one
 * timestep sharded by tile / row band over the GPUs of a box, every GPU wanting every coarse value:
 * coarsen_surface_c3072.py:29-76 gathers the sub-tiles in a later Beam stage).  No counterpart in the
 * reference's native interface.  Every coarse value the call produces - outs[v][i] and delp_coarse[i] - is stored
 * at its own address AND at that address + replica_offsets[r] bytes for r < n_replicas: with the outputs placed
 * in a buffer that is mapped at the same layout on every GPU (symmetric / peer memory over NVLink), the offsets
 * are peer_base - own_base and the kernel's stores ARE the all-gather; the caller only adds a barrier between
 * the ranks before the values are read.
 *   replica_kind CG_REPLICA_PEERS:     one NVLink store per value and offset, n_replicas <= CG_MAX_REPLICAS
 *   replica_kind CG_REPLICA_MULTICAST: replica_offsets[0] = multicast_base - own_base of an NVLS multicast mapping
 *                                      of the buffer; every value is stored ONCE (multimem.st) and the NVSwitch
 *                                      writes it into every GPU's copy, this GPU's included
 * The four coarse cells a CTA owns in a row go out as one 16-byte store: outs[v], delp_coarse and the offsets must
 * be multiples of 16 bytes (CG_EINVAL).  Served by the pair-thread kernel only (f = 8, nx % 32 == 0; CG_ENOTSUP
 * otherwise, nothing is launched). */
#define CG_MAX_REPLICAS 7
#define CG_REPLICA_PEERS 0
#define CG_REPLICA_MULTICAST 1
int cg_pressure_level_average_replicated(const cg_field *fields, float *const *outs, int nfields,
                                         const float *delp, const float *area, float *delp_coarse,
                                         long n_outer, int nz, int ny, int nx, int f,
                                         int mask_below, double toa,
                                         int replica_kind, int n_replicas, const long long *replica_offsets,
                                         void *workspace, long workspace_bytes, void *stream);

/* The exchange step of that sharded path as a stand-alone NVLS broadcast, no counterpart in the reference: every 16
 * bytes of [src, src + nbytes) are stored once (multimem.st) to src + mc_offset, the multicast mapping of the
 * symmetric buffer that src lies in, and the NVSwitch writes them into every GPU's copy.  With every rank
 * broadcasting its own slot this is an all-gather in one small kernel per rank plus a barrier; pointer, offset and
 * size are multiples of 16. */
int cg_multicast_store(const void *src, long long mc_offset, long nbytes, void *stream);

/* budgets.py:62-77 area_above_fine_surface: block sum of area where the coarse
 * log-midpoint pressure (upsampled) <= fine surface pressure. */
int cg_area_above_fine_surface(const float *delp, const float *delp_coarse, const float *area,
                               float *out, long n_outer, int nz, int ny, int nx, int f,
                               double toa, void *stream);

/* No counterpart in the reference.
 * cg_area_above_fine_surface without any allocation, for callers that may not allocate
 * (CUDA graph capture, serving loops).  Same arguments and result; its two temporaries live
 * in a file-scope device array of cg_static_scratch_bytes() = 16 MiB that every GPU loading
 * the library holds until exit.  CG_EINVAL when
 *   align256(4 n_outer ny nx) + align256(4 n_outer (nz + 1) (ny/f) (nx/f)) exceeds it
 * (a 6-tile C384 timestep needs 8.0 MiB).  All calls share the array (it keeps the last
 * call's contents, which are overwritten before use): calls that overlap on the GPU race. */
long cg_static_scratch_bytes(void);
int cg_area_above_fine_surface_static(const float *delp, const float *delp_coarse,
                                      const float *area, float *out, long n_outer, int nz,
                                      int ny, int nx, int f, double toa, void *stream);

/* ---- restart-coarsening recipes (external/vcm/vcm/cubedsphere/coarsen_restarts.py) ----------- */

/* Element-wise steps between the block kernels, on same-shape dense arrays of one dtype
 * (masks are 0 / 1 in that dtype).  `b` may be NULL: the scalar takes its place.
 *   CG_EW_MUL      out = a * b                  delp * area (:374), area * vfrac (:1160)
 *   CG_EW_ISCLOSE  out = np.isclose(a, b)       xarray_utils.isclose (:1097, :1107, :1109, :1352)
 *   CG_EW_AND      out = a && b                 is_dominant_surface_type & is_dominant_vtype (:1156)
 *   CG_EW_WHERE    out = c ? a : b              DataArray.where(mask) with b = NaN, xr.where (:1176, :1352)
 *   CG_EW_GT / LT  out = a > b, a < b           :1176, :1360, :1394
 *   CG_EW_BLEND    out = a*b + (1 - a)*c        blend (:624-639)
 *   CG_EW_FILLNA   out = isnan(a) ? b : a       .fillna(0.0) (:1242, :1254)
 *   CG_EW_SUB / DIV out = a - b, a / b          compute_blending_weights (:519-536) */
#define CG_EW_MUL 0
#define CG_EW_ISCLOSE 1
#define CG_EW_AND 2
#define CG_EW_WHERE 3
#define CG_EW_GT 4
#define CG_EW_LT 5
#define CG_EW_BLEND 6
#define CG_EW_FILLNA 7
#define CG_EW_SUB 8
#define CG_EW_DIV 9
int cg_elementwise(int op, const void *a, const void *b, const void *c, double scalar, void *out,
                   long n, int dtype, void *stream);

/* calc/thermo/vertically_dependent.py:189-208 surface_pressure_from_delp:
 * (n_outer, nz, h) -> (n_outer, h), sum over z (top-down) + toa. */
int cg_surface_pressure(const void *delp, void *out, double toa, long n_outer, int nz, long h,
                        int dtype, void *stream);
/* the same with xarray's default skipna=True of DataArray.sum (NaN thicknesses count as zero): what
 * surface_pressure_from_delp returns for xarray inputs (vertically_dependent.py:203) */
int cg_surface_pressure_nansum(const void *delp, void *out, double toa, long n_outer, int nz, long h,
                               int dtype, void *stream);
/* calc/thermo/vertically_dependent.py:153-179 pressure_at_midpoint_log: (n_outer, nz, h) -> same shape,
 * delp / diff(log(p_interface)) with the interfaces of cg_pressure_at_interface (sequential cumsum from toa). */
int cg_pressure_at_midpoint_log(const void *delp, void *out, double toa, long n_outer, int nz, long h,
                                int dtype, void *stream);
/* cubedsphere/coarsen.py:54-80 shift_edge_var_to_center: (n_outer, n, inner) -> (n_outer, n-1, inner),
 * out[o,j,x] = 0.5 * (in[o,j+1,x] + in[o,j,x]) along the staggered dim. */
int cg_shift_to_center(const void *in, void *out, long n_outer, int n, long inner, int dtype, void *stream);

/* coarsen_restarts.py:519-581 (compute_blending_weights with pressure_at_midpoint_log and
 * surface_pressure_from_delp of the coarse delp): delp_coarse (n_outer, nz, h), ps_min
 * (n_outer, h) = block minimum of the fine surface pressure; out (n_outer, nz, h):
 *   w = pfull > sigma*ps_min ? (ps - pfull) / (ps - sigma*ps_min) : 1 */
int cg_blending_weights(const void *delp_coarse, const void *ps_min, void *out, double toa,
                        double sigma, long n_outer, int nz, long h, int dtype, void *stream);

/* coarsen_restarts.py:916-938 _impose_hydrostatic_balance (height_at_interface, hydrostatic_dz,
 * dz_and_top_to_phis of vertically_dependent.py:69-100, 211-232, 182-186) per coarse column:
 * dz, temp, sphum, delp, dz_out (n_outer, nz, h); phis, phis_out (n_outer, h). */
int cg_hydrostatic_balance(const void *dz, const void *phis, const void *temp, const void *sphum,
                           const void *delp, void *dz_out, void *phis_out, double toa,
                           long n_outer, int nz, long h, int dtype, void *stream);

#ifdef __cplusplus
}
#endif
#endif
