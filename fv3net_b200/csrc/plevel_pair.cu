// Pair-thread streaming pressure-level coarse-graining for sm_100a: f = 8, iv = 1, kord = 1, nx % 32 == 0.
// (regridz.py:115-147 + coarsen.py:183, budgets.py:49-57; mappm.f90:10-126, :614-851, :854-931)
//
// Same tandem walk as plevel_stream.cu (source layers in order, coarse interfaces met on the way close one
// target level and open the next; the pressure-only geometry is formed once per column and level and shared by
// all fields through shared-memory records), re-cut so that every lane works:
//   * A CTA owns 4 x-adjacent coarse cells = 32 fine columns = 16 column PAIRS x 8 rows x nz levels; one CTA
//     of 16 warps per SM.
//   * A field thread is one work item (column pair, field): 16 V items flattened over the field warps
//     (V = 26 -> 13 full warps).  It carries the rolling PPM state of its two columns as f32x2 operands (one
//     instruction, two columns) and owns a PRIVATE accumulator column acc[target level][item] in shared memory:
//     its read-modify-writes never meet another thread's, and bank = lane whatever the level.  The four pairs of
//     a cell are summed once at the end in a fixed order (bitwise reproducible).
//   * Three geometry warps (lanes over the 32 columns, warp g takes level 3 chunk + g) write the records and
//     issue the TMA loads: 13 + 3 warps = 4 per scheduler.
//   * The walk is unrolled by the 3 levels of a pipeline stage with the rolling state in 3-slot rings indexed
//     at compile time: no register moves between steps.
//   * Slope and parabola limiters are written for the instruction mix of sm_100a: |.|-minimum with the sign
//     spliced back (FMNMX, LOP3) and 0 / 1 masks (FSET) instead of chains of half-rate 3-input min / max.
// Every fine input element is read from HBM exactly once per launch.
#include "plevel_common.cuh"
#include <cmath>
#include <cstring>
#include <mutex>

namespace cg {
namespace pairk {
using namespace plv;

constexpr int NCELL = 4;             // coarse cells per CTA
constexpr int NCOL = 32;             // fine columns per CTA row = lanes of a geometry warp
constexpr int NPAIR = NCOL / 2;      // column pairs per CTA row
constexpr int KC = 3;                // levels per pipeline stage = geometry warps
constexpr int NGW = KC;
constexpr int NSTAGE = 3;
constexpr int BOX_BYTES = KC * 128;  // one source, one stage: 3 rows of 32 columns (odd row count: 8 swizzle phases over 8 sources)
constexpr int MAXFW = 13;            // field warps: 16 warps with the geometry warps, 128 registers each
constexpr int MAXV = MAXFW * 32 / NPAIR;     // 26 output fields per launch
constexpr int MAXSRC = 2 * MAXV;
// geometry record of one step, structure of arrays over the 16 pairs: group g of pair p at g * GROUP_BYTES + 16 p
//   0: E1a E1b E2a E2b | 1: Fa Fb Ga Gb | 2: -Ha -Hb dpwa dpwb | 3: exa exb eya eyb | 4: eza ezb -ewa -ewb
//   5: offa offb infos flag   (off = accumulator byte offset of the level the first event closes; infos = info_a |
//      info_b << 8, info = number of coarse interfaces in the layer; flag = some column has further events)
constexpr int GROUP_BYTES = NPAIR * 16;
constexpr int NGROUP = 6;
constexpr int REC_LEVEL = NGROUP * GROUP_BYTES;
constexpr int REC_BYTES = KC * REC_LEVEL;
constexpr int MAXE = 2;              // listed further events per (column, layer)
constexpr int INFO_SLOW = 255;
constexpr int OVA_BYTES = KC * NCOL * MAXE * 16;     // (x, y, z, -w) per event
constexpr int OVB_BYTES = KC * NCOL * MAXE * 4;      // accumulator byte offset per event
constexpr int TAILMAX = 16;          // listed further events of the layer below the surface: (x, offset) per event
constexpr int TAIL_BYTES = NCOL * TAILMAX * 8;
constexpr int EDGE_WORDS = 24;       // per column: one-sided cubic coefficients, top (0..8) and bottom (12..20)
enum { CTX_NZ, CTX_MASK, CTX_TOA, CTX_REC, CTX_OVA, CTX_OVB, CTX_EDGE, CTX_TAIL, CTX_CELL0, CTX_CELLB, CTX_ACC, CTX_PE2,
       CTX_RDP2, CTX_ASTRIDE, CTX_WORDS };

struct Params {
    CUtensorMap maps[MAXSRC];
    float *outs[MAXV];
    int fa[MAXV], fb[MAXV];          // TMA sources of the parents (fb = fa for a plain field)
    const float *area;               // (n_outer, ny, nx)
    const float *psurf;              // (n_outer, ny, nx) sequential-cumsum surface pressure, mask_below only
    const float *delp_coarse;        // (n_outer, nz, ny/8, nx/8)
    const float *delp;               // (n_outer, nz, ny, nx), read by the geometry warps
    float fc1[MAXV], fc2[MAXV], fc3[MAXV];   // q = (a - c1 b) (c2 b + c3): plain (0,0,1), product (0,1,0), diff (1,0,1/scale)
    int any_derived;
    int nsrc, nfields, nz, ny, nx, mask_below;
    int nfw;                         // field warps = ceil(16 nfields / 32)
    float toa;
    int *flags;                      // per CTA: the fast variant met a non-finite field value (NaN-safe variant reruns the CTA)
};

// ---- shared-memory carve-up --------------------------------------------------------------------
struct Layout {
    int ring, stage_bytes, bars, cell0, cell_bytes;
    int pe2, rdp2, den;              // within a cell block
    int rec, ova, ovb, tail, ctx, edge, colw, colps, acc, acc_stride, total;
};

__host__ __device__ inline Layout make_layout(int nsrc, int nz, int nfw)
{
    Layout L;
    int o = 0;
    L.stage_bytes = nsrc * BOX_BYTES;
    L.ring = o;  o += NSTAGE * L.stage_bytes;    o = (o + 127) & ~127;
    L.bars = o;  o += 128;
    L.cell0 = o;
    int w = 0;
    L.pe2 = w;   w += (nz + 2) * 4;              w = (w + 15) & ~15;
    L.rdp2 = w;  w += (nz + 1) * 4;              w = (w + 15) & ~15;
    L.den = w;   w += (nz + 1) * 4;
    L.cell_bytes = (w + 127) & ~127;
    o += NCELL * L.cell_bytes;
    L.rec = o;   o += NSTAGE * REC_BYTES;
    L.ova = o;   o += NSTAGE * OVA_BYTES;
    L.ovb = o;   o += NSTAGE * OVB_BYTES;
    L.tail = o;  o += TAIL_BYTES;
    L.ctx = o;   o += 128;
    L.edge = o;  o += NCOL * EDGE_WORDS * 4;
    L.colw = o;  o += 8 * NCOL * 4;              // [row][column of the CTA]
    L.colps = o; o += 8 * NCOL * 4;
    L.acc_stride = nfw * 32 * 4;                 // bytes per accumulator slot: one float per work item
    L.acc = o;   o += (nz + 1) * L.acc_stride;
    L.total = o;
    return L;
}

