// ms_kde.cu -- per-vertex Gaussian KDE densities as a tiled fp64 pair-sum on B200.
//
// Replaces (LLNL/MemSurfer): kde_2d / kde_3d memsurfer/src/TriMesh_kde.cpp:344-392 (the
// O(V*n_ids) double loop: two virtual calls + one libm exp per pair), the normalisations of
// TriMesh::kde TriMesh_kde.cpp:504-524, GaussianKernel2D/3D DensityKernels.hpp:44-84 and
// DistanceSquared / DistancePeriodicXYSquared DistanceKernels.hpp:75-136 (fused in).
//
// Design (not a contraction -> no tensor cores; bound = the fp64 pipe, 64 DFMA/clk/SM):
//   * coordinates are pre-scaled by sqrt(-efactor*log2(e)*512) once per field, so the pair term is
//     2^-(r/512) with r = dx^2+dy^2[+dz^2]: no per-pair multiply by efactor; sfactor is applied
//     once per target;
//   * software exp2 in 6 fp64 instructions: round-to-nearest split of r with the 1.5*2^52 trick
//     (n = round(r), f = r - n in [-1/2,1/2]), 2^-(f/512) by a degree-2 near-minimax polynomial
//     (1.3e-11 relative; every term is positive, so this also bounds the sum), 2^-((n&511)/512) from a
//     512-entry shared-memory table, 2^-(n>>9) by an integer subtraction from the table value's exponent
//     field, all folded into the accumulating DFMA;
//   * two pair-sum kernels share that core:
//       k_kde_pairs   all V x n_ids pairs (the reference's sum).  Single-wrap minimum image as
//                     min(|dx|, boxw-|dx|): the reference adds boxw to the smaller coordinate when
//                     |dx| >= boxw/2, and (|dx|-boxw)^2 == min(|dx|, boxw-|dx|)^2;
//       k_kde_groups  pairs inside the window box(64 targets) (+) disc(cutoff_sigmas*sigma) of a fine cell list
//                     (dropped relative tail mass <= exp(-cutoff_sigmas^2/2)); the periodic image shift is
//                     applied once per staged batch, so there is no per-pair wrap arithmetic at all; sources at
//                     least 4.5 sigma away from the box are summed in float32 on the packed fp32 pipe + MUFU.EX2
//                     (their share of a sum is < 1e-4: kde_split_batch / kde_far_sum2, with an fp64 redo guard);
//   * source points are staged into shared memory by the TMA engine (cp.async.bulk + mbarrier,
//     double buffered); one broadcast LDS.128 per source feeds the lanes' DFMA chains.
#include "ms_common.cuh"
#include <mutex>
#include <math.h>
#include <stdlib.h>

#define KDE_STAGE_BYTES 16384    // all-pairs: bytes per shared-memory stage (1024 double2 / 512 double4)
#define KDE_MAGIC 6755399441055744.0   // 1.5 * 2^52
#define KDE_TAB_LOG 9
#define KDE_TAB (1 << KDE_TAB_LOG)
#define KDE_RMAX 2147483000.0    // the scaled r must stay below 2^31 (host check or in-kernel clamp)

__device__ const double c_exp2_tab[KDE_TAB] = {
#include "ms_exp2_table.inc"
};
// 2^(-f/512) on [-1/2,1/2] as 1 + f*(C1 + f*C2): near-minimax, relative error 1.3e-11 (tools/gen_exp2_table.py).
// The 512-entry table buys one DFMA per pair: with 256 entries the same accuracy class needs degree 3.
#define KDE_C1 (-0x1.62e43144a09e5p-10)
#define KDE_C2 (0x1.ebfbdff816ce7p-21)

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    uint32_t ok = 0;
    while (!ok) {
        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                     : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    }
}

// acc += 2^-(r/256), r in [0, 2^31)
__device__ __forceinline__ void kde_term(double r, const double *__restrict__ tab, double &acc) {
    const double t = r + KDE_MAGIC;
    const int n = __double2loint(t);
    const double f = r - (t - KDE_MAGIC);             // in [-1/2, 1/2]
    double p = fma(KDE_C2, f, KDE_C1);
    p = fma(p, f, 1.0);
    const double T = tab[n & (KDE_TAB - 1)];          // (1/2, 1]
    const int e = min(n >> KDE_TAB_LOG, 1021);                  // beyond 2^-1021 the term is held at ~1e-308 (i.e. dropped)
    const double sc = __hiloint2double(__double2hiint(T) - (e << 20), __double2loint(T));
    acc = fma(sc, p, acc);
}

template <int DIM> struct SrcT;
template <> struct SrcT<2> { typedef double2 type; };
template <> struct SrcT<3> { typedef double4 type; };

// all threads of the CTA: accumulate cnt staged sources into R register targets
template <int DIM, bool WRAP, bool CLAMP, int R>
__device__ __forceinline__ void kde_accumulate(const typename SrcT<DIM>::type *__restrict__ T, int cnt,
                                               const double *__restrict__ tab, const double (&tx)[R], const double (&ty)[R],
                                               const double (&tz)[R], double (&acc)[R], double wx, double wy) {
    typedef typename SrcT<DIM>::type S;
#pragma unroll 2
    for (int i = 0; i < cnt; i++) {
        const S sp = T[i];
#pragma unroll
        for (int r = 0; r < R; r++) {
            double dx = sp.x - tx[r], dy = sp.y - ty[r];
            if (WRAP) {
                // plain compare-select (fmin() would add NaN-quieting instructions per pair)
                // (only the square is used, so the sign of the kept value does not matter)
                const double bx = wx - fabs(dx), by = wy - fabs(dy);
                dx = fabs(dx) < bx ? dx : bx; dy = fabs(dy) < by ? dy : by;
            }
            double r2 = fma(dy, dy, dx * dx);
            if (DIM == 3) { const double dz = ((const double*)&sp)[2] - tz[r]; r2 = fma(dz, dz, r2); }
            if (CLAMP) r2 = r2 < KDE_RMAX ? r2 : KDE_RMAX;
            kde_term(r2, tab, acc[r]);
        }
    }
}

// ------------------------------------------------------------------ staging of scaled coordinates
// dst[i] = scale * vertex (ids ? ids[i] : i)
template <int DIM>
__global__ void __launch_bounds__(256) k_kde_stage(const float *__restrict__ verts, const int32_t *__restrict__ ids, int64_t n,
                                                   double scale, typename SrcT<DIM>::type *dst) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int64_t s = ids ? (int64_t)__ldg(ids + i) : i;
    typename SrcT<DIM>::type o;
    o.x = ms_ld_f(verts + 3*s) * scale; o.y = ms_ld_f(verts + 3*s + 1) * scale;
    if (DIM == 3) { ((double*)&o)[2] = ms_ld_f(verts + 3*s + 2) * scale; ((double*)&o)[3] = 0.0; }
    dst[i] = o;
}

// ------------------------------------------------------------------ all pairs
// grid.x: target blocks (THREADS*R targets each); grid.y: source splits (tiles [y*tps,(y+1)*tps)).
// partial[y*nv + j] = sum over the split's sources of 2^-(scaled distance^2 / 256).
template <int DIM, bool PERIODIC, bool CLAMP, int THREADS, int R>
__global__ void __launch_bounds__(THREADS) k_kde_pairs(const typename SrcT<DIM>::type *__restrict__ tgt, int64_t nv,
                                                       const typename SrcT<DIM>::type *__restrict__ src, int64_t ns,
                                                       int ntiles, int tiles_per_split,
                                                       double wx, double wy, double *__restrict__ partial) {
    typedef typename SrcT<DIM>::type S;
    constexpr int TILE = KDE_STAGE_BYTES / (int)sizeof(S);
    __shared__ __align__(128) S tile[2][TILE];
    __shared__ __align__(16) double tab[KDE_TAB];
    __shared__ __align__(8) uint64_t bar[2];
    const int tid = threadIdx.x;
    const int t_begin = blockIdx.y * tiles_per_split;
    const int t_end = min(ntiles, t_begin + tiles_per_split);
    const int64_t j0 = (int64_t)blockIdx.x * (THREADS * R) + tid;
    for (int i = tid; i < KDE_TAB; i += THREADS) tab[i] = c_exp2_tab[i];

    double tx[R], ty[R], tz[R], acc[R];
#pragma unroll
    for (int r = 0; r < R; r++) {
        const int64_t j = j0 + (int64_t)r * THREADS;
        S q; q.x = 0.0; q.y = 0.0; if (DIM == 3) { ((double*)&q)[2] = 0.0; }
        if (j < nv) q = tgt[j];
        tx[r] = q.x; ty[r] = q.y; tz[r] = DIM == 3 ? ((double*)&q)[2] : 0.0; acc[r] = 0.0;
    }
    if (tid == 0) { mbar_init(&bar[0], 1); mbar_init(&bar[1], 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
    __syncthreads();
    auto tile_count = [&](int t) { return (int)min((int64_t)TILE, ns - (int64_t)t * TILE); };
    if (tid == 0 && t_begin < t_end) {
        const uint32_t bytes = (uint32_t)tile_count(t_begin) * (uint32_t)sizeof(S);
        mbar_expect_tx(&bar[0], bytes); bulk_g2s(tile[0], src + (size_t)t_begin * TILE, bytes, &bar[0]);
    }
    uint32_t phase = 0u;                  // bit s = parity to wait for on stage s
    for (int t = t_begin; t < t_end; t++) {
        const int s = (t - t_begin) & 1;
        if (tid == 0 && t + 1 < t_end) {   // buffer s^1 was released by the __syncthreads closing iteration t-1
            const uint32_t bytes = (uint32_t)tile_count(t + 1) * (uint32_t)sizeof(S);
            mbar_expect_tx(&bar[s ^ 1], bytes);
            bulk_g2s(tile[s ^ 1], src + (size_t)(t + 1) * TILE, bytes, &bar[s ^ 1]);
        }
        mbar_wait(&bar[s], (phase >> s) & 1u); phase ^= 1u << s;
        kde_accumulate<DIM, PERIODIC, CLAMP, R>(tile[s], tile_count(t), tab, tx, ty, tz, acc, wx, wy);
        __syncthreads();
    }
#pragma unroll
    for (int r = 0; r < R; r++) {
        const int64_t j = j0 + (int64_t)r * THREADS;
        if (j < nv) partial[(size_t)blockIdx.y * nv + j] = acc[r];
    }
}

// ------------------------------------------------------------------ cell list (cut-off variant)
struct KdeGrid {
    double b0x, b0y;          // scaled origin of the grid
    double inv_cx, inv_cy;    // 1 / scaled cell size per axis
    double wx, wy;            // scaled extent (the period when periodic): ncx*cell_x, ncy*cell_y
    int ncx, ncy, periodic;
};

template <int DIM>
__device__ __forceinline__ int kde_cell_of(const KdeGrid &g, const typename SrcT<DIM>::type &p) {
    int cx = (int)floor((p.x - g.b0x) * g.inv_cx), cy = (int)floor((p.y - g.b0y) * g.inv_cy);
    if (g.periodic) {            // points on / beyond a face fold back: the cells tile the period exactly
        cx %= g.ncx; if (cx < 0) cx += g.ncx;
        cy %= g.ncy; if (cy < 0) cy += g.ncy;
    } else {
        cx = max(0, min(g.ncx - 1, cx)); cy = max(0, min(g.ncy - 1, cy));
    }
    return cy * g.ncx + cx;
}

// bin id of every point: cell (labels == NULL) or label*ncell + cell; points with a negative label
// are no sources (bin -1)
template <int DIM>
__global__ void __launch_bounds__(256) k_kde_cell_count(const typename SrcT<DIM>::type *__restrict__ pts, int64_t n, KdeGrid g,
                                                        const int32_t *__restrict__ labels, int nlabels, int32_t *cell, int32_t *cnt) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int c = kde_cell_of<DIM>(g, pts[i]);
    if (labels) { const int l = __ldg(labels + i); c = (l >= 0 && l < nlabels) ? l * (g.ncx * g.ncy) + c : -1; }
    cell[i] = c;
    if (c >= 0) atomicAdd(cnt + c, 1);
}
__global__ void __launch_bounds__(256) k_kde_cell_scatter(int64_t n, const int32_t *__restrict__ cell, const int32_t *__restrict__ start,
                                                          int32_t *cursor, int32_t *perm) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int c = cell[i];
    if (c >= 0) perm[start[c] + atomicAdd(cursor + c, 1)] = (int32_t)i;
}
// ascending element order inside every cell => the summation order (and the result) is reproducible.
// One LANE per cell for the common small cells (serial insertion sort of <= SORT_SMALL indices); the
// rare bigger cells of a warp's 32 are then sorted one at a time by the whole warp: every lane ranks its
// elements against the whole cell (counting sort by comparison, in registers); cells with more than 256
// elements fall back to a serial sort.
#define SORT_K 8
#define SORT_SMALL 16
__global__ void __launch_bounds__(128) k_kde_cell_sort(int ncell, const int32_t *__restrict__ start, int32_t *perm) {
    const int c0 = blockIdx.x * blockDim.x + threadIdx.x, lane = threadIdx.x & 31;
    int b0 = 0, n0 = 0;
    if (c0 < ncell) { b0 = start[c0]; n0 = start[c0 + 1] - b0; }
    if (n0 > 1 && n0 <= SORT_SMALL) {
        for (int i = b0 + 1; i < b0 + n0; i++) {
            const int32_t key = perm[i]; int j = i - 1;
            while (j >= b0 && perm[j] > key) { perm[j + 1] = perm[j]; j--; }
            perm[j + 1] = key;
        }
    }
    unsigned big = __ballot_sync(0xffffffffu, n0 > SORT_SMALL);
    while (big) {
        const int src = __ffs(big) - 1; big &= big - 1;
        const int b = __shfl_sync(0xffffffffu, b0, src), n = __shfl_sync(0xffffffffu, n0, src), e = b + n;
        if (n > 32 * SORT_K) {            // crowded cell (strongly clustered input): serial insertion sort, correct but slow
            if (lane == 0)
                for (int i = b + 1; i < e; i++) {
                    const int32_t key = perm[i]; int j = i - 1;
                    while (j >= b && perm[j] > key) { perm[j + 1] = perm[j]; j--; }
                    perm[j + 1] = key;
                }
            __syncwarp();
            continue;
        }
        int32_t v[SORT_K]; int rank[SORT_K];
#pragma unroll
        for (int k = 0; k < SORT_K; k++) { const int i = lane + 32 * k; v[k] = i < n ? perm[b + i] : 0x7fffffff; rank[k] = 0; }
        const int rounds = (n + 31) >> 5;
#pragma unroll
        for (int k2 = 0; k2 < SORT_K; k2++) {
            if (k2 < rounds) {
                for (int l = 0; l < 32; l++) {
                    const int32_t o = __shfl_sync(0xffffffffu, v[k2], l);
#pragma unroll
                    for (int k = 0; k < SORT_K; k++) rank[k] += o < v[k];
                }
            }
        }
        __syncwarp();
#pragma unroll
        for (int k = 0; k < SORT_K; k++) if (lane + 32 * k < n) perm[b + rank[k]] = v[k];      // indices are distinct
        __syncwarp();
    }
}

// sorted[i] = point perm[i], folded into the canonical period [b0, b0+w) when the grid is periodic so
// that its coordinates agree with the cell it was binned into
template <int DIM>
__global__ void __launch_bounds__(256) k_kde_cell_gather(const typename SrcT<DIM>::type *__restrict__ pts, int64_t n, KdeGrid g,
                                                         const int32_t *__restrict__ nbinned, const int32_t *__restrict__ perm,
                                                         typename SrcT<DIM>::type *sorted) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n || i >= (int64_t)__ldg(nbinned)) return;
    typename SrcT<DIM>::type p = pts[perm[i]];
    if (g.periodic) {
        const int cx = (int)floor((p.x - g.b0x) * g.inv_cx), cy = (int)floor((p.y - g.b0y) * g.inv_cy);
        const int kx = (cx >= 0 ? cx : cx - g.ncx + 1) / g.ncx, ky = (cy >= 0 ? cy : cy - g.ncy + 1) / g.ncy;
        p.x -= (double)kx * g.wx; p.y -= (double)ky * g.wy;
    }
    sorted[i] = p;
}

// ------------------------------------------------------------------ target groups (cut-off variant)
// Sources live in FINE cells (pitch ~rcut/8) sorted by (label, cell).  Targets are sorted by
// (strip of B fine rows, fine column, index) and cut into GROUPS of 64 consecutive targets of one
// strip: every lane slot is used (the fixed cell/tile versions wasted 12-25 % of the lanes).  The
// window of a group is the Minkowski sum box(group) (+) disc(rcut) at the fine-cell pitch: per fine row
// one contiguous range of the sorted source array whose half-width follows the circle.
// Every WARP is an autonomous worker: it draws (group, label, half-of-the-rows) items from an atomic
// counter, keeps its 64 targets in registers (2 per lane), plans its own batches (lane = row piece:
// two offset loads + a warp scan), stages them with cp.async.bulk into its private double buffer
// (one mbarrier per stage, no CTA-wide barrier anywhere) and adds its sums with one fp64 atomicAdd per
// target: with exactly two addends per output (the two halves) onto zero the result is independent of
// their order, i.e. bit-reproducible.  The 256-entry exp2 table is replicated 16x (entry e of copy k
// at double index e*16+k, lane l reads copy l&15): every lane of a half-warp owns its bank pair, so
// the table LDS.64 costs 2 wavefronts instead of ~6.5 (ncu on the first cell-list kernel: 4.1-way conflicts).
// CTA shape: registers beat warps.  No pipe is saturated (fp64 ~40 %, XU ~20 %, issue ~70 %): the loop is bound by
// dependent chains, so what counts is how many pairs a warp keeps in flight.  16 warps per SM with 128 registers each
// and 16 sources x 2 targets unrolled beat 24 warps with 80 registers (same box, ms per frame of the 8 launches):
//   12 x 2, unroll 8: 7.62 | 8 x 2, unroll 8: 6.88 | 8 x 2, unroll 16: 6.77 | 8 x 2, unroll 32: 6.99 | 16 x 1: 6.89
//   6 x 2: 7.29 | 12 x 1: 7.27 | 7 x 2: 8.08 | 10 x 2: 7.83 | 20 x 1: 7.86 | stages of 96 / 160 / 192 sources: 6.82 / 6.87 / 7.56
#ifndef GR_NW
#define GR_NW 8                  // warps per CTA
#endif
#ifndef GR_MINB
#define GR_MINB 2                // CTAs per SM
#endif
#ifndef GR_UNROLL
#define GR_UNROLL 16
#endif
#define GR_THREADS (GR_NW * 32)
#ifndef GR_CAP
#define GR_CAP 128               // sources per stage (per warp)
#endif
#ifndef GR_TGT
#define GR_TGT 64                // targets per group, GR_TGT/32 per lane.  Same box, ms/step of the launch series: 32: 9.40, 64: 8.12,
                                 // 96: 8.73, 128: 9.37 -- more targets share the per-batch work (planning, copies, the near/far
                                 // split) but widen the window (+11 % pairs at 64, +19 % at 96)
#endif
#define GR_TAB_BYTES (KDE_TAB * 16 * 8)

#define KDE_MAXGAP 47
#ifndef KDE_FAR_SIGMAS
#define KDE_FAR_SIGMAS 4.5
#endif
#define KDE_FAR_SHARE 1e-4      // largest fp32 part of a sum, relative to its fp64 part, that is accepted (else: fp64 redo)
struct KdeTiling {
    KdeGrid gs;               // fine source grid
    int B, nty, mf, wmax;     // fine rows per strip, strips, window reach in fine rows, widest group (columns)
    double rc2, cxs, cys;     // scaled cut-off radius^2 and fine cell pitch
    double ri2;               // scaled radius^2 beyond which a pair is summed in fp32 (0: no far field)
    float hx[KDE_MAXGAP + 1]; // half width of the window at `gap` full rows from the strip, rounded UP (< 0: the row is outside)
    double gscale;            // factor on the stored group x extents (1 unless the binning is shared between sigmas)
    double oscale;            // factor on every sum before it is added to the output (the kernel's sfactor, or 1)
};
struct KdeGroup { int t0, t1, strip, ca, cb, pad0, pad1, pad2; double xlo, xhi; };   // targets [t0,t1), fine columns [ca,cb], x extent

// The window of the box [xlo,xhi] x strip TY is cut, row by row, into contiguous ranges of the sorted source array:
// piece q = nparts*row + part, part 1 = what wrapped around the period.  code = 3*(sx+1)+(sy+1): the image shift
// (sx*wx, sy*wy) to add to the targets.
__device__ __forceinline__ void kde_group_piece(const KdeTiling &T, int TY, double xlo, double xhi, int nparts, int q,
                                                const int32_t *__restrict__ sstart, int &off, int &cnt, int &code) {
    off = 0; cnt = 0; code = 4;
    const KdeGrid &g = T.gs;
    const int rows = T.B + 2 * T.mf;
    // nparts = 2 only for groups whose window crosses the period in x (kde_window_parts): twice the pieces, half of them empty
    const int part = nparts == 2 ? q & 1 : 0, ri = nparts == 2 ? q >> 1 : q;
    if (ri >= rows) return;
    const int gap = ri < T.mf ? T.mf - 1 - ri : (ri >= T.mf + T.B ? ri - T.mf - T.B : 0);   // full fine rows between row and strip
    double hx;
    if (gap <= KDE_MAXGAP) hx = (double)T.hx[gap];
    else { const double gy = (double)gap * T.cys, h2 = T.rc2 - gy * gy; hx = h2 > 0.0 ? sqrt(h2) : -1.0; }
    if (hx < 0.0) return;
    int row = TY * T.B - T.mf + ri, sy = 0;
    if (row < 0) { if (!g.periodic) return; row += g.ncy; sy = 1; }
    else if (row >= g.ncy) { if (!g.periodic) return; row -= g.ncy; sy = -1; }
    int xa = (int)floor((xlo - hx - g.b0x) * g.inv_cx), xb = (int)floor((xhi + hx - g.b0x) * g.inv_cx), sx = 0;    // unwrapped columns
    if (!g.periodic) {
        if (part) return;
        xa = max(xa, 0); xb = min(xb, g.ncx - 1);
    } else if (xb < 0) {
        if (part) return;
        xa += g.ncx; xb += g.ncx; sx = 1;
    } else if (xa >= g.ncx) {
        if (part) return;
        xa -= g.ncx; xb -= g.ncx; sx = -1;
    } else if (xa < 0) {
        if (part == 0) { xa += g.ncx; xb = g.ncx - 1; sx = 1; } else { xa = 0; }
    } else if (xb >= g.ncx) {
        if (part == 0) { xb = g.ncx - 1; } else { xa = 0; xb -= g.ncx; sx = -1; }
    } else if (part) return;
    if (xa > xb) return;
    off = __ldg(sstart + row * g.ncx + xa);
    cnt = __ldg(sstart + row * g.ncx + xb + 1) - off;
    code = 3 * (sx + 1) + (sy + 1);
}
// does the window of the box [xlo, xhi] reach across the period in x (then every range may come in two parts)?
__device__ __forceinline__ int kde_window_parts(const KdeTiling &T, double xlo, double xhi) {
    if (!T.gs.periodic) return 1;
    const double hx = (double)T.hx[0];
    const int xa = (int)floor((xlo - hx - T.gs.b0x) * T.gs.inv_cx), xb = (int)floor((xhi + hx - T.gs.b0x) * T.gs.inv_cx);
    return (xa < 0 || xb >= T.gs.ncx) ? 2 : 1;
}

// 2^-(r/256) = sc * p with the lane-private table copy (tbase = shared address of the table + 8*(lane&15)).
// !ECLAMP: the table holds T'[i] = T[i] with (i << 12) pre-added to its high word, so that the exponent
// shift is ONE integer multiply-add: hi(T') - (n << 12) = hi(T) - ((n >> 8) << 20)  (n < 2^18 guaranteed
// by the host: the window keeps r below 256*1000).  Per pair: LOP3 + IMAD + LDS.64 + IMAD.
template <bool ECLAMP>
__device__ __forceinline__ double kde_scale_t(int n, uint32_t tbase) {
    int lo, hi;
    asm("{ .reg .b32 i, a; and.b32 i, %2, %4; mad.lo.u32 a, i, 128, %3; ld.shared.v2.b32 {%0, %1}, [a]; }"
        : "=r"(lo), "=r"(hi) : "r"(n), "r"(tbase), "n"(KDE_TAB - 1));
    if (ECLAMP) hi -= min(n >> KDE_TAB_LOG, 1021) << 20;
    else        hi -= n << (20 - KDE_TAB_LOG);
    return __hiloint2double(hi, lo);
}
template <bool ECLAMP>
__device__ __forceinline__ void kde_term_t(double r, uint32_t tbase, double &acc) {
    const double t = r + KDE_MAGIC;
    const int n = __double2loint(t);
    const double f = r - (t - KDE_MAGIC);
    double p = fma(KDE_C2, f, KDE_C1);
    p = fma(p, f, 1.0);
    acc = fma(kde_scale_t<ECLAMP>(n, tbase), p, acc);
}

template <int DIM, bool CLAMP, int R>
__device__ __forceinline__ void kde_accumulate_t(const typename SrcT<DIM>::type *__restrict__ S, int cnt, uint32_t tabl,
                                                 const double (&tx)[R], const double (&ty)[R], const double (&tz)[R], double (&acc)[R]) {
    constexpr int kUnroll = GR_UNROLL;
#pragma unroll kUnroll
    for (int i = 0; i < cnt; i++) {
        const typename SrcT<DIM>::type sp = S[i];
#pragma unroll
        for (int r = 0; r < R; r++) {
            const double dx = sp.x - tx[r], dy = sp.y - ty[r];
            double r2 = fma(dy, dy, dx * dx);
            if (DIM == 3) { const double dz = ((const double*)&sp)[2] - tz[r]; r2 = fma(dz, dz, r2); }
            if (CLAMP) r2 = r2 < KDE_RMAX ? r2 : KDE_RMAX;
            kde_term_t<CLAMP>(r2, tabl, acc[r]);
        }
    }
}

// ---- groups of 64 consecutive targets per strip
__global__ void __launch_bounds__(128) k_kde_group_count(const int32_t *__restrict__ tstart, int ncx, int nty, int32_t *ng) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s < nty) ng[s] = (tstart[(s + 1) * ncx] - tstart[s * ncx] + GR_TGT - 1) / GR_TGT;
}
__global__ void __launch_bounds__(128) k_kde_group_fill(const int32_t *__restrict__ tstart, const int32_t *__restrict__ tcell,
                                                        const int32_t *__restrict__ tperm, int ncx, int nty,
                                                        const int32_t *__restrict__ gstart, const double *__restrict__ tx, int xstride,
                                                        KdeGroup *groups) {
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= gstart[nty]) return;
    int lo = 0, hi = nty;                       // largest s with gstart[s] <= g
    while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (gstart[mid] <= g) lo = mid; else hi = mid; }
    const int s = lo, b = tstart[s * ncx], e = tstart[(s + 1) * ncx];
    KdeGroup G;
    G.t0 = b + (g - gstart[s]) * GR_TGT; G.t1 = min(G.t0 + GR_TGT, e); G.strip = s;
    G.ca = tcell[tperm[G.t0]] % ncx; G.cb = tcell[tperm[G.t1 - 1]] % ncx;
    G.pad0 = G.pad1 = G.pad2 = 0;
    double xlo = tx[(size_t)G.t0 * xstride], xhi = xlo;
    for (int j = G.t0 + 1; j < G.t1; j++) { const double x = tx[(size_t)j * xstride]; xlo = fmin(xlo, x); xhi = fmax(xhi, x); }
    G.xlo = xlo; G.xhi = xhi;
    groups[g] = G;
}

// ---- per-warp batch planner: the warp caches 32 consecutive pieces (lane = piece) and hands out runs of
// them that fit one stage and share one image shift; a piece larger than a stage is sliced
struct KdePlanState { int pbase, pos, intra, c_off, c_cnt, c_code, nparts; };
struct KdeBatch { int total, code, take_off, take_cnt, dst; };

template <int CAP>
__device__ __forceinline__ bool kde_plan_next(KdePlanState &st, const KdeTiling &T, int TY, double xlo, double xhi, int q_hi,
                                              const int32_t *__restrict__ sstart, int lane, KdeBatch &b) {
    for (;;) {
        if (st.pos >= 32) {
            st.pbase += 32; st.pos = 0; st.intra = 0;
            if (st.pbase >= q_hi) return false;
            kde_group_piece(T, TY, xlo, xhi, st.nparts, st.pbase + lane, sstart, st.c_off, st.c_cnt, st.c_code);
            if (st.pbase + lane >= q_hi) st.c_cnt = 0;
        }
        const int eff = lane < st.pos ? 0 : (lane == st.pos ? st.c_cnt - st.intra : st.c_cnt);
        const unsigned ne = __ballot_sync(0xffffffffu, eff > 0);
        if (ne == 0u) { st.pos = 32; continue; }
        const int first = __ffs(ne) - 1;
        const int code = __shfl_sync(0xffffffffu, st.c_code, first);
        int pre = eff;                                   // inclusive prefix sum of the remaining piece sizes
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int v = __shfl_up_sync(0xffffffffu, pre, o); if (lane >= o) pre += v; }
        // the batch: the pieces of one image shift, in order, until the stage is full -- the piece that crosses the end
        // of the stage is cut there and resumed in the next batch, so every batch but the last of a shift is full
        const unsigned other = __ballot_sync(0xffffffffu, eff > 0 && st.c_code != code);
        const int nsame = other ? __ffs(other) - 1 : 32;
        const unsigned over = __ballot_sync(0xffffffffu, lane < nsame && eff > 0 && pre > CAP);
        b.code = code; b.dst = pre - eff;
        b.take_off = st.c_off + (lane == st.pos ? st.intra : 0);
        if (over == 0u) {
            b.take_cnt = lane < nsame ? eff : 0;
            b.total = __shfl_sync(0xffffffffu, pre, nsame - 1);
            st.pos = nsame; st.intra = 0;
        } else {
            const int cut = __ffs(over) - 1;
            const int part = CAP - (pre - eff);       // what of the cut piece still fits (>= 0 on lane `cut`)
            b.take_cnt = lane < cut ? eff : (lane == cut ? part : 0);
            b.total = CAP;
            const int part_c = __shfl_sync(0xffffffffu, part, cut);
            st.intra = (cut == st.pos ? st.intra : 0) + part_c;
            st.pos = cut;
        }
        return true;
    }
}

// lane 0 issues the copies of every lane's piece (a cp.async.bulk is a one-thread instruction: spread over
// the lanes the compiler serialises it anyway and leaves the warp split for the rest of the loop)
template <class S>
__device__ __forceinline__ void kde_issue_batch(const KdeBatch &b, S *stage, const S *__restrict__ src, uint64_t *bar, int lane) {
    unsigned m = __ballot_sync(0xffffffffu, b.take_cnt > 0);
    if (lane == 0) mbar_expect_tx(bar, (uint32_t)b.total * (uint32_t)sizeof(S));
    while (m) {
        const int i = __ffs(m) - 1; m &= m - 1;
        const int off = __shfl_sync(0xffffffffu, b.take_off, i), cnt = __shfl_sync(0xffffffffu, b.take_cnt, i);
        const int dst = __shfl_sync(0xffffffffu, b.dst, i);
        if (lane == 0) bulk_g2s(stage + dst, src + off, (uint32_t)cnt * (uint32_t)sizeof(S), bar);
    }
    __syncwarp();
}

// ---- far field in fp32.  A source that is at least r_in = 4.5 sigma away from every point of the group's box gives
// terms below exp(-10.1) = 4e-5 of a near one, and all such terms together carry ~1e-4 of a sum: fp32 is enough for
// them -- IF the coordinates are small.  Per staged batch the warp therefore (lane-parallel, four sources per lane)
// measures every source against the box and rewrites the stage in place: the near sources compacted at its head as
// they are, the far ones behind them as float coordinates relative to the centre of the group and scaled by
// 1/sqrt(512) (one fp64 subtraction per SOURCE, not per pair: relative error of a term ~6e-6, of the sum ~2e-11).
// The pair loops then run over the two lists: 10 DFMA-class + 7 other instructions per near pair, FADD/FFMA +
// MUFU.EX2 = 8 issue slots per far pair.  Returns the number of near sources.
// sources per stage: a 3-D source is 32 bytes, so half as many keep the stage (and the CTAs per SM) what they are in 2-D
template <int DIM> struct GrCap { static constexpr int value = DIM == 3 ? GR_CAP / 2 : GR_CAP; };
template <int DIM> struct FarT;
template <> struct FarT<2> { typedef float2 type; };
template <> struct FarT<3> { typedef float4 type; };
template <int DIM>
__device__ __forceinline__ int kde_split_batch(typename SrcT<DIM>::type *Bs, int cnt, float hwx, float hwy, float ri2f,
                                               double cx, double cy, float FS, int lane) {
    typedef typename SrcT<DIM>::type S;
    typedef typename FarT<DIM>::type F;
    constexpr int K = GrCap<DIM>::value / 32;
    S v[K];
    float fx[K], fy[K];
    unsigned mn[K], mf[K];
    int nnear = 0;
#pragma unroll
    for (int k = 0; k < K; k++) {
        const bool valid = lane + 32 * k < cnt;
        bool isfar = false;
        if (valid) {
            v[k] = Bs[lane + 32 * k];
            fx[k] = (float)(v[k].x - cx); fy[k] = (float)(v[k].y - cy);
            const float ax = fmaxf(fabsf(fx[k]) - hwx, 0.f), ay = fmaxf(fabsf(fy[k]) - hwy, 0.f);
            isfar = fmaf(ay, ay, ax * ax) >= ri2f;
        }
        mn[k] = __ballot_sync(0xffffffffu, valid && !isfar);
        mf[k] = __ballot_sync(0xffffffffu, valid && isfar);
        nnear += __popc(mn[k]);
    }
    __syncwarp();                                  // every load above is done before the first store below
    F *Fs = (F*)(Bs + nnear);
    const unsigned lt = (1u << lane) - 1u;
    int nb = 0, fb = 0;
#pragma unroll
    for (int k = 0; k < K; k++) {
        if ((mn[k] >> lane) & 1u) Bs[nb + __popc(mn[k] & lt)] = v[k];
        else if ((mf[k] >> lane) & 1u) {
            const int j = fb + __popc(mf[k] & lt);
            if (DIM == 2) {                        // two sources per 16 bytes: {x0, x1, y0, y1} for the packed FADD2/FFMA2 loop
                float *q = (float*)Fs + 4 * (j >> 1) + (j & 1);
                q[0] = fx[k] * FS; q[2] = fy[k] * FS;
            } else {
                F f;
                f.x = fx[k] * FS; f.y = fy[k] * FS;
                ((float*)&f)[2] = (float)((const double*)&v[k])[2] * FS; ((float*)&f)[3] = 0.f;
                Fs[j] = f;
            }
        }
        nb += __popc(mn[k]); fb += __popc(mf[k]);
    }
    if (DIM == 2 && (fb & 1) && lane == 0) {       // odd count: the last pair is completed by a source at infinity (term 0)
        float *q = (float*)Fs + 4 * (fb >> 1) + 1;
        q[0] = 1e18f; q[2] = 0.f;
    }
    __syncwarp();
    return nnear;
}
// 2-D far field, two sources per step on the packed fp32 pipe: FADD2 x2, FMUL2, FFMA2, 2 x MUFU.EX2, FADD2 per TWO pairs
// (the scalar loop spends 8 issue slots per pair, of which the MUFU is one: the XU pipe sustains one per 8 cycles, so
// the scalar loop keeps the issue port as busy as the XU -- this one leaves half of the port to the other warps)
template <int R>
__device__ __forceinline__ void kde_far_sum2(const float4 *Fp, int cnt, const float (&ftx)[R], const float (&fty)[R], float (&facc)[R]) {
    float2 ntx[R], nty[R], a2[R];
#pragma unroll
    for (int r = 0; r < R; r++) { ntx[r] = make_float2(-ftx[r], -ftx[r]); nty[r] = make_float2(-fty[r], -fty[r]); a2[r] = make_float2(0.f, 0.f); }
    const int np = (cnt + 1) >> 1;
#pragma unroll 4
    for (int i = 0; i < np; i++) {
        const float4 sp = Fp[i];
#pragma unroll
        for (int r = 0; r < R; r++) {
            const float2 dx = __fadd2_rn(make_float2(sp.x, sp.y), ntx[r]), dy = __fadd2_rn(make_float2(sp.z, sp.w), nty[r]);
            const float2 r2 = __ffma2_rn(dy, dy, __fmul2_rn(dx, dx));
            float2 e;
            asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e.x) : "f"(-r2.x));
            asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e.y) : "f"(-r2.y));
            a2[r] = __fadd2_rn(a2[r], e);
        }
    }
#pragma unroll
    for (int r = 0; r < R; r++) facc[r] += a2[r].x + a2[r].y;
}
template <int DIM, int R>
__device__ __forceinline__ void kde_far_sum(const typename FarT<DIM>::type *Fs, int cnt, const float (&ftx)[R], const float (&fty)[R],
                                            const float (&ftz)[R], float (&facc)[R]) {
    typedef typename FarT<DIM>::type F;
#pragma unroll 8
    for (int i = 0; i < cnt; i++) {
        const F sp = Fs[i];
#pragma unroll
        for (int r = 0; r < R; r++) {
            const float dx = sp.x - ftx[r], dy = sp.y - fty[r];
            float r2 = fmaf(dy, dy, dx * dx);
            if (DIM == 3) { const float dz = ((const float*)&sp)[2] - ftz[r]; r2 = fmaf(dz, dz, r2); }
            float e;
            asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(-r2));
            facc[r] += e;
        }
    }
}

template <int DIM, bool CLAMP>
__global__ void __launch_bounds__(GR_THREADS, GR_MINB) k_kde_groups(const typename SrcT<DIM>::type *__restrict__ tgt, const int32_t *__restrict__ tstart,
                                                                   const int32_t *__restrict__ tperm, const KdeGroup *__restrict__ groups,
                                                                   const int32_t *__restrict__ gtotal,
                                                                   const typename SrcT<DIM>::type *__restrict__ src, const int32_t *__restrict__ sstart_all,
                                                                   int nlabels, int64_t out_stride, KdeTiling T, unsigned *counter,
                                                                   double *__restrict__ out_all) {
    typedef typename SrcT<DIM>::type S;
    constexpr int CAP = GrCap<DIM>::value;
    extern __shared__ __align__(128) unsigned char gr_smem[];
    double *tab = (double*)gr_smem;                                             // [KDE_TAB][16]
    S *buf_all = (S*)(gr_smem + GR_TAB_BYTES);                                  // [GR_NW][2][CAP]
    uint64_t *bar_all = (uint64_t*)(gr_smem + GR_TAB_BYTES + (size_t)GR_NW * 2 * CAP * sizeof(S));   // [GR_NW][2]
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    for (int i = tid; i < KDE_TAB * 16; i += GR_THREADS) {
        const double v = c_exp2_tab[i >> 4];
        tab[i] = CLAMP ? v : __hiloint2double(__double2hiint(v) + ((i >> 4) << (20 - KDE_TAB_LOG)), __double2loint(v));
    }
    S *buf = buf_all + (size_t)warp * 2 * CAP;
    uint64_t *bar = bar_all + warp * 2;
    if (lane == 0) { mbar_init(&bar[0], 1); mbar_init(&bar[1], 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
    __syncthreads();
    const uint32_t tabl = smem_u32(tab) + 8u * (uint32_t)(lane & 15);
    const int ncell = T.gs.ncx * T.gs.ncy;
    const int nrows = T.B + 2 * T.mf;
    const int r_mid = nrows / 2;                 // an item takes the rows [0, r_mid) or [r_mid, nrows) of its window
#ifndef GR_HALVES
#define GR_HALVES 2
#endif
    const int nitems = __ldg(gtotal) * nlabels * GR_HALVES;
    const bool far = T.ri2 > 0.0;
    uint32_t phase = 0u;                       // bit s = parity to wait for on stage s

    for (;;) {
        int item = 0;
        if (lane == 0) item = (int)atomicAdd(counter, 1u);
        item = __shfl_sync(0xffffffffu, item, 0);
        if (item >= nitems) break;
        const int half = item % GR_HALVES, lab = (item / GR_HALVES) % nlabels, g = (item / GR_HALVES) / nlabels;
        const int4 gi = __ldg((const int4*)(groups + g));            // t0, t1, strip, ca
        const int g_cb = __ldg(&groups[g].cb);
        const double g_xlo = __ldg(&groups[g].xlo) * T.gscale, g_xhi = __ldg(&groups[g].xhi) * T.gscale;
        const int32_t *__restrict__ sstart = sstart_all + (size_t)lab * ncell;
        const int r_lo = half ? r_mid : 0, r_hi = (half || GR_HALVES == 1) ? nrows : r_mid;
        for (int pa = gi.w; pa <= g_cb; pa += T.wmax) {               // one pass unless the group is wider than the period allows
            const int pb = min(g_cb, pa + T.wmax - 1);
            int a0 = gi.x, a1 = gi.y;
            double xlo = g_xlo, xhi = g_xhi;
            if (pa != gi.w || pb != g_cb) {
                a0 = max(a0, __ldg(tstart + gi.z * T.gs.ncx + pa));
                a1 = min(a1, __ldg(tstart + gi.z * T.gs.ncx + pb + 1));
                xlo = T.gs.b0x + (double)pa * T.cxs; xhi = T.gs.b0x + (double)(pb + 1) * T.cxs;
            }
            if (a0 >= a1) continue;
            // GR_R targets per lane
            constexpr int GR_R = GR_TGT / 32;
            double tx[GR_R], ty[GR_R], tz[GR_R], acc[GR_R];
            // far field: coordinates relative to the centre of the group, scaled by 1/sqrt(512) so that the fp32 term is
            // ex2(-(dx^2 + dy^2)); the centre moves with the periodic image of a batch, the targets do not
            const double gcx = 0.5 * (xlo + xhi), gcy = T.gs.b0y + ((double)gi.z + 0.5) * (double)T.B * T.cys;
            const float FS = 0.044194173824159216f;                  // 1 / sqrt(512)
            float ftx[GR_R], fty[GR_R], ftz[GR_R], facc[GR_R];
#pragma unroll
            for (int r = 0; r < GR_R; r++) {
                const int j = a0 + lane + 32 * r;
                S q; q.x = 0.0; q.y = 0.0; if (DIM == 3) { ((double*)&q)[2] = 0.0; }
                if (j < a1) q = tgt[j];
                tx[r] = q.x; ty[r] = q.y; tz[r] = DIM == 3 ? ((double*)&q)[2] : 0.0;
                acc[r] = 0.0; facc[r] = 0.f;
                ftx[r] = (float)(tx[r] - gcx) * FS; fty[r] = (float)(ty[r] - gcy) * FS; ftz[r] = (float)tz[r] * FS;
            }
            // the distance of a source to the box is taken in fp32 from the centre-relative coordinates: the margin covers
            // the rounding, a source within it simply stays in the fp64 list
            const float hwx = (float)(0.5 * (xhi - xlo)), hwy = (float)(0.5 * (double)T.B * T.cys), ri2f = (float)T.ri2 * 1.0001f;
            const int nparts = kde_window_parts(T, xlo, xhi);
            // The fp32 part of a sum is trusted only while it is a small part: if it exceeds KDE_FAR_SHARE of the fp64
            // part on any target (a target with few or no near sources: sparse ids, the rim of a cluster), the whole
            // item is summed again in fp64.  Bound on the result: KDE_FAR_SHARE x the fp32 sum's error (~5e-6) < 1e-9.
            bool use_far = far;
          again:
            KdePlanState st; st.pbase = nparts * r_lo - 32; st.pos = 32; st.intra = 0; st.c_off = st.c_cnt = 0; st.c_code = 4;
            st.nparts = nparts;
            const int q_hi = nparts * r_hi;
            KdeBatch cur, nxt;
            int s = 0;
            bool have = kde_plan_next<CAP>(st, T, gi.z, xlo, xhi, q_hi, sstart, lane, cur);
            if (have) kde_issue_batch<S>(cur, buf, src, &bar[0], lane);
            while (have) {
                const bool more = kde_plan_next<CAP>(st, T, gi.z, xlo, xhi, q_hi, sstart, lane, nxt);
                if (more) kde_issue_batch<S>(nxt, buf + (s ^ 1) * CAP, src, &bar[s ^ 1], lane);    // stage s^1 was drained in the previous round
                mbar_wait(&bar[s], (phase >> s) & 1u); phase ^= 1u << s;
                __syncwarp();                  // the lanes leave the try_wait loop one by one: reconverge before the pair loop
                const int cdx = (cur.code * 11) >> 5, cdy = cur.code - 3 * cdx;           // code / 3, code % 3 for code < 9
                const double shx = cdx == 1 ? 0.0 : (cdx ? T.gs.wx : -T.gs.wx), shy = cdy == 1 ? 0.0 : (cdy ? T.gs.wy : -T.gs.wy);
                S *Bs = buf + s * CAP;
                int nnear = cur.total;
                if (use_far) nnear = kde_split_batch<DIM>(Bs, cur.total, hwx, hwy, ri2f, gcx + shx, gcy + shy, FS, lane);
                {
                    double sx[GR_R], sy[GR_R];
#pragma unroll
                    for (int r = 0; r < GR_R; r++) { sx[r] = tx[r] + shx; sy[r] = ty[r] + shy; }
                    kde_accumulate_t<DIM, CLAMP, GR_R>(Bs, nnear, tabl, sx, sy, tz, acc);
                }
                if (use_far) {
                    if (DIM == 2) kde_far_sum2<GR_R>((const float4*)(Bs + nnear), cur.total - nnear, ftx, fty, facc);
                    else kde_far_sum<DIM, GR_R>((const typename FarT<DIM>::type*)(Bs + nnear), cur.total - nnear, ftx, fty, ftz, facc);
                    __syncwarp();
                    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");      // the next bulk copy into this stage follows generic stores
                } else __syncwarp();
                s ^= 1; have = more; cur = nxt;
            }
            if (use_far) {
                bool redo = false;
#pragma unroll
                for (int r = 0; r < GR_R; r++) redo |= a0 + lane + 32 * r < a1 && (double)facc[r] > KDE_FAR_SHARE * acc[r];
                if (__any_sync(0xffffffffu, redo)) {
#pragma unroll
                    for (int r = 0; r < GR_R; r++) { acc[r] = 0.0; facc[r] = 0.f; }
                    use_far = false;
                    goto again;
                }
            }
#pragma unroll
            for (int r = 0; r < GR_R; r++) {
                const int j = a0 + lane + 32 * r;
                if (j < a1) atomicAdd(out_all + (size_t)lab * out_stride + __ldg(tperm + j), (acc[r] + (double)facc[r]) * T.oscale);
            }
        }
    }
}

// pairs the group kernel evaluates: sum over groups of (#targets) x (#sources of every label in the window);
// total[0] = all pairs, total[1] = the near-field (fp64) ones
__global__ void __launch_bounds__(128) k_kde_groups_count(const KdeGroup *__restrict__ groups, const int32_t *__restrict__ gtotal,
                                                          const int32_t *__restrict__ tstart, const int32_t *__restrict__ sstart_all,
                                                          const double *__restrict__ src, int sstride,
                                                          int nlabels, KdeTiling T, unsigned long long *total) {
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    const int ncell = T.gs.ncx * T.gs.ncy, nrows = T.B + 2 * T.mf;
    unsigned long long mine = 0ull, near_ = 0ull;
    if (g < gtotal[0]) {
        const KdeGroup G = groups[g];
        for (int pa = G.ca; pa <= G.cb; pa += T.wmax) {
            const int pb = min(G.cb, pa + T.wmax - 1);
            int a0 = G.t0, a1 = G.t1;
            double xlo = G.xlo * T.gscale, xhi = G.xhi * T.gscale;
            if (pa != G.ca || pb != G.cb) {
                a0 = max(a0, tstart[G.strip * T.gs.ncx + pa]); a1 = min(a1, tstart[G.strip * T.gs.ncx + pb + 1]);
                xlo = T.gs.b0x + (double)pa * T.cxs; xhi = T.gs.b0x + (double)(pb + 1) * T.cxs;
            }
            if (a0 >= a1) continue;
            unsigned long long nsrc = 0ull, nnear = 0ull;
            const double ylo = T.gs.b0y + (double)G.strip * (double)T.B * T.cys, yhi = ylo + (double)T.B * T.cys;
            for (int l = 0; l < nlabels; l++) {
                int off, cnt, code;
                for (int q = 0; q < 2 * nrows; q++) {
                    kde_group_piece(T, G.strip, xlo, xhi, 2, q, sstart_all + (size_t)l * ncell, off, cnt, code);
                    nsrc += (unsigned long long)cnt;
                    if (T.ri2 > 0.0) {               // the same test as kde_split_batch
                        const double shx = (double)(code / 3 - 1) * T.gs.wx, shy = (double)(code % 3 - 1) * T.gs.wy;
                        for (int i = off; i < off + cnt; i++) {
                            const double x = src[(size_t)i * sstride] * T.gscale, y = src[(size_t)i * sstride + 1] * T.gscale;   // staged at the reference sigma's scale
                            const double dx = fmax(fmax(xlo + shx - x, x - xhi - shx), 0.0), dy = fmax(fmax(ylo + shy - y, y - yhi - shy), 0.0);
                            nnear += fma(dy, dy, dx * dx) >= T.ri2 ? 0ull : 1ull;
                        }
                    } else nnear += (unsigned long long)cnt;
                }
            }
            mine += nsrc * (unsigned long long)(a1 - a0);
            near_ += nnear * (unsigned long long)(a1 - a0);
        }
    }
    for (int o = 16; o > 0; o >>= 1) { mine += __shfl_xor_sync(0xffffffffu, mine, o); near_ += __shfl_xor_sync(0xffffffffu, near_, o); }
    if ((threadIdx.x & 31) == 0 && mine) { atomicAdd(total, mine); atomicAdd(total + 1, near_); }
}

// ------------------------------------------------------------------ normalisation (TriMesh_kde.cpp:504-524)
#define RED_T 256

__global__ void __launch_bounds__(RED_T) k_kde_combine(const double *__restrict__ partial, int64_t nv, int nsplit,
                                                       double factor, double *__restrict__ dens, double *__restrict__ bsum) {
    __shared__ double sm[RED_T];
    const int64_t j = (int64_t)blockIdx.x * RED_T + threadIdx.x;
    double d = 0.0;
    if (j < nv) {
        for (int y = 0; y < nsplit; y++) d += partial[(size_t)y * nv + j];
        d *= factor;
        dens[j] = d;
    }
    sm[threadIdx.x] = d;
    __syncthreads();
    for (int o = RED_T / 2; o > 0; o >>= 1) { if (threadIdx.x < o) sm[threadIdx.x] += sm[threadIdx.x + o]; __syncthreads(); }
    if (threadIdx.x == 0) bsum[blockIdx.x] = sm[0];
}
// get_counts: every block re-reduces the nb block sums in the same fixed order (a few KB from L2;
// deterministic, and no separate single-block reduction launch), then scales its 256 values
__global__ void __launch_bounds__(RED_T) k_kde_emit(const double *__restrict__ dens, int64_t nv, const double *__restrict__ bsum, int64_t nb,
                                                    double npoints, double *__restrict__ o64, float *__restrict__ o32) {
    __shared__ double sm[RED_T];
    double scale = 1.0;
    if (bsum) {
        double s = 0.0;
        for (int64_t i = threadIdx.x; i < nb; i += RED_T) s += bsum[i];
        sm[threadIdx.x] = s;
        __syncthreads();
        for (int o = RED_T / 2; o > 0; o >>= 1) { if (threadIdx.x < o) sm[threadIdx.x] += sm[threadIdx.x + o]; __syncthreads(); }
        scale = npoints / sm[0];
    }
    const int64_t j = (int64_t)blockIdx.x * RED_T + threadIdx.x;
    if (j >= nv) return;
    double d = dens[j];
    if (bsum) d *= scale;
    if (o64) o64[j] = d;
    if (o32) o32[j] = (float)d;
}

// stable partition by label (all-pairs, label-fused): flag -> exclusive scan -> scatter, one label at a time
__global__ void __launch_bounds__(256) k_kde_label_flag(const int32_t *__restrict__ labels, int64_t n, int l, int32_t *flag) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) flag[i] = __ldg(labels + i) == l;
}
__global__ void __launch_bounds__(256) k_kde_label_scatter(const int32_t *__restrict__ labels, int64_t n, int l, const int32_t *__restrict__ pos,
                                                           int32_t *start, int32_t *perm) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int base = start[l];
    if (__ldg(labels + i) == l) perm[base + pos[i]] = (int32_t)i;
    if (i == 0) start[l + 1] = base + pos[n];
}

// ------------------------------------------------------------------ label-fused normalisation
// dens[j] = factor * sum_{l in mask} raw[l*nv + j]  (+ block sums for the get_counts total)
__global__ void __launch_bounds__(RED_T) k_kde_select(const double *__restrict__ raw, int64_t nv, int nlabels, unsigned long long mask,
                                                      double factor, double *__restrict__ dens, double *__restrict__ bsum) {
    __shared__ double sm[RED_T];
    const int64_t j = (int64_t)blockIdx.x * RED_T + threadIdx.x;
    double d = 0.0;
    if (j < nv) {
        for (int l = 0; l < nlabels; l++) if ((mask >> l) & 1ull) d += raw[(size_t)l * nv + j];
        d *= factor;
        dens[j] = d;
    }
    sm[threadIdx.x] = d;
    __syncthreads();
    for (int o = RED_T / 2; o > 0; o >>= 1) { if (threadIdx.x < o) sm[threadIdx.x] += sm[threadIdx.x + o]; __syncthreads(); }
    if (threadIdx.x == 0) bsum[blockIdx.x] = sm[0];
}

// ------------------------------------------------------------------ host side
struct KdePlan { int64_t ns; int threads, R, nsplit; int64_t blocks_x; };
static KdePlan kde_plan(int64_t nv, int64_t ns) {
    KdePlan p;
    p.ns = ns;
    const bool big = nv >= (int64_t)ms_sm_count() * 128 * 4 * 2;
    p.threads = big ? 128 : 64; p.R = big ? 4 : 1;
    p.blocks_x = (nv + p.threads * p.R - 1) / (p.threads * p.R);
    // enough CTAs for >= 16 full waves (6 resident CTAs per SM) so that the last, partial wave costs
    // a few percent at most; the splits are combined by k_kde_combine
    const int64_t ntiles = (p.ns + 511) / 512;                      // smallest tile (3-D)
    int64_t want = ((int64_t)ms_sm_count() * 6 * 16 + p.blocks_x - 1) / (p.blocks_x > 0 ? p.blocks_x : 1);
    if (want < 1) want = 1; if (want > 64) want = 64; if (want > ntiles) want = ntiles > 0 ? ntiles : 1;
    p.nsplit = (int)want;
    return p;
}
static int64_t kde_max_cells(int64_t nv) { return nv / 3 + 4096; }     // fine source cells of the tile plan: <= 16 per ~56-target tile
#define KDE_MAX_LABELS 64
#define KDE_MAX_STRIPS 65536
static int64_t kde_max_groups(int64_t nv) { return nv / GR_TGT + KDE_MAX_STRIPS + 8; }

static size_t kde_ws_bytes(int64_t nv, int64_t ns, int nlabels) {
    const KdePlan p = kde_plan(nv, ns);
    const size_t nc = (size_t)kde_max_cells(nv) * (size_t)nlabels + 2;
    return 2 * ms_align((size_t)nv * 32) + 2 * ms_align((size_t)ns * 32)         // staged + cell-sorted targets, sources
         + ms_align((size_t)p.nsplit * nv * 8) + ms_align((size_t)nv * 8) + ms_align(((size_t)nv / RED_T + 2) * 8)
         + 2 * ms_align((size_t)(nv + 1) * 4) + 2 * ms_align((size_t)(ns + 1) * 4) // bin ids, permutations
         + 4 * ms_align(nc * 4) + ms_scan_ws_bytes((int64_t)nc) + 8192
         + 2 * ms_align((KDE_MAX_STRIPS + 8) * 4) + ms_align((size_t)kde_max_groups(nv) * sizeof(KdeGroup));
}
extern "C" size_t ms_kde_workspace_bytes(int64_t nv, int64_t nids) { return kde_ws_bytes(nv, nids > 0 ? nids : nv, 1); }
extern "C" size_t ms_kde_labels_workspace_bytes(int64_t nv, int nlabels) { return kde_ws_bytes(nv, nv, nlabels < 1 ? 1 : nlabels); }

template <int DIM, bool PER, bool CLAMP>
static int kde_run_pairs(const KdePlan &p, int64_t nv, double wx, double wy, const void *tgt_, const void *src_, double *partial, cudaStream_t s) {
    typedef typename SrcT<DIM>::type S;
    const S *tgt = (const S*)tgt_, *src = (const S*)src_;
    const int tile = KDE_STAGE_BYTES / (int)sizeof(S);
    const int ntiles = (int)((p.ns + tile - 1) / tile);
    const int tps = (ntiles + p.nsplit - 1) / p.nsplit;
    dim3 grid((unsigned)p.blocks_x, (unsigned)p.nsplit);
    MS_K("kde_pairs", s);
    if (p.R == 4) k_kde_pairs<DIM, PER, CLAMP, 128, 4><<<grid, 128, 0, s>>>(tgt, nv, src, p.ns, ntiles, tps, wx, wy, partial);
    else          k_kde_pairs<DIM, PER, CLAMP, 64, 1><<<grid, 64, 0, s>>>(tgt, nv, src, p.ns, ntiles, tps, wx, wy, partial);
    MS_LAUNCH_CHECK();
    return 0;
}
static int kde_pairs_dispatch(int type, int periodic, bool clamp, const KdePlan &p, int64_t nv, double wx, double wy,
                              const void *tgt, const void *src, double *partial, cudaStream_t s) {
    if (type == 2) {
        if (periodic) return clamp ? kde_run_pairs<2, true, true>(p, nv, wx, wy, tgt, src, partial, s)
                                   : kde_run_pairs<2, true, false>(p, nv, wx, wy, tgt, src, partial, s);
        return kde_run_pairs<2, false, true>(p, nv, wx, wy, tgt, src, partial, s);
    }
    return periodic ? kde_run_pairs<3, true, true>(p, nv, wx, wy, tgt, src, partial, s)
                    : kde_run_pairs<3, false, true>(p, nv, wx, wy, tgt, src, partial, s);
}

// sort n staged points into bins (cell, or label*ncell + cell): start (nbins+1), perm, sorted
template <int DIM>
static int kde_bin(const void *pts_, int64_t n, const KdeGrid &g, const int32_t *labels, int nlabels, int32_t *cell, int32_t *cnt,
                   int32_t *cursor, int32_t *start, int32_t *perm, void *sorted_, void *scan_ws, cudaStream_t s) {
    typedef typename SrcT<DIM>::type S;
    const S *pts = (const S*)pts_; S *sorted = (S*)sorted_;
    const int nbins = g.ncx * g.ncy * (labels ? nlabels : 1);
    MS_CUDA(cudaMemsetAsync(cnt, 0, sizeof(int32_t) * (nbins + 1), s));
    MS_CUDA(cudaMemsetAsync(cursor, 0, sizeof(int32_t) * (nbins + 1), s));
    MS_K("kde_cell_count", s); k_kde_cell_count<DIM><<<ms_blocks(n, 256), 256, 0, s>>>(pts, n, g, labels, nlabels, cell, cnt);
    if (ms_exclusive_scan_i32(cnt, start, nbins, scan_ws, s)) return 2;
    MS_K("kde_cell_scatter", s); k_kde_cell_scatter<<<ms_blocks(n, 256), 256, 0, s>>>(n, cell, start, cursor, perm);
    MS_K("kde_cell_sort", s); k_kde_cell_sort<<<ms_blocks((int64_t)nbins, 128), 128, 0, s>>>(nbins, start, perm);
    MS_K("kde_cell_gather", s); k_kde_cell_gather<DIM><<<ms_blocks(n, 256), 256, 0, s>>>(pts, n, g, start + nbins, perm, sorted);
    MS_LAUNCH_CHECK();
    return 0;
}

// Sources at least KDE_FAR_SIGMAS sigma away from the box of a group are summed in fp32 (kde_split_batch / kde_far_sum):
// at 4.5 sigma they carry exp(-4.5^2 / 2) = 4e-5 of a sum and an fp32 term on group-local coordinates is good to ~6e-6
// (random) / 2.4e-7 (ex2.approx, possibly biased), i.e. <= 2e-10 of the result against the 1e-9 bar; measured against
// the fp64 oracle on the bench frame: 1.8e-11 (5.0 sigma: 5.8e-12 = the fp64 path's own deviation, 6 % slower).
// Only when the cut-off leaves a rim of at least one sigma beyond it.
// half widths per gap (rounded up to float: the window may only grow), once on the host instead of a sqrt per piece
static void kde_fill_half_widths(KdeTiling &T) {
    for (int gap = 0; gap <= KDE_MAXGAP; gap++) {
        const double gy = (double)gap * T.cys, h2 = T.rc2 - gy * gy;
        T.hx[gap] = h2 > 0.0 ? nextafterf((float)sqrt(h2), INFINITY) : -1.f;
    }
}
static double kde_far_radius2(double cutoff_sigmas, double sigma_scaled) {
    if (cutoff_sigmas < KDE_FAR_SIGMAS + 1.0) return 0.0;
    const double r = KDE_FAR_SIGMAS * sigma_scaled;
    return r * r;
}

struct KdeCtx {
    KdePlan p; bool cells, tiles; KdeGrid g; KdeTiling tl; int type, periodic, nsplit, nlabels; bool clamp;
    double wx, wy; float sfactor;
    void *tgt, *tgt_sorted, *src, *src_sorted; const void *ssorted; const int32_t *sst;
    double *partial, *dens, *bsum; int32_t *tperm, *tstart, *sstart, *tcell, *ng, *gstart; KdeGroup *groups; int64_t max_groups;
    unsigned long long *counter; void *scan_ws; int32_t *sperm; double scale;
};

// validation, staging of the scaled coordinates, choice between all pairs and the cell list, binning.
// Sources: ids (nids > 0), every vertex (ids == NULL, labels == NULL), or every vertex grouped by label.
static int kde_front(const float *verts, int64_t nv, const int32_t *ids, int64_t nids, const int32_t *labels, int nlabels,
                     int type, int periodic, const float *box6, float sigma, double cutoff_sigmas,
                     void *ws, size_t ws_bytes, cudaStream_t s, KdeCtx &c) {
    MS_REQUIRE(type == 2 || type == 3, "ms_kde: invalid density type %d (should be 2 (2D) or 3 (3D); geodesic type 1 is not on the GPU path)", type);
    MS_REQUIRE(nv > 0 && nids >= 0 && sigma > 0.0f, "ms_kde: bad arguments nv=%lld nids=%lld sigma=%g", (long long)nv, (long long)nids, (double)sigma);
    MS_REQUIRE(nv < ((int64_t)1 << 31) && nids < ((int64_t)1 << 31), "ms_kde: too many points");
    MS_REQUIRE(!periodic || box6 != NULL, "ms_kde: periodic KDE needs the 6-float bbox");
    MS_REQUIRE(!labels || (nlabels >= 1 && nlabels <= KDE_MAX_LABELS), "ms_kde_labels: need 1 <= nlabels <= %d", KDE_MAX_LABELS);
    if (nids == 0) ids = NULL;
    if (!labels) nlabels = 1;
    const int64_t ns = ids ? nids : nv;
    MS_REQUIRE(ws_bytes >= kde_ws_bytes(nv, ns, nlabels), "ms_kde: workspace too small");
    const KdePlan p = kde_plan(nv, ns);
    c.p = p; c.type = type; c.periodic = periodic; c.nlabels = nlabels;
    // constants rounded through float32 exactly as DensityKernels.hpp:47-48 stores them
    const double sg = (double)sigma;
    const float efactor = (float)(-1.0 / (2.0 * sg * sg));
    c.sfactor = type == 3 ? (float)(1.0 / pow(sqrt(2.0 * M_PI) * sg, 3.0)) : (float)(1.0 / (2.0 * M_PI * sg * sg));
    const double scale = sqrt(-(double)efactor * 1.4426950408889634074 * (double)KDE_TAB);
    double wx = 0.0, wy = 0.0;
    if (box6) {      // float32 widths as DistanceKernels.hpp:46-68 forms them
        wx = (double)(float)(box6[3] - box6[0]) * scale; wy = (double)(float)(box6[4] - box6[1]) * scale;
    }
    MS_REQUIRE(!periodic || (wx > 0.0 && wy > 0.0), "ms_kde: periodic bbox has no extent");
    c.wx = wx; c.wy = wy;
    MsArena A(ws, ws_bytes);
    const size_t nc = (size_t)kde_max_cells(nv) * (size_t)nlabels + 2;
    void *tgt = A.take<char>((size_t)nv * 32), *tgt_sorted = A.take<char>((size_t)nv * 32);
    void *src = A.take<char>((size_t)ns * 32), *src_sorted = A.take<char>((size_t)ns * 32);
    c.partial = A.take<double>((size_t)p.nsplit * nv); c.dens = A.take<double>(nv);
    c.bsum = A.take<double>(nv / RED_T + 2);
    int32_t *tcell = A.take<int32_t>(nv + 1), *tperm = A.take<int32_t>(nv), *scell = A.take<int32_t>(ns + 1), *sperm = A.take<int32_t>(ns);
    int32_t *cnt = A.take<int32_t>(nc), *cursor = A.take<int32_t>(nc), *tstart = A.take<int32_t>(nc), *sstart = A.take<int32_t>(nc);
    void *scan_ws = A.take<char>(ms_scan_ws_bytes((int64_t)nc)); c.scan_ws = scan_ws;
    c.counter = (unsigned long long*)A.take<unsigned long long>(4);
    c.ng = A.take<int32_t>(KDE_MAX_STRIPS + 8); c.gstart = A.take<int32_t>(KDE_MAX_STRIPS + 8);
    c.max_groups = kde_max_groups(nv); c.groups = A.take<KdeGroup>((size_t)c.max_groups); c.tcell = tcell;

    MS_K("kde_stage", s);
    if (type == 2) k_kde_stage<2><<<ms_blocks(nv, 256), 256, 0, s>>>(verts, NULL, nv, scale, (double2*)tgt);
    else           k_kde_stage<3><<<ms_blocks(nv, 256), 256, 0, s>>>(verts, NULL, nv, scale, (double4*)tgt);
    if (ids) {
        MS_K("kde_stage", s);
        if (type == 2) k_kde_stage<2><<<ms_blocks(ns, 256), 256, 0, s>>>(verts, ids, ns, scale, (double2*)src);
        else           k_kde_stage<3><<<ms_blocks(ns, 256), 256, 0, s>>>(verts, ids, ns, scale, (double4*)src);
    } else src = tgt;
    MS_LAUNCH_CHECK();
    c.tgt = tgt; c.src = src; c.tgt_sorted = tgt_sorted; c.src_sorted = src_sorted; c.tperm = tperm; c.tstart = tstart; c.sstart = sstart;
    c.sperm = sperm; c.scale = scale;

    // ---- cell-list plan: needs the grid extent (box6); periodic grids must tile the period
    bool cells = false; KdeGrid g; double rwin = 0.0;
    g.ncx = g.ncy = 1; g.periodic = periodic ? 1 : 0; g.b0x = g.b0y = 0.0; g.inv_cx = g.inv_cy = 0.0; g.wx = wx; g.wy = wy;
    c.tiles = false;
    if (cutoff_sigmas > 0.0 && box6 != NULL && wx > 0.0 && wy > 0.0) {
        // ---- group plan: strips of B fine rows, ~square groups of GR_TGT targets, fine cells of ~rcut/16
        const double rcut = cutoff_sigmas * sg * scale;
        const double tside = sqrt((double)GR_TGT * wx * wy / (double)nv);
        const double ntyd = floor(wy / tside);
        if (ntyd >= 1.0 && ntyd <= (double)KDE_MAX_STRIPS) {
            KdeTiling T;
            T.nty = (int)ntyd;
            const double tsy = wy / T.nty;
            int B = (int)lround(tsy / (rcut / 16.0));
            B = B < 1 ? 1 : (B > 6 ? 6 : B);
            double ncxd = 0.0;
            for (;; B--) {                                 // fine cells ~square; their count must fit the workspace
                ncxd = floor(wx / (tsy / B)); if (ncxd < 1.0) ncxd = 1.0;
                if (B == 1 || ncxd * ntyd * B <= (double)kde_max_cells(nv)) break;
            }
            KdeGrid gs = g;
            gs.ncx = (int)ncxd; gs.ncy = T.nty * B;
            T.B = B; T.cxs = wx / gs.ncx; T.cys = tsy / B; T.rc2 = rcut * rcut; T.gscale = 1.0; T.oscale = 1.0;
            T.ri2 = kde_far_radius2(cutoff_sigmas, sg * scale);
            kde_fill_half_widths(T);
            T.mf = (int)ceil(rcut / T.cys);
            const int mxmax = (int)ceil(rcut / T.cxs);
            gs.b0x = (double)box6[0] * scale; gs.b0y = (double)box6[1] * scale;
            gs.inv_cx = 1.0 / T.cxs; gs.inv_cy = 1.0 / T.cys;
            T.gs = gs;
            T.wmax = periodic ? gs.ncx - 2 * mxmax - 1 : (1 << 30);      // widest box whose window does not meet itself
            const bool fits = ncxd * ntyd * B <= (double)kde_max_cells(nv) && T.mf <= 4096
                           && (!periodic || (T.wmax >= 1 && B + 2 * T.mf <= gs.ncy));
            if (fits) {
                c.tiles = cells = true; c.tl = T;
                g = gs; g.ncy = T.nty; g.inv_cy = 1.0 / tsy;            // the target grid: fine columns x strips
                const double ex = (double)(2 * mxmax + 2) * T.cxs + 2.0 * tside, ey = (B + T.mf + 1) * T.cys;
                rwin = ex * ex + ey * ey;
                // the exponent shift n>>8 is not clamped in the kernel unless CLAMP: keep it far below 1021
                if (rwin >= (double)KDE_TAB * 1000.0) rwin = 2.0 * KDE_RMAX;
            }
        }
    }
    c.cells = cells; c.nsplit = cells ? 1 : p.nsplit;
    if (!cells) {
        // r = scaled distance^2 must stay below 2^31 for the integer split of the exponent: periodic 2-D
        // pairs are bounded by the half box (checked here), everything else clamps in the kernel
        c.clamp = type == 3 || !periodic || !(0.25 * (wx * wx + wy * wy) < KDE_RMAX);
        c.g = g;
        if (!labels) return 0;
        // all pairs per label: stable partition of the staged vertices by label (sstart = nlabels+1 offsets)
        MS_CUDA(cudaMemsetAsync(sstart, 0, sizeof(int32_t), s));
        for (int l = 0; l < nlabels; l++) {
            MS_K("kde_label_flag", s); k_kde_label_flag<<<ms_blocks(nv, 256), 256, 0, s>>>(labels, nv, l, scell);
            if (ms_exclusive_scan_i32(scell, tcell, nv, scan_ws, s)) return 2;      // tcell (nv+1 <= workspace slack) holds positions
            MS_K("kde_label_scatter", s); k_kde_label_scatter<<<ms_blocks(nv, 256), 256, 0, s>>>(labels, nv, l, tcell, sstart, sperm);
        }
        KdeGrid g1 = g; g1.periodic = 0;
        MS_K("kde_cell_gather", s);
        if (type == 2) k_kde_cell_gather<2><<<ms_blocks(nv, 256), 256, 0, s>>>((const double2*)src, nv, g1, sstart + nlabels, sperm, (double2*)src_sorted);
        else           k_kde_cell_gather<3><<<ms_blocks(nv, 256), 256, 0, s>>>((const double4*)src, nv, g1, sstart + nlabels, sperm, (double4*)src_sorted);
        MS_LAUNCH_CHECK();
        c.ssorted = src_sorted; c.sst = sstart;
        return 0;
    } else c.clamp = type == 3 || !periodic || !(rwin < KDE_RMAX);
    c.g = g;
    int rc;
    if (cells) {
        if (type == 2) rc = kde_bin<2>(tgt, nv, g, NULL, 1, tcell, cnt, cursor, tstart, tperm, tgt_sorted, scan_ws, s);
        else           rc = kde_bin<3>(tgt, nv, g, NULL, 1, tcell, cnt, cursor, tstart, tperm, tgt_sorted, scan_ws, s);
        if (rc) return rc;
        c.ssorted = tgt_sorted; c.sst = tstart;
    }
    if (ids || labels || c.tiles) {
        const KdeGrid &gsrc = c.tiles ? c.tl.gs : g;      // tiles: the sources always live on the fine grid
        if (type == 2) rc = kde_bin<2>(src, ns, gsrc, labels, nlabels, scell, cnt, cursor, sstart, sperm, src_sorted, scan_ws, s);
        else           rc = kde_bin<3>(src, ns, gsrc, labels, nlabels, scell, cnt, cursor, sstart, sperm, src_sorted, scan_ws, s);
        if (rc) return rc;
        c.ssorted = src_sorted; c.sst = sstart;
    }
    return 0;
}

// groups of 64 consecutive targets per strip: counts -> offsets -> descriptors
static int kde_build_groups(const KdeCtx &c, cudaStream_t s) {
    const KdeTiling &T = c.tl;
    MS_K("kde_group_count", s); k_kde_group_count<<<ms_blocks(T.nty, 128), 128, 0, s>>>(c.tstart, T.gs.ncx, T.nty, c.ng);
    if (ms_exclusive_scan_i32(c.ng, c.gstart, T.nty, c.scan_ws, s)) return 2;
    MS_K("kde_group_fill", s); k_kde_group_fill<<<ms_blocks(c.max_groups, 128), 128, 0, s>>>(c.tstart, c.tcell, c.tperm, T.gs.ncx, T.nty, c.gstart,
                                                                                         (const double*)c.tgt_sorted, c.type == 2 ? 2 : 4, c.groups);
    MS_LAUNCH_CHECK();
    return 0;
}

template <int DIM, bool CLAMP>
static int kde_run_groups(const KdeCtx &c, int64_t nv, double *out, cudaStream_t s, bool build = true) {
    typedef typename SrcT<DIM>::type S;
    const size_t smem = GR_TAB_BYTES + (size_t)GR_NW * 2 * GrCap<DIM>::value * sizeof(S) + (size_t)GR_NW * 2 * sizeof(uint64_t);
    // set on every call (idempotent, a few microseconds): no per-process cache, any device, any thread.
    // Two 2-D CTAs (64 KB table + 48 KB of stages each) only fit with the whole 228 KB carved out as shared memory.
    MS_CUDA(cudaFuncSetAttribute(k_kde_groups<DIM, CLAMP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    MS_CUDA(cudaFuncSetAttribute(k_kde_groups<DIM, CLAMP>, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
    int per_sm = 0;
    MS_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_kde_groups<DIM, CLAMP>, GR_THREADS, smem));
    MS_REQUIRE(per_sm >= 1, "ms_kde: k_kde_groups does not fit an SM (%zu bytes of shared memory)", smem);
    const int sms = ms_sm_count();
    const KdeTiling &T = c.tl;
    if (build && kde_build_groups(c, s)) return 2;
    unsigned *item_counter = (unsigned*)(c.counter + 2);
    MS_CUDA(cudaMemsetAsync(item_counter, 0, sizeof(unsigned), s));
    MS_CUDA(cudaMemsetAsync(out, 0, sizeof(double) * (size_t)nv * (size_t)c.nlabels, s));
    MS_K("kde_groups", s);
    k_kde_groups<DIM, CLAMP><<<sms * per_sm, GR_THREADS, smem, s>>>((const S*)c.tgt_sorted, c.tstart, c.tperm, c.groups, c.gstart + T.nty,
                                                                     (const S*)c.ssorted, c.sst, c.nlabels, nv, T, item_counter, out);
    MS_LAUNCH_CHECK();
    return 0;
}

static int kde_launch_cells(const KdeCtx &c, int64_t nv, double *out, cudaStream_t s, bool build = true) {
    if (c.type == 2) return c.clamp ? kde_run_groups<2, true>(c, nv, out, s, build) : kde_run_groups<2, false>(c, nv, out, s, build);
    return kde_run_groups<3, true>(c, nv, out, s, build);
}

extern "C" int ms_kde(const float *verts, int64_t nv, const int32_t *ids, int64_t nids,
                      int type, int periodic, const float *box6, float sigma, int get_counts,
                      double cutoff_sigmas, double *density_f64, float *density_f32,
                      void *ws, size_t ws_bytes, void *stream) {
    cudaStream_t s = (cudaStream_t)stream;
    KdeCtx c;
    int rc = kde_front(verts, nv, ids, nids, NULL, 1, type, periodic, box6, sigma, cutoff_sigmas, ws, ws_bytes, s, c);
    if (rc) return rc;
    if (c.cells) rc = kde_launch_cells(c, nv, c.partial, s);
    else         rc = kde_pairs_dispatch(type, periodic, c.clamp, c.p, nv, c.wx, c.wy, c.tgt, c.src, c.partial, s);
    if (rc) return rc;
    const int nb = ms_blocks(nv, RED_T);
    MS_K("kde_combine", s); k_kde_combine<<<nb, RED_T, 0, s>>>(c.partial, nv, c.nsplit, (double)c.sfactor * (1.0 / (double)nv), c.dens, c.bsum);
    MS_K("kde_emit", s); k_kde_emit<<<nb, RED_T, 0, s>>>(c.dens, nv, get_counts ? c.bsum : NULL, nb, (double)c.p.ns, density_f64, density_f32);
    MS_K(NULL, s);
    MS_LAUNCH_CHECK();
    return 0;
}

extern "C" int ms_kde_labels(const float *verts, int64_t nv, const int32_t *labels, int nlabels,
                             int type, int periodic, const float *box6, float sigma, double cutoff_sigmas,
                             double *raw, void *ws, size_t ws_bytes, void *stream) {
    cudaStream_t s = (cudaStream_t)stream;
    MS_REQUIRE(labels != NULL && raw != NULL, "ms_kde_labels: labels and raw are required");
    KdeCtx c;
    int rc = kde_front(verts, nv, NULL, 0, labels, nlabels, type, periodic, box6, sigma, cutoff_sigmas, ws, ws_bytes, s, c);
    if (rc) return rc;
    if (c.cells) {
        c.tl.oscale = (double)c.sfactor;
        rc = kde_launch_cells(c, nv, raw, s);
        if (rc) return rc;
        MS_K(NULL, s);
        MS_LAUNCH_CHECK();
        return 0;
    }
    // all pairs: the label-sorted source array is cut into one range per label (offsets read back)
    int32_t h[KDE_MAX_LABELS + 1];
    MS_CUDA(cudaMemcpyAsync(h, c.sstart, sizeof(int32_t) * (nlabels + 1), cudaMemcpyDeviceToHost, s));
    MS_CUDA(cudaStreamSynchronize(s));
    const size_t esz = type == 2 ? sizeof(double2) : sizeof(double4);
    for (int l = 0; l < nlabels; l++) {
        const int64_t nsl = h[l + 1] - h[l];
        double *out = raw + (size_t)l * nv;
        if (nsl <= 0) { MS_CUDA(cudaMemsetAsync(out, 0, sizeof(double) * nv, s)); continue; }
        KdePlan p = kde_plan(nv, nsl);
        if (p.nsplit > c.p.nsplit) p.nsplit = c.p.nsplit;          // partial buffer was sized for the whole source set
        rc = kde_pairs_dispatch(type, periodic, c.clamp, p, nv, c.wx, c.wy, c.tgt, (const char*)c.src_sorted + (size_t)h[l] * esz, c.partial, s);
        if (rc) return rc;
        MS_K("kde_combine", s); k_kde_combine<<<ms_blocks(nv, RED_T), RED_T, 0, s>>>(c.partial, nv, p.nsplit, (double)c.sfactor, out, c.bsum);
    }
    MS_K(NULL, s);
    MS_LAUNCH_CHECK();
    return 0;
}

// ------------------------------------------------------------------ several sigmas on one binning
// sorted[i] = fold(scale * vertex perm[i]): the staged + cell-sorted coordinates of another sigma on the SAME
// permutation (bins are scale-invariant: only the grid is rescaled)
template <int DIM>
__global__ void __launch_bounds__(256) k_kde_restage(const float *__restrict__ verts, const int32_t *__restrict__ perm, int64_t n,
                                                     const int32_t *__restrict__ nbinned, double scale_ref, KdeGrid gref,
                                                     double scale, double wx, double wy, typename SrcT<DIM>::type *sorted) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n || (nbinned && i >= (int64_t)__ldg(nbinned))) return;
    const int64_t v = (int64_t)__ldg(perm + i);
    const double x = ms_ld_f(verts + 3*v), y = ms_ld_f(verts + 3*v + 1);
    typename SrcT<DIM>::type p;
    p.x = x * scale; p.y = y * scale;
    if (DIM == 3) { ((double*)&p)[2] = ms_ld_f(verts + 3*v + 2) * scale; ((double*)&p)[3] = 0.0; }
    if (gref.periodic) {      // the period count is decided on the REFERENCE staging, exactly as the binning decided it
        const double xr = __dmul_rn(x, scale_ref), yr = __dmul_rn(y, scale_ref);      // rounded products, as k_kde_stage stored them (no FMA contraction)
        const int cx = (int)floor((xr - gref.b0x) * gref.inv_cx), cy = (int)floor((yr - gref.b0y) * gref.inv_cy);
        const int kx = (cx >= 0 ? cx : cx - gref.ncx + 1) / gref.ncx, ky = (cy >= 0 ? cy : cy - gref.ncy + 1) / gref.ncy;
        p.x -= (double)kx * wx; p.y -= (double)ky * wy;
    }
    sorted[i] = p;
}

extern "C" size_t ms_kde_labels_multi_workspace_bytes(int64_t nv, int nlabels) { return ms_kde_labels_workspace_bytes(nv, nlabels); }

// raw (nsigma, nlabels, nv): per-label pair sums for every sigma.  With a cut-off whose group plan fits every
// sigma the vertices are binned ONCE (on the grid of the smallest sigma) and only re-staged per sigma;
// otherwise this is a loop over ms_kde_labels.
extern "C" int ms_kde_evaluated_pairs(const float *verts, int64_t nv, const int32_t *ids, int64_t nids, const int32_t *labels, int nlabels,
                                      int type, int periodic, const float *box6, float sigma, double cutoff_sigmas,
                                      double *pairs, void *ws, size_t ws_bytes, void *stream);
// Optional tail of every sigma inside the multi call: the normalised fields of `nsets` label sets are written (and, if
// asked for, sent to page-locked host memory on a copy stream) as soon as that sigma's pair sums are queued -- the
// downloads of one sigma then run under the pair kernel of the next instead of after the last one.
struct KdeEmit {
    int nsets; const uint64_t *masks; const double *nsel; int get_counts;
    double *partials, *o64; float *o32, *host32; cudaStream_t copy;
};
static int kde_emit_sigma(const KdeEmit &E, const double *raw_i, int i, int64_t nv, int nlabels, cudaStream_t s);

static int kde_labels_multi_impl(const float *verts, int64_t nv, const int32_t *labels, int nlabels,
                                 int type, int periodic, const float *box6, const float *sigmas, int nsigma, double cutoff_sigmas,
                                 double *raw, double *pairs, void *ws, size_t ws_bytes, void *stream, const KdeEmit *emit = NULL) {
    cudaStream_t s = (cudaStream_t)stream;
    MS_REQUIRE(labels != NULL && (raw != NULL || pairs != NULL) && sigmas != NULL && nsigma >= 1, "ms_kde_labels_multi: labels, sigmas and an output are required");
    const size_t per = (size_t)nlabels * (size_t)nv;
    int ref = 0;
    for (int i = 1; i < nsigma; i++) if (sigmas[i] < sigmas[ref]) ref = i;
    bool shared = nsigma > 1 && cutoff_sigmas > 0.0;
    KdeCtx c;
    KdeTiling Ts[64]; bool clamps[64]; double scales[64]; float sfac[64];
    if (shared && nsigma <= 64) {
        int rc = kde_front(verts, nv, NULL, 0, labels, nlabels, type, periodic, box6, sigmas[ref], cutoff_sigmas, ws, ws_bytes, s, c);
        if (rc) return rc;
        shared = c.tiles;
        for (int i = 0; shared && i < nsigma; i++) {
            MS_REQUIRE(sigmas[i] > 0.0f, "ms_kde_labels_multi: bad sigma %g", (double)sigmas[i]);
            const double sg = (double)sigmas[i];
            const float efactor = (float)(-1.0 / (2.0 * sg * sg));
            sfac[i] = type == 3 ? (float)(1.0 / pow(sqrt(2.0 * M_PI) * sg, 3.0)) : (float)(1.0 / (2.0 * M_PI * sg * sg));
            scales[i] = sqrt(-(double)efactor * 1.4426950408889634074 * (double)KDE_TAB);
            const double rho = scales[i] / c.scale, rcut = cutoff_sigmas * sg * scales[i];
            KdeTiling T = c.tl;
            T.gs.b0x = (double)box6[0] * scales[i]; T.gs.b0y = (double)box6[1] * scales[i];
            T.gs.wx = (double)(float)(box6[3] - box6[0]) * scales[i]; T.gs.wy = (double)(float)(box6[4] - box6[1]) * scales[i];
            T.cxs = T.gs.wx / T.gs.ncx; T.cys = T.gs.wy / T.gs.ncy;
            T.gs.inv_cx = 1.0 / T.cxs; T.gs.inv_cy = 1.0 / T.cys;
            T.rc2 = rcut * rcut;
            T.ri2 = kde_far_radius2(cutoff_sigmas, sg * scales[i]);
            kde_fill_half_widths(T);
            T.mf = (int)ceil(rcut / T.cys);
            const int mxmax = (int)ceil(rcut / T.cxs);
            T.wmax = periodic ? T.gs.ncx - 2 * mxmax - 1 : (1 << 30);
            if (T.mf > 4096 || (periodic && (T.wmax < 1 || T.B + 2 * T.mf > T.gs.ncy))) { shared = false; break; }
            const double tside = T.cys * T.B;
            const double ex = (double)(2 * mxmax + 2) * T.cxs + 2.0 * tside, ey = (T.B + T.mf + 1) * T.cys;
            const double rwin = ex * ex + ey * ey;
            clamps[i] = type == 3 || !periodic || !(rwin < (double)KDE_TAB * 1000.0);
            Ts[i] = T;
            (void)rho;
        }
    } else shared = false;
    if (!shared) {
        for (int i = 0; i < nsigma; i++) {
            double two[2] = { 0.0, 0.0 };
            const int rc = pairs ? ms_kde_evaluated_pairs(verts, nv, NULL, 0, labels, nlabels, type, periodic, box6, sigmas[i], cutoff_sigmas, two, ws, ws_bytes, stream)
                                 : ms_kde_labels(verts, nv, labels, nlabels, type, periodic, box6, sigmas[i], cutoff_sigmas, raw + (size_t)i * per, ws, ws_bytes, stream);
            if (rc) return rc;
            if (pairs) { pairs[i] = two[0]; pairs[nsigma + i] = two[1]; }
            else if (emit && kde_emit_sigma(*emit, raw + (size_t)i * per, i, nv, nlabels, s)) return 1;
        }
        return 0;
    }
    if (kde_build_groups(c, s)) return 2;
    const int nbins_s = c.tl.gs.ncx * c.tl.gs.ncy * nlabels;
    const KdeGrid gs_ref = c.tl.gs, gt_ref = c.g;
    if (pairs) {             // diagnostic: the pairs every sigma's launch evaluates on this shared binning
        for (int i = 0; i < nsigma; i++) {
            c.tl = Ts[i]; c.tl.gscale = scales[i] / c.scale;
            MS_CUDA(cudaMemsetAsync(c.counter, 0, 2 * sizeof(unsigned long long), s));
            MS_K("kde_cells_count", s); k_kde_groups_count<<<ms_blocks(c.max_groups, 128), 128, 0, s>>>(c.groups, c.gstart + c.tl.nty, c.tstart, c.sst, (const double*)c.ssorted, c.type == 2 ? 2 : 4, c.nlabels, c.tl, c.counter);
            MS_K(NULL, s);
            MS_LAUNCH_CHECK();
            unsigned long long h[2] = { 0ull, 0ull };
            MS_CUDA(cudaMemcpyAsync(h, c.counter, sizeof(h), cudaMemcpyDeviceToHost, s));
            MS_CUDA(cudaStreamSynchronize(s));
            pairs[i] = (double)h[0]; pairs[nsigma + i] = (double)h[1];
        }
        return 0;
    }
    for (int k = 0; k < nsigma; k++) {
        const int i = k == 0 ? ref : (k <= ref ? k - 1 : k);           // the reference sigma first: its coordinates are staged
        if (i != ref || k != 0) {
            const KdeGrid gref = gs_ref, gref_t = gt_ref;      // every array folds on the grid it was binned on
            const double wxi = Ts[i].gs.wx, wyi = Ts[i].gs.wy;
            MS_K("kde_restage", s);
            if (type == 2) {
                k_kde_restage<2><<<ms_blocks(nv, 256), 256, 0, s>>>(verts, c.tperm, nv, NULL, c.scale, gref_t, scales[i], wxi, wyi, (double2*)c.tgt_sorted);
                k_kde_restage<2><<<ms_blocks(nv, 256), 256, 0, s>>>(verts, c.sperm, nv, c.sstart + nbins_s, c.scale, gref, scales[i], wxi, wyi, (double2*)c.src_sorted);
            } else {
                k_kde_restage<3><<<ms_blocks(nv, 256), 256, 0, s>>>(verts, c.tperm, nv, NULL, c.scale, gref_t, scales[i], wxi, wyi, (double4*)c.tgt_sorted);
                k_kde_restage<3><<<ms_blocks(nv, 256), 256, 0, s>>>(verts, c.sperm, nv, c.sstart + nbins_s, c.scale, gref, scales[i], wxi, wyi, (double4*)c.src_sorted);
            }
            MS_LAUNCH_CHECK();
        }
        c.tl = Ts[i]; c.clamp = clamps[i];
        // group x extents were measured on the reference staging: they scale with the coordinates
        c.tl.gscale = scales[i] / c.scale;
        double *out = raw + (size_t)i * per;
        c.tl.oscale = (double)sfac[i];                // the groups kernel applies the kernel's sfactor itself
        const int rc = kde_launch_cells(c, nv, out, s, false);
        if (rc) return rc;
        if (emit && kde_emit_sigma(*emit, out, i, nv, nlabels, s)) return 1;
    }
    MS_K(NULL, s);
    MS_LAUNCH_CHECK();
    return 0;
}

extern "C" int ms_kde_labels_multi(const float *verts, int64_t nv, const int32_t *labels, int nlabels,
                                   int type, int periodic, const float *box6, const float *sigmas, int nsigma, double cutoff_sigmas,
                                   double *raw, void *ws, size_t ws_bytes, void *stream) {
    MS_REQUIRE(raw != NULL, "ms_kde_labels_multi: raw is required");
    return kde_labels_multi_impl(verts, nv, labels, nlabels, type, periodic, box6, sigmas, nsigma, cutoff_sigmas, raw, NULL, ws, ws_bytes, stream);
}
extern "C" int ms_kde_labels_multi_pairs(const float *verts, int64_t nv, const int32_t *labels, int nlabels,
                                         int type, int periodic, const float *box6, const float *sigmas, int nsigma, double cutoff_sigmas,
                                         double *pairs, void *ws, size_t ws_bytes, void *stream) {
    MS_REQUIRE(pairs != NULL, "ms_kde_labels_multi_pairs: pairs is required");
    return kde_labels_multi_impl(verts, nv, labels, nlabels, type, periodic, box6, sigmas, nsigma, cutoff_sigmas, NULL, pairs, ws, ws_bytes, stream);
}

extern "C" size_t ms_kde_select_workspace_bytes(int64_t nv) { return ms_align((size_t)nv * 8) + ms_align(((size_t)nv / RED_T + 2) * 8) + 1024; }
extern "C" int ms_kde_select(const double *raw, int64_t nv, int nlabels, const uint8_t *select, double n_selected, int get_counts,
                             double *density_f64, float *density_f32, void *ws, size_t ws_bytes, void *stream) {
    cudaStream_t s = (cudaStream_t)stream;
    MS_REQUIRE(raw && select && nv > 0 && nlabels >= 1 && nlabels <= KDE_MAX_LABELS, "ms_kde_select: bad arguments");
    MS_REQUIRE(ws_bytes >= ms_kde_select_workspace_bytes(nv), "ms_kde_select: workspace too small");
    unsigned long long mask = 0ull;
    for (int l = 0; l < nlabels; l++) if (select[l]) mask |= 1ull << l;
    MsArena A(ws, ws_bytes);
    double *dens = A.take<double>(nv), *bsum = A.take<double>(nv / RED_T + 2);
    const int nb = ms_blocks(nv, RED_T);
    MS_K("kde_select", s); k_kde_select<<<nb, RED_T, 0, s>>>(raw, nv, nlabels, mask, 1.0 / (double)nv, dens, bsum);
    MS_K("kde_emit", s); k_kde_emit<<<nb, RED_T, 0, s>>>(dens, nv, get_counts ? bsum : NULL, nb, n_selected, density_f64, density_f32);
    MS_K(NULL, s);
    MS_LAUNCH_CHECK();
    return 0;
}

// ------------------------------------------------------------------ ms_kde_select without a workspace
// No counterpart in the reference: added for callers that cannot allocate
// (CUDA graph capture, tight serving loops).  Same two kernels as ms_kde_select; their scratch (nv densities +
// one partial sum per block of RED_T) lives in a file-scope device array, which every GPU that loads the library
// holds until exit (MS_KDE_SELECT_STATIC_MAX_NV vertices: 8 MiB + 32 KiB).  Nothing in it has to start at zero --
// both kernels overwrite what they read -- but all calls on a device share it: calls that overlap on the GPU race.
static __device__ __align__(256) double g_kde_select_scratch[MS_KDE_SELECT_STATIC_MAX_NV + MS_KDE_SELECT_STATIC_MAX_NV / RED_T + 2];
#define KDE_STATIC_MAX_DEVICES 64
static double *g_kde_select_scratch_ptr[KDE_STATIC_MAX_DEVICES];      // device address of the array, per device
static std::mutex g_kde_select_scratch_lock;

extern "C" int ms_kde_select_static(const double *raw, int64_t nv, int nlabels, const uint8_t *select, double n_selected, int get_counts,
                                    double *density_f64, float *density_f32, void *stream) {
    cudaStream_t s = (cudaStream_t)stream;
    MS_REQUIRE(raw && select && nv > 0 && nlabels >= 1 && nlabels <= KDE_MAX_LABELS, "ms_kde_select_static: bad arguments");
    MS_REQUIRE(nv <= (int64_t)MS_KDE_SELECT_STATIC_MAX_NV,
               "ms_kde_select_static: %lld vertices exceed the static scratch (MS_KDE_SELECT_STATIC_MAX_NV = %lld); use ms_kde_select with a workspace",
               (long long)nv, (long long)MS_KDE_SELECT_STATIC_MAX_NV);
    int dev = -1;
    MS_CUDA(cudaGetDevice(&dev));
    MS_REQUIRE(dev >= 0 && dev < KDE_STATIC_MAX_DEVICES, "ms_kde_select_static: device %d out of range", dev);
    double *scratch = NULL;
    {
        // in host code the array's NAME is a host-side stand-in, not a device address: ask the runtime once per device
        std::lock_guard<std::mutex> hold(g_kde_select_scratch_lock);
        if (!g_kde_select_scratch_ptr[dev]) {
            double *ptr = NULL;
            MS_CUDA(cudaGetSymbolAddress((void**)&ptr, g_kde_select_scratch));
            MS_REQUIRE(ptr != NULL, "ms_kde_select_static: no device address for the scratch array");
            g_kde_select_scratch_ptr[dev] = ptr;
        }
        scratch = g_kde_select_scratch_ptr[dev];
    }
    unsigned long long mask = 0ull;
    for (int l = 0; l < nlabels; l++) if (select[l]) mask |= 1ull << l;
    double *dens = scratch, *bsum = scratch + MS_KDE_SELECT_STATIC_MAX_NV;
    const int nb = ms_blocks(nv, RED_T);
    MS_K("kde_select", s); k_kde_select<<<nb, RED_T, 0, s>>>(raw, nv, nlabels, mask, 1.0 / (double)nv, dens, bsum);
    MS_K("kde_emit", s); k_kde_emit<<<nb, RED_T, 0, s>>>(dens, nv, get_counts ? bsum : NULL, nb, n_selected, density_f64, density_f32);
    MS_K(NULL, s);
    MS_LAUNCH_CHECK();
    return 0;
}

// ------------------------------------------------------------------ all fields of a call in two launches
// Membrane.compute_densities (membrane.py:289-297) asks for one field per (sigma, label set); each is
//   density = sum_{l in set} raw[sigma][l] / nv          (TriMesh_kde.cpp:504-512)
//   get_counts: density *= n_selected / sum_j density    (TriMesh_kde.cpp:515-524)
// The sum over all vertices only depends on the per-(sigma, label) totals of raw: ms_kde_label_totals reduces every
// row of raw once (fixed order: KDE_TOT_BLOCKS partial sums per row), ms_kde_fields then writes ANY number of
// fields in one launch -- instead of two launches per field.
#define KDE_TOT_BLOCKS 256
#define KDE_MAX_FIELDS 64
__global__ void __launch_bounds__(RED_T) k_kde_label_totals(const double *__restrict__ raw, int64_t nv, double *__restrict__ partials) {
    __shared__ double sm[RED_T];
    const double *__restrict__ x = raw + (size_t)blockIdx.y * nv;
    double s = 0.0;
    for (int64_t i = (int64_t)blockIdx.x * RED_T + threadIdx.x; i < nv; i += (int64_t)KDE_TOT_BLOCKS * RED_T) s += x[i];
    sm[threadIdx.x] = s;
    __syncthreads();
    for (int o = RED_T / 2; o > 0; o >>= 1) { if (threadIdx.x < o) sm[threadIdx.x] += sm[threadIdx.x + o]; __syncthreads(); }
    if (threadIdx.x == 0) partials[(size_t)blockIdx.y * KDE_TOT_BLOCKS + blockIdx.x] = sm[0];
}
struct KdeFieldSet { int row0[KDE_MAX_FIELDS]; unsigned long long mask[KDE_MAX_FIELDS]; double nsel[KDE_MAX_FIELDS]; };
#define FIELD_PER_T 4
__global__ void __launch_bounds__(RED_T) k_kde_fields(const double *__restrict__ raw, const double *__restrict__ partials, int64_t nv, int nlabels,
                                                      KdeFieldSet F, int get_counts, double *__restrict__ o64, float *__restrict__ o32) {
    __shared__ double sm[RED_T];
    const int f = blockIdx.y;
    const int row0 = F.row0[f];
    const unsigned long long mask = F.mask[f];
    const double inv_nv = 1.0 / (double)nv;
    double scale = 1.0;
    if (get_counts) {             // sum_j density_j = (sum of the selected rows' totals) / nv, re-reduced in a fixed order by every block
        double t = 0.0;
        for (int l = 0; l < nlabels; l++) if ((mask >> l) & 1ull) t += partials[(size_t)(row0 + l) * KDE_TOT_BLOCKS + threadIdx.x];
        sm[threadIdx.x] = t;
        __syncthreads();
        for (int o = RED_T / 2; o > 0; o >>= 1) { if (threadIdx.x < o) sm[threadIdx.x] += sm[threadIdx.x + o]; __syncthreads(); }
        scale = F.nsel[f] / (sm[0] * inv_nv);
    }
#pragma unroll
    for (int k = 0; k < FIELD_PER_T; k++) {
        const int64_t j = ((int64_t)blockIdx.x * FIELD_PER_T + k) * RED_T + threadIdx.x;
        if (j >= nv) break;
        double d = 0.0;
        for (int l = 0; l < nlabels; l++) if ((mask >> l) & 1ull) d += raw[(size_t)(row0 + l) * nv + j];
        d *= inv_nv;
        if (get_counts) d *= scale;
        if (o64) o64[(size_t)f * nv + j] = d;
        if (o32) o32[(size_t)f * nv + j] = (float)d;
    }
}

static int kde_emit_sigma(const KdeEmit &E, const double *raw_i, int i, int64_t nv, int nlabels, cudaStream_t s) {
    double *part = E.partials + (size_t)i * nlabels * KDE_TOT_BLOCKS;
    MS_K("kde_label_totals", s); k_kde_label_totals<<<dim3(KDE_TOT_BLOCKS, (unsigned)nlabels), RED_T, 0, s>>>(raw_i, nv, part);
    KdeFieldSet F;
    for (int f = 0; f < E.nsets; f++) { F.row0[f] = 0; F.mask[f] = E.masks[f]; F.nsel[f] = E.nsel[f]; }
    const size_t off = (size_t)i * E.nsets * (size_t)nv;
    MS_K("kde_fields", s);
    k_kde_fields<<<dim3((unsigned)ms_blocks(nv, RED_T * FIELD_PER_T), (unsigned)E.nsets), RED_T, 0, s>>>(
        raw_i, part, nv, nlabels, F, E.get_counts, E.o64 ? E.o64 + off : NULL, E.o32 ? E.o32 + off : NULL);
    MS_K(NULL, s);
    MS_LAUNCH_CHECK();
    if (E.host32) {               // download on the copy stream, ordered after the fields kernel by an event
        cudaEvent_t ev;
        MS_CUDA(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
        MS_CUDA(cudaEventRecord(ev, s));
        MS_CUDA(cudaStreamWaitEvent(E.copy, ev, 0));
        MS_CUDA(cudaMemcpyAsync(E.host32 + off, E.o32 + off, sizeof(float) * (size_t)E.nsets * (size_t)nv, cudaMemcpyDeviceToHost, E.copy));
        MS_CUDA(cudaEventDestroy(ev));
    }
    return 0;
}

extern "C" int ms_kde_labels_multi_fields(const float *verts, int64_t nv, const int32_t *labels, int nlabels,
                                          int type, int periodic, const float *box6, const float *sigmas, int nsigma, double cutoff_sigmas,
                                          double *raw, int nsets, const uint64_t *masks, const double *n_selected, int get_counts,
                                          double *partials, double *out_f64, float *out_f32, float *host_f32, void *copy_stream,
                                          void *ws, size_t ws_bytes, void *stream) {
    MS_REQUIRE(raw && masks && n_selected && partials && out_f32, "ms_kde_labels_multi_fields: raw, masks, n_selected, partials and out_f32 are required");
    MS_REQUIRE(nsets >= 1 && nsets <= KDE_MAX_FIELDS && nlabels >= 1 && nlabels <= KDE_MAX_LABELS, "ms_kde_labels_multi_fields: 1 <= nsets <= %d", KDE_MAX_FIELDS);
    for (int f = 0; f < nsets; f++)
        MS_REQUIRE(masks[f] != 0ull && n_selected[f] > 0.0, "ms_kde_labels_multi_fields: set %d selects no source label", f);
    MS_REQUIRE(host_f32 == NULL || (copy_stream != NULL && copy_stream != stream), "ms_kde_labels_multi_fields: downloads need a copy stream of their own");
    KdeEmit E;
    E.nsets = nsets; E.masks = masks; E.nsel = n_selected; E.get_counts = get_counts; E.partials = partials;
    E.o64 = out_f64; E.o32 = out_f32; E.host32 = host_f32; E.copy = (cudaStream_t)copy_stream;
    return kde_labels_multi_impl(verts, nv, labels, nlabels, type, periodic, box6, sigmas, nsigma, cutoff_sigmas, raw, NULL, ws, ws_bytes, stream, &E);
}

extern "C" size_t ms_kde_label_totals_bytes(int64_t nrows) { return (size_t)(nrows < 1 ? 1 : nrows) * KDE_TOT_BLOCKS * sizeof(double); }
extern "C" int ms_kde_label_totals(const double *raw, int64_t nrows, int64_t nv, double *partials, void *stream) {
    cudaStream_t s = (cudaStream_t)stream;
    MS_REQUIRE(raw && partials && nrows >= 1 && nrows <= 65535 && nv > 0, "ms_kde_label_totals: bad arguments");
    MS_K("kde_label_totals", s); k_kde_label_totals<<<dim3(KDE_TOT_BLOCKS, (unsigned)nrows), RED_T, 0, s>>>(raw, nv, partials);
    MS_K(NULL, s);
    MS_LAUNCH_CHECK();
    return 0;
}
extern "C" int ms_kde_fields(const double *raw, const double *partials, int64_t nv, int nlabels, int nfields,
                             const int32_t *row0, const uint64_t *masks, const double *n_selected, int get_counts,
                             double *out_f64, float *out_f32, void *stream) {
    cudaStream_t s = (cudaStream_t)stream;
    MS_REQUIRE(raw && partials && row0 && masks && n_selected && nv > 0 && nlabels >= 1 && nlabels <= KDE_MAX_LABELS,
               "ms_kde_fields: bad arguments");
    MS_REQUIRE(nfields >= 1 && nfields <= KDE_MAX_FIELDS, "ms_kde_fields: 1 <= nfields <= %d", KDE_MAX_FIELDS);
    KdeFieldSet F;
    for (int f = 0; f < nfields; f++) {
        MS_REQUIRE(row0[f] >= 0 && masks[f] != 0ull && n_selected[f] > 0.0, "ms_kde_fields: field %d selects no source label", f);
        F.row0[f] = row0[f]; F.mask[f] = masks[f]; F.nsel[f] = n_selected[f];
    }
    MS_K("kde_fields", s);
    k_kde_fields<<<dim3((unsigned)ms_blocks(nv, RED_T * FIELD_PER_T), (unsigned)nfields), RED_T, 0, s>>>(raw, partials, nv, nlabels, F, get_counts, out_f64, out_f32);
    MS_K(NULL, s);
    MS_LAUNCH_CHECK();
    return 0;
}

extern "C" int ms_kde_evaluated_pairs(const float *verts, int64_t nv, const int32_t *ids, int64_t nids, const int32_t *labels, int nlabels,
                                      int type, int periodic, const float *box6, float sigma, double cutoff_sigmas,
                                      double *pairs, void *ws, size_t ws_bytes, void *stream) {
    cudaStream_t s = (cudaStream_t)stream;
    MS_REQUIRE(pairs != NULL, "ms_kde_evaluated_pairs: need an output");
    KdeCtx c;
    int rc = kde_front(verts, nv, ids, nids, labels, nlabels, type, periodic, box6, sigma, cutoff_sigmas, ws, ws_bytes, s, c);
    if (rc) return rc;
    if (!c.cells) { MS_CUDA(cudaStreamSynchronize(s)); pairs[0] = pairs[1] = (double)nv * (double)c.p.ns; return 0; }
    MS_CUDA(cudaMemsetAsync(c.counter, 0, 2 * sizeof(unsigned long long), s));
    if (kde_build_groups(c, s)) return 2;
    MS_K("kde_cells_count", s); k_kde_groups_count<<<ms_blocks(c.max_groups, 128), 128, 0, s>>>(c.groups, c.gstart + c.tl.nty, c.tstart, c.sst, (const double*)c.ssorted, c.type == 2 ? 2 : 4, c.nlabels, c.tl, c.counter);
    MS_K(NULL, s);
    MS_LAUNCH_CHECK();
    unsigned long long h[2] = { 0ull, 0ull };
    MS_CUDA(cudaMemcpyAsync(h, c.counter, sizeof(h), cudaMemcpyDeviceToHost, s));
    MS_CUDA(cudaStreamSynchronize(s));
    pairs[0] = (double)h[0]; pairs[1] = (double)h[1];
    return 0;
}
