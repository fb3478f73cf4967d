// ms_knn.cu -- point-normal estimation, stages 1-3: periodic ghost images, exact k-nearest
// neighbours through a uniform xy cell list (one warp per query), PCA normal per point.
//
// Replaces (LLNL/MemSurfer): PointSet::set_periodic memsurfer/src/PointSet.cpp:94-147 and the
// CGAL::pca_estimate_normals call of PointSet::need_normals PointSet.cpp:163-165 (kd-tree
// (k+1)-NN + 3x3 covariance + eigen-solve; CGAL 4.13 itself is not in the tree -- algorithm as
// in SURVEY.md Appendix A.1).
//
// Parity notes: ghosts are materialised explicitly, in the reference's order and with its
// float32 rounding, so neighbour INDICES are comparable bit for bit.  Squared distances are formed
// in fp64 as dx*dx + dy*dy + dz*dz without FMA contraction (this file is built --fmad=false);
// ties break by ascending index.
#include "ms_common.cuh"
#include <float.h>
#include <string.h>

// ------------------------------------------------------------------ stage 1: ghosts
struct GhostCuts { float boxw0, boxw1, l0, l1, r0, r1; };

__device__ __forceinline__ int ghost_flags(const float *__restrict__ p, const GhostCuts &c) {
    const float x = __ldg(p), y = __ldg(p + 1);
    const int lx = x < c.l0, ly = y < c.l1, rx = x > c.r0, ry = y > c.r1;
    return lx | (rx << 1) | (ly << 2) | (ry << 3);
}

__global__ void __launch_bounds__(256) k_ghost_count(const float *__restrict__ pts, int64_t n, GhostCuts c, int32_t *cnt) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int f = ghost_flags(pts + 3*i, c);
    const int hx = (f & 3) != 0, hy = (f & 12) != 0;
    cnt[i] = hx + hy + (hx & hy);
}

__global__ void __launch_bounds__(256) k_ghost_fill(const float *__restrict__ pts, int64_t n, GhostCuts c,
                                                    const int32_t *__restrict__ off, float *out, int32_t *src) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float x = pts[3*i], y = pts[3*i+1], z = pts[3*i+2];
    out[3*i] = x; out[3*i+1] = y; out[3*i+2] = z; src[i] = (int32_t)i;
    const int f = ghost_flags(pts + 3*i, c);
    const int lx = f & 1, rx = (f >> 1) & 1, ly = (f >> 2) & 1, ry = (f >> 3) & 1;
    if (!f) return;
    // create_point(const float&...) narrows the double sums to float (PointSet.cpp:26-28,133-142)
    const float gx = lx ? (float)((double)x + (double)c.boxw0) : (float)((double)x - (double)c.boxw0);
    const float gy = ly ? (float)((double)y + (double)c.boxw1) : (float)((double)y - (double)c.boxw1);
    int64_t g = n + off[i];
    if (lx | rx) { out[3*g] = gx; out[3*g+1] = y;  out[3*g+2] = z; src[g] = (int32_t)i; g++; }
    if (ly | ry) { out[3*g] = x;  out[3*g+1] = gy; out[3*g+2] = z; src[g] = (int32_t)i; g++; }
    if ((lx | rx) && (ly | ry)) { out[3*g] = gx; out[3*g+1] = gy; out[3*g+2] = z; src[g] = (int32_t)i; }
}

__global__ void __launch_bounds__(256) k_copy_points(const float *__restrict__ pts, int64_t n, float *out, int32_t *src) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    out[3*i] = pts[3*i]; out[3*i+1] = pts[3*i+1]; out[3*i+2] = pts[3*i+2]; src[i] = (int32_t)i;
}

extern "C" int64_t ms_pointset_capacity(int64_t n, int periodic) { return periodic ? 4 * n : n; }

// number of x / y coordinates outside the box: what Membrane.fit_points_to_box_xy (membrane.py:100-121) would move
__global__ void __launch_bounds__(256) k_count_outside(const float *__restrict__ pts, int64_t n, float b0x, float b0y, float b1x, float b1y,
                                                       unsigned long long *count) {
    unsigned c = 0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const float x = pts[3*i], y = pts[3*i+1];
        c += (x < b0x) + (x > b1x) + (y < b0y) + (y > b1y);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(count, (unsigned long long)c);
}
extern "C" int ms_points_outside_box(const float *points, int64_t n, const float *box6, int64_t *count_dev, void *stream) {
    cudaStream_t s = (cudaStream_t)stream;
    MS_REQUIRE(points && n > 0 && box6 && count_dev, "ms_points_outside_box: bad arguments");
    MS_CUDA(cudaMemsetAsync(count_dev, 0, sizeof(int64_t), s));
    MS_K("count_outside", s);
    k_count_outside<<<ms_sm_count() * 4, 256, 0, s>>>(points, n, box6[0], box6[1], box6[3], box6[4], (unsigned long long*)count_dev);
    MS_K(NULL, s);
    MS_LAUNCH_CHECK();
    return 0;
}
extern "C" size_t ms_ghosts_workspace_bytes(int64_t n) {
    return 2 * ms_align((size_t)(n + 1) * sizeof(int32_t)) + ms_scan_ws_bytes(n) + 1024;
}

// info[MS_INFO_M] = n + total ghosts (or n); the other diagnostic slots are cleared
__global__ void k_ghost_total(const int32_t *__restrict__ off_n, int64_t n, int64_t *info) {
    info[MS_INFO_M] = n + (off_n ? (int64_t)*off_n : 0);
    info[MS_INFO_UNORIENTED] = 0; info[MS_INFO_KNN_TIES] = 0; info[MS_INFO_MST_ROUNDS] = 0; info[MS_INFO_ERROR] = 0; info[MS_INFO_KNN_CANDS] = 0;
}

extern "C" int ms_ghosts(const float *points, int64_t n, int periodic, const float *box6, float blayer,
                         float *points_all, int32_t *ghost_src, int64_t *info,
                         void *ws, size_t ws_bytes, void *stream) {
    cudaStream_t s = (cudaStream_t)stream;
    MS_REQUIRE(n > 0 && info, "ms_ghosts: bad arguments");
    if (!periodic) {
        MS_K("copy_points", s); k_copy_points<<<ms_blocks(n, 256), 256, 0, s>>>(points, n, points_all, ghost_src);
        MS_K("ghost_total", s); k_ghost_total<<<1, 1, 0, s>>>(NULL, n, info);
        MS_K(NULL, s);
        MS_LAUNCH_CHECK();
        return 0;
    }
    MS_REQUIRE(box6, "ms_ghosts: periodic point set needs the box");
    MS_REQUIRE(blayer > 0.0f && blayer < 1.0f, "PointSet::set_periodic(): Invalid thickness of boundary layer (should be 0 < t < 1)! got %g", (double)blayer);
    MS_REQUIRE(ws_bytes >= ms_ghosts_workspace_bytes(n), "ms_ghosts: workspace too small");
    // float32 arithmetic exactly as PointSet.cpp:114-118 (volatile: keep each rounding step)
    volatile float bw0 = box6[3] - box6[0], bw1 = box6[4] - box6[1];
    volatile float t0 = blayer * bw0, t1 = blayer * bw1;
    GhostCuts c; c.boxw0 = bw0; c.boxw1 = bw1;
    c.l0 = box6[0] + t0; c.l1 = box6[1] + t1; c.r0 = box6[3] - t0; c.r1 = box6[4] - t1;
    MsArena A(ws, ws_bytes);
    int32_t *cnt = A.take<int32_t>(n + 1), *off = A.take<int32_t>(n + 1);
    void *scan_ws = A.take<char>(ms_scan_ws_bytes(n));
    MS_K("ghost_count", s); k_ghost_count<<<ms_blocks(n, 256), 256, 0, s>>>(points, n, c, cnt);
    if (ms_exclusive_scan_i32(cnt, off, n, scan_ws, s)) return 2;
    MS_K("ghost_fill", s); k_ghost_fill<<<ms_blocks(n, 256), 256, 0, s>>>(points, n, c, off, points_all, ghost_src);
    MS_K("ghost_total", s); k_ghost_total<<<1, 1, 0, s>>>(off + n, n, info);
    MS_K(NULL, s);
    MS_LAUNCH_CHECK();
    return 0;
}

// ------------------------------------------------------------------ stage 2: cell list + exact kNN
struct Grid { float lox, loy; double h, inv_h; int nx, ny, ok, pad; };

__device__ __forceinline__ unsigned f2ord(float f) { unsigned u = __float_as_uint(f); return (u & 0x80000000u) ? ~u : (u | 0x80000000u); }
__device__ __forceinline__ float ord2f(unsigned u) { return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u); }

__device__ __forceinline__ int64_t ms_count(const int64_t *info, int64_t cap) { return info ? min(__ldg(info + MS_INFO_M), cap) : cap; }

// bb[0..3] = ordered-uint (min x, min y, max x, max y)
__global__ void __launch_bounds__(256) k_bbox(const float *__restrict__ pts, int64_t m_cap, const int64_t *__restrict__ info, unsigned *bb) {
    const int64_t m = ms_count(info, m_cap);
    unsigned mnx = 0xffffffffu, mny = 0xffffffffu, mxx = 0u, mxy = 0u;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < m; i += (int64_t)gridDim.x * blockDim.x) {
        const unsigned x = f2ord(pts[3*i]), y = f2ord(pts[3*i+1]);
        mnx = min(mnx, x); mny = min(mny, y); mxx = max(mxx, x); mxy = max(mxy, y);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        mnx = min(mnx, __shfl_xor_sync(0xffffffffu, mnx, o)); mny = min(mny, __shfl_xor_sync(0xffffffffu, mny, o));
        mxx = max(mxx, __shfl_xor_sync(0xffffffffu, mxx, o)); mxy = max(mxy, __shfl_xor_sync(0xffffffffu, mxy, o));
    }
    if ((threadIdx.x & 31) == 0) { atomicMin(bb, mnx); atomicMin(bb + 1, mny); atomicMax(bb + 2, mxx); atomicMax(bb + 3, mxy); }
}

__device__ __forceinline__ int cell_of(const Grid &g, float x, float y, int &cx, int &cy) {
    cx = (int)(((double)x - (double)g.lox) * g.inv_h); cy = (int)(((double)y - (double)g.loy) * g.inv_h);
    cx = max(0, min(g.nx - 1, cx)); cy = max(0, min(g.ny - 1, cy));
    return cy * g.nx + cx;
}

// the cell list's geometry from the measured extent and the point count: one thread, on the device, so that
// the host never waits for either (ppc = target points per cell; h_min = smallest admissible cell side)
__global__ void k_make_grid(const unsigned *__restrict__ bb, float4 box4, int has_box, int64_t m_cap, int64_t *info,
                            double ppc, double h_min, int64_t ncell_max, Grid *gout) {
    const int64_t m = info ? min(info[MS_INFO_M], m_cap) : m_cap;
    float lo[2], hi[2];
    if (has_box) { lo[0] = box4.x; lo[1] = box4.y; hi[0] = box4.z; hi[1] = box4.w; }
    else { lo[0] = ord2f(bb[0]); lo[1] = ord2f(bb[1]); hi[0] = ord2f(bb[2]); hi[1] = ord2f(bb[3]); }
    double wx = (double)hi[0] - (double)lo[0], wy = (double)hi[1] - (double)lo[1];
    Grid g;
    g.ok = (wx == wx && wy == wy && wx < 1e30 && wy < 1e30) ? 1 : 0;        // non-finite coordinates: every query gives up
    if (!g.ok) { wx = 1.0; wy = 1.0; lo[0] = lo[1] = 0.f; if (info) info[MS_INFO_ERROR] |= MS_ERR_NONFINITE; }
    if (!(wx > 0.0)) wx = 1e-6; if (!(wy > 0.0)) wy = 1e-6;
    double h = sqrt(wx * wy * ppc / (double)(m > 0 ? m : 1));
    if (h < h_min) h = h_min;
    for (;;) {
        const double nxd = floor(wx / h) + 1.0, nyd = floor(wy / h) + 1.0;
        if (nxd * nyd <= (double)(ncell_max - 1)) { g.nx = (int)nxd; g.ny = (int)nyd; break; }
        h *= 1.5;
    }
    g.lox = lo[0]; g.loy = lo[1]; g.h = h; g.inv_h = 1.0 / h;
    *gout = g;
}

__global__ void __launch_bounds__(256) k_cell_count(const float *__restrict__ pts, int64_t m_cap, const int64_t *__restrict__ info,
                                                    const Grid *__restrict__ gp, int32_t *cell, int32_t *cnt) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= ms_count(info, m_cap)) return;
    const Grid g = *gp;
    int cx, cy; const int c = cell_of(g, pts[3*i], pts[3*i+1], cx, cy);
    cell[i] = c; atomicAdd(cnt + c, 1);
}

// sorted[p] = (x, y, z, original index as int bits) grouped by cell
__global__ void __launch_bounds__(256) k_cell_scatter(const float *__restrict__ pts, int64_t m_cap, const int64_t *__restrict__ info,
                                                      const int32_t *__restrict__ cell,
                                                      const int32_t *__restrict__ start, int32_t *cursor, float4 *sorted) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= ms_count(info, m_cap)) return;
    const int c = cell[i];
    const int p = start[c] + atomicAdd(cursor + c, 1);
    sorted[p] = make_float4(pts[3*i], pts[3*i+1], pts[3*i+2], __int_as_float((int)i));
}

__device__ __forceinline__ bool key_lt(double d, int i, double e, int j) { return d < e || (d == e && i < j); }

// One THREAD per query (queries in cell order, so a warp works on neighbouring cells): the K best
// (d2, index) so far live in registers as a sorted list; a candidate that beats the current worst is
// inserted by a branch-free compare-exchange chain.  K >= k1 is a compile-time capacity; the search
// stops when the K-th distance lies inside the explored square, so the first k1 entries are exact.
// The block's neighbour lists go through shared memory so that rows are written with coalescing.
#define KNNT_T 128
template <int K>
__device__ __forceinline__ void knn_scan(const float4 *__restrict__ sorted, int pb, int pe, double qx, double qy, double qz,
                                         double (&bd)[K], int (&bi)[K]) {
    for (int p = pb; p < pe; p++) {
        const float4 c = __ldg(sorted + p);
        const double dx = qx - (double)c.x, dy = qy - (double)c.y, dz = qz - (double)c.z;
        double cd = dx * dx; cd += dy * dy; cd += dz * dz;
        int ci = __float_as_int(c.w);
        if (!key_lt(cd, ci, bd[K - 1], bi[K - 1])) continue;
#pragma unroll
        for (int j = 0; j < K; j++) {
            const bool lt = key_lt(cd, ci, bd[j], bi[j]);
            const double td = lt ? bd[j] : cd; const int ti = lt ? bi[j] : ci;
            bd[j] = lt ? cd : bd[j]; bi[j] = lt ? ci : bi[j];
            cd = td; ci = ti;
        }
    }
}

// Same scan ordered by the squared distance ALONE (one fp64 compare per list slot, all independent, instead of
// the serial two-compare exchange chain above).  Exact as long as no two distances tie; `tie` is raised
// whenever a tie could have mattered -- a rejected candidate equal to the current worst, or a pushed-out worst
// equal to its neighbour -- and the caller then repeats the query with knn_scan (ties are decided by index).
template <int K>
__device__ __forceinline__ void knn_scan_fast(const float4 *__restrict__ sorted, int pb, int pe, double qx, double qy, double qz,
                                              double (&bd)[K], int (&bi)[K], bool &tie) {
    for (int p = pb; p < pe; p++) {
        const float4 c = __ldg(sorted + p);
        const double dx = qx - (double)c.x, dy = qy - (double)c.y, dz = qz - (double)c.z;
        double cd = dx * dx; cd += dy * dy; cd += dz * dz;
        const int ci = __float_as_int(c.w);
        if (!(cd < bd[K - 1])) { tie |= (cd == bd[K - 1]); continue; }
        if (K > 1) tie |= (bd[K - 1] == bd[K - 2]) && bd[K - 1] < DBL_MAX;
        // new b[j] = (c < b[j]) ? ((c < b[j-1]) ? b[j-1] : c) : b[j]   (the list is ascending, so the tests are monotone)
        double last_d = cd; int last_i = ci; bool prev = false;
#pragma unroll
        for (int j = 0; j < K; j++) {
            const double cur_d = bd[j]; const int cur_i = bi[j];
            const bool lt = cd < cur_d;
            bd[j] = lt ? (prev ? last_d : cd) : cur_d;
            bi[j] = lt ? (prev ? last_i : ci) : cur_i;
            last_d = cur_d; last_i = cur_i; prev = lt;
        }
    }
}

// ring-by-ring search around cell (cx,cy) until the K-th distance lies inside the explored square
template <int K, bool FAST>
__device__ __forceinline__ unsigned knn_search(const float4 *__restrict__ sorted, const Grid &g, const int32_t *__restrict__ start,
                                           int cx, int cy, double qx, double qy, double qz, double (&bd)[K], int (&bi)[K], bool &tie,
                                           int r0 = 0) {
    const double rx = (qx - (double)g.lox), ry = (qy - (double)g.loy);
    unsigned ncand = 0;
    for (int r = r0; ; r++) {
        for (int yy = cy - r; yy <= cy + r; yy++) {
            if (yy < 0 || yy >= g.ny) continue;
            const bool full_row = (yy == cy - r) || (yy == cy + r);
            const int nseg = full_row ? 1 : 2;
            for (int sg = 0; sg < nseg; sg++) {
                int x0, x1;
                if (full_row) { x0 = cx - r; x1 = cx + r; } else { x0 = x1 = (sg == 0 ? cx - r : cx + r); }
                x0 = max(x0, 0); x1 = min(x1, g.nx - 1);
                if (x0 > x1) continue;
                const int pb = __ldg(start + yy * g.nx + x0), pe = __ldg(start + yy * g.nx + x1 + 1);
                ncand += (unsigned)(pe - pb);
                if (FAST) knn_scan_fast<K>(sorted, pb, pe, qx, qy, qz, bd, bi, tie);
                else      knn_scan<K>(sorted, pb, pe, qx, qy, qz, bd, bi);
            }
        }
        if (cx - r <= 0 && cy - r <= 0 && cx + r >= g.nx - 1 && cy + r >= g.ny - 1) break;
        // distance to the unexplored region; a side on which the window already reaches the grid edge has none
        const double b0 = cx - r > 0 ? rx - (double)(cx - r) * g.h : DBL_MAX;
        const double b1 = cx + r < g.nx - 1 ? (double)(cx + r + 1) * g.h - rx : DBL_MAX;
        const double b2 = cy - r > 0 ? ry - (double)(cy - r) * g.h : DBL_MAX;
        const double b3 = cy + r < g.ny - 1 ? (double)(cy + r + 1) * g.h - ry : DBL_MAX;
        const double bm = fmin(fmin(b0, b1), fmin(b2, b3));
        if (bm > 0.0 && bd[K - 1] < DBL_MAX && sqrt(bd[K - 1]) < bm * (1.0 - 1e-9)) break;
    }
    return ncand;
}

template <int K>
__global__ void __launch_bounds__(KNNT_T) k_knn_thread(const float4 *__restrict__ sorted, int64_t m_cap, const int64_t *__restrict__ info,
                                                       const Grid *__restrict__ gp,
                                                       const int32_t *__restrict__ start, int k1, int32_t *__restrict__ nbr,
                                                       int64_t *__restrict__ info_out) {
    extern __shared__ int32_t sm_nbr[];                   // KNNT_T * k1 neighbour ids + KNNT_T query ids
    int32_t *sm_q = sm_nbr + KNNT_T * k1;
    const int64_t m = ms_count(info, m_cap);
    if ((int64_t)blockIdx.x * KNNT_T >= m) return;         // capacity-sized launch: blocks beyond the point count leave at once
    const Grid g = *gp;
    const int64_t q = (int64_t)blockIdx.x * KNNT_T + threadIdx.x;
    double bd[K]; int bi[K];
#pragma unroll
    for (int j = 0; j < K; j++) { bd[j] = DBL_MAX; bi[j] = 0x7fffffff; }
    int qi = -1;
    unsigned ncand = 0;
    if (q < m && g.ok) {
        const float4 me = sorted[q];
        qi = __float_as_int(me.w);
        const double qx = me.x, qy = me.y, qz = me.z;
        int cx, cy; cell_of(g, me.x, me.y, cx, cy);
        bool tie = false;
        ncand = knn_search<K, true>(sorted, g, start, cx, cy, qx, qy, qz, bd, bi, tie);
#pragma unroll
        for (int j = 1; j < K; j++) tie |= (bd[j] == bd[j - 1]) && bd[j] < DBL_MAX;      // ties inside the kept list
        if (tie) {             // rare: repeat with the exact (distance, index) order
            if (info_out) atomicAdd((unsigned long long*)(info_out + MS_INFO_KNN_TIES), 1ull);
#pragma unroll
            for (int j = 0; j < K; j++) { bd[j] = DBL_MAX; bi[j] = 0x7fffffff; }
            ncand += knn_search<K, false>(sorted, g, start, cx, cy, qx, qy, qz, bd, bi, tie);
        }
    }
    if (info_out) {             // diagnostic: candidate distances evaluated (one atomic per warp)
        unsigned long long c = ncand;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
        if ((threadIdx.x & 31) == 0 && c) atomicAdd((unsigned long long*)(info_out + MS_INFO_KNN_CANDS), c);
    }
    if (q < m && !g.ok) qi = __float_as_int(sorted[q].w);      // non-finite input: the row is all -1
    sm_q[threadIdx.x] = qi;
#pragma unroll
    for (int j = 0; j < K; j++) if (j < k1) sm_nbr[threadIdx.x * k1 + j] = (bi[j] == 0x7fffffff) ? -1 : bi[j];
    __syncthreads();
    for (int e = threadIdx.x; e < KNNT_T * k1; e += KNNT_T) {
        const int row = e / k1, col = e - row * k1;
        const int id = sm_q[row];
        if (id >= 0) nbr[(int64_t)id * k1 + col] = sm_nbr[e];
    }
}

// ---- the same search with the 3 x 3 block of cells pre-screened in float32 (r2).
// In k_knn_thread the warp runs the K-slot exchange chain (~6 instructions per slot) whenever ANY of its lanes accepts
// a candidate, i.e. for nearly every one of the ~63 candidates of the 3 x 3 block, although a lane only needs ~K of
// them.  Here a thread first writes the float32 distances of the block's candidates to shared memory, finds from a
// 32-bin histogram a threshold below which at least K of them lie (K + ~2 on average), and only those go through the
// exact fp64 distance and the exchange chain -- every lane walking to ITS next candidate before the warp runs the
// chain once for all of them: ~25 chain passes per warp instead of ~63.
// Exactness: if x is one of the true K nearest then at most K - 1 candidates are closer than x; at least K candidates
// have a float32 distance <= T, so one of them is not closer than x, and with a relative error eps (2.4e-7) of the
// float32 distance d32(x) <= T (1 + eps)^2: the margin 1e-5 keeps every true neighbour.  The chain and its tie flags
// are those of knn_scan_fast; a query whose block holds fewer than K or more than KNNS_CAP candidates, and the ring
// expansion beyond the block, take the code of k_knn_thread.
#define KNNS_CAP 96
template <int K>
__global__ void __launch_bounds__(KNNT_T) k_knn_select(const float4 *__restrict__ sorted, int64_t m_cap, const int64_t *__restrict__ info,
                                                       const Grid *__restrict__ gp,
                                                       const int32_t *__restrict__ start, int k1, int32_t *__restrict__ nbr,
                                                       int64_t *__restrict__ info_out) {
    extern __shared__ int32_t sm_nbr[];                   // first: float d32[KNNS_CAP][KNNT_T] + uint8 hist[32][KNNT_T]; then the output staging
    float *smd = (float*)sm_nbr;
    unsigned char *smh = (unsigned char*)(smd + KNNS_CAP * KNNT_T);
    const int tid = threadIdx.x;
    const int64_t m = ms_count(info, m_cap);
    if ((int64_t)blockIdx.x * KNNT_T >= m) return;         // capacity-sized launch: blocks beyond the point count leave at once
    const Grid g = *gp;
    const int64_t q = (int64_t)blockIdx.x * KNNT_T + tid;
    double bd[K]; int bi[K];
#pragma unroll
    for (int j = 0; j < K; j++) { bd[j] = DBL_MAX; bi[j] = 0x7fffffff; }
    int qi = -1;
    unsigned ncand = 0;
    const bool live = q < m && g.ok;
    float4 me = make_float4(0.f, 0.f, 0.f, 0.f);
    int cx = 0, cy = 0, n = 0, pb0 = 0, pb1 = 0, pb2 = 0, l0 = 0, l1 = 0;
    bool tie = false, screened = false;
    float thr = -1.f;
    int nsel = 0;
    if (live) {
        me = sorted[q];
        qi = __float_as_int(me.w);
        cell_of(g, me.x, me.y, cx, cy);
        const int x0 = max(cx - 1, 0), x1 = min(cx + 1, g.nx - 1);
        int len[3] = { 0, 0, 0 }, pb[3] = { 0, 0, 0 };
#pragma unroll
        for (int r = 0; r < 3; r++) {
            const int yy = cy - 1 + r;
            if (yy < 0 || yy >= g.ny) continue;
            pb[r] = __ldg(start + yy * g.nx + x0);
            len[r] = __ldg(start + yy * g.nx + x1 + 1) - pb[r];
        }
        pb0 = pb[0]; pb1 = pb[1]; pb2 = pb[2]; l0 = len[0]; l1 = len[0] + len[1];
        n = l1 + len[2];
        screened = n >= K && n <= KNNS_CAP;
        if (screened) {
            float dmax = 0.f;
            int i = 0;
#pragma unroll
            for (int r = 0; r < 3; r++)
                for (int p = pb[r]; p < pb[r] + len[r]; p++, i++) {
                    const float4 c = __ldg(sorted + p);
                    const float dx = c.x - me.x, dy = c.y - me.y, dz = c.z - me.z;
                    const float d = fmaf(dz, dz, fmaf(dy, dy, dx * dx));
                    smd[i * KNNT_T + tid] = d;
                    dmax = fmaxf(dmax, d);
                }
            screened = dmax > 0.f && dmax < 1e30f;
            if (screened) {
#pragma unroll
                for (int b = 0; b < 32; b++) smh[b * KNNT_T + tid] = 0;
                const float scale = 32.f / dmax;
                for (int j = 0; j < n; j++) {
                    const int b = min(31, (int)(smd[j * KNNT_T + tid] * scale));
                    smh[b * KNNT_T + tid]++;
                }
                int cum = 0, bstar = 31;
#pragma unroll
                for (int b = 0; b < 32; b++) {
                    cum += smh[b * KNNT_T + tid];
                    if (cum >= K && bstar == 31) bstar = b;
                }
                // every candidate of the bins 0..bstar: d * scale < bstar + 1 in float arithmetic, i.e. d below this edge
                thr = bstar == 31 ? dmax * 1.00002f : ((float)(bstar + 1) / scale) * 1.00002f;
                // the candidates below the threshold, compacted in place as indices (slot nsel <= j: its distance was read)
                int *sel = (int*)smd;
                for (int j = 0; j < n; j++)
                    if (smd[j * KNNT_T + tid] <= thr) { sel[nsel * KNNT_T + tid] = j; nsel++; }
            }
        }
    }
    // ---- exact pass over the screened candidates, the warp in step
    {
        const double qx = me.x, qy = me.y, qz = me.z;
        const int *sel = (const int*)smd;
        const int nn = screened ? nsel : 0;
        for (int it = 0; ; it++) {
            const bool have = it < nn;
            if (!__any_sync(0xffffffffu, have)) break;
            if (have) {
                const int i = sel[it * KNNT_T + tid];
                const int p = i < l0 ? pb0 + i : (i < l1 ? pb1 + (i - l0) : pb2 + (i - l1));
                const float4 c = __ldg(sorted + p);
                const double dx = qx - (double)c.x, dy = qy - (double)c.y, dz = qz - (double)c.z;
                double cd = dx * dx; cd += dy * dy; cd += dz * dz;
                const int ci = __float_as_int(c.w);
                if (!(cd < bd[K - 1])) tie |= (cd == bd[K - 1]);
                else {
                    if (K > 1) tie |= (bd[K - 1] == bd[K - 2]) && bd[K - 1] < DBL_MAX;
                    double last_d = cd; int last_i = ci; bool prev = false;
#pragma unroll
                    for (int j = 0; j < K; j++) {
                        const double cur_d = bd[j]; const int cur_i = bi[j];
                        const bool lt = cd < cur_d;
                        bd[j] = lt ? (prev ? last_d : cd) : cur_d;
                        bi[j] = lt ? (prev ? last_i : ci) : cur_i;
                        last_d = cur_d; last_i = cur_i; prev = lt;
                    }
                }
            }
        }
        if (live) {
            if (screened) {
                ncand = (unsigned)n;
                // is the K-th distance inside the explored 3 x 3 block?  (the test of knn_search after ring 1)
                const double rx = (qx - (double)g.lox), ry = (qy - (double)g.loy);
                const bool whole = cx - 1 <= 0 && cy - 1 <= 0 && cx + 1 >= g.nx - 1 && cy + 1 >= g.ny - 1;
                const double b0 = cx - 1 > 0 ? rx - (double)(cx - 1) * g.h : DBL_MAX;
                const double b1 = cx + 1 < g.nx - 1 ? (double)(cx + 2) * g.h - rx : DBL_MAX;
                const double b2 = cy - 1 > 0 ? ry - (double)(cy - 1) * g.h : DBL_MAX;
                const double b3 = cy + 1 < g.ny - 1 ? (double)(cy + 2) * g.h - ry : DBL_MAX;
                const double bm = fmin(fmin(b0, b1), fmin(b2, b3));
                const bool done = whole || (bm > 0.0 && bd[K - 1] < DBL_MAX && sqrt(bd[K - 1]) < bm * (1.0 - 1e-9));
                if (!done) ncand += knn_search<K, true>(sorted, g, start, cx, cy, qx, qy, qz, bd, bi, tie, 2);
            } else {
                ncand = knn_search<K, true>(sorted, g, start, cx, cy, qx, qy, qz, bd, bi, tie);
            }
#pragma unroll
            for (int j = 1; j < K; j++) tie |= (bd[j] == bd[j - 1]) && bd[j] < DBL_MAX;      // ties inside the kept list
            if (tie) {             // rare: repeat with the exact (distance, index) order
                if (info_out) atomicAdd((unsigned long long*)(info_out + MS_INFO_KNN_TIES), 1ull);
#pragma unroll
                for (int j = 0; j < K; j++) { bd[j] = DBL_MAX; bi[j] = 0x7fffffff; }
                ncand += knn_search<K, false>(sorted, g, start, cx, cy, qx, qy, qz, bd, bi, tie);
            }
        }
    }
    if (info_out) {             // diagnostic: candidate distances evaluated (one atomic per warp)
        unsigned long long c = ncand;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
        if ((tid & 31) == 0 && c) atomicAdd((unsigned long long*)(info_out + MS_INFO_KNN_CANDS), c);
    }
    if (q < m && !g.ok) qi = __float_as_int(sorted[q].w);      // non-finite input: the row is all -1
    __syncthreads();                                            // the distances are dead: the block's rows take their place
    int32_t *sm_q = sm_nbr + KNNT_T * k1;
    sm_q[tid] = qi;
#pragma unroll
    for (int j = 0; j < K; j++) if (j < k1) sm_nbr[tid * k1 + j] = (bi[j] == 0x7fffffff) ? -1 : bi[j];
    __syncthreads();
    const float inv_k1 = 1.0f / (float)k1;
    for (int e = tid; e < KNNT_T * k1; e += KNNT_T) {
        int row = (int)(((float)e + 0.5f) * inv_k1);          // e / k1 for e < 4096 (exact: the quotient is far from an integer edge)
        int col = e - row * k1;
        if (col < 0) { row--; col += k1; } else if (col >= k1) { row++; col -= k1; }
        const int id = sm_q[row];
        if (id >= 0) nbr[(int64_t)id * k1 + col] = sm_nbr[e];
    }
}

template <int K>
static int knn_select_launch(const float4 *sorted, int64_t m_cap, const int64_t *info, const Grid *g, const int32_t *start, int k1, int32_t *nbr,
                             int64_t *info_out, cudaStream_t s) {
    const size_t smem = (size_t)KNNS_CAP * KNNT_T * sizeof(float) + 32 * KNNT_T;      // >= the KNNT_T * (k1 + 1) ints of the output staging
    MS_CUDA(cudaFuncSetAttribute(k_knn_select<K>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_knn_select<K><<<ms_blocks(m_cap, KNNT_T), KNNT_T, smem, s>>>(sorted, m_cap, info, g, start, k1, nbr, info_out);
    return 0;
}

template <int K>
static void knn_thread_launch(const float4 *sorted, int64_t m_cap, const int64_t *info, const Grid *g, const int32_t *start, int k1, int32_t *nbr,
                              int64_t *info_out, cudaStream_t s) {
    const size_t smem = (size_t)KNNT_T * (k1 + 1) * sizeof(int32_t);
    k_knn_thread<K><<<ms_blocks(m_cap, KNNT_T), KNNT_T, smem, s>>>(sorted, m_cap, info, g, start, k1, nbr, info_out);
}

extern "C" size_t ms_knn_workspace_bytes(int64_t m, int k1) {
    (void)k1;
    const size_t ncell = (size_t)m + 2048;
    return ms_align((size_t)m * 4) + 3 * ms_align(ncell * 4) + ms_align((size_t)m * 16) + ms_scan_ws_bytes((int64_t)ncell) + 4096;
}

// the cell list shared by ms_knn and ms_overlapping_pairs: extent (measured, or box4 given by the caller) -> grid on
// the device -> counting sort of the points into `sorted` (x, y, z, index).  Nothing is read back.
struct CellList { int32_t *start; float4 *sorted; Grid *grid; unsigned *bb; };
static int build_cell_list(const float *points, int64_t m_cap, int64_t *info, const float *box4_host, double ppc, double h_min,
                           void *ws, size_t ws_bytes, cudaStream_t s, CellList &cl) {
    const int64_t ncell_max = m_cap + 2048;
    MsArena A(ws, ws_bytes);
    int32_t *cell = A.take<int32_t>(m_cap);
    int32_t *cnt = A.take<int32_t>(ncell_max), *start = A.take<int32_t>(ncell_max), *cursor = A.take<int32_t>(ncell_max);
    float4 *sorted = A.take<float4>(m_cap);
    void *scan_ws = A.take<char>(ms_scan_ws_bytes(ncell_max));
    unsigned *bb = (unsigned*)A.take<unsigned>(64);          // [0..3] extent, [8..11] overlap result, [32..] Grid
    Grid *grid = (Grid*)(bb + 32);
    float4 b4 = make_float4(0.f, 0.f, 0.f, 0.f);
    if (box4_host) b4 = make_float4(box4_host[0], box4_host[1], box4_host[2], box4_host[3]);
    else {
        MS_CUDA(cudaMemsetAsync(bb, 0xff, 2 * sizeof(unsigned), s));        // min x, min y (ordered uints)
        MS_CUDA(cudaMemsetAsync(bb + 2, 0, 2 * sizeof(unsigned), s));       // max x, max y
        MS_K("bbox", s); k_bbox<<<ms_sm_count() * 4, 256, 0, s>>>(points, m_cap, info, bb);
    }
    MS_K("make_grid", s); k_make_grid<<<1, 1, 0, s>>>(bb, b4, box4_host != NULL, m_cap, info, ppc, h_min, ncell_max, grid);
    MS_CUDA(cudaMemsetAsync(cnt, 0, sizeof(int32_t) * ncell_max, s));
    MS_CUDA(cudaMemsetAsync(cursor, 0, sizeof(int32_t) * ncell_max, s));
    MS_K("cell_count", s); k_cell_count<<<ms_blocks(m_cap, 256), 256, 0, s>>>(points, m_cap, info, grid, cell, cnt);
    if (ms_exclusive_scan_i32(cnt, start, ncell_max - 1, scan_ws, s)) return 2;      // cells beyond nx*ny are empty
    MS_K("cell_scatter", s); k_cell_scatter<<<ms_blocks(m_cap, 256), 256, 0, s>>>(points, m_cap, info, cell, start, cursor, sorted);
    MS_LAUNCH_CHECK();
    cl.start = start; cl.sorted = sorted; cl.grid = grid; cl.bb = bb;
    return 0;
}

extern "C" int ms_knn(const float *points_all, int64_t m_cap, int64_t *info, int k1, int32_t *nbr,
                      void *ws, size_t ws_bytes, void *stream) {
    cudaStream_t s = (cudaStream_t)stream;
    MS_REQUIRE(m_cap > 0 && k1 >= 1 && k1 <= 32, "ms_knn: need m > 0 and 1 <= k+1 <= 32 (got m=%lld, k+1=%d)", (long long)m_cap, k1);
    MS_REQUIRE(m_cap < ((int64_t)1 << 31) - 4096, "ms_knn: too many points");
    MS_REQUIRE(ws_bytes >= ms_knn_workspace_bytes(m_cap, k1), "ms_knn: workspace too small");
    double ppc = 0.37 * (double)k1; if (ppc < 2.0) ppc = 2.0;
    CellList cl;
    if (build_cell_list(points_all, m_cap, info, NULL, ppc, 0.0, ws, ws_bytes, s, cl)) return 2;
    MS_K("knn", s);
    if (k1 <= 8) knn_thread_launch<8>(cl.sorted, m_cap, info, cl.grid, cl.start, k1, nbr, info, s);
    else if (k1 <= 12) knn_thread_launch<12>(cl.sorted, m_cap, info, cl.grid, cl.start, k1, nbr, info, s);
    else if (k1 <= 16) knn_thread_launch<16>(cl.sorted, m_cap, info, cl.grid, cl.start, k1, nbr, info, s);
#ifndef KNN_NO_SELECT
    else if (k1 <= 19) { if (knn_select_launch<19>(cl.sorted, m_cap, info, cl.grid, cl.start, k1, nbr, info, s)) return 2; }
    else if (k1 <= 24) { if (knn_select_launch<24>(cl.sorted, m_cap, info, cl.grid, cl.start, k1, nbr, info, s)) return 2; }
#else
    else if (k1 <= 19) knn_thread_launch<19>(cl.sorted, m_cap, info, cl.grid, cl.start, k1, nbr, info, s);
    else if (k1 <= 24) knn_thread_launch<24>(cl.sorted, m_cap, info, cl.grid, cl.start, k1, nbr, info, s);
#endif
    else knn_thread_launch<32>(cl.sorted, m_cap, info, cl.grid, cl.start, k1, nbr, info, s);
    MS_K(NULL, s);
    MS_LAUNCH_CHECK();
    return 0;
}

// ------------------------------------------------------------------ overlapping points
// Replaces utils.test_overlapping_points memsurfer/utils.py:91-159 (scipy pdist over ALL pairs, O(N^2) memory:
// unusable beyond ~5e4 points) for the detection part: pairs closer than dthreshold can only sit in the
// same or an adjacent cell of the kNN cell list (cell side >= dthreshold), so one thread per point scans its
// 3x3 cells.  pairs (max_pairs,2) receives (i<j) index pairs in no particular order; *npairs their total
// count (may exceed max_pairs); *min_above = smallest distance > dthreshold seen among the scanned
// candidates (the reference's move radius, utils.py:129; +inf if none).
__global__ void __launch_bounds__(128) k_overlap(const float4 *__restrict__ sorted, int64_t m, const Grid *__restrict__ gp, const int32_t *__restrict__ start,
                                                 double thr, int32_t *pairs, unsigned long long max_pairs,
                                                 unsigned long long *npairs, unsigned long long *min_above_bits) {
    const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const Grid g = *gp;
    double mine = 1e300;
    if (p < m && g.ok) {
        const float4 q = __ldg(sorted + p);
        const int qi = __float_as_int(q.w);
        int cx, cy; cell_of(g, q.x, q.y, cx, cy);
        for (int yy = max(cy - 1, 0); yy <= min(cy + 1, g.ny - 1); yy++)
            for (int xx = max(cx - 1, 0); xx <= min(cx + 1, g.nx - 1); xx++) {
                const int c = yy * g.nx + xx;
                const int e1 = __ldg(start + c + 1);
                for (int e = __ldg(start + c); e < e1; e++) {
                    const float4 o = __ldg(sorted + e);
                    const int oi = __float_as_int(o.w);
                    if (oi <= qi) continue;
                    const double dx = (double)q.x - (double)o.x, dy = (double)q.y - (double)o.y, dz = (double)q.z - (double)o.z;
                    double d2 = dx * dx; d2 += dy * dy; d2 += dz * dz;
                    const double d = sqrt(d2);
                    if (d < thr) {
                        const unsigned long long at = atomicAdd(npairs, 1ull);
                        if (at < max_pairs) { pairs[2 * at] = qi; pairs[2 * at + 1] = oi; }
                    } else if (d > thr && d < mine) mine = d;
                }
            }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mine = fmin(mine, __shfl_xor_sync(0xffffffffu, mine, o));
    if ((threadIdx.x & 31) == 0 && mine < 1e300) atomicMin(min_above_bits, (unsigned long long)__double_as_longlong(mine));   // positive doubles order as integers
}
__global__ void k_overlap_init(unsigned long long *res) { res[0] = 0ull; res[1] = 0x7ff0000000000000ull; }

extern "C" size_t ms_overlap_workspace_bytes(int64_t n) { return ms_knn_workspace_bytes(n, 19); }

// box4 (host, may be NULL = measure the extent on the device) = xy extent to lay the cell list over
// {xmin, ymin, xmax, ymax} (points outside are clamped into the border cells, which costs speed, not pairs);
// result_dev (device, 2 x u64) receives {pair count, bits of the smallest distance above the threshold}.
static int overlap_launch(const float *points, int64_t m, double dthreshold, const float *box4, int32_t *pairs, int64_t max_pairs,
                          unsigned long long *result_dev, void *ws, size_t ws_bytes, cudaStream_t s, unsigned long long **res_out) {
    MS_REQUIRE(points && m > 0 && dthreshold >= 0.0 && max_pairs >= 0 && (pairs || max_pairs == 0), "ms_overlapping_pairs: bad arguments");
    MS_REQUIRE(m < ((int64_t)1 << 31) - 4096, "ms_overlapping_pairs: too many points");
    MS_REQUIRE(ws_bytes >= ms_overlap_workspace_bytes(m), "ms_overlapping_pairs: workspace too small");
    CellList cl;
    if (build_cell_list(points, m, NULL, box4, 2.0, dthreshold, ws, ws_bytes, s, cl)) return 2;
    unsigned long long *res = result_dev ? result_dev : (unsigned long long*)(cl.bb + 8);   // [0] = pair count, [1] = min distance bits
    MS_K("overlap_init", s); k_overlap_init<<<1, 1, 0, s>>>(res);
    MS_K("overlap", s); k_overlap<<<ms_blocks(m, 128), 128, 0, s>>>(cl.sorted, m, cl.grid, cl.start, dthreshold, pairs, (unsigned long long)max_pairs, res, res + 1);
    MS_K(NULL, s);
    MS_LAUNCH_CHECK();
    if (res_out) *res_out = res;
    return 0;
}

extern "C" int ms_overlapping_pairs_async(const float *points, int64_t m, double dthreshold, const float *box4, int32_t *pairs,
                                          int64_t max_pairs, unsigned long long *result_dev, void *ws, size_t ws_bytes, void *stream) {
    MS_REQUIRE(result_dev != NULL, "ms_overlapping_pairs_async: result_dev is required");
    return overlap_launch(points, m, dthreshold, box4, pairs, max_pairs, result_dev, ws, ws_bytes, (cudaStream_t)stream, NULL);
}

extern "C" int ms_overlapping_pairs(const float *points, int64_t m, double dthreshold, int32_t *pairs, int64_t max_pairs,
                                    int64_t *npairs_out, double *min_above_out, void *ws, size_t ws_bytes, void *stream) {
    cudaStream_t s = (cudaStream_t)stream;
    MS_REQUIRE(npairs_out != NULL, "ms_overlapping_pairs: npairs is required");
    unsigned long long *res = NULL;
    const int rc = overlap_launch(points, m, dthreshold, NULL, pairs, max_pairs, NULL, ws, ws_bytes, s, &res);
    if (rc) return rc;
    unsigned long long h2[2];
    MS_CUDA(cudaMemcpyAsync(h2, res, sizeof(h2), cudaMemcpyDeviceToHost, s));
    MS_CUDA(cudaStreamSynchronize(s));
    *npairs_out = (int64_t)h2[0];
    if (min_above_out) memcpy(min_above_out, &h2[1], 8);
    return 0;
}

// ------------------------------------------------------------------ stage 3: PCA normals
// One thread per point: centroid and covariance in neighbour order, cyclic Jacobi (CGAL's internal
// 3x3 solver when CGAL_EIGEN3_ENABLED is off, setup.py:217), smallest-eigenvalue eigenvector.
__global__ void __launch_bounds__(128) k_pca(const float *__restrict__ pts, int64_t m_cap, const int64_t *__restrict__ info,
                                             const int32_t *__restrict__ nbr, int k1, double *__restrict__ normals) {
    const int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= ms_count(info, m_cap)) return;
    double c0 = 0.0, c1 = 0.0, c2 = 0.0; int cnt = 0;
    for (int j = 0; j < k1; j++) {
        const int i = __ldg(nbr + q * k1 + j); if (i < 0) continue;
        c0 += ms_ld_f(pts + 3*(size_t)i); c1 += ms_ld_f(pts + 3*(size_t)i + 1); c2 += ms_ld_f(pts + 3*(size_t)i + 2); cnt++;
    }
    c0 /= cnt; c1 /= cnt; c2 /= cnt;
    double a00 = 0, a01 = 0, a02 = 0, a11 = 0, a12 = 0, a22 = 0;
    for (int j = 0; j < k1; j++) {
        const int i = __ldg(nbr + q * k1 + j); if (i < 0) continue;
        const double dx = ms_ld_f(pts + 3*(size_t)i) - c0, dy = ms_ld_f(pts + 3*(size_t)i + 1) - c1, dz = ms_ld_f(pts + 3*(size_t)i + 2) - c2;
        a00 += dx*dx; a01 += dx*dy; a02 += dx*dz; a11 += dy*dy; a12 += dy*dz; a22 += dz*dz;
    }
    double a[3][3] = { { a00, a01, a02 }, { a01, a11, a12 }, { a02, a12, a22 } };
    double v[3][3] = { { 1, 0, 0 }, { 0, 1, 0 }, { 0, 0, 1 } };
    for (int sweep = 0; sweep < 64; sweep++) {
        const double offn = a[0][1]*a[0][1] + a[0][2]*a[0][2] + a[1][2]*a[1][2];
        if (offn == 0.0) break;
#pragma unroll
        for (int p = 0; p < 2; p++) {
#pragma unroll
            for (int qq = p + 1; qq < 3; qq++) {
                if (a[p][qq] == 0.0) continue;
                const double theta = (a[qq][qq] - a[p][p]) / (2.0 * a[p][qq]);
                const double t = (theta >= 0.0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
                const double cs = 1.0 / sqrt(t * t + 1.0), sn = t * cs;
                const double apq = a[p][qq];
                a[p][p] -= t * apq; a[qq][qq] += t * apq; a[p][qq] = a[qq][p] = 0.0;
                const int r = 3 - p - qq;
                const double arp = a[r][p], arq = a[r][qq];
                a[r][p] = a[p][r] = cs * arp - sn * arq;
                a[r][qq] = a[qq][r] = sn * arp + cs * arq;
#pragma unroll
                for (int i = 0; i < 3; i++) {
                    const double vip = v[i][p], viq = v[i][qq];
                    v[i][p] = cs * vip - sn * viq;
                    v[i][qq] = sn * vip + cs * viq;
                }
            }
        }
    }
    double n0, n1, n2;
    if (a[0][0] == a[1][1] && a[1][1] == a[2][2]) { n0 = 0.0; n1 = 0.0; n2 = 1.0; }
    else {
        double e = a[0][0]; n0 = v[0][0]; n1 = v[1][0]; n2 = v[2][0];
        if (a[1][1] < e) { e = a[1][1]; n0 = v[0][1]; n1 = v[1][1]; n2 = v[2][1]; }
        if (a[2][2] < e) { e = a[2][2]; n0 = v[0][2]; n1 = v[1][2]; n2 = v[2][2]; }
        const double l = sqrt(n0*n0 + n1*n1 + n2*n2);
        n0 /= l; n1 /= l; n2 /= l;
    }
    normals[3*q] = n0; normals[3*q+1] = n1; normals[3*q+2] = n2;
}

extern "C" int ms_pca_normals(const float *points_all, int64_t m_cap, const int64_t *info, const int32_t *nbr, int k1,
                              double *normals, void *stream) {
    MS_REQUIRE(m_cap > 0 && k1 >= 1, "ms_pca_normals: bad arguments");
    MS_K("pca", (cudaStream_t)stream); k_pca<<<ms_blocks(m_cap, 128), 128, 0, (cudaStream_t)stream>>>(points_all, m_cap, info, nbr, k1, normals);
    MS_K(NULL, (cudaStream_t)stream);
    MS_LAUNCH_CHECK();
    return 0;
}
