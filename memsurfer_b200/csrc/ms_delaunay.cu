// ms_delaunay.cu -- periodic 2-D Delaunay triangulation on B200 (first slice).
//
// Replaces (LLNL/MemSurfer): TriMesh::periodicDelaunay memsurfer/src/TriMesh_cgal.cpp:160-281
// (CGAL::Periodic_2_Delaunay_triangulation_2, converted to the 9-sheeted covering; the loop at :212-262 keeps
// every face with at least one vertex in the original sheet as three PeriodicVertex{origVidx, offx, offy}).
//
// Design: no incremental insertion, no global data structure.  The periodic images the faces around the
// original sheet can touch are the ghost points ms_ghosts already makes (0.2 box widths per side), the
// candidate neighbours of a point are its k nearest ones from ms_knn (sorted by distance).  One thread per
// ORIGINAL point builds its Voronoi cell by clipping a square with the bisector of each neighbour in turn and
// stops at the security radius (a neighbour farther than twice the cell's circumradius cannot cut it; the
// list is sorted, so neither can any later one).  The generators of consecutive cell edges, taken
// counter-clockwise, are the point's Delaunay neighbours in angular order: every cell vertex is one
// triangle.  A triangle is written by its smallest point index (ghosts come after the originals, so that is
// always an original vertex): count pass -> exclusive scan -> fill pass, i.e. a deterministic face order.
// A point whose list ends before the security radius is reached scans all points for the remaining cutters
// (rare: very uneven density); a cell that cannot be closed inside the ghost layer is counted in
// *n_incomplete (the caller falls back to its host producer).
//
// Arithmetic: fp64 on float32 inputs, compiled without FMA contraction.  Whether a bisector cuts off a cell vertex is
// the in-circle predicate of four points; when the fp64 value of that test is within 1e-10 of zero (relative to its
// terms) it is re-evaluated in double-double arithmetic on the exact coordinate differences, the four points taken in
// one canonical (lexicographic) order so that every cell computes bit-identical values for a shared quadruple, and a
// ZERO (|det| <= 2^-96 of its terms) -- four or more cocircular points -- is decided by the symbolic perturbation CGAL's Delaunay_triangulation_2 uses
// (side_of_oriented_circle with `perturb`: points ordered lexicographically, the largest one is pushed off the
// circle first [recalled, CGAL 4.13; CGAL itself is not in the tree]).  The answer is a function of the four points
// only, so the four cells that share a degenerate Voronoi vertex agree and the stars always glue to a triangulation.
#include "ms_common.cuh"

// ---- double-double helpers (explicit fma(): this file is built --fmad=false, which only stops contraction)
struct dd { double hi, lo; };
__device__ __forceinline__ dd dd_two_sum(double a, double b) { const double s = a + b, bb = s - a; dd r; r.hi = s; r.lo = (a - (s - bb)) + (b - bb); return r; }
__device__ __forceinline__ dd dd_two_prod(double a, double b) { dd r; r.hi = a * b; r.lo = fma(a, b, -r.hi); return r; }
__device__ __forceinline__ dd dd_add(dd a, dd b) {
    dd s = dd_two_sum(a.hi, b.hi); const dd t = dd_two_sum(a.lo, b.lo);
    s.lo += t.hi; s = dd_two_sum(s.hi, s.lo); s.lo += t.lo; return dd_two_sum(s.hi, s.lo);
}
__device__ __forceinline__ dd dd_neg(dd a) { a.hi = -a.hi; a.lo = -a.lo; return a; }
__device__ __forceinline__ dd dd_mul(dd a, dd b) {
    dd p = dd_two_prod(a.hi, b.hi);
    p.lo += a.hi * b.lo + a.lo * b.hi;
    return dd_two_sum(p.hi, p.lo);
}
__device__ __forceinline__ int dd_sign(dd a) { return a.hi > 0.0 ? 1 : (a.hi < 0.0 ? -1 : (a.lo > 0.0 ? 1 : (a.lo < 0.0 ? -1 : 0))); }
// sign of the orientation of (a, b, c): > 0 counter-clockwise
__device__ __forceinline__ int dl_orient(double ax, double ay, double bx, double by, double cx, double cy) {
    const double ux = bx - ax, uy = by - ay, vx = cx - ax, vy = cy - ay;     // exact (float32-valued inputs)
    return dd_sign(dd_add(dd_two_prod(ux, vy), dd_neg(dd_two_prod(uy, vx))));
}
// sign of the in-circle determinant: > 0 iff d lies inside the circle through a, b, c (counter-clockwise).
// The three products carry up to ~108 bits for 26-bit differences, two more than a double-double holds, so a value
// within 2^-96 of the terms' magnitude is reported as ZERO (the caller then perturbs symbolically): a true non-zero
// value that small is as good as cocircular, and the caller evaluates every quadruple in ONE canonical order, so
// whatever is decided here is decided identically in all four cells that share the quadruple.
__device__ __forceinline__ int dl_incircle(double ax, double ay, double bx, double by, double cx, double cy, double dx, double dy) {
    const double x0 = ax - dx, y0 = ay - dy, x1 = bx - dx, y1 = by - dy, x2 = cx - dx, y2 = cy - dy;   // exact
    const dd l0 = dd_add(dd_two_prod(x0, x0), dd_two_prod(y0, y0));
    const dd l1 = dd_add(dd_two_prod(x1, x1), dd_two_prod(y1, y1));
    const dd l2 = dd_add(dd_two_prod(x2, x2), dd_two_prod(y2, y2));
    const dd m0 = dd_add(dd_two_prod(x1, y2), dd_neg(dd_two_prod(y1, x2)));
    const dd m1 = dd_add(dd_two_prod(x2, y0), dd_neg(dd_two_prod(y2, x0)));
    const dd m2 = dd_add(dd_two_prod(x0, y1), dd_neg(dd_two_prod(y0, x1)));
    const dd t0 = dd_mul(l0, m0), t1 = dd_mul(l1, m1), t2 = dd_mul(l2, m2);
    const dd det = dd_add(dd_add(t0, t1), t2);
    const double mag = fabs(t0.hi) + fabs(t1.hi) + fabs(t2.hi);
    if (fabs(det.hi) <= 1.2621774483536189e-29 * mag) return 0;                    // 2^-96
    return dd_sign(det);
}
// The point set: n original points followed by their periodic images.  An image is stored as the float32 sum
// point + k * width, i.e. ROUNDED -- two cells that see the same four points through different images would then
// evaluate the predicates on slightly different quadruples and could disagree on a nearly cocircular one (a lattice
// whose spacing is not a power of two).  Every coordinate is therefore rebuilt as source + k * width in fp64, which
// is exact for float32 operands: all the images of a configuration are exact translates of each other.
struct DlPts { const float *pts; const int32_t *gsrc; int64_t n; double wx, wy; };
__device__ __forceinline__ void dl_xy(const DlPts &P, int64_t i, double &x, double &y) {
    x = (double)P.pts[3*i]; y = (double)P.pts[3*i + 1];
    if (i >= P.n) {
        const int64_t o = P.gsrc[i];
        const double ox = (double)P.pts[3*o], oy = (double)P.pts[3*o + 1];
        x = ox + rint((x - ox) / P.wx) * P.wx; y = oy + rint((y - oy) / P.wy) * P.wy;
    }
}
// lexicographic (x, y, index) order of two points
__device__ __forceinline__ bool dl_lex_less(const double (&X)[4], const double (&Y)[4], const int64_t (&id)[4], int i, int j) {
    if (X[i] != X[j]) return X[i] < X[j];
    if (Y[i] != Y[j]) return Y[i] < Y[j];
    return id[i] < id[j];
}
// is q inside the circle through (p0, p1, p2), counter-clockwise and not collinear -- ties decided by perturbation
__device__ __forceinline__ bool dl_inside_sos(const DlPts &P, int64_t p0, int64_t p1, int64_t p2, int64_t q) {
    const int64_t id[4] = { p0, p1, p2, q };
    double X[4], Y[4];
#pragma unroll
    for (int i = 0; i < 4; i++) dl_xy(P, id[i], X[i], Y[i]);
    // canonical order of the quadruple: lexicographic in the exact coordinates (the same in every image of it)
    int o[4] = { 0, 1, 2, 3 };
    int parity = 1;
#pragma unroll
    for (int a = 0; a < 3; a++)
#pragma unroll
        for (int b = 0; b < 3 - a; b++)
            if (dl_lex_less(X, Y, id, o[b + 1], o[b])) { const int t = o[b]; o[b] = o[b + 1]; o[b + 1] = t; parity = -parity; }
    // the lifted 4 x 4 determinant is antisymmetric in its four points
    const int s = parity * dl_incircle(X[o[0]], Y[o[0]], X[o[1]], Y[o[1]], X[o[2]], Y[o[2]], X[o[3]], Y[o[3]]);
    if (s != 0) return s > 0;
    // cocircular: walk the four points from the lexicographically largest down (CGAL: 2 iterations are enough)
    for (int i = 3; i > 0; i--) {
        const int w = o[i];
        if (w == 3) return false;                                   // ON_NEGATIVE_SIDE: outside
        int r = 0;
        if (w == 2) r = dl_orient(X[0], Y[0], X[1], Y[1], X[3], Y[3]);
        else if (w == 1) r = dl_orient(X[0], Y[0], X[3], Y[3], X[2], Y[2]);
        else r = dl_orient(X[3], Y[3], X[1], Y[1], X[2], Y[2]);
        if (r != 0) return r > 0;
    }
    return false;
}

#define DL_MAXV 24          // cell vertices kept per thread (a Voronoi cell has ~6)
// security radius: a point farther than twice the cell's radius cannot cut the cell.  The margin keeps the points ON
// that circle (bisector through the farthest cell vertex: the diagonal neighbour of a lattice point) in the list even
// when rounding puts them just beyond it -- the symbolic perturbation may still count them as cutters, and the cell
// across the diagonal will have asked the same question.  (|s| <= 1e-10 in dl_clip corresponds to 2e-10 here.)
#define DL_SECURE (4.0 * (1.0 + 1e-8))

struct DlCell { double x[DL_MAXV], y[DL_MAXV]; int g[DL_MAXV]; int n; };

// clip the convex cell (counter-clockwise) with the half-plane {v : v.d <= |d|^2/2}; the new edge gets generator q.
// Cell vertex i is the circumcentre of (p, g[i-1], g[i]) (counter-clockwise): it is cut off iff q lies inside that
// circle.  The fp64 value decides unless it is within 1e-10 (relative) of zero; then the robust predicate does.
__device__ __forceinline__ bool dl_clip(DlCell &c, double dx, double dy, int q, const DlPts &P, int64_t p) {
    const double h = 0.5 * (dx * dx + dy * dy);
    double s[DL_MAXV];
    unsigned outm = 0u;
    for (int i = 0; i < c.n; i++) {
        const double tx = c.x[i] * dx, ty = c.y[i] * dy;
        s[i] = tx + ty - h;
        bool out = s[i] > 0.0;
        const int ga = c.g[i == 0 ? c.n - 1 : i - 1], gb = c.g[i];
        if (ga >= 0 && gb >= 0 && ga != gb && fabs(s[i]) <= 1e-10 * (fabs(tx) + fabs(ty) + h))
            out = dl_inside_sos(P, p, ga, gb, q);
        outm |= (unsigned)out << i;
    }
    if (!outm) return true;
    DlCell o; o.n = 0;
    for (int i = 0; i < c.n; i++) {
        const int j = i + 1 == c.n ? 0 : i + 1;
        const bool in_i = !((outm >> i) & 1u), in_j = !((outm >> j) & 1u);
        if (in_i) {
            if (o.n >= DL_MAXV) return false;
            o.x[o.n] = c.x[i]; o.y[o.n] = c.y[i];
            o.g[o.n] = in_j ? c.g[i] : c.g[i];            // edge i survives (whole or up to the cut)
            o.n++;
        }
        if (in_i != in_j) {
            double t = s[i] / (s[i] - s[j]);                  // a predicate overruled by the robust test leaves t outside [0,1]
            t = (t >= 0.0 && t <= 1.0) ? t : (in_i ? 1.0 : 0.0);
            if (o.n >= DL_MAXV) return false;
            o.x[o.n] = c.x[i] + t * (c.x[j] - c.x[i]); o.y[o.n] = c.y[i] + t * (c.y[j] - c.y[i]);
            // leaving the half-plane: the next edge runs along the bisector (generator q);
            // entering it: the rest of edge i continues with its own generator
            o.g[o.n] = in_i ? q : c.g[i];
            o.n++;
        }
    }
    c = o;
    return true;
}

__device__ __forceinline__ double dl_radius2(const DlCell &c) {
    double r2 = 0.0;
    for (int i = 0; i < c.n; i++) { const double v = c.x[i] * c.x[i] + c.y[i] * c.y[i]; r2 = v > r2 ? v : r2; }
    return r2;
}

// the Voronoi cell of point p; returns the number of cell vertices (0: cannot be closed).
// When the sorted neighbour list ends before the security radius is reached (very uneven density) the cell
// clipped so far is a superset of the true one, so only points within twice its radius can still cut it: the
// thread scans ALL m points for those (order does not matter: the cell is an intersection of half-planes).
// reach2 = (width of the ghost layer)^2: a cell whose security disc leaves the ghost layer could be cut by a
// periodic image that was never made -- reported as not closed.
__device__ __forceinline__ void dl_square(DlCell &c, double R) {
    c.n = 4;
    c.x[0] = -R; c.y[0] = -R; c.x[1] = R; c.y[1] = -R; c.x[2] = R; c.y[2] = R; c.x[3] = -R; c.y[3] = R;
    c.g[0] = c.g[1] = c.g[2] = c.g[3] = -1;
}
__device__ __forceinline__ int dl_cell(const DlPts &P, const int32_t *__restrict__ nbr, int k1, int64_t p, int64_t m,
                                       double reach2, DlCell &c) {
    const double px = (double)P.pts[3*p], py = (double)P.pts[3*p + 1];       // p < n: an original point
    const int32_t *row = nbr + p * k1;
    // fast path: start from a square that contains every bisector of the list (half side = distance to the last neighbour)
    double far2 = 0.0;
    {
        const int32_t q = row[k1 - 1];
        double qx, qy; dl_xy(P, q, qx, qy);
        const double dx = qx - px, dy = qy - py;
        far2 = dx * dx + dy * dy;
    }
    dl_square(c, sqrt(far2));
    bool complete = false;
    double r2 = 2.0 * far2;
    for (int j = 0; j < k1; j++) {
        const int32_t q = row[j];
        if ((int64_t)q == p) continue;
        double qx, qy; dl_xy(P, q, qx, qy);
        const double dx = qx - px, dy = qy - py;
        if (dx * dx + dy * dy > DL_SECURE * r2) { complete = true; break; }
        if (!dl_clip(c, dx, dy, q, P, p)) return 0;
        r2 = dl_radius2(c);
    }
    bool open = false;
    for (int i = 0; i < c.n; i++) open |= c.g[i] < 0;
    // the list ended before the security radius was reached, or part of the starting square is still there (all
    // listed neighbours on one side: the edge of a cluster, the rim of a void): start again from the largest square a
    // closable cell fits in (half side = half the ghost layer), clip with the whole list, then scan ALL points for
    // the cutters beyond it (order does not matter: the cell is an intersection of half-planes)
    if ((!complete && !(far2 > 4.0 * r2)) || open) {
        dl_square(c, 0.5001 * sqrt(reach2));
        r2 = dl_radius2(c);
        for (int j = 0; j < k1; j++) {
            const int32_t q = row[j];
            if ((int64_t)q == p) continue;
            double qx, qy; dl_xy(P, q, qx, qy);
            const double dx = qx - px, dy = qy - py;
            if (!dl_clip(c, dx, dy, q, P, p)) return 0;
        }
        r2 = dl_radius2(c);
        for (int64_t q = 0; q < m; q++) {
            if (q == p) continue;
            double qx, qy; dl_xy(P, q, qx, qy);
            const double dx = qx - px, dy = qy - py;
            const double d2 = dx * dx + dy * dy;
            if (d2 > DL_SECURE * r2 || d2 < far2) continue;             // cannot cut / closer than the list's last entry: already used
            if (d2 == far2) {                                      // the last entry itself (or a tie with it): clip a point only once
                bool used = false;
                for (int j = 0; j < k1; j++) used |= (int64_t)row[j] == q;
                if (used) continue;
            }
            if (!dl_clip(c, dx, dy, (int)q, P, p)) return 0;
            r2 = dl_radius2(c);
        }
    }
    if (4.0 * r2 > reach2) return 0;
    for (int i = 0; i < c.n; i++) if (c.g[i] < 0) return 0;
    return c.n;
}

// FILL == false: cnt[p] = faces owned by p (p is their smallest index), incomplete points counted;
// FILL == true: faces written to dfaces/doffs from offset start[p]
template <bool FILL>
__global__ void __launch_bounds__(128) k_delaunay_star(const float *__restrict__ pts, int64_t n, const int32_t *__restrict__ gsrc,
                                                       const int32_t *__restrict__ nbr, int k1, int64_t m, double reach2, double wx, double wy,
                                                       const int32_t *__restrict__ start, int32_t *__restrict__ cnt, int32_t *__restrict__ incomplete,
                                                       int32_t *__restrict__ dfaces, int8_t *__restrict__ doffs) {
    const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    DlCell c;
    DlPts P; P.pts = pts; P.gsrc = gsrc; P.n = n; P.wx = wx; P.wy = wy;
    const int nv = p < n ? dl_cell(P, nbr, k1, p, m, reach2, c) : 0;
    if (!FILL) {
        // corners of all stars: 6n on a torus (Euler); less when a cocircular quadruple was left without a diagonal
        const int tot = __reduce_add_sync(0xffffffffu, nv);
        if ((threadIdx.x & 31) == 0) atomicAdd(incomplete + 1, tot);
        if (p >= n) return;
        if (nv == 0) { atomicAdd(incomplete, 1); cnt[p] = 0; return; }
    }
    if (p >= n) return;
    int w = FILL ? start[p] : 0, mine = 0;
    for (int i = 0; i < nv; i++) {
        // cell vertex i sits between edge i-1 and edge i: triangle (p, g[i-1], g[i]), counter-clockwise
        const int a = c.g[i == 0 ? nv - 1 : i - 1], b = c.g[i];
        if (a == b || (int64_t)a < p || (int64_t)b < p) continue;
        if (FILL) {
            const int64_t v[3] = { p, a, b };
            for (int k = 0; k < 3; k++) {
                const int64_t o = gsrc[v[k]];
                dfaces[3 * (int64_t)w + k] = (int32_t)o;
                const double ox = ((double)pts[3*v[k]] - (double)pts[3*o]) / wx, oy = ((double)pts[3*v[k] + 1] - (double)pts[3*o + 1]) / wy;
                doffs[6 * (int64_t)w + 2*k] = (int8_t)rint(ox);
                doffs[6 * (int64_t)w + 2*k + 1] = (int8_t)rint(oy);
            }
            w++;
        }
        mine++;
    }
    if (!FILL) cnt[p] = mine;
}

extern "C" size_t ms_delaunay_periodic_workspace_bytes(int64_t n) {
    return 2 * ms_align((size_t)(n + 1) * sizeof(int32_t)) + ms_scan_ws_bytes(n) + 1024;
}

extern "C" int ms_delaunay_periodic(const float *points_all, int64_t m, int64_t n, const int32_t *ghost_src,
                                    const int32_t *nbr, int k1, double box_wx, double box_wy, double ghost_layer,
                                    int32_t *dfaces, int8_t *doffs, int64_t cap, int64_t *nfaces, int64_t *n_incomplete,
                                    void *ws, size_t ws_bytes, void *stream) {
    cudaStream_t s = (cudaStream_t)stream;
    MS_REQUIRE(points_all && ghost_src && nbr && nfaces && n_incomplete, "ms_delaunay_periodic: bad arguments");
    MS_REQUIRE(n > 0 && n <= m && k1 >= 4 && k1 <= 32 && m >= k1, "ms_delaunay_periodic: bad sizes (n=%lld, m=%lld, k+1=%d)", (long long)n, (long long)m, k1);
    MS_REQUIRE(box_wx > 0.0 && box_wy > 0.0, "ms_delaunay_periodic: empty box");
    MS_REQUIRE(ws_bytes >= ms_delaunay_periodic_workspace_bytes(n), "ms_delaunay_periodic: workspace too small");
    const double reach2 = ghost_layer * ghost_layer;
    MsArena A(ws, ws_bytes);
    int32_t *cnt = A.take<int32_t>(n + 1), *start = A.take<int32_t>(n + 1);
    void *scan_ws = A.take<char>(ms_scan_ws_bytes(n));
    int32_t *flag = (int32_t*)A.take<int32_t>(8);
    MS_CUDA(cudaMemsetAsync(flag, 0, 8 * sizeof(int32_t), s));
    MS_K("delaunay_count", s);
    k_delaunay_star<false><<<ms_blocks(n, 128), 128, 0, s>>>(points_all, n, ghost_src, nbr, k1, m, reach2, box_wx, box_wy, NULL, cnt, flag, NULL, NULL);
    if (ms_exclusive_scan_i32(cnt, start, n, scan_ws, s)) return 2;
    MS_K(NULL, s);
    MS_LAUNCH_CHECK();
    int32_t h[3] = { 0, 0, 0 };
    MS_CUDA(cudaMemcpyAsync(&h[0], start + n, sizeof(int32_t), cudaMemcpyDeviceToHost, s));
    MS_CUDA(cudaMemcpyAsync(&h[1], flag, 2 * sizeof(int32_t), cudaMemcpyDeviceToHost, s));
    MS_CUDA(cudaStreamSynchronize(s));
    *nfaces = h[0]; *n_incomplete = h[1];
    if (h[1] == 0 && (int64_t)h[2] != 6 * n) {      // degenerate input (exactly cocircular points): not a triangulation
        const int64_t off = (int64_t)h[2] - 6 * n;      // star corners missing (< 0) or in excess
        *n_incomplete = -(off < 0 ? -off : off);
        return 0;
    }
    if (h[1] != 0 || dfaces == NULL) return 0;
    MS_REQUIRE((int64_t)h[0] <= cap, "ms_delaunay_periodic: %d faces do not fit the output (%lld)", h[0], (long long)cap);
    MS_K("delaunay_fill", s);
    k_delaunay_star<true><<<ms_blocks(n, 128), 128, 0, s>>>(points_all, n, ghost_src, nbr, k1, m, reach2, box_wx, box_wy, start, NULL, NULL, dfaces, doffs);
    MS_K(NULL, s);
    MS_LAUNCH_CHECK();
    return 0;
}

// ------------------------------------------------------------------ split for the trim step
// TriMesh::trim_periodicDelaunay (TriMesh.cpp:493-575) keeps the faces whose three corners lie in the original
// sheet as they are (mFaces) and re-indexes the others onto duplicated vertices.  The first group is ~99 % of a
// large mesh: it is compacted here, in order (flag -> exclusive scan -> scatter); the few faces that cross the
// box go to the host-side bookkeeping as (vertex ids, offsets), also in order.
__global__ void __launch_bounds__(256) k_split_flag(const int8_t *__restrict__ doffs, int64_t nf, int32_t *flag) {
    const int64_t f = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= nf) return;
    int any = 0;
#pragma unroll
    for (int k = 0; k < 6; k++) any |= doffs[6*f + k];
    flag[f] = any == 0;
}
__global__ void __launch_bounds__(256) k_split_scatter(const int32_t *__restrict__ dfaces, const int8_t *__restrict__ doffs, int64_t nf,
                                                       const int32_t *__restrict__ flag, const int32_t *__restrict__ pos,
                                                       uint32_t *__restrict__ full, int32_t *__restrict__ pfaces, int8_t *__restrict__ poffs) {
    const int64_t f = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= nf) return;
    if (flag[f]) {
        const int64_t o = pos[f];
#pragma unroll
        for (int k = 0; k < 3; k++) full[3*o + k] = (uint32_t)dfaces[3*f + k];
    } else {
        const int64_t o = f - pos[f];
#pragma unroll
        for (int k = 0; k < 3; k++) pfaces[3*o + k] = dfaces[3*f + k];
#pragma unroll
        for (int k = 0; k < 6; k++) poffs[6*o + k] = doffs[6*f + k];
    }
}

extern "C" size_t ms_delaunay_split_workspace_bytes(int64_t nf) {
    return 2 * ms_align((size_t)(nf + 1) * sizeof(int32_t)) + ms_scan_ws_bytes(nf) + 1024;
}

// full (nf,3) u32, pfaces (nf,3) i32, poffs (nf,3,2) i8 sized for the worst case; *nfull = faces written to `full`
extern "C" int ms_delaunay_split(const int32_t *dfaces, const int8_t *doffs, int64_t nf, uint32_t *full, int32_t *pfaces, int8_t *poffs,
                                 int64_t *nfull, void *ws, size_t ws_bytes, void *stream) {
    cudaStream_t s = (cudaStream_t)stream;
    MS_REQUIRE(dfaces && doffs && full && pfaces && poffs && nfull && nf >= 0, "ms_delaunay_split: bad arguments");
    MS_REQUIRE(ws_bytes >= ms_delaunay_split_workspace_bytes(nf), "ms_delaunay_split: workspace too small");
    *nfull = 0;
    if (nf == 0) return 0;
    MsArena A(ws, ws_bytes);
    int32_t *flag = A.take<int32_t>(nf + 1), *pos = A.take<int32_t>(nf + 1);
    void *scan_ws = A.take<char>(ms_scan_ws_bytes(nf));
    MS_K("split_flag", s); k_split_flag<<<ms_blocks(nf, 256), 256, 0, s>>>(doffs, nf, flag);
    if (ms_exclusive_scan_i32(flag, pos, nf, scan_ws, s)) return 2;
    MS_K("split_scatter", s); k_split_scatter<<<ms_blocks(nf, 256), 256, 0, s>>>(dfaces, doffs, nf, flag, pos, full, pfaces, poffs);
    MS_K(NULL, s);
    MS_LAUNCH_CHECK();
    int32_t h = 0;
    MS_CUDA(cudaMemcpyAsync(&h, pos + nf, sizeof(int32_t), cudaMemcpyDeviceToHost, s));
    MS_CUDA(cudaStreamSynchronize(s));
    *nfull = h;
    return 0;
}

// ------------------------------------------------------------------ duplicated vertices of the crossing faces
// Second step of TriMesh::trim_periodicDelaunay (TriMesh.cpp:513-560): every corner of a crossing face that is a periodic
// image (offset != 0) is re-indexed onto a DUPLICATED vertex; the reference finds or appends (orig, offx, offy) with a
// linear std::find, so duplicates are numbered in order of FIRST APPEARANCE in the (face, corner) scan.  Here: every
// image corner inserts its key into an open-addressing hash table with atomicMin on the corner's position (the first
// appearance wins whatever the thread order), the first appearances are flagged and scanned, and every corner looks
// its number up.  Coordinates as TriMesh.cpp:547-549: float(double(v) + double(float(off) * boxw)).
__device__ __forceinline__ unsigned long long dup_key(int v, int ox, int oy) { return (unsigned long long)(unsigned)v * 9ull + (unsigned long long)((ox + 1) * 3 + (oy + 1)); }
__device__ __forceinline__ uint32_t dup_hash(unsigned long long k) { k ^= k >> 33; k *= 0xff51afd7ed558ccdull; k ^= k >> 33; return (uint32_t)k; }
#define DUP_EMPTY 0xffffffffffffffffull
// table slot: key (u64) and the smallest corner position that carries it (u32 in a u64 for atomicMin)
__global__ void __launch_bounds__(256) k_dup_insert(const int32_t *__restrict__ pf, const int8_t *__restrict__ po, int64_t ncorner,
                                                    unsigned long long *keys, unsigned long long *firstpos, uint32_t tmask, int32_t *slot_of) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= ncorner) return;
    const int ox = po[2*i], oy = po[2*i + 1];
    if (ox == 0 && oy == 0) { slot_of[i] = -1; return; }
    const unsigned long long k = dup_key(pf[i], ox, oy);
    uint32_t h = dup_hash(k) & tmask;
    for (;;) {
        const unsigned long long prev = atomicCAS(keys + h, DUP_EMPTY, k);
        if (prev == DUP_EMPTY || prev == k) break;
        h = (h + 1) & tmask;
    }
    atomicMin(firstpos + h, (unsigned long long)i);
    slot_of[i] = (int32_t)h;
}
__global__ void __launch_bounds__(256) k_dup_flag(int64_t ncorner, const int32_t *__restrict__ slot_of, const unsigned long long *__restrict__ firstpos, int32_t *flag) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= ncorner) return;
    const int s = slot_of[i];
    flag[i] = s >= 0 && firstpos[s] == (unsigned long long)i;
}
__global__ void __launch_bounds__(256) k_dup_emit(const float *__restrict__ cverts, int dim, int64_t nv, const int32_t *__restrict__ pf, const int8_t *__restrict__ po,
                                                  int64_t ncorner, const int32_t *__restrict__ slot_of, const unsigned long long *__restrict__ firstpos,
                                                  const int32_t *__restrict__ rank, float bwx, float bwy,
                                                  uint32_t *__restrict__ tfaces, uint32_t *__restrict__ dup_ids, float *__restrict__ dup_verts) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= ncorner) return;
    const int s = slot_of[i];
    const int v = pf[i];
    if (s < 0) { tfaces[i] = (uint32_t)v; return; }
    const int64_t first = (int64_t)firstpos[s];
    const int d = rank[first];
    tfaces[i] = (uint32_t)(nv + d);
    if (first == i) {
        dup_ids[d] = (uint32_t)v;
        const int ox = po[2*i], oy = po[2*i + 1];
        for (int c = 0; c < dim; c++) {
            float x = cverts[i * dim + c];
            if (c == 0 && ox != 0) x = (float)((double)x + (double)((float)ox * bwx));
            if (c == 1 && oy != 0) x = (float)((double)x + (double)((float)oy * bwy));
            dup_verts[(int64_t)d * dim + c] = x;
        }
    }
}
static int64_t dup_table_size(int64_t ncorner) { int64_t t = 1024; while (t < 4 * ncorner) t <<= 1; return t; }
extern "C" size_t ms_delaunay_duplicates_workspace_bytes(int64_t nfaces) {
    const int64_t nc = 3 * nfaces;
    return 2 * ms_align((size_t)dup_table_size(nc) * 8) + 3 * ms_align((size_t)(nc + 1) * 4) + ms_scan_ws_bytes(nc) + 1024;
}
// corner_verts (3 nf, dim) f32 = this mesh's coordinates of the ORIGINAL vertex of every corner (dim 2 or 3; a few thousand rows:
// cheaper to gather than to ship the whole vertex array), nv = number of original vertices, pfaces (nf,3) i32 / poffs (nf,3,2) i8 =
// the crossing faces from ms_delaunay_split, box widths as float32 (bbox1 - bbox0 of THIS mesh); outputs: tfaces (nf,3) u32 (image corners -> nv + duplicate number), dup_ids (3 nf) u32,
// dup_verts (3 nf, dim) f32; *ndup (host) = number of duplicated vertices.  Syncs the stream once.
extern "C" int ms_delaunay_duplicates(const float *corner_verts, int64_t nv, int dim, const int32_t *pfaces, const int8_t *poffs, int64_t nf,
                                      float box_wx, float box_wy, uint32_t *tfaces, uint32_t *dup_ids, float *dup_verts, int64_t *ndup,
                                      void *ws, size_t ws_bytes, void *stream) {
    cudaStream_t s = (cudaStream_t)stream;
    MS_REQUIRE((dim == 2 || dim == 3) && nf >= 0 && ndup, "ms_delaunay_duplicates: bad arguments");
    MS_REQUIRE(ws_bytes >= ms_delaunay_duplicates_workspace_bytes(nf), "ms_delaunay_duplicates: workspace too small");
    *ndup = 0;
    if (nf == 0) return 0;
    MS_REQUIRE(corner_verts && pfaces && poffs && tfaces && dup_ids && dup_verts, "ms_delaunay_duplicates: bad arguments");
    const int64_t nc = 3 * nf, tsz = dup_table_size(nc);
    MS_REQUIRE(tsz < ((int64_t)1 << 31), "ms_delaunay_duplicates: too many crossing faces");
    MsArena A(ws, ws_bytes);
    unsigned long long *keys = A.take<unsigned long long>(tsz), *firstpos = A.take<unsigned long long>(tsz);
    int32_t *slot_of = A.take<int32_t>(nc + 1), *flag = A.take<int32_t>(nc + 1), *rank = A.take<int32_t>(nc + 1);
    void *scan_ws = A.take<char>(ms_scan_ws_bytes(nc));
    MS_CUDA(cudaMemsetAsync(keys, 0xff, (size_t)tsz * 8, s));
    MS_CUDA(cudaMemsetAsync(firstpos, 0xff, (size_t)tsz * 8, s));
    MS_K("dup_insert", s); k_dup_insert<<<ms_blocks(nc, 256), 256, 0, s>>>(pfaces, poffs, nc, keys, firstpos, (uint32_t)(tsz - 1), slot_of);
    MS_K("dup_flag", s); k_dup_flag<<<ms_blocks(nc, 256), 256, 0, s>>>(nc, slot_of, firstpos, flag);
    if (ms_exclusive_scan_i32(flag, rank, nc, scan_ws, s)) return 2;
    MS_K("dup_emit", s); k_dup_emit<<<ms_blocks(nc, 256), 256, 0, s>>>(corner_verts, dim, nv, pfaces, poffs, nc, slot_of, firstpos, rank, box_wx, box_wy,
                                                                     tfaces, dup_ids, dup_verts);
    MS_K(NULL, s);
    MS_LAUNCH_CHECK();
    int32_t h = 0;
    MS_CUDA(cudaMemcpyAsync(&h, rank + nc, sizeof(int32_t), cudaMemcpyDeviceToHost, s));
    MS_CUDA(cudaStreamSynchronize(s));
    *ndup = h;
    return 0;
}
