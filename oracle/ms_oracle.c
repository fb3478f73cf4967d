/* ms_oracle.c -- TEST INFRASTRUCTURE ONLY.
 *
 * CPU restatement (fp64 arithmetic on float32-valued inputs) of MemSurfer's
 * per-frame membrane-property hot path.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this library; the product (memsurfer_b200)
 * never does.  Citations are path:line under /root/reference.
 *
 * Parity pinning:
 *   - mesh normals / point areas / KDE (A6, A7, A13-A16): PINNED against the
 *     reference's own C++ compiled in place (oracle/_ref/libmsref.so, float32)
 *     by tests/test_oracle.py.
 *   - point normals (A3, A4: CGAL 4.13 pca_estimate_normals/mst_orient_normals)
 *     and curvatures (A9, A10: VTK >= 9.1 vtkCurvatures): third-party code that
 *     is absent from the tree and from this image -> restated from the
 *     published algorithms; PARITY UNPINNED.
 *
 * Build: make -C oracle   (gcc -O2 -ffp-contract=off: no FMA contraction, so
 * sums are formed as the x86-64 baseline build of the reference would).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include <float.h>

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

/* ------------------------------------------------------------------------- */
/* A2: PointSet::set_periodic  memsurfer/src/PointSet.cpp:94-147              */
/* Ghost images of the points inside the boundary layer, appended per original */
/* point in the order x-image, y-image, xy-image.  All arithmetic is float32   */
/* as in the reference (TypeFunction = float, Types.hpp:26; create_point takes */
/* const float&, PointSet.cpp:26-28; p[0] is a double holding a float value,   */
/* boxw a float: the sum is formed in double and narrowed to float by the      */
/* create_point call).  box = {b0x,b0y,b0z,b1x,b1y,b1z}.                        */
/* out may be NULL to count only.  Returns the number of ghosts.              */
/* ------------------------------------------------------------------------- */
int64_t orc_ghosts(const float *pts, int64_t n, const float *box, float thickness,
                   float *out /* (ng,3) */, int32_t *src /* (ng,) original index, may be NULL */) {
    const float boxw[2] = { box[3] - box[0], box[4] - box[1] };
    const float lcut[2] = { box[0] + thickness * boxw[0], box[1] + thickness * boxw[1] };
    const float rcut[2] = { box[3] - thickness * boxw[0], box[4] - thickness * boxw[1] };
    int64_t g = 0;
    for (int64_t i = 0; i < n; i++) {
        const double px = pts[3*i], py = pts[3*i+1], pz = pts[3*i+2];
        const int lx = px < lcut[0], ly = py < lcut[1];
        const int rx = px > rcut[0], ry = py > rcut[1];
        if (!lx && !rx && !ly && !ry) continue;
#define ORC_PUSH(X, Y) do { if (out) { out[3*g] = (float)(X); out[3*g+1] = (float)(Y); out[3*g+2] = (float)pz; } \
                            if (src) src[g] = (int32_t)i; g++; } while (0)
        if (lx)      ORC_PUSH(px + boxw[0], py);
        else if (rx) ORC_PUSH(px - boxw[0], py);
        if (ly)      ORC_PUSH(px, py + boxw[1]);
        else if (ry) ORC_PUSH(px, py - boxw[1]);
        if (lx && ly)      ORC_PUSH(px + boxw[0], py + boxw[1]);
        else if (lx && ry) ORC_PUSH(px + boxw[0], py - boxw[1]);
        else if (rx && ly) ORC_PUSH(px - boxw[0], py + boxw[1]);
        else if (rx && ry) ORC_PUSH(px - boxw[0], py - boxw[1]);
#undef ORC_PUSH
    }
    return g;
}

/* ------------------------------------------------------------------------- */
/* A6: TriMesh::need_normals (static)  memsurfer/src/TriMesh.cpp:99-133        */
/* verts (nv,3) f32 [periodic: verts ++ dup verts], faces (nf,3) u32           */
/* [periodic: faces ++ trimmed faces] as the wrapper TriMesh.hpp:356-370       */
/* concatenates them; caller keeps the first nv_orig rows.                    */
/* fp64 throughout; the reference's float(1.0/(l2a*l2c)) scale is kept double. */
/* Zero vector normalises to (1,0,0)  memsurfer/src/Vec.h:796-813.             */
/* ------------------------------------------------------------------------- */
void orc_mesh_normals(const float *verts, int64_t nv, const uint32_t *faces, int64_t nf,
                      double *pn /* (nv,3) */) {
    memset(pn, 0, sizeof(double) * 3 * (size_t)nv);
    for (int64_t i = 0; i < nf; i++) {
        const uint32_t f0 = faces[3*i], f1 = faces[3*i+1], f2 = faces[3*i+2];
        double p0[3], p1[3], p2[3], a[3], b[3], c[3], fn[3];
        for (int d = 0; d < 3; d++) { p0[d] = verts[3*f0+d]; p1[d] = verts[3*f1+d]; p2[d] = verts[3*f2+d]; }
        for (int d = 0; d < 3; d++) { a[d] = p0[d]-p1[d]; b[d] = p1[d]-p2[d]; c[d] = p2[d]-p0[d]; }
        const double l2a = a[0]*a[0]+a[1]*a[1]+a[2]*a[2];
        const double l2b = b[0]*b[0]+b[1]*b[1]+b[2]*b[2];
        const double l2c = c[0]*c[0]+c[1]*c[1]+c[2]*c[2];
        if (l2a == 0.0 || l2b == 0.0 || l2c == 0.0) continue;
        fn[0] = a[1]*b[2] - a[2]*b[1];
        fn[1] = a[2]*b[0] - a[0]*b[2];
        fn[2] = a[0]*b[1] - a[1]*b[0];
        const double s0 = 1.0 / (l2a * l2c), s1 = 1.0 / (l2b * l2a), s2 = 1.0 / (l2c * l2b);
        for (int d = 0; d < 3; d++) {
            pn[3*f0+d] += fn[d] * s0;
            pn[3*f1+d] += fn[d] * s1;
            pn[3*f2+d] += fn[d] * s2;
        }
    }
    for (int64_t i = 0; i < nv; i++) {
        double *v = pn + 3*i;
        const double l = sqrt(v[0]*v[0] + v[1]*v[1] + v[2]*v[2]);
        if (l <= 0.0) { v[0] = 1.0; v[1] = 0.0; v[2] = 0.0; }
        else { const double r = 1.0 / l; v[0] *= r; v[1] *= r; v[2] *= r; }
    }
}

/* ------------------------------------------------------------------------- */
/* A7: TriMesh::need_pointareas (static)  memsurfer/src/TriMesh.cpp:140-190    */
/* Mixed Voronoi / obtuse-corner vertex areas.                                */
/* ------------------------------------------------------------------------- */
static inline double dot3(const double *a, const double *b) { return a[0]*b[0] + a[1]*b[1] + a[2]*b[2]; }
static inline void cross3(const double *a, const double *b, double *o) {
    o[0] = a[1]*b[2] - a[2]*b[1]; o[1] = a[2]*b[0] - a[0]*b[2]; o[2] = a[0]*b[1] - a[1]*b[0];
}

void orc_corner_areas(const double e[3][3], double ca[3]) {
    double cr[3];
    cross3(e[0], e[1], cr);
    const double area = 0.5 * sqrt(dot3(cr, cr));
    const double l2[3] = { dot3(e[0], e[0]), dot3(e[1], e[1]), dot3(e[2], e[2]) };
    const double ew[3] = { l2[0] * (l2[1] + l2[2] - l2[0]),
                           l2[1] * (l2[2] + l2[0] - l2[1]),
                           l2[2] * (l2[0] + l2[1] - l2[2]) };
    if (ew[0] <= 0.0) {
        ca[1] = -0.25 * l2[2] * area / dot3(e[0], e[2]);
        ca[2] = -0.25 * l2[1] * area / dot3(e[0], e[1]);
        ca[0] = area - ca[1] - ca[2];
    } else if (ew[1] <= 0.0) {
        ca[2] = -0.25 * l2[0] * area / dot3(e[1], e[0]);
        ca[0] = -0.25 * l2[2] * area / dot3(e[1], e[2]);
        ca[1] = area - ca[2] - ca[0];
    } else if (ew[2] <= 0.0) {
        ca[0] = -0.25 * l2[1] * area / dot3(e[2], e[1]);
        ca[1] = -0.25 * l2[0] * area / dot3(e[2], e[0]);
        ca[2] = area - ca[0] - ca[1];
    } else {
        const double ewscale = 0.5 * area / (ew[0] + ew[1] + ew[2]);
        for (int j = 0; j < 3; j++) ca[j] = ewscale * (ew[(j+1)%3] + ew[(j+2)%3]);
    }
}

void orc_pointareas(const float *verts, int64_t nv, const uint32_t *faces, int64_t nf,
                    double *areas /* (nv,) */) {
    memset(areas, 0, sizeof(double) * (size_t)nv);
    for (int64_t i = 0; i < nf; i++) {
        const uint32_t f[3] = { faces[3*i], faces[3*i+1], faces[3*i+2] };
        double e[3][3], ca[3];
        for (int d = 0; d < 3; d++) {
            e[0][d] = (double)verts[3*f[2]+d] - (double)verts[3*f[1]+d];
            e[1][d] = (double)verts[3*f[0]+d] - (double)verts[3*f[2]+d];
            e[2][d] = (double)verts[3*f[1]+d] - (double)verts[3*f[0]+d];
        }
        orc_corner_areas(e, ca);
        areas[f[0]] += ca[0]; areas[f[1]] += ca[1]; areas[f[2]] += ca[2];
    }
}

/* ------------------------------------------------------------------------- */
/* A8-A10: TriMesh.compute_curvatures  memsurfer/trimesh.py:284-316            */
/*   mesh assembly + half-box face filter: memsurfer/trimesh.py:441-500        */
/*   vtkCurvatures (VTK >= 9.1; NOT in the tree) mean = dihedral-angle,        */
/*   Gauss = angle deficit: restated from SURVEY.md Appendix A.3 [recalled].   */
/*   PARITY UNPINNED (no VTK here).                                            */
/* verts/faces are the concatenated periodic arrays; xw,yw = half box widths   */
/* (pass < 0 to disable the filter = non-periodic mesh).  Outputs (nv,).       */
/* ------------------------------------------------------------------------- */
static void tri_normal_dir(const double *v1, const double *v2, const double *v3, double *n) {
    /* vtkTriangle::ComputeNormalDirection: (v3-v2) x (v1-v2) */
    double a[3] = { v3[0]-v2[0], v3[1]-v2[1], v3[2]-v2[2] };
    double b[3] = { v1[0]-v2[0], v1[1]-v2[1], v1[2]-v2[2] };
    cross3(a, b, n);
}
static void tri_unit_normal(const double *v1, const double *v2, const double *v3, double *n) {
    tri_normal_dir(v1, v2, v3, n);
    const double l = sqrt(dot3(n, n));
    if (l != 0.0) { n[0] /= l; n[1] /= l; n[2] /= l; }
}
static double tri_area(const double *v1, const double *v2, const double *v3) {
    double n[3]; tri_normal_dir(v1, v2, v3, n);
    return 0.5 * sqrt(dot3(n, n));
}
static double angle_between(const double *u, const double *v) {
    double c[3]; cross3(u, v, c);
    return atan2(sqrt(dot3(c, c)), dot3(u, v));
}

void orc_face_keep(const float *verts, const uint32_t *faces, int64_t nf, double xw, double yw,
                   uint8_t *keep) {
    for (int64_t f = 0; f < nf; f++) {
        int per = 0;
        if (xw >= 0.0) {
            for (int i = 0; i < 3; i++) {
                const float *p = verts + 3 * (size_t)faces[3*f+i];
                const float *q = verts + 3 * (size_t)faces[3*f+(i+1)%3];
                per |= (fabs((double)p[0] - (double)q[0]) > xw) || (fabs((double)p[1] - (double)q[1]) > yw);
            }
        }
        keep[f] = !per;
    }
}

int orc_curvatures(const float *verts, int64_t nv, const uint32_t *faces, int64_t nf,
                   double xw, double yw, double *H, double *K) {
    uint8_t *keep = (uint8_t*)malloc((size_t)nf + 1);
    int64_t *off = (int64_t*)calloc((size_t)nv + 1, sizeof(int64_t));
    int32_t *cnt = (int32_t*)calloc((size_t)nv + 1, sizeof(int32_t));
    double *dA = (double*)calloc((size_t)nv + 1, sizeof(double));
    if (!keep || !off || !cnt || !dA) return 1;
    orc_face_keep(verts, faces, nf, xw, yw, keep);
    /* vertex -> incident kept faces (ascending face index) */
    for (int64_t f = 0; f < nf; f++) if (keep[f]) for (int i = 0; i < 3; i++) off[faces[3*f+i] + 1]++;
    for (int64_t v = 0; v < nv; v++) off[v+1] += off[v];
    int64_t *adj = (int64_t*)malloc(sizeof(int64_t) * (size_t)(off[nv] + 1));
    int64_t *cur = (int64_t*)malloc(sizeof(int64_t) * (size_t)(nv + 1));
    memcpy(cur, off, sizeof(int64_t) * (size_t)nv);
    for (int64_t f = 0; f < nf; f++) if (keep[f]) for (int i = 0; i < 3; i++) adj[cur[faces[3*f+i]]++] = f;

    for (int64_t v = 0; v < nv; v++) { H[v] = 0.0; K[v] = 2.0 * M_PI; }

#define P(idx, arr) do { const float *q_ = verts + 3*(size_t)(idx); arr[0]=q_[0]; arr[1]=q_[1]; arr[2]=q_[2]; } while (0)
    for (int64_t f = 0; f < nf; f++) {
        if (!keep[f]) continue;
        const uint32_t *t = faces + 3*f;
        /* ---- mean (A9) ---- */
        for (int v = 0; v < 3; v++) {
            const uint32_t vl = t[v], vr = t[(v+1)%3], vo = t[(v+2)%3];
            int64_t nb = -1; int nnb = 0;
            for (int64_t j = off[vl]; j < off[vl+1]; j++) {
                const int64_t g = adj[j];
                if (g == f) continue;
                const uint32_t *u = faces + 3*g;
                if (u[0] == vr || u[1] == vr || u[2] == vr) { nb = g; nnb++; }
            }
            if (nnb != 1 || nb <= f) continue;
            double ore[3], end[3], oth[3], nf_[3], nn_[3], e[3], a0[3], a1[3], a2[3], cr[3];
            P(vl, ore); P(vr, end); P(vo, oth);
            tri_unit_normal(ore, end, oth, nf_);
            const double Af0 = tri_area(ore, end, oth);
            const uint32_t *u = faces + 3*nb;
            P(u[0], a0); P(u[1], a1); P(u[2], a2);
            tri_unit_normal(a0, a1, a2, nn_);
            const double Af = Af0 + tri_area(a0, a1, a2);
            e[0] = end[0]-ore[0]; e[1] = end[1]-ore[1]; e[2] = end[2]-ore[2];
            const double len = sqrt(dot3(e, e));
            if (len != 0.0) { e[0] /= len; e[1] /= len; e[2] /= len; }
            const double cs = dot3(nf_, nn_);
            cross3(nf_, nn_, cr);
            const double sn = dot3(cr, e);
            double Hf = (sn != 0.0 || cs != 0.0) ? len * atan2(sn, cs) : 0.0;
            if (Af != 0.0) Hf = Hf / Af * 3.0;
            H[vl] += Hf; H[vr] += Hf; cnt[vl]++; cnt[vr]++;
        }
        /* ---- Gauss (A10) ---- */
        {
            double v0[3], v1[3], v2[3], e0[3], e1[3], e2[3];
            P(t[0], v0); P(t[1], v1); P(t[2], v2);
            for (int d = 0; d < 3; d++) { e0[d] = v1[d]-v0[d]; e1[d] = v2[d]-v1[d]; e2[d] = v0[d]-v2[d]; }
            const double al0 = M_PI - angle_between(e1, e2);
            const double al1 = M_PI - angle_between(e2, e0);
            const double al2 = M_PI - angle_between(e0, e1);
            const double A = tri_area(v0, v1, v2);
            dA[t[0]] += A; dA[t[1]] += A; dA[t[2]] += A;
            K[t[0]] -= al1; K[t[1]] -= al2; K[t[2]] -= al0;
        }
    }
#undef P
    for (int64_t v = 0; v < nv; v++) {
        H[v] = cnt[v] > 0 ? 0.5 * H[v] / (double)cnt[v] : 0.0;
        K[v] = dA[v] > 0.0 ? 3.0 * K[v] / dA[v] : 0.0;
    }
    free(keep); free(off); free(cnt); free(dA); free(adj); free(cur);
    return 0;
}

/* ------------------------------------------------------------------------- */
/* A13-A16: kde_2d / kde_3d + TriMesh::kde normalisation                       */
/*   memsurfer/src/TriMesh_kde.cpp:344-392 (pair loops), :504-524 (norms)      */
/*   GaussianKernel2D/3D   memsurfer/src/DensityKernels.hpp:44-84              */
/*   DistanceSquared / DistancePeriodicXYSquared  DistanceKernels.hpp:75-136   */
/* type 2 = 2-D kernel on (x,y); type 3 = 3-D kernel on (x,y,z), z never       */
/* wrapped.  Constants efactor, sfactor and boxw are ROUNDED THROUGH float32   */
/* exactly as the reference stores them (DensityKernels.hpp:47-48,             */
/* DistanceKernels.hpp:84,101); all pair arithmetic and sums are fp64.         */
/* bbox: 6 floats {b0x,b0y,b0z,b1x,b1y,b1z} (trimesh.py:346-347 passes bboxl). */
/* ids: source vertex ids (nids == 0 -> all vertices).  Targets: all vertices. */
/* ------------------------------------------------------------------------- */
void orc_kde_constants(float sigma, int type, float *efactor, float *sfactor) {
    const double s = (double)sigma;
    *efactor = (float)(-1.0 / (2.0 * s * s));
    if (type == 3) *sfactor = (float)(1.0 / pow(sqrt(2.0 * M_PI) * s, 3.0));
    else           *sfactor = (float)(1.0 / (2.0 * M_PI * s * s));
}

static inline double orc_dist2(const float *a, const float *b, int type, int periodic, const double *boxw) {
    double ax = a[0], ay = a[1], bx = b[0], by = b[1];
    if (periodic) {
        if (fabs(ax - bx) >= 0.5 * boxw[0]) { if (ax < bx) ax += boxw[0]; else bx += boxw[0]; }
        if (fabs(ay - by) >= 0.5 * boxw[1]) { if (ay < by) ay += boxw[1]; else by += boxw[1]; }
    }
    double d = (ax - bx) * (ax - bx) + (ay - by) * (ay - by);
    if (type == 3) { const double dz = (double)a[2] - (double)b[2]; d += dz * dz; }
    return d;
}

/* raw (un-normalised) pair sums: density[j] = sum_i sfactor*exp(d2*efactor) */
void orc_kde_raw(const float *verts, int64_t nv, const int32_t *ids, int64_t nids,
                 int type, int periodic, const float *bbox, float sigma, double *density) {
    float ef, sf; orc_kde_constants(sigma, type, &ef, &sf);
    const double efactor = ef, sfactor = sf;
    double boxw[2] = { 0.0, 0.0 };
    if (periodic) { boxw[0] = (double)(float)(bbox[3] - bbox[0]); boxw[1] = (double)(float)(bbox[4] - bbox[1]); }
    const int64_t ns = nids > 0 ? nids : nv;
#pragma omp parallel for schedule(static)
    for (int64_t j = 0; j < nv; j++) {
        double acc = 0.0;
        for (int64_t i = 0; i < ns; i++) {
            const int64_t s = nids > 0 ? ids[i] : i;
            acc += sfactor * exp(orc_dist2(verts + 3*s, verts + 3*j, type, periodic, boxw) * efactor);
        }
        density[j] = acc;
    }
}

void orc_kde(const float *verts, int64_t nv, const int32_t *ids, int64_t nids,
             int type, int periodic, const float *bbox, float sigma, int get_counts,
             double *density) {
    orc_kde_raw(verts, nv, ids, nids, type, periodic, bbox, sigma, density);
    const double norm1 = 1.0 / (double)nv;                 /* TriMesh_kde.cpp:504-512 */
    for (int64_t j = 0; j < nv; j++) density[j] *= norm1;
    if (get_counts) {                                      /* TriMesh_kde.cpp:515-524 */
        double dsm = 0.0;
        for (int64_t j = 0; j < nv; j++) dsm += density[j];
        const double norm2 = (double)(nids > 0 ? nids : nv) / dsm;
        for (int64_t j = 0; j < nv; j++) density[j] *= norm2;
    }
}

/* ------------------------------------------------------------------------- */
/* A3 (search part): (k+1)-nearest neighbours incl. self, sorted by            */
/* (squared distance, index).  CGAL 4.13 Orthogonal_k_neighbor_search is       */
/* exact (eps = 0) and NOT in the tree [recalled, SURVEY.md Appendix A.1];     */
/* call site memsurfer/src/PointSet.cpp:163-165.  Squared distance is formed   */
/* in double from the float32-valued coordinates, accumulated x, y, z, no FMA. */
/* Exact-tie order between distinct points is not pinned by CGAL; here ties    */
/* break by ascending index.  Brute force O(M^2) -- small cases only.          */
/* ------------------------------------------------------------------------- */
static inline double orc_d2(const float *p, const float *q) {
    const double dx = (double)p[0] - (double)q[0];
    const double dy = (double)p[1] - (double)q[1];
    const double dz = (double)p[2] - (double)q[2];
    double d = 0.0; d += dx * dx; d += dy * dy; d += dz * dz;
    return d;
}

static inline void orc_topk_insert(double *bd, int32_t *bi, int k1, int *cnt, double d, int32_t idx) {
    /* keep bd/bi sorted ascending by (d, idx); capacity k1 */
    int n = *cnt;
    if (n == k1) {
        if (d > bd[n-1] || (d == bd[n-1] && idx > bi[n-1])) return;
        n--;
    }
    int p = n;
    while (p > 0 && (bd[p-1] > d || (bd[p-1] == d && bi[p-1] > idx))) { bd[p] = bd[p-1]; bi[p] = bi[p-1]; p--; }
    bd[p] = d; bi[p] = idx;
    *cnt = n + 1;
}

void orc_knn_brute(const float *pts, int64_t m, int k1, int32_t *nbr /* (m,k1) */, double *nbr_d2 /* (m,k1) or NULL */) {
#pragma omp parallel for schedule(static)
    for (int64_t q = 0; q < m; q++) {
        double bd[64]; int32_t bi[64]; int cnt = 0;
        for (int64_t i = 0; i < m; i++)
            orc_topk_insert(bd, bi, k1, &cnt, orc_d2(pts + 3*q, pts + 3*i), (int32_t)i);
        for (int j = 0; j < k1; j++) {
            nbr[q*k1 + j] = j < cnt ? bi[j] : -1;
            if (nbr_d2) nbr_d2[q*k1 + j] = j < cnt ? bd[j] : -1.0;
        }
    }
}

/* Same result through a uniform xy grid (exact: a ring is accepted only when  */
/* the k1-th best distance is <= the distance to the unexplored region).       */
/* Used for larger parity cases and as the timed CPU restatement.              */
void orc_knn_grid(const float *pts, int64_t m, int k1, int32_t *nbr, double *nbr_d2) {
    float lo[2] = { pts[0], pts[1] }, hi[2] = { pts[0], pts[1] };
    for (int64_t i = 0; i < m; i++) for (int d = 0; d < 2; d++) {
        if (pts[3*i+d] < lo[d]) lo[d] = pts[3*i+d];
        if (pts[3*i+d] > hi[d]) hi[d] = pts[3*i+d];
    }
    const double wx = (double)hi[0] - lo[0], wy = (double)hi[1] - lo[1];
    double h = sqrt((wx > 0 ? wx : 1.0) * (wy > 0 ? wy : 1.0) * 4.0 / (double)m);   /* ~4 pts / cell */
    if (!(h > 0.0)) h = 1.0;
    const int64_t nx = (int64_t)(wx / h) + 1, ny = (int64_t)(wy / h) + 1;
    int64_t *start = (int64_t*)calloc((size_t)(nx*ny + 1), sizeof(int64_t));
    int32_t *order = (int32_t*)malloc(sizeof(int32_t) * (size_t)m);
    int64_t *cell = (int64_t*)malloc(sizeof(int64_t) * (size_t)m);
    for (int64_t i = 0; i < m; i++) {
        int64_t cx = (int64_t)(((double)pts[3*i] - lo[0]) / h), cy = (int64_t)(((double)pts[3*i+1] - lo[1]) / h);
        if (cx >= nx) cx = nx - 1; if (cy >= ny) cy = ny - 1;
        cell[i] = cy * nx + cx; start[cell[i] + 1]++;
    }
    for (int64_t c = 0; c < nx*ny; c++) start[c+1] += start[c];
    int64_t *cur = (int64_t*)malloc(sizeof(int64_t) * (size_t)(nx*ny + 1));
    memcpy(cur, start, sizeof(int64_t) * (size_t)(nx*ny));
    for (int64_t i = 0; i < m; i++) order[cur[cell[i]]++] = (int32_t)i;
#pragma omp parallel for schedule(dynamic, 256)
    for (int64_t q = 0; q < m; q++) {
        double bd[64]; int32_t bi[64]; int cnt = 0;
        const double qx = (double)pts[3*q] - lo[0], qy = (double)pts[3*q+1] - lo[1];
        const int64_t cx = cell[q] % nx, cy = cell[q] / nx;
        for (int64_t r = 0; ; r++) {
            for (int64_t yy = cy - r; yy <= cy + r; yy++) {
                if (yy < 0 || yy >= ny) continue;
                const int64_t step = (yy == cy - r || yy == cy + r) ? 1 : 2 * r;
                for (int64_t xx = cx - r; xx <= cx + r; xx += (step > 0 ? step : 1)) {
                    if (xx < 0 || xx >= nx) continue;
                    for (int64_t s = start[yy*nx + xx]; s < start[yy*nx + xx + 1]; s++)
                        orc_topk_insert(bd, bi, k1, &cnt, orc_d2(pts + 3*q, pts + 3*(int64_t)order[s]), order[s]);
                }
            }
            /* distance from q to the boundary of the explored window */
            double bx0 = cx - r > 0 ? qx - (double)(cx - r) * h : 1e300, bx1 = cx + r < nx - 1 ? (double)(cx + r + 1) * h - qx : 1e300;
            double by0 = cy - r > 0 ? qy - (double)(cy - r) * h : 1e300, by1 = cy + r < ny - 1 ? (double)(cy + r + 1) * h - qy : 1e300;
            double bmin = bx0; if (bx1 < bmin) bmin = bx1; if (by0 < bmin) bmin = by0; if (by1 < bmin) bmin = by1;
            const int covers_all = (cx - r <= 0 && cy - r <= 0 && cx + r >= nx - 1 && cy + r >= ny - 1);
            if (covers_all) break;
            if (cnt == k1 && bmin > 0.0 && sqrt(bd[k1-1]) < bmin * (1.0 - 1e-9)) break;
        }
        for (int j = 0; j < k1; j++) {
            nbr[q*k1 + j] = j < cnt ? bi[j] : -1;
            if (nbr_d2) nbr_d2[q*k1 + j] = j < cnt ? bd[j] : -1.0;
        }
    }
    free(start); free(order); free(cell); free(cur);
}

/* ------------------------------------------------------------------------- */
/* A3 (fit part): CGAL::pca_estimate_normals  [recalled, Appendix A.1]         */
/* centroid of the k1 neighbours, covariance sum (p-c)(p-c)^T (not divided),   */
/* symmetric 3x3 eigen-decomposition by cyclic Jacobi (CGAL's internal solver  */
/* when CGAL_EIGEN3_ENABLED is not defined; setup.py:217), normal = unit       */
/* eigenvector of the smallest eigenvalue; all eigenvalues equal -> (0,0,1).   */
/* Sign is arbitrary here (fixed by orc_mst_orient).                          */
/* ------------------------------------------------------------------------- */
void orc_eig3_smallest(const double c[6] /* xx,xy,xz,yy,yz,zz */, double n[3], double ev[3]) {
    double a[3][3] = { { c[0], c[1], c[2] }, { c[1], c[3], c[4] }, { c[2], c[4], c[5] } };
    double v[3][3] = { { 1, 0, 0 }, { 0, 1, 0 }, { 0, 0, 1 } };
    for (int sweep = 0; sweep < 64; sweep++) {
        const double offn = a[0][1]*a[0][1] + a[0][2]*a[0][2] + a[1][2]*a[1][2];
        if (offn == 0.0) break;
        for (int p = 0; p < 2; p++) for (int q = p + 1; q < 3; q++) {
            if (a[p][q] == 0.0) continue;
            const double theta = (a[q][q] - a[p][p]) / (2.0 * a[p][q]);
            const double t = (theta >= 0.0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
            const double cs = 1.0 / sqrt(t * t + 1.0), sn = t * cs;
            const double apq = a[p][q];
            a[p][p] -= t * apq; a[q][q] += t * apq; a[p][q] = a[q][p] = 0.0;
            const int r = 3 - p - q;
            const double arp = a[r][p], arq = a[r][q];
            a[r][p] = a[p][r] = cs * arp - sn * arq;
            a[r][q] = a[q][r] = sn * arp + cs * arq;
            for (int i = 0; i < 3; i++) {
                const double vip = v[i][p], viq = v[i][q];
                v[i][p] = cs * vip - sn * viq;
                v[i][q] = sn * vip + cs * viq;
            }
        }
    }
    int m = 0;
    if (a[1][1] < a[m][m]) m = 1;
    if (a[2][2] < a[m][m]) m = 2;
    ev[0] = a[0][0]; ev[1] = a[1][1]; ev[2] = a[2][2];
    if (a[0][0] == a[1][1] && a[1][1] == a[2][2]) { n[0] = 0; n[1] = 0; n[2] = 1; return; }
    const double l = sqrt(v[0][m]*v[0][m] + v[1][m]*v[1][m] + v[2][m]*v[2][m]);
    n[0] = v[0][m] / l; n[1] = v[1][m] / l; n[2] = v[2][m] / l;
}

void orc_pca_normals(const float *pts, int64_t m, const int32_t *nbr, int k1, double *normals /* (m,3) */) {
#pragma omp parallel for schedule(static)
    for (int64_t q = 0; q < m; q++) {
        double c[3] = { 0, 0, 0 }, cov[6] = { 0, 0, 0, 0, 0, 0 }, ev[3];
        int cnt = 0;
        for (int j = 0; j < k1; j++) { const int32_t i = nbr[q*k1 + j]; if (i < 0) continue;
            c[0] += pts[3*(int64_t)i]; c[1] += pts[3*(int64_t)i+1]; c[2] += pts[3*(int64_t)i+2]; cnt++; }
        c[0] /= cnt; c[1] /= cnt; c[2] /= cnt;
        for (int j = 0; j < k1; j++) { const int32_t i = nbr[q*k1 + j]; if (i < 0) continue;
            const double dx = pts[3*(int64_t)i] - c[0], dy = pts[3*(int64_t)i+1] - c[1], dz = pts[3*(int64_t)i+2] - c[2];
            cov[0] += dx*dx; cov[1] += dx*dy; cov[2] += dx*dz; cov[3] += dy*dy; cov[4] += dy*dz; cov[5] += dz*dz; }
        orc_eig3_smallest(cov, normals + 3*q, ev);
    }
}

/* ------------------------------------------------------------------------- */
/* A4: CGAL::mst_orient_normals  [recalled, Appendix A.2]; call site           */
/* memsurfer/src/PointSet.cpp:171-174.  Source = first max-z point, flipped to */
/* n_z >= 0.  Riemannian graph: edge {a<b} iff b in (k+1)-NN(a), weight        */
/* float(max(0, 1-|na.nb|)).  Minimum spanning tree of the source's component, */
/* then parent->child propagation: child flipped iff dot(parent, child) < 0.   */
/* The MST is unique when weights are distinct; for exact float ties boost's   */
/* Prim order is not pinned -- here ties break by the total order (w, a, slot of b in NN(a))    */
/* (Kruskal), which the CUDA Boruvka path uses too.  normals updated in place; */
/* oriented[i] = 1 for points reached from the source.  Returns #unoriented.   */
/* ------------------------------------------------------------------------- */
typedef struct { float w; int32_t a, b; int64_t eid; } orc_edge;
static int orc_edge_cmp(const void *x, const void *y) {
    const orc_edge *p = (const orc_edge*)x, *q = (const orc_edge*)y;
    if (p->w != q->w) return p->w < q->w ? -1 : 1;
    return p->eid < q->eid ? -1 : (p->eid > q->eid);   /* eid = a*k1 + slot of b in a's list */
}
static int32_t orc_find(int32_t *par, int32_t x) {
    while (par[x] != x) { par[x] = par[par[x]]; x = par[x]; }
    return x;
}

int64_t orc_mst_orient(const float *pts, int64_t m, const int32_t *nbr, int k1,
                       double *normals, uint8_t *oriented) {
    int64_t ne = 0;
    orc_edge *E = (orc_edge*)malloc(sizeof(orc_edge) * (size_t)(m * k1 + 1));
    for (int64_t a = 0; a < m; a++) for (int j = 0; j < k1; j++) {
        const int32_t b = nbr[a*k1 + j];
        if (b <= a) continue;
        const double *na = normals + 3*a, *nb = normals + 3*(int64_t)b;
        double dot = na[0]*nb[0] + na[1]*nb[1] + na[2]*nb[2];
        double w = 1.0 - fabs(dot); if (w < 0.0) w = 0.0;
        E[ne].w = (float)w; E[ne].a = (int32_t)a; E[ne].b = b; E[ne].eid = a * k1 + j; ne++;
    }
    qsort(E, (size_t)ne, sizeof(orc_edge), orc_edge_cmp);
    int32_t *par = (int32_t*)malloc(sizeof(int32_t) * (size_t)m);
    for (int64_t i = 0; i < m; i++) par[i] = (int32_t)i;
    /* tree adjacency as edge lists */
    int32_t *deg = (int32_t*)calloc((size_t)m + 1, sizeof(int32_t));
    int64_t nt = 0;
    orc_edge *T = (orc_edge*)malloc(sizeof(orc_edge) * (size_t)(m + 1));
    for (int64_t e = 0; e < ne; e++) {
        const int32_t ra = orc_find(par, E[e].a), rb = orc_find(par, E[e].b);
        if (ra == rb) continue;
        par[ra] = rb; T[nt++] = E[e]; deg[E[e].a]++; deg[E[e].b]++;
    }
    int64_t *off = (int64_t*)calloc((size_t)m + 1, sizeof(int64_t));
    for (int64_t i = 0; i < m; i++) off[i+1] = off[i] + deg[i];
    int32_t *adj = (int32_t*)malloc(sizeof(int32_t) * (size_t)(2 * nt + 1));
    memset(deg, 0, sizeof(int32_t) * (size_t)m);
    for (int64_t e = 0; e < nt; e++) {
        adj[off[T[e].a] + deg[T[e].a]++] = T[e].b;
        adj[off[T[e].b] + deg[T[e].b]++] = T[e].a;
    }
    /* source */
    int64_t src = 0;
    for (int64_t i = 1; i < m; i++) if (pts[3*i+2] > pts[3*src+2]) src = i;
    memset(oriented, 0, (size_t)m);
    if (normals[3*src+2] < 0.0) { normals[3*src] = -normals[3*src]; normals[3*src+1] = -normals[3*src+1]; normals[3*src+2] = -normals[3*src+2]; }
    int32_t *queue = (int32_t*)malloc(sizeof(int32_t) * (size_t)m);
    int64_t qh = 0, qt = 0; queue[qt++] = (int32_t)src; oriented[src] = 1;
    while (qh < qt) {
        const int32_t s = queue[qh++];
        for (int64_t j = off[s]; j < off[s+1]; j++) {
            const int32_t t = adj[j];
            if (oriented[t]) continue;
            double *ns = normals + 3*(int64_t)s, *ntv = normals + 3*(int64_t)t;
            const double d = ns[0]*ntv[0] + ns[1]*ntv[1] + ns[2]*ntv[2];
            if (d < 0.0) { ntv[0] = -ntv[0]; ntv[1] = -ntv[1]; ntv[2] = -ntv[2]; }
            oriented[t] = 1; queue[qt++] = t;
        }
    }
    free(E); free(par); free(deg); free(T); free(off); free(adj); free(queue);
    return m - qt;
}

/* raw pair sums for a SUBSET of targets (parity checks at sizes where all V^2 pairs are too slow on a CPU) */
void orc_kde_raw_targets(const float *verts, int64_t nv, const int32_t *ids, int64_t nids,
                         const int32_t *targets, int64_t nt,
                         int type, int periodic, const float *bbox, float sigma, double *density /* (nt,) */) {
    float ef, sf; orc_kde_constants(sigma, type, &ef, &sf);
    const double efactor = ef, sfactor = sf;
    double boxw[2] = { 0.0, 0.0 };
    if (periodic) { boxw[0] = (double)(float)(bbox[3] - bbox[0]); boxw[1] = (double)(float)(bbox[4] - bbox[1]); }
    const int64_t ns = nids > 0 ? nids : nv;
#pragma omp parallel for schedule(static)
    for (int64_t t = 0; t < nt; t++) {
        const int64_t j = targets[t];
        double acc = 0.0;
        for (int64_t i = 0; i < ns; i++) {
            const int64_t s = nids > 0 ? ids[i] : i;
            acc += sfactor * exp(orc_dist2(verts + 3*s, verts + 3*j, type, periodic, boxw) * efactor);
        }
        density[t] = acc;
    }
}

/* ------------------------------------------------------------------ closest point on a mesh
 * TriMesh::project_on_surface (memsurfer/src/TriMesh_cgal.cpp:295-363) asks CGAL's AABB tree for the
 * closest point of the mesh's own faces to every query and converts it to barycentric coordinates with
 * TriMesh::Point2Bary (memsurfer/src/TriMesh.cpp:28-42); TriMesh::distance_to_other_mesh
 * (TriMesh_cgal.cpp:390-425) turns that into a distance.  CGAL is not in the tree: the closest point is
 * restated as a sequential scan of all faces with the standard region test (Ericson 5.1.5), fp64, the
 * first (smallest id) face winning ties.                                                              */
static void orc_closest_on_triangle(const double *p, const double *a, const double *b, const double *c, double *o) {
    double ab[3], ac[3], ap[3], bp[3], cp[3];
    for (int k = 0; k < 3; k++) { ab[k] = b[k] - a[k]; ac[k] = c[k] - a[k]; ap[k] = p[k] - a[k]; }
    const double d1 = dot3(ab, ap), d2 = dot3(ac, ap);
    if (d1 <= 0.0 && d2 <= 0.0) { for (int k = 0; k < 3; k++) o[k] = a[k]; return; }
    for (int k = 0; k < 3; k++) bp[k] = p[k] - b[k];
    const double d3 = dot3(ab, bp), d4 = dot3(ac, bp);
    if (d3 >= 0.0 && d4 <= d3) { for (int k = 0; k < 3; k++) o[k] = b[k]; return; }
    const double vc = d1 * d4 - d3 * d2;
    if (vc <= 0.0 && d1 >= 0.0 && d3 <= 0.0) {
        const double v = d1 / (d1 - d3);
        for (int k = 0; k < 3; k++) o[k] = a[k] + v * ab[k];
        return;
    }
    for (int k = 0; k < 3; k++) cp[k] = p[k] - c[k];
    const double d5 = dot3(ab, cp), d6 = dot3(ac, cp);
    if (d6 >= 0.0 && d5 <= d6) { for (int k = 0; k < 3; k++) o[k] = c[k]; return; }
    const double vb = d5 * d2 - d1 * d6;
    if (vb <= 0.0 && d2 >= 0.0 && d6 <= 0.0) {
        const double w = d2 / (d2 - d6);
        for (int k = 0; k < 3; k++) o[k] = a[k] + w * ac[k];
        return;
    }
    const double va = d3 * d6 - d5 * d4;
    if (va <= 0.0 && (d4 - d3) >= 0.0 && (d5 - d6) >= 0.0) {
        const double w = (d4 - d3) / ((d4 - d3) + (d5 - d6));
        for (int k = 0; k < 3; k++) o[k] = b[k] + w * (c[k] - b[k]);
        return;
    }
    const double denom = 1.0 / (va + vb + vc);
    const double v = vb * denom, w = vc * denom;
    for (int k = 0; k < 3; k++) o[k] = a[k] + ab[k] * v + ac[k] * w;
}

void orc_closest_point(const float *verts, int64_t nv, const uint32_t *faces, int64_t nf, const float *queries, int64_t nq,
                       int32_t *face_id, double *bary, double *dist) {
    (void)nv;
    for (int64_t i = 0; i < nq; i++) {
        const double p[3] = { queries[3*i], queries[3*i+1], queries[3*i+2] };
        double best = 1e300, bc[3] = { 0, 0, 0 }; int64_t bf = -1;
        for (int64_t f = 0; f < nf; f++) {
            double a[3], b[3], c[3], q[3];
            for (int k = 0; k < 3; k++) { a[k] = verts[3*(int64_t)faces[3*f] + k]; b[k] = verts[3*(int64_t)faces[3*f+1] + k]; c[k] = verts[3*(int64_t)faces[3*f+2] + k]; }
            orc_closest_on_triangle(p, a, b, c, q);
            const double dq[3] = { p[0] - q[0], p[1] - q[1], p[2] - q[2] };
            const double dd = dot3(dq, dq);
            if (dd < best) { best = dd; bf = f; bc[0] = q[0]; bc[1] = q[1]; bc[2] = q[2]; }
        }
        double u = 0.0, v = 0.0, w = 0.0;
        if (bf >= 0) {          /* TriMesh::Point2Bary, TriMesh.cpp:28-42 */
            double a[3], v0[3], v1[3], v2[3];
            for (int k = 0; k < 3; k++) {
                a[k] = verts[3*(int64_t)faces[3*bf] + k];
                v0[k] = (double)verts[3*(int64_t)faces[3*bf+1] + k] - a[k]; v1[k] = (double)verts[3*(int64_t)faces[3*bf+2] + k] - a[k]; v2[k] = bc[k] - a[k];
            }
            const double d00 = dot3(v0, v0), d01 = dot3(v0, v1), d11 = dot3(v1, v1), d20 = dot3(v2, v0), d21 = dot3(v2, v1);
            const double denom = d00 * d11 - d01 * d01;
            v = (d11 * d20 - d01 * d21) / denom;
            w = (d00 * d21 - d01 * d20) / denom;
            u = 1.0 - v - w;
        }
        if (face_id) face_id[i] = (int32_t)bf;
        if (bary) { bary[3*i] = u; bary[3*i+1] = v; bary[3*i+2] = w; }
        if (dist) dist[i] = bf >= 0 ? sqrt(best) : 0.0;
    }
}

/* ---------------------------------------------------------------- vertex neighbours / shells
 * orc_need_neighbors follows TriMesh::need_neighbors (memsurfer/src/TriMesh.cpp:197-255): the sequential face
 * loop appends the two other vertices of corner j of face i to the list of faces[i][j] unless present.
 * off (nv+1) offsets and lists (cap entries); returns the total length or -1 when cap is too small.       */
int64_t orc_need_neighbors(const uint32_t *faces, int64_t nf, int64_t nv, int32_t *off, int32_t *lists, int64_t cap) {
    int32_t *cnt = (int32_t*)calloc((size_t)nv + 1, sizeof(int32_t));
    int32_t *room = (int32_t*)calloc((size_t)nv + 1, sizeof(int32_t));
    for (int64_t i = 0; i < 3 * nf; i++) if ((int64_t)faces[i] < nv) room[faces[i]] += 2;
    int64_t *start = (int64_t*)malloc(((size_t)nv + 1) * sizeof(int64_t));
    start[0] = 0;
    for (int64_t v = 0; v < nv; v++) start[v + 1] = start[v] + room[v];
    int32_t *tmp = (int32_t*)malloc(((size_t)start[nv] + 1) * sizeof(int32_t));
    for (int64_t i = 0; i < nf; i++)
        for (int j = 0; j < 3; j++) {
            const int64_t me = faces[3*i + j];
            if (me >= nv) continue;
            int32_t *lst = tmp + start[me];
            for (int k = 1; k <= 2; k++) {
                const int32_t u = (int32_t)faces[3*i + (j + k) % 3];
                int seen = 0;
                for (int q = 0; q < cnt[me]; q++) if (lst[q] == u) { seen = 1; break; }
                if (!seen) lst[cnt[me]++] = u;
            }
        }
    int64_t tot = 0;
    for (int64_t v = 0; v < nv; v++) { off[v] = (int32_t)tot; tot += cnt[v]; }
    off[nv] = (int32_t)tot;
    if (tot <= cap)
        for (int64_t v = 0; v < nv; v++) memcpy(lists + off[v], tmp + start[v], sizeof(int32_t) * (size_t)cnt[v]);
    free(cnt); free(room); free(start); free(tmp);
    return tot <= cap ? tot : -1;
}

/* orc_shells follows the FIFO breadth-first loop of TriMesh.compute_shells (memsurfer/trimesh.py:392-404):
 * reference vertices start at 0 and enter the queue in the given order; an unvisited neighbour gets
 * 1 + the shell of the vertex that reaches it first.  Unreached vertices stay at -1.                      */
void orc_shells(const int32_t *off, const int32_t *lists, int64_t nv, const int32_t *ref, int64_t nref, int32_t *shells) {
    int32_t *queue = (int32_t*)malloc(((size_t)nv + (size_t)nref + 1) * sizeof(int32_t));
    int64_t head = 0, tail = 0;
    for (int64_t v = 0; v < nv; v++) shells[v] = -1;
    for (int64_t i = 0; i < nref; i++) shells[ref[i]] = 0;
    for (int64_t i = 0; i < nref; i++) queue[tail++] = ref[i];
    while (head < tail) {
        const int32_t cur = queue[head++];
        for (int32_t j = off[cur]; j < off[cur + 1]; j++) {
            const int32_t u = lists[j];
            if (shells[u] == -1) { shells[u] = 1 + shells[cur]; queue[tail++] = u; }
        }
    }
    free(queue);
}

/* ---------------------------------------------------------------- geodesic KDE, type 1
 * orc_geodesics follows TriMesh::compute_geodesics_fw (memsurfer/src/TriMesh_kde.cpp:196-332, the non-PDIST
 * branch): D = "infinity" (FLT_MAX there, DBL_MAX here), 0 on the diagonal; every face edge gets the square
 * root of the (periodic-xy, 3-D) squared distance of its end points; then Floyd-Warshall in k, i, j order.
 * The symmetric write of :287 is redundant (both candidates are the same two addends).  fp64 throughout.   */
void orc_geodesics(const float *verts, int64_t nv, const uint32_t *faces, int64_t nf, int periodic, const float *bbox, double *D) {
    double boxw[2] = { 0.0, 0.0 };
    if (periodic) { boxw[0] = (double)(float)(bbox[3] - bbox[0]); boxw[1] = (double)(float)(bbox[4] - bbox[1]); }
    for (int64_t i = 0; i < nv * nv; i++) D[i] = DBL_MAX;
    for (int64_t i = 0; i < nv; i++) D[i * nv + i] = 0.0;
    for (int64_t f = 0; f < nf; f++)
        for (int d = 0; d < 3; d++) {
            const int64_t a = faces[3*f + d], b = faces[3*f + (d == 2 ? 0 : d + 1)];
            const double val = sqrt(orc_dist2(verts + 3*a, verts + 3*b, 3, periodic, boxw));
            D[a * nv + b] = val; D[b * nv + a] = val;
        }
    for (int64_t k = 0; k < nv; k++) {
        const double *Dk = D + k * nv;
#pragma omp parallel for schedule(static)
        for (int64_t i = 0; i < nv; i++) {
            double *Di = D + i * nv;
            const double dik = Di[k];
            if (i == k || dik == DBL_MAX) continue;
            for (int64_t j = 0; j < nv; j++) if (dik + Dk[j] < Di[j]) Di[j] = dik + Dk[j];
        }
    }
}

/* orc_kde_geodesic follows kde_2m (TriMesh_kde.cpp:394-437) + the normalisation of TriMesh::kde (:504-524):
 * all vertices: density[j] = sum_i k(D[i][j]^2); selected ids: density[j] = sum_i k(D[ids[i]][j]) -- the
 * reference does NOT square the distance in that branch (:432), reproduced as is.  k = GaussianKernel2D.   */
void orc_kde_geodesic(const double *D, int64_t nv, const int32_t *ids, int64_t nids, float sigma, int get_counts, double *density) {
    float ef, sf; orc_kde_constants(sigma, 2, &ef, &sf);
    const double efactor = ef, sfactor = sf;
#pragma omp parallel for schedule(static)
    for (int64_t j = 0; j < nv; j++) {
        double acc = 0.0;
        if (nids == 0) for (int64_t i = 0; i < nv; i++) { const double d = D[i * nv + j]; acc += sfactor * exp(d * d * efactor); }
        else for (int64_t i = 0; i < nids; i++) acc += sfactor * exp(D[(int64_t)ids[i] * nv + j] * efactor);
        density[j] = acc;
    }
    const double norm1 = 1.0 / (double)nv;
    for (int64_t j = 0; j < nv; j++) density[j] *= norm1;
    if (get_counts) {
        double dsm = 0.0;
        for (int64_t j = 0; j < nv; j++) dsm += density[j];
        const double norm2 = (double)(nids > 0 ? nids : nv) / dsm;
        for (int64_t j = 0; j < nv; j++) density[j] *= norm2;
    }
}
