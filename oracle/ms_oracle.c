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
