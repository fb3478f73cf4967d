// TEST INFRASTRUCTURE ONLY (oracle/): extern "C" driver around the reference's
// own, unmodified C++ (compiled in place from /root/reference/memsurfer/src by
// oracle/Makefile -> oracle/_ref/libmsref.so).  Gives the float32 "truth" for
// rows A6, A7, A13-A16 of SURVEY.md section 8 and the timed CPU baseline
// (bench.py --impl reference, cpu_baseline.kind = "reference").
// Calls only the reference's public API:
//   TriMesh(float*,n,d)            memsurfer/src/TriMesh.hpp:229
//   set_faces(uint32_t*,n,d)       memsurfer/src/TriMesh.hpp:248
//   need_normals(bool)             memsurfer/src/TriMesh.hpp:346
//   need_pointareas(bool)          memsurfer/src/TriMesh.hpp:379
//   kde(name,type,get_counts,...)  memsurfer/src/TriMesh_kde.cpp:443
// trim_periodicDelaunay (TriMesh.hpp:214) and its input mDelaunayFaces (TriMesh.hpp:104) are private and are only fed by
// the CGAL path, which cannot compile here: open the class for this driver.  The standard
// headers were already included by ref_stub.h, so the macro only touches the reference's own headers.
#define private public
#define protected public
#include "TriMesh.hpp"
#undef private
#undef protected
#include "DensityKernels.hpp"
#include "DistanceKernels.hpp"
#include <cstring>

extern "C" {

// verts: (nv,3) f32, faces: (nf,3) u32 -- for a periodic mesh pass the
// concatenation verts ++ dup_verts, faces ++ trimmed_faces (what the wrapper
// at TriMesh.hpp:356-370 builds) and read back the first nv_orig rows.
int msref_need_normals(const float *verts, int nv, const uint32_t *faces, int nf, float *out_normals) {
    try {
        TriMesh m(const_cast<float*>(verts), nv, 3);
        m.set_faces(const_cast<uint32_t*>(faces), nf, 3);
        std::vector<TypeFunction> r = m.need_normals(false);
        std::memcpy(out_normals, r.data(), sizeof(float)*r.size());
        return 0;
    } catch (const std::exception &e) { std::cerr << e.what(); return 1; }
}

int msref_need_pointareas(const float *verts, int nv, const uint32_t *faces, int nf, float *out_areas) {
    try {
        TriMesh m(const_cast<float*>(verts), nv, 3);
        m.set_faces(const_cast<uint32_t*>(faces), nf, 3);
        const std::vector<TypeFunction> &r = m.need_pointareas(false);
        std::memcpy(out_areas, r.data(), sizeof(float)*r.size());
        return 0;
    } catch (const std::exception &e) { std::cerr << e.what(); return 1; }
}

// TriMesh::need_neighbors() (TriMesh.cpp:197-255): linearised [counts..., lists...]; returns its length
// (-1: out too small, -2: exception).
long long msref_need_neighbors(const float *verts, int nv, const uint32_t *faces, int nf, int32_t *out, long long cap) {
    try {
        TriMesh m(const_cast<float*>(verts), nv, 3);
        m.set_faces(const_cast<uint32_t*>(faces), nf, 3);
        const std::vector<TypeIndexI> r = m.need_neighbors();
        if ((long long)r.size() > cap) return -1;
        std::memcpy(out, r.data(), sizeof(int32_t) * r.size());
        return (long long)r.size();
    } catch (const std::exception &e) { std::cerr << e.what(); return -2; }
}

// type 1 (geodesic, TriMesh_kde.cpp:484-498): Floyd-Warshall over the edges of `faces` (the caller passes the
// mesh's faces followed by its periodic faces, which is what the periodic branch concatenates), then kde_2m.
int msref_kde_geodesic(const float *verts, int nv, const uint32_t *faces, int nf, int get_counts, float sigma,
                       int periodic, const float *bbox, int nbbox, const int32_t *ids, int nids, float *out_density) {
    try {
        TriMesh m(const_cast<float*>(verts), nv, 3);
        m.set_faces(const_cast<uint32_t*>(faces), nf, 3);
        std::vector<TypeIndexI> vids(ids, ids + nids);
        const std::vector<TypeFunction> *r = 0;
        GaussianKernel2D gk(sigma);
        if (periodic) { DistancePeriodicXYSquared dk(const_cast<float*>(bbox), nbbox);
                        r = &m.kde("d", 1, get_counts != 0, gk, dk, vids, false); }
        else          { DistanceSquared dk;
                        r = &m.kde("d", 1, get_counts != 0, gk, dk, vids, false); }
        std::memcpy(out_density, r->data(), sizeof(float)*r->size());
        return 0;
    } catch (const std::exception &e) { std::cerr << e.what(); return 1; }
}

// type: 2 (kde_2d, GaussianKernel2D) or 3 (kde_3d, GaussianKernel3D), as
// memsurfer/trimesh.py:337-349 picks them.  periodic != 0 ->
// DistancePeriodicXYSquared(bbox, nbbox) else DistanceSquared().
int msref_kde(const float *verts, int nv, int type, int get_counts, float sigma,
              int periodic, const float *bbox, int nbbox,
              const int32_t *ids, int nids, float *out_density) {
    try {
        TriMesh m(const_cast<float*>(verts), nv, 3);
        std::vector<TypeIndexI> vids(ids, ids + nids);
        const std::vector<TypeFunction> *r = 0;
        if (type == 3) {
            GaussianKernel3D gk(sigma);
            if (periodic) { DistancePeriodicXYSquared dk(const_cast<float*>(bbox), nbbox);
                            r = &m.kde("d", type, get_counts != 0, gk, dk, vids, false); }
            else          { DistanceSquared dk;
                            r = &m.kde("d", type, get_counts != 0, gk, dk, vids, false); }
        } else {
            GaussianKernel2D gk(sigma);
            if (periodic) { DistancePeriodicXYSquared dk(const_cast<float*>(bbox), nbbox);
                            r = &m.kde("d", type, get_counts != 0, gk, dk, vids, false); }
            else          { DistanceSquared dk;
                            r = &m.kde("d", type, get_counts != 0, gk, dk, vids, false); }
        }
        std::memcpy(out_density, r->data(), sizeof(float)*r->size());
        return 0;
    } catch (const std::exception &e) { std::cerr << e.what(); return 1; }
}

// TriMesh::trim_periodicDelaunay (TriMesh.cpp:493-575) on a given periodic Delaunay: dfaces (nf,3) original vertex ids,
// doffs (nf,3,2) periodic offsets in {-1,0,1}; the mesh's own bbox (n = 4 for 2-D, 6 for 3-D vertices).  Outputs sized
// for the worst case by the caller: faces / pfaces / tfaces (nf,3), dup_ids (3 nf), dup_verts (3 nf, dim).
// counts = {#faces, #periodic faces (= #trimmed faces), #duplicated vertices}.
int msref_trim_periodic(const float *verts, int nv, int dim, const float *bbox, int nbbox,
                        const int32_t *dfaces, const int8_t *doffs, int nf,
                        int32_t *faces, int32_t *pfaces, int32_t *tfaces, int32_t *dup_ids, float *dup_verts, long long *counts) {
    try {
        TriMesh m(const_cast<float*>(verts), nv, dim);
        m.set_periodic();
        m.set_bbox(const_cast<float*>(bbox), nbbox);
        m.mDelaunayFaces.resize(nf);
        for (int f = 0; f < nf; f++) {
            m.mDelaunayFaces[f].resize(3);
            for (int d = 0; d < 3; d++)
                m.mDelaunayFaces[f][d] = PeriodicVertex((TypeIndex)dfaces[3*f+d], doffs[6*f+2*d], doffs[6*f+2*d+1]);
        }
        m.trim_periodicDelaunay(false);
        const std::vector<TypeIndexI> a = m.get_faces(), b = m.periodic_faces(false), c = m.trimmed_faces(false), d = m.duplicate_ids();
        const std::vector<TypeFunction> e = m.duplicated_vertices(false);
        std::memcpy(faces, a.data(), sizeof(int32_t) * a.size());
        std::memcpy(pfaces, b.data(), sizeof(int32_t) * b.size());
        std::memcpy(tfaces, c.data(), sizeof(int32_t) * c.size());
        std::memcpy(dup_ids, d.data(), sizeof(int32_t) * d.size());
        std::memcpy(dup_verts, e.data(), sizeof(float) * e.size());
        counts[0] = (long long)a.size() / 3; counts[1] = (long long)b.size() / 3; counts[2] = (long long)d.size();
        return (b.size() == c.size() && e.size() == d.size() * (size_t)dim) ? 0 : 2;
    } catch (const std::exception &e) { std::cerr << e.what(); return 1; }
}

// TriMesh::wrap_vertices (TriMesh.cpp:462-491): x,y (dim = 2) folded into [box0, box1), in place on a copy.
int msref_wrap_vertices(const float *verts, int nv, int dim, const float *bbox, int nbbox, int wrap_dim, float *out) {
    try {
        TriMesh m(const_cast<float*>(verts), nv, dim);
        m.set_periodic();
        m.set_bbox(const_cast<float*>(bbox), nbbox);
        m.wrap_vertices((uint8_t)wrap_dim);
        const std::vector<TypeFunction> v = m.get_vertices();
        std::memcpy(out, v.data(), sizeof(float) * v.size());
        return 0;
    } catch (const std::exception &e) { std::cerr << e.what(); return 1; }
}

} // extern "C"
