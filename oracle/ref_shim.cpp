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
#include "TriMesh.hpp"
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

} // extern "C"
