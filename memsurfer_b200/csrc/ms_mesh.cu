// ms_mesh.cu -- TriMesh geometry on B200: vertex->face CSR build and ONE fused per-vertex
// gather kernel for vertex normals, mixed Voronoi areas, mean and Gaussian curvature.
//
// Replaces (LLNL/MemSurfer): TriMesh::need_normals memsurfer/src/TriMesh.cpp:99-133,
// TriMesh::need_pointareas TriMesh.cpp:140-190, their periodic wrappers TriMesh.hpp:346-407,
// and the vtkCurvatures calls of memsurfer/trimesh.py:284-316 (face filter trimesh.py:484-499).
//
// Design: HBM-bound streaming work.  Instead of the reference's scatter (face -> 3 vertex
// accumulators; atomics under OpenMP) every ORIGINAL vertex owns one thread that walks its
// incident (face,corner) list in ascending order -- the same order in which the reference's
// sequential face loop adds to that vertex -- so results are deterministic and agree with a
// sequential fp64 evaluation bit for bit (this file is compiled with --fmad=false).
// Duplicate (periodic image) vertices only ever serve as gather sources; their rows are never
// produced (the reference computes and then drops them, TriMesh.hpp:368-369, trimesh.py:310-311).
#include "ms_common.cuh"

#define MESH_T 128

// ------------------------------------------------------------------ topology
__global__ void __launch_bounds__(256) k_face_degree(const float *__restrict__ verts, const uint32_t *__restrict__ faces,
                                                     int64_t nf, int64_t nv, double xw, double yw,
                                                     int32_t *deg, uint8_t *keep) {
    const int64_t f = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= nf) return;
    uint32_t t[3] = { faces[3*f], faces[3*f+1], faces[3*f+2] };
    int per = 0;
    if (xw >= 0.0) {
        double x[3], y[3];
#pragma unroll
        for (int i = 0; i < 3; i++) { x[i] = ms_ld_f(verts + 3*(size_t)t[i]); y[i] = ms_ld_f(verts + 3*(size_t)t[i] + 1); }
#pragma unroll
        for (int i = 0; i < 3; i++) per |= (fabs(x[i] - x[(i+1)%3]) > xw) || (fabs(y[i] - y[(i+1)%3]) > yw);
    }
    keep[f] = !per;
#pragma unroll
    for (int i = 0; i < 3; i++) if ((int64_t)t[i] < nv) atomicAdd(deg + t[i], 1);
}

__global__ void __launch_bounds__(256) k_face_fill(const uint32_t *__restrict__ faces, int64_t nf, int64_t nv,
                                                   const int32_t *__restrict__ off, int32_t *cursor, int32_t *adj) {
    const int64_t f = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= nf) return;
#pragma unroll
    for (int i = 0; i < 3; i++) {
        const uint32_t v = faces[3*f+i];
        if ((int64_t)v < nv) { const int slot = atomicAdd(cursor + v, 1); adj[off[v] + slot] = (int32_t)(f * 4 + i); }
    }
}

// ascending (face,corner) order inside every vertex list => deterministic gather order
__global__ void __launch_bounds__(256) k_sort_lists(int64_t nv, const int32_t *__restrict__ off, int32_t *adj) {
    const int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= nv) return;
    const int b = off[v], e = off[v+1];
    for (int i = b + 1; i < e; i++) {
        const int32_t key = adj[i]; int j = i - 1;
        while (j >= b && adj[j] > key) { adj[j+1] = adj[j]; j--; }
        adj[j+1] = key;
    }
}

extern "C" size_t ms_mesh_topology_workspace_bytes(int64_t nv, int64_t nf) {
    (void)nf;
    return 2 * ms_align((size_t)(nv + 1) * sizeof(int32_t)) + ms_scan_ws_bytes(nv) + 1024;
}

extern "C" int ms_mesh_topology(const float *verts_all, int64_t nv_all, int64_t nv,
                                const uint32_t *faces_all, int64_t nf, double half_xw, double half_yw,
                                int32_t *csr_off, int32_t *csr_adj, uint8_t *keep,
                                void *ws, size_t ws_bytes, void *stream) {
    cudaStream_t s = (cudaStream_t)stream;
    MS_REQUIRE(nv >= 0 && nv <= nv_all && nf >= 0, "ms_mesh_topology: bad sizes nv=%lld nv_all=%lld nf=%lld", (long long)nv, (long long)nv_all, (long long)nf);
    MS_REQUIRE(nf < ((int64_t)1 << 29), "ms_mesh_topology: too many faces");
    MS_REQUIRE(ws_bytes >= ms_mesh_topology_workspace_bytes(nv, nf), "ms_mesh_topology: workspace too small");
    MsArena A(ws, ws_bytes);
    int32_t *deg = A.take<int32_t>(nv + 1), *cursor = A.take<int32_t>(nv + 1);
    void *scan_ws = A.take<char>(ms_scan_ws_bytes(nv));
    MS_CUDA(cudaMemsetAsync(deg, 0, sizeof(int32_t) * (nv + 1), s));
    MS_CUDA(cudaMemsetAsync(cursor, 0, sizeof(int32_t) * (nv + 1), s));
    if (nf > 0) { MS_K("face_degree", s); k_face_degree<<<ms_blocks(nf, 256), 256, 0, s>>>(verts_all, faces_all, nf, nv, half_xw, half_yw, deg, keep); }
    if (ms_exclusive_scan_i32(deg, csr_off, nv, scan_ws, s)) return 2;
    if (nf > 0) { MS_K("face_fill", s); k_face_fill<<<ms_blocks(nf, 256), 256, 0, s>>>(faces_all, nf, nv, csr_off, cursor, csr_adj); }
    if (nv > 0) { MS_K("sort_lists", s); k_sort_lists<<<ms_blocks(nv, 256), 256, 0, s>>>(nv, csr_off, csr_adj); }
    MS_K(NULL, s);
    MS_LAUNCH_CHECK();
    return 0;
}

// ------------------------------------------------------------------ fused properties
struct D3 { double x, y, z; };
__device__ __forceinline__ D3 ldv(const float *__restrict__ verts, uint32_t i) {
    const float *p = verts + 3 * (size_t)i;
    D3 r; r.x = (double)__ldg(p); r.y = (double)__ldg(p + 1); r.z = (double)__ldg(p + 2); return r;
}
__device__ __forceinline__ D3 sub(const D3 &a, const D3 &b) { D3 r; r.x = a.x - b.x; r.y = a.y - b.y; r.z = a.z - b.z; return r; }
__device__ __forceinline__ double dot(const D3 &a, const D3 &b) { return a.x*b.x + a.y*b.y + a.z*b.z; }
__device__ __forceinline__ D3 cross(const D3 &a, const D3 &b) {
    D3 r; r.x = a.y*b.z - a.z*b.y; r.y = a.z*b.x - a.x*b.z; r.z = a.x*b.y - a.y*b.x; return r;
}
// vtkTriangle::ComputeNormalDirection(v1,v2,v3) = (v3-v2) x (v1-v2)
__device__ __forceinline__ D3 tri_dir(const D3 &v1, const D3 &v2, const D3 &v3) { return cross(sub(v3, v2), sub(v1, v2)); }
__device__ __forceinline__ D3 unit_or_same(D3 n) {
    const double l = sqrt(dot(n, n));
    if (l != 0.0) { n.x /= l; n.y /= l; n.z /= l; }
    return n;
}
__device__ __forceinline__ double angle_between(const D3 &u, const D3 &v) {
    const D3 c = cross(u, v);
    return atan2(sqrt(dot(c, c)), dot(u, v));
}

// count list entries of vertex list [b,e) (other than face g) whose face is kept and contains u;
// returns the count and the last such face in nb
__device__ __forceinline__ int edge_neighbours(const uint32_t *__restrict__ faces, const int32_t *__restrict__ adj,
                                               const uint8_t *__restrict__ keep, int b, int e, int g, uint32_t u, int &nb) {
    int cnt = 0;
    for (int j = b; j < e; j++) {
        const int h = __ldg(adj + j) >> 2;
        if (h == g || !__ldg(keep + h)) continue;
        const uint32_t *t = faces + 3 * (size_t)h;
        if (__ldg(t) == u || __ldg(t + 1) == u || __ldg(t + 2) == u) { nb = h; cnt++; }
    }
    return cnt;
}

__global__ void __launch_bounds__(MESH_T) k_mesh_properties(
        const float *__restrict__ verts, const uint32_t *__restrict__ faces, int64_t nv,
        const int32_t *__restrict__ off, const int32_t *__restrict__ adj, const uint8_t *__restrict__ keep,
        double *__restrict__ vn64, float *__restrict__ vn32, double *__restrict__ pa64, float *__restrict__ pa32,
        double *__restrict__ Hout, double *__restrict__ Kout) {
    const int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= nv) return;
    const double PI = 3.14159265358979323846;
    const bool want_curv = (Hout != nullptr) || (Kout != nullptr);
    D3 pn; pn.x = pn.y = pn.z = 0.0;
    double parea = 0.0, Hsum = 0.0, Ksum = 2.0 * PI, dA = 0.0;
    int hcnt = 0;
    const int b = __ldg(off + v), e = __ldg(off + v + 1);
    for (int j = b; j < e; j++) {
        const int packed = __ldg(adj + j);
        const int g = packed >> 2, ci = packed & 3;
        const uint32_t t[3] = { __ldg(faces + 3*(size_t)g), __ldg(faces + 3*(size_t)g + 1), __ldg(faces + 3*(size_t)g + 2) };
        const D3 p[3] = { ldv(verts, t[0]), ldv(verts, t[1]), ldv(verts, t[2]) };
        // ---- vertex normal contribution (TriMesh.cpp:110-127)
        {
            const D3 a = sub(p[0], p[1]), bb = sub(p[1], p[2]), c = sub(p[2], p[0]);
            const double l2a = dot(a, a), l2b = dot(bb, bb), l2c = dot(c, c);
            if (l2a != 0.0 && l2b != 0.0 && l2c != 0.0) {
                const D3 fn = cross(a, bb);
                const double sc = ci == 0 ? 1.0 / (l2a * l2c) : (ci == 1 ? 1.0 / (l2b * l2a) : 1.0 / (l2c * l2b));
                pn.x += fn.x * sc; pn.y += fn.y * sc; pn.z += fn.z * sc;
            }
        }
        // ---- mixed Voronoi corner area (TriMesh.cpp:152-188)
        {
            const D3 ed[3] = { sub(p[2], p[1]), sub(p[0], p[2]), sub(p[1], p[0]) };
            const D3 cr = cross(ed[0], ed[1]);
            const double area = 0.5 * sqrt(dot(cr, cr));
            const double l2[3] = { dot(ed[0], ed[0]), dot(ed[1], ed[1]), dot(ed[2], ed[2]) };
            const double ew[3] = { l2[0] * (l2[1] + l2[2] - l2[0]), l2[1] * (l2[2] + l2[0] - l2[1]), l2[2] * (l2[0] + l2[1] - l2[2]) };
            double ca[3];
            if (ew[0] <= 0.0) {
                ca[1] = -0.25 * l2[2] * area / dot(ed[0], ed[2]);
                ca[2] = -0.25 * l2[1] * area / dot(ed[0], ed[1]);
                ca[0] = area - ca[1] - ca[2];
            } else if (ew[1] <= 0.0) {
                ca[2] = -0.25 * l2[0] * area / dot(ed[1], ed[0]);
                ca[0] = -0.25 * l2[2] * area / dot(ed[1], ed[2]);
                ca[1] = area - ca[2] - ca[0];
            } else if (ew[2] <= 0.0) {
                ca[0] = -0.25 * l2[1] * area / dot(ed[2], ed[1]);
                ca[1] = -0.25 * l2[0] * area / dot(ed[2], ed[0]);
                ca[2] = area - ca[0] - ca[1];
            } else {
                const double ewscale = 0.5 * area / (ew[0] + ew[1] + ew[2]);
                ca[0] = ewscale * (ew[1] + ew[2]); ca[1] = ewscale * (ew[2] + ew[0]); ca[2] = ewscale * (ew[0] + ew[1]);
            }
            parea += ci == 0 ? ca[0] : (ci == 1 ? ca[1] : ca[2]);
        }
        if (!want_curv || !__ldg(keep + g)) continue;
        // ---- mean curvature: edges of g that touch v, in ascending edge index (vtkCurvatures mean)
        {
            const int e0 = ci == 0 ? 0 : ci - 1, e1 = ci == 0 ? 2 : ci;   // {ci,(ci+2)%3} sorted
#pragma unroll 1
            for (int k = 0; k < 2; k++) {
                const int ed = k == 0 ? e0 : e1;
                const int il = ed, ir = (ed + 1) % 3, io = (ed + 2) % 3;
                const uint32_t u = (il == ci) ? t[ir] : t[il];
                int nb = -1;
                const int cnt = edge_neighbours(faces, adj, keep, b, e, g, u, nb);
                if (cnt != 1 || nb <= g) continue;
                const D3 ore = p[il], end = p[ir], oth = p[io];
                const D3 d0 = tri_dir(ore, end, oth);
                const D3 nf = unit_or_same(d0);
                const double Af0 = 0.5 * sqrt(dot(d0, d0));
                const uint32_t *tn = faces + 3 * (size_t)nb;
                const D3 a0 = ldv(verts, __ldg(tn)), a1 = ldv(verts, __ldg(tn + 1)), a2 = ldv(verts, __ldg(tn + 2));
                const D3 d1 = tri_dir(a0, a1, a2);
                const D3 nn = unit_or_same(d1);
                const double Af = Af0 + 0.5 * sqrt(dot(d1, d1));
                D3 ev = sub(end, ore);
                const double len = sqrt(dot(ev, ev));
                if (len != 0.0) { ev.x /= len; ev.y /= len; ev.z /= len; }
                const double cs = dot(nf, nn);
                const double sn = dot(cross(nf, nn), ev);
                double Hf = (sn != 0.0 || cs != 0.0) ? len * atan2(sn, cs) : 0.0;
                if (Af != 0.0) Hf = Hf / Af * 3.0;
                Hsum += Hf; hcnt++;
            }
        }
        // ---- Gaussian curvature: interior angle at v and face area (vtkCurvatures Gauss)
        {
            const D3 g0 = sub(p[1], p[0]), g1 = sub(p[2], p[1]), g2 = sub(p[0], p[2]);
            const double al = ci == 0 ? PI - angle_between(g2, g0) : (ci == 1 ? PI - angle_between(g0, g1) : PI - angle_between(g1, g2));
            const D3 dn = tri_dir(p[0], p[1], p[2]);
            dA += 0.5 * sqrt(dot(dn, dn));
            Ksum -= al;
        }
    }
    {
        const double l = sqrt(dot(pn, pn));
        if (l <= 0.0) { pn.x = 1.0; pn.y = 0.0; pn.z = 0.0; }
        else { const double r = 1.0 / l; pn.x *= r; pn.y *= r; pn.z *= r; }
    }
    if (vn64) { vn64[3*v] = pn.x; vn64[3*v+1] = pn.y; vn64[3*v+2] = pn.z; }
    if (vn32) { vn32[3*v] = (float)pn.x; vn32[3*v+1] = (float)pn.y; vn32[3*v+2] = (float)pn.z; }
    if (pa64) pa64[v] = parea;
    if (pa32) pa32[v] = (float)parea;
    if (Hout) Hout[v] = hcnt > 0 ? 0.5 * Hsum / (double)hcnt : 0.0;
    if (Kout) Kout[v] = dA > 0.0 ? 3.0 * Ksum / dA : 0.0;
}

extern "C" int ms_mesh_properties(const float *verts_all, int64_t nv_all, int64_t nv,
                                  const uint32_t *faces_all, int64_t nf,
                                  const int32_t *csr_off, const int32_t *csr_adj, const uint8_t *keep,
                                  double *vnormals_f64, float *vnormals_f32,
                                  double *pareas_f64, float *pareas_f32,
                                  double *mean_curv, double *gaus_curv, void *stream) {
    MS_REQUIRE(nv >= 0 && nv <= nv_all && nf >= 0, "ms_mesh_properties: bad sizes");
    MS_REQUIRE(csr_off && csr_adj && keep, "ms_mesh_properties: topology missing (call ms_mesh_topology first)");
    if (nv == 0) return 0;
    MS_K("mesh_properties", (cudaStream_t)stream); k_mesh_properties<<<ms_blocks(nv, MESH_T), MESH_T, 0, (cudaStream_t)stream>>>(
        verts_all, faces_all, nv, csr_off, csr_adj, keep,
        vnormals_f64, vnormals_f32, pareas_f64, pareas_f32, mean_curv, gaus_curv);
    MS_K(NULL, (cudaStream_t)stream);
    MS_LAUNCH_CHECK();
    return 0;
}
