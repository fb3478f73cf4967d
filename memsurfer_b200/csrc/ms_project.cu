// ms_project.cu -- closest point on a triangle mesh for a set of query points, with the face id and
// barycentric coordinates of the closest point and the distance (leaflet thickness).
//
// Replaces (LLNL/MemSurfer): TriMesh::project_on_surface
// memsurfer/src/TriMesh_cgal.cpp:295-363 (CGAL AABB tree over the mesh's own faces mFaces,
// closest_point_and_primitive per query, barycentric coordinates by TriMesh::Point2Bary
// memsurfer/src/TriMesh.cpp:28-42) and TriMesh::distance_to_other_mesh TriMesh_cgal.cpp:390-425
// (vertices of one mesh projected on the other, Euclidean distance, no periodic wrap).
//
// B200 design: instead of a pointer-chasing AABB tree, a uniform xy grid of triangle lists (membrane
// meshes are height fields: a few triangles per cell); one thread per query walks square rings of
// cells outwards and stops when the best squared distance is below the squared xy distance of the
// next ring.  The closest point on a triangle is Ericson's region test in fp64 without FMA
// contraction (this file is built --fmad=false) and ties break by the smaller face id, so the result
// is independent of the order of the lists and bit-identical to a sequential scan of all faces
// (oracle/ms_oracle.c orc_closest_point).
#include "ms_common.cuh"
#include <math.h>

struct ProjGrid { double x0, y0, inv_c, cell; int ncx, ncy; };

struct D3p { double x, y, z; };
__device__ __forceinline__ D3p p_sub(const D3p &a, const D3p &b) { return { a.x - b.x, a.y - b.y, a.z - b.z }; }
__device__ __forceinline__ double p_dot(const D3p &a, const D3p &b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ D3p p_ld(const float *__restrict__ v, int64_t i) {
    return { (double)__ldg(v + 3 * i), (double)__ldg(v + 3 * i + 1), (double)__ldg(v + 3 * i + 2) };
}

// closest point of triangle (a,b,c) to p (Ericson, Real-Time Collision Detection 5.1.5)
__device__ __forceinline__ D3p closest_on_triangle(const D3p &p, const D3p &a, const D3p &b, const D3p &c) {
    const D3p ab = p_sub(b, a), ac = p_sub(c, a), ap = p_sub(p, a);
    const double d1 = p_dot(ab, ap), d2 = p_dot(ac, ap);
    if (d1 <= 0.0 && d2 <= 0.0) return a;
    const D3p bp = p_sub(p, b);
    const double d3 = p_dot(ab, bp), d4 = p_dot(ac, bp);
    if (d3 >= 0.0 && d4 <= d3) return b;
    const double vc = d1 * d4 - d3 * d2;
    if (vc <= 0.0 && d1 >= 0.0 && d3 <= 0.0) {
        const double v = d1 / (d1 - d3);
        return { a.x + v * ab.x, a.y + v * ab.y, a.z + v * ab.z };
    }
    const D3p cp = p_sub(p, c);
    const double d5 = p_dot(ab, cp), d6 = p_dot(ac, cp);
    if (d6 >= 0.0 && d5 <= d6) return c;
    const double vb = d5 * d2 - d1 * d6;
    if (vb <= 0.0 && d2 >= 0.0 && d6 <= 0.0) {
        const double w = d2 / (d2 - d6);
        return { a.x + w * ac.x, a.y + w * ac.y, a.z + w * ac.z };
    }
    const double va = d3 * d6 - d5 * d4;
    if (va <= 0.0 && (d4 - d3) >= 0.0 && (d5 - d6) >= 0.0) {
        const double w = (d4 - d3) / ((d4 - d3) + (d5 - d6));
        return { b.x + w * (c.x - b.x), b.y + w * (c.y - b.y), b.z + w * (c.z - b.z) };
    }
    const double denom = 1.0 / (va + vb + vc);
    const double v = vb * denom, w = vc * denom;
    return { a.x + ab.x * v + ac.x * w, a.y + ab.y * v + ac.y * w, a.z + ab.z * v + ac.z * w };
}

__device__ __forceinline__ void face_cells(const float *__restrict__ verts, const uint32_t *__restrict__ faces, int64_t f, const ProjGrid &g,
                                           int &cx0, int &cx1, int &cy0, int &cy1) {
    double xl = 1e300, xh = -1e300, yl = 1e300, yh = -1e300;
#pragma unroll
    for (int k = 0; k < 3; k++) {
        const int64_t v = (int64_t)__ldg(faces + 3 * f + k);
        const double x = (double)__ldg(verts + 3 * v), y = (double)__ldg(verts + 3 * v + 1);
        xl = fmin(xl, x); xh = fmax(xh, x); yl = fmin(yl, y); yh = fmax(yh, y);
    }
    cx0 = max(0, min(g.ncx - 1, (int)floor((xl - g.x0) * g.inv_c))); cx1 = max(0, min(g.ncx - 1, (int)floor((xh - g.x0) * g.inv_c)));
    cy0 = max(0, min(g.ncy - 1, (int)floor((yl - g.y0) * g.inv_c))); cy1 = max(0, min(g.ncy - 1, (int)floor((yh - g.y0) * g.inv_c)));
}

__global__ void __launch_bounds__(256) k_proj_count(const float *__restrict__ verts, const uint32_t *__restrict__ faces, int64_t nf, ProjGrid g, int32_t *cnt) {
    const int64_t f = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= nf) return;
    int cx0, cx1, cy0, cy1;
    face_cells(verts, faces, f, g, cx0, cx1, cy0, cy1);
    for (int cy = cy0; cy <= cy1; cy++) for (int cx = cx0; cx <= cx1; cx++) atomicAdd(cnt + cy * g.ncx + cx, 1);
}
__global__ void __launch_bounds__(256) k_proj_fill(const float *__restrict__ verts, const uint32_t *__restrict__ faces, int64_t nf, ProjGrid g,
                                                   const int32_t *__restrict__ start, int32_t *cursor, int32_t *list, int64_t cap) {
    const int64_t f = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= nf) return;
    int cx0, cx1, cy0, cy1;
    face_cells(verts, faces, f, g, cx0, cx1, cy0, cy1);
    for (int cy = cy0; cy <= cy1; cy++) for (int cx = cx0; cx <= cx1; cx++) {
        const int c = cy * g.ncx + cx;
        const int64_t at = (int64_t)start[c] + atomicAdd(cursor + c, 1);
        if (at < cap) list[at] = (int32_t)f;
    }
}

__global__ void __launch_bounds__(128) k_proj_query(const float *__restrict__ verts, const uint32_t *__restrict__ faces, ProjGrid g,
                                                    const int32_t *__restrict__ start, const int32_t *__restrict__ list,
                                                    const float *__restrict__ queries, int64_t nq,
                                                    int32_t *__restrict__ face_id, double *__restrict__ bary, double *__restrict__ d64, float *__restrict__ d32) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nq) return;
    const D3p p = p_ld(queries, i);
    const int qx = max(0, min(g.ncx - 1, (int)floor((p.x - g.x0) * g.inv_c)));
    const int qy = max(0, min(g.ncy - 1, (int)floor((p.y - g.y0) * g.inv_c)));
    double best = 1e300; int bf = -1; D3p bc = { 0.0, 0.0, 0.0 };
    const int rmax = max(g.ncx, g.ncy);
    for (int r = 0; r <= rmax; r++) {
        // every cell of ring r+1 is at least r cells away from p in x or in y
        if (r > 0) { const double lb = (double)(r - 1) * g.cell; if (bf >= 0 && best <= lb * lb) break; }
        const int y0 = qy - r, y1 = qy + r, x0 = qx - r, x1 = qx + r;
        auto visit = [&](int cx, int cy) {
            const int c = cy * g.ncx + cx;
            const int e1 = __ldg(start + c + 1);
            for (int e = __ldg(start + c); e < e1; e++) {
                const int f = __ldg(list + e);
                const D3p a = p_ld(verts, (int64_t)__ldg(faces + 3 * (int64_t)f)), b = p_ld(verts, (int64_t)__ldg(faces + 3 * (int64_t)f + 1)),
                          cc = p_ld(verts, (int64_t)__ldg(faces + 3 * (int64_t)f + 2));
                const D3p q = closest_on_triangle(p, a, b, cc);
                const D3p dq = p_sub(p, q);
                const double dd = p_dot(dq, dq);
                if (dd < best || (dd == best && f < bf)) { best = dd; bf = f; bc = q; }
            }
        };
        for (int cy = max(y0, 0); cy <= min(y1, g.ncy - 1); cy++) {
            if (cy == y0 || cy == y1) {
                for (int cx = max(x0, 0); cx <= min(x1, g.ncx - 1); cx++) visit(cx, cy);
            } else {
                if (x0 >= 0) visit(x0, cy);
                if (x1 < g.ncx) visit(x1, cy);
            }
        }
    }
    double u = 0.0, v = 0.0, w = 0.0;
    if (bf >= 0) {      // TriMesh::Point2Bary (TriMesh.cpp:28-42)
        const D3p a = p_ld(verts, (int64_t)__ldg(faces + 3 * (int64_t)bf)), b = p_ld(verts, (int64_t)__ldg(faces + 3 * (int64_t)bf + 1)),
                  c = p_ld(verts, (int64_t)__ldg(faces + 3 * (int64_t)bf + 2));
        const D3p v0 = p_sub(b, a), v1 = p_sub(c, a), v2 = p_sub(bc, a);
        const double d00 = p_dot(v0, v0), d01 = p_dot(v0, v1), d11 = p_dot(v1, v1), d20 = p_dot(v2, v0), d21 = p_dot(v2, v1);
        const double denom = d00 * d11 - d01 * d01;
        v = (d11 * d20 - d01 * d21) / denom;
        w = (d00 * d21 - d01 * d20) / denom;
        u = 1.0 - v - w;
    }
    if (face_id) face_id[i] = bf;
    if (bary) { bary[3 * i] = u; bary[3 * i + 1] = v; bary[3 * i + 2] = w; }
    const double d = bf >= 0 ? sqrt(best) : 0.0;
    if (d64) d64[i] = d;
    if (d32) d32[i] = (float)d;
}

// ------------------------------------------------------------------ host side
static void proj_grid(const float *box4, int64_t nf, ProjGrid &g) {
    const double wx = fmax((double)box4[2] - (double)box4[0], 1e-30), wy = fmax((double)box4[3] - (double)box4[1], 1e-30);
    double cell = sqrt(wx * wy / fmax(1.0, (double)nf / 2.0));          // ~2 faces per cell before overlap duplication
    const double maxc = 4.0e6;
    if ((wx / cell) * (wy / cell) > maxc) cell = sqrt(wx * wy / maxc);
    g.x0 = box4[0]; g.y0 = box4[1]; g.cell = cell; g.inv_c = 1.0 / cell;
    g.ncx = (int)fmax(1.0, ceil(wx / cell)); g.ncy = (int)fmax(1.0, ceil(wy / cell));
}
static int64_t proj_list_cap(int64_t nf) { return 12 * nf + 1024; }

extern "C" size_t ms_closest_point_workspace_bytes(int64_t nv, int64_t nf) {
    (void)nv;
    const size_t nc = 4000000 + 4096 + 2;
    return 3 * ms_align(nc * 4) + ms_align((size_t)proj_list_cap(nf) * 4) + ms_scan_ws_bytes((int64_t)nc) + 4096;
}

extern "C" int ms_closest_point(const float *verts, int64_t nv, const uint32_t *faces, int64_t nf, const float *box4,
                                const float *queries, int64_t nq, int32_t *face_id, double *bary, double *dist_f64, float *dist_f32,
                                void *ws, size_t ws_bytes, void *stream) {
    cudaStream_t s = (cudaStream_t)stream;
    MS_REQUIRE(verts && faces && box4 && queries, "ms_closest_point: null input");
    MS_REQUIRE(nv > 0 && nf > 0 && nq >= 0 && nf < ((int64_t)1 << 31), "ms_closest_point: bad sizes nv=%lld nf=%lld nq=%lld", (long long)nv, (long long)nf, (long long)nq);
    MS_REQUIRE(ws_bytes >= ms_closest_point_workspace_bytes(nv, nf), "ms_closest_point: workspace too small");
    if (nq == 0) return 0;
    ProjGrid g; proj_grid(box4, nf, g);
    const int64_t ncell = (int64_t)g.ncx * g.ncy;
    MS_REQUIRE(ncell <= 4000000 + 4096, "ms_closest_point: grid too large");
    MsArena A(ws, ws_bytes);
    const size_t nc = 4000000 + 4096 + 2;
    int32_t *cnt = A.take<int32_t>(nc), *start = A.take<int32_t>(nc), *cursor = A.take<int32_t>(nc);
    const int64_t cap = proj_list_cap(nf);
    int32_t *list = A.take<int32_t>((size_t)cap);
    void *scan_ws = A.take<char>(ms_scan_ws_bytes((int64_t)nc));
    MS_CUDA(cudaMemsetAsync(cnt, 0, sizeof(int32_t) * (ncell + 1), s));
    MS_CUDA(cudaMemsetAsync(cursor, 0, sizeof(int32_t) * (ncell + 1), s));
    MS_K("proj_count", s); k_proj_count<<<ms_blocks(nf, 256), 256, 0, s>>>(verts, faces, nf, g, cnt);
    if (ms_exclusive_scan_i32(cnt, start, ncell, scan_ws, s)) return 2;
    int32_t total = 0;
    MS_CUDA(cudaMemcpyAsync(&total, start + ncell, sizeof(int32_t), cudaMemcpyDeviceToHost, s));
    MS_CUDA(cudaStreamSynchronize(s));
    MS_REQUIRE((int64_t)total <= cap, "ms_closest_point: %d (face, cell) entries exceed the list capacity %lld (triangles much larger than the mean)", total, (long long)cap);
    MS_K("proj_fill", s); k_proj_fill<<<ms_blocks(nf, 256), 256, 0, s>>>(verts, faces, nf, g, start, cursor, list, cap);
    MS_K("proj_query", s); k_proj_query<<<ms_blocks(nq, 128), 128, 0, s>>>(verts, faces, g, start, list, queries, nq, face_id, bary, dist_f64, dist_f32);
    MS_K(NULL, s);
    MS_LAUNCH_CHECK();
    return 0;
}
