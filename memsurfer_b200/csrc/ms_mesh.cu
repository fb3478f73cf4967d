// ms_mesh.cu -- TriMesh geometry on B200: vertex->face CSR build, a corner-parallel pass for the atan2 terms and an
// eight-lanes-per-vertex ordered gather for vertex normals, mixed Voronoi areas, mean and Gaussian curvature.
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

// ------------------------------------------------------------------ properties: corner pass + vertex gather
// Data flow (algorithmic bytes: 84 B/vertex): vertices are padded ONCE to float4 (one LDG.128 per
// gather); the only per-face state that goes through memory is what is expensive to recompute -- the atan2
// results: the three corner angle terms of the Gaussian curvature and the (up to three) dihedral edge terms of
// the mean curvature, 48 bytes per face.  Everything else (face normal, corner weights, mixed Voronoi corner
// areas, triangle area: cross products, a square root and a division) is recomputed per (vertex, incident face)
// from the L2-resident float4 vertices -- the 128-byte face records of the first version cost 11x the
// algorithmic DRAM traffic.
struct D3 { double x, y, z; };
// random gathers go straight to L2 (ld.cg): through L1 they were measured 2.8x slower in the orientation kernel
__device__ __forceinline__ D3 ldv4(const float4 *__restrict__ v4, uint32_t i) {
    const float4 p = __ldcg(v4 + i);
    D3 r; r.x = (double)p.x; r.y = (double)p.y; r.z = (double)p.z; return r;
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

__global__ void __launch_bounds__(256) k_pad_verts(const float *__restrict__ verts, int64_t n, float4 *__restrict__ v4) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) v4[i] = make_float4(verts[3*i], verts[3*i+1], verts[3*i+2], 0.f);
}

// One thread per CORNER / half-edge (g, i) -- 3F threads, at most two atan2 each:
//  * al[3g+i] = pi - angle at corner i (the term vtkCurvatures::GetGaussCurvature subtracts at that vertex);
//  * He[3g+i] = mean-curvature term of the edge (t[i], t[i+1]) of a kept face g (vtkCurvatures::GetMeanCurvature):
//    if exactly one other kept face n shares the edge and n > g, Hf = 3 * len * dihedral / (A_g + A_n) and
//    hflag[3g+i] = 1; both end points pick it up in the gather.  The neighbour search walks the incident list of
//    an end point that is an original vertex (every face on the edge is in that list); edges between two
//    duplicated vertices feed dropped rows only.  The neighbour's unit normal and area are recomputed from its
//    own three vertices (same operations as for g: the result does not depend on who computes it).
// 12 blocks of 128 threads per SM (40 registers, a few spilled words): the pass waits on dependent gathers, so resident
// warps count for more than registers (0.339 -> 0.307 ms per frame; 16 blocks: the same)
#ifndef CT_MINB
#define CT_MINB 12
#endif
__global__ void __launch_bounds__(128, CT_MINB) k_corner_terms(const float4 *__restrict__ v4, const uint32_t *__restrict__ faces, int64_t nf, int64_t nv,
                                                      const int32_t *__restrict__ off, const int32_t *__restrict__ adj,
                                                      const uint8_t *__restrict__ keep, int want_mean,
                                                      double *__restrict__ al, double *__restrict__ He, uint8_t *__restrict__ hflag) {
    const int64_t he = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (he >= 3 * nf) return;
    const double PI = 3.14159265358979323846;
    const int64_t g = he / 3;
    const int ed = (int)(he - 3 * g);
    const uint32_t t[3] = { __ldg(faces + 3*g), __ldg(faces + 3*g + 1), __ldg(faces + 3*g + 2) };
    const D3 p[3] = { ldv4(v4, t[0]), ldv4(v4, t[1]), ldv4(v4, t[2]) };
    {   // corner angle: corner 0 -> (g2, g0), 1 -> (g0, g1), 2 -> (g1, g2) with g_k = p[k+1] - p[k]
        const D3 u = ed == 0 ? sub(p[0], p[2]) : (ed == 1 ? sub(p[1], p[0]) : sub(p[2], p[1]));
        const D3 v = ed == 0 ? sub(p[1], p[0]) : (ed == 1 ? sub(p[2], p[1]) : sub(p[0], p[2]));
        al[he] = PI - angle_between(u, v);
    }
    uint8_t flag = 0;
    if (want_mean && __ldg(keep + g)) {
        // (no dynamic indexing: p[] and t[] stay in registers)
        const uint32_t vl = ed == 0 ? t[0] : (ed == 1 ? t[1] : t[2]);
        const uint32_t vr = ed == 0 ? t[1] : (ed == 1 ? t[2] : t[0]);
        uint32_t a = vl, u = vr;                       // walk a's list, look for u
        if ((int64_t)a >= nv) { a = vr; u = vl; }
        if ((int64_t)a < nv) {
            const int b = __ldg(off + a), e = __ldg(off + a + 1);
            int cnt = 0, nb = -1;
            for (int j = b; j < e; j++) {
                const int packed = __ldg(adj + j);
                const int h = packed >> 2, c = packed & 3;            // a sits at corner c of face h
                if (h == (int)g || !__ldg(keep + h)) continue;
                const uint32_t *th = faces + 3 * (size_t)h;
                if (__ldg(th + (c == 2 ? 0 : c + 1)) == u || __ldg(th + (c == 0 ? 2 : c - 1)) == u) { nb = h; cnt++; }
            }
            if (cnt == 1 && nb > (int)g) {
                const D3 ore = ed == 0 ? p[0] : (ed == 1 ? p[1] : p[2]);
                const D3 end = ed == 0 ? p[1] : (ed == 1 ? p[2] : p[0]);
                const D3 oth = ed == 0 ? p[2] : (ed == 1 ? p[0] : p[1]);
                const D3 d0 = tri_dir(ore, end, oth);
                const D3 nfu = unit_or_same(d0);
                const double Af0 = 0.5 * sqrt(dot(d0, d0));
                const uint32_t *tn = faces + 3 * (size_t)nb;
                const D3 q0 = ldv4(v4, __ldg(tn)), q1 = ldv4(v4, __ldg(tn + 1)), q2 = ldv4(v4, __ldg(tn + 2));
                const D3 dn = tri_dir(q0, q1, q2);
                const D3 nn = unit_or_same(dn);
                const double Af = Af0 + 0.5 * sqrt(dot(dn, dn));
                D3 ev = sub(end, ore);
                const double len = sqrt(dot(ev, ev));
                if (len != 0.0) { ev.x /= len; ev.y /= len; ev.z /= len; }
                const double cs = dot(nfu, nn);
                const double sn = dot(cross(nfu, nn), ev);
                double Hf = (sn != 0.0 || cs != 0.0) ? len * atan2(sn, cs) : 0.0;
                if (Af != 0.0) Hf = Hf / Af * 3.0;
                He[he] = Hf;
                flag = 1;
            }
        }
    }
    hflag[he] = flag;
}

// Vertex gather.  A block owns VG_V consecutive ORIGINAL vertices, i.e. one contiguous range of the CSR incidence
// array.  Phase 1: one thread per incident (face, corner) ENTRY recomputes what that corner contributes (dense: every
// lane has an entry, whatever the vertex degrees) and leaves eight doubles in shared memory.  Phase 2: thread
// (vertex i, quantity q) adds up quantity q over the entries of vertex i in ascending order -- the order in which the
// reference's sequential face loop adds to that vertex, so the sums are bit-equal to the sequential oracle.
#define VG_T 128
#define VG_V 16           // vertices per block (x VG_Q quantities = VG_T summing threads)
#define VG_Q 8            // doubles per entry: pn.x, pn.y, pn.z, corner area, triangle area, angle term, H term 0, H term 1
// 12 blocks per SM (40 registers): 0.43 -> 0.33 ms per frame; 14 blocks spill into the summing loop (0.41)
#ifndef VG_MINB
#define VG_MINB 12
#endif
__global__ void __launch_bounds__(VG_T, VG_MINB) k_vertex_gather8(
        const float4 *__restrict__ v4, const uint32_t *__restrict__ faces, int64_t nv,
        const int32_t *__restrict__ off, const int32_t *__restrict__ adj, const uint8_t *__restrict__ keep,
        const double *__restrict__ al, const double *__restrict__ He, const uint8_t *__restrict__ hflag,
        double *__restrict__ vn64, float *__restrict__ vn32, double *__restrict__ pa64, float *__restrict__ pa32,
        double *__restrict__ Hout, double *__restrict__ Kout) {
    __shared__ double sm[VG_T][VG_Q + 1];                 // +1: the summing threads read a column, 9 doubles apart
    __shared__ int sm_h[VG_T];
    __shared__ int sm_off[VG_V + 1];
    const double PI = 3.14159265358979323846;
    const int tid = threadIdx.x;
    const int64_t v0 = (int64_t)blockIdx.x * VG_V;
    const int nvb = (int)min((int64_t)VG_V, nv - v0);
    if (tid <= VG_V) sm_off[tid] = __ldg(off + v0 + min(tid, nvb));
    __syncthreads();
    const int eb = sm_off[0], ee = sm_off[VG_V];
    const int vi = tid / VG_Q, q = tid % VG_Q;            // phase 2 role: vertex vi of the block, quantity q
    const int my_b = sm_off[vi], my_e = sm_off[vi + 1];
    // running sum: q = 0..2 normal, 3 point area, 4 dA, 5 K (starts at 2 pi, terms are subtracted), 6 H, 7 number of H terms
    double acc = q == 5 ? 2.0 * PI : 0.0;
    int hcnt = 0;
    for (int r0 = eb; r0 < ee; r0 += VG_T) {
        const int j = r0 + tid;
        double c[VG_Q] = { 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0 };
        int myh = 0;
        if (j < ee) {
            const int packed = __ldg(adj + j);
            const int g = packed >> 2, ci = packed & 3;
            const uint32_t t[3] = { __ldcg(faces + 3*(size_t)g), __ldcg(faces + 3*(size_t)g + 1), __ldcg(faces + 3*(size_t)g + 2) };
            const D3 p[3] = { ldv4(v4, t[0]), ldv4(v4, t[1]), ldv4(v4, t[2]) };
            {   // ---- vertex normal contribution (TriMesh.cpp:110-127): fn = a x b, weight 1 / (|e|^2 |e'|^2) of the corner's edges
                const D3 a = sub(p[0], p[1]), bb = sub(p[1], p[2]), cc = sub(p[2], p[0]);
                const double l2a = dot(a, a), l2b = dot(bb, bb), l2c = dot(cc, cc);
                if (l2a != 0.0 && l2b != 0.0 && l2c != 0.0) {
                    const D3 fn = cross(a, bb);
                    const double w = ci == 0 ? 1.0 / (l2a * l2c) : (ci == 1 ? 1.0 / (l2b * l2a) : 1.0 / (l2c * l2b));
                    c[0] = fn.x * w; c[1] = fn.y * w; c[2] = fn.z * w;
                }
            }
            {   // ---- mixed Voronoi corner area (TriMesh.cpp:152-188)
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
                c[3] = ci == 0 ? ca[0] : (ci == 1 ? ca[1] : ca[2]);
            }
            if (__ldcg(keep + g)) {   // ---- curvature terms of kept faces (vtkCurvatures)
                const D3 dn = tri_dir(p[0], p[1], p[2]);
                c[4] = 0.5 * sqrt(dot(dn, dn));
                c[5] = __ldcg(al + 3*(size_t)g + ci);
                // edges of g that touch this vertex, in ascending edge index: {ci, (ci+2)%3} sorted
                const int e0 = ci == 0 ? 0 : ci - 1, e1 = ci == 0 ? 2 : ci;
                if (__ldcg(hflag + 3*(size_t)g + e0)) { c[6] = __ldcg(He + 3*(size_t)g + e0); myh++; }
                if (__ldcg(hflag + 3*(size_t)g + e1)) { c[7] = __ldcg(He + 3*(size_t)g + e1); myh++; }
            }
        }
#pragma unroll
        for (int k = 0; k < VG_Q; k++) sm[tid][k] = c[k];
        sm_h[tid] = myh;
        __syncthreads();
        const int jb = max(my_b, r0) - r0, je = min(my_e, r0 + VG_T) - r0;      // this vertex's entries inside the round
        if (q <= 4) { for (int i = jb; i < je; i++) acc += sm[i][q]; }
        else if (q == 5) { for (int i = jb; i < je; i++) acc -= sm[i][5]; }
        else if (q == 6) { for (int i = jb; i < je; i++) { acc += sm[i][6]; acc += sm[i][7]; } }
        else { for (int i = jb; i < je; i++) hcnt += sm_h[i]; }
        __syncthreads();
    }
    // hand the sums of a vertex (8 consecutive lanes) to its first lane
    const int base = (tid & 31) / VG_Q * VG_Q;
    double r[7];
#pragma unroll
    for (int k = 0; k < 7; k++) r[k] = __shfl_sync(0xffffffffu, acc, base + k);
    hcnt = __shfl_sync(0xffffffffu, hcnt, base + 7);
    const int64_t v = v0 + vi;
    if (vi >= nvb || q != 0) return;
    double pnx = r[0], pny = r[1], pnz = r[2];
    {
        const double l = sqrt(pnx*pnx + pny*pny + pnz*pnz);
        if (l <= 0.0) { pnx = 1.0; pny = 0.0; pnz = 0.0; }
        else { const double qq = 1.0 / l; pnx *= qq; pny *= qq; pnz *= qq; }
    }
    if (vn64) { vn64[3*v] = pnx; vn64[3*v+1] = pny; vn64[3*v+2] = pnz; }
    if (vn32) { vn32[3*v] = (float)pnx; vn32[3*v+1] = (float)pny; vn32[3*v+2] = (float)pnz; }
    if (pa64) pa64[v] = r[3];
    if (pa32) pa32[v] = (float)r[3];
    if (Hout) Hout[v] = hcnt > 0 ? 0.5 * r[6] / (double)hcnt : 0.0;
    if (Kout) Kout[v] = r[4] > 0.0 ? 3.0 * r[5] / r[4] : 0.0;
}

extern "C" size_t ms_mesh_properties_workspace_bytes(int64_t nv, int64_t nf) {
    // float4 vertices for up to 3 nf + nv rows are never needed: nv_all <= nv + 3 nf (every duplicated vertex is used by a face)
    return ms_align((size_t)(nv + 3 * nf + 1) * sizeof(float4)) + 2 * ms_align((size_t)(3 * nf + 1) * sizeof(double)) + ms_align((size_t)3 * nf + 1) + 1024;
}

extern "C" int ms_mesh_properties(const float *verts_all, int64_t nv_all, int64_t nv,
                                  const uint32_t *faces_all, int64_t nf,
                                  const int32_t *csr_off, const int32_t *csr_adj, const uint8_t *keep,
                                  double *vnormals_f64, float *vnormals_f32,
                                  double *pareas_f64, float *pareas_f32,
                                  double *mean_curv, double *gaus_curv,
                                  void *ws, size_t ws_bytes, void *stream) {
    cudaStream_t s = (cudaStream_t)stream;
    MS_REQUIRE(nv >= 0 && nv <= nv_all && nf >= 0, "ms_mesh_properties: bad sizes");
    MS_REQUIRE(nv_all <= nv + 3 * nf, "ms_mesh_properties: %lld vertices beyond the first %lld, but %lld faces can reference %lld at most",
               (long long)(nv_all - nv), (long long)nv, (long long)nf, (long long)(3 * nf));
    MS_REQUIRE(csr_off && csr_adj && keep, "ms_mesh_properties: topology missing (call ms_mesh_topology first)");
    MS_REQUIRE(ws_bytes >= ms_mesh_properties_workspace_bytes(nv, nf), "ms_mesh_properties: workspace too small");
    if (nv == 0) return 0;
    MsArena A(ws, ws_bytes);
    float4 *v4 = A.take<float4>(nv + 3 * nf + 1);
    double *al = A.take<double>(3 * nf + 1), *He = A.take<double>(3 * nf + 1);
    uint8_t *hflag = A.take<uint8_t>(3 * nf + 1);
    MS_K("pad_verts", s); k_pad_verts<<<ms_blocks(nv_all, 256), 256, 0, s>>>(verts_all, nv_all, v4);
    if (nf > 0) {
        MS_K("corner_terms", s);
        k_corner_terms<<<ms_blocks(3 * nf, 128), 128, 0, s>>>(v4, faces_all, nf, nv, csr_off, csr_adj, keep, mean_curv != NULL, al, He, hflag);
    }
    MS_K("vertex_gather8", s);
    k_vertex_gather8<<<ms_blocks(nv, VG_V), VG_T, 0, s>>>(v4, faces_all, nv, csr_off, csr_adj, keep, al, He, hflag,
                                                                     vnormals_f64, vnormals_f32, pareas_f64, pareas_f32, mean_curv, gaus_curv);
    MS_K(NULL, s);
    MS_LAUNCH_CHECK();
    return 0;
}

// ------------------------------------------------------------------ vertex neighbours and shells
// TriMesh::need_neighbors (TriMesh.cpp:197-255): the sequential face loop appends, for corner j of face i, the
// two other vertices of the face to the list of vertex faces[i][j] unless already present.  The list of a
// vertex is therefore its incident (face, corner) pairs in ascending order, each contributing (next, next-next),
// de-duplicated by first appearance -- reproduced here by one thread per vertex over the sorted CSR list.
__global__ void __launch_bounds__(256) k_corner_degree(const uint32_t *__restrict__ faces, int64_t nf, int64_t nv, int32_t *deg) {
    const int64_t f = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= nf) return;
#pragma unroll
    for (int i = 0; i < 3; i++) { const uint32_t v = faces[3*f+i]; if ((int64_t)v < nv) atomicAdd(deg + v, 1); }
}

// FILL == false: cnt[v] = number of distinct neighbours; FILL == true: write them to nbr[noff[v]..]
template <bool FILL>
__global__ void __launch_bounds__(128) k_vertex_neighbors(const uint32_t *__restrict__ faces, int64_t nv,
                                                          const int32_t *__restrict__ off, const int32_t *__restrict__ adj,
                                                          const int32_t *__restrict__ noff, int32_t *__restrict__ cnt, int32_t *__restrict__ nbr) {
    const int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= nv) return;
    const int b = off[v], e = off[v + 1];
    int32_t *mine = FILL ? nbr + noff[v] : NULL;
    int n = 0;
    for (int j = b; j < e; j++) {
        const int packed = adj[j];
        const int64_t f = packed >> 2; const int c = packed & 3;
        for (int k = 1; k <= 2; k++) {
            const int32_t u = (int32_t)faces[3*f + (c + k) % 3];
            bool seen = false;
            if (FILL) { for (int q = 0; q < n; q++) seen |= mine[q] == u; }
            else {
                // first appearance among the entries before (j, k): rescan them (lists are ~6 faces long)
                for (int jj = b; jj <= j && !seen; jj++) {
                    const int p2 = adj[jj]; const int64_t f2 = p2 >> 2; const int c2 = p2 & 3;
                    const int kend = jj < j ? 2 : k - 1;
                    for (int kk = 1; kk <= kend; kk++) seen |= (int32_t)faces[3*f2 + (c2 + kk) % 3] == u;
                }
            }
            if (!seen) { if (FILL) mine[n] = u; n++; }
        }
    }
    if (!FILL) cnt[v] = n;
}

extern "C" size_t ms_mesh_neighbors_workspace_bytes(int64_t nv, int64_t nf) {
    return 3 * ms_align((size_t)(nv + 1) * sizeof(int32_t)) + ms_align((size_t)(3 * nf + 1) * sizeof(int32_t)) + ms_scan_ws_bytes(nv) + 1024;
}

// nbr_off: (nv+1) offsets, nbr: up to 6*nf entries (2 per corner), *total_out = nbr_off[nv] (synchronises the stream)
extern "C" int ms_mesh_neighbors(const uint32_t *faces, int64_t nf, int64_t nv, int32_t *nbr_off, int32_t *nbr, int64_t *total_out,
                                 void *ws, size_t ws_bytes, void *stream) {
    cudaStream_t s = (cudaStream_t)stream;
    MS_REQUIRE(nv >= 0 && nf >= 0 && nf < ((int64_t)1 << 29) && nbr_off && total_out, "ms_mesh_neighbors: bad arguments");
    MS_REQUIRE(ws_bytes >= ms_mesh_neighbors_workspace_bytes(nv, nf), "ms_mesh_neighbors: workspace too small");
    MsArena A(ws, ws_bytes);
    int32_t *deg = A.take<int32_t>(nv + 1), *cursor = A.take<int32_t>(nv + 1), *off = A.take<int32_t>(nv + 1);
    int32_t *adj = A.take<int32_t>(3 * nf + 1);
    void *scan_ws = A.take<char>(ms_scan_ws_bytes(nv));
    MS_CUDA(cudaMemsetAsync(deg, 0, sizeof(int32_t) * (nv + 1), s));
    MS_CUDA(cudaMemsetAsync(cursor, 0, sizeof(int32_t) * (nv + 1), s));
    if (nf > 0) { MS_K("corner_degree", s); k_corner_degree<<<ms_blocks(nf, 256), 256, 0, s>>>(faces, nf, nv, deg); }
    if (ms_exclusive_scan_i32(deg, off, nv, scan_ws, s)) return 2;
    if (nf > 0) { MS_K("face_fill", s); k_face_fill<<<ms_blocks(nf, 256), 256, 0, s>>>(faces, nf, nv, off, cursor, adj); }
    if (nv > 0) {
        MS_K("sort_lists", s); k_sort_lists<<<ms_blocks(nv, 256), 256, 0, s>>>(nv, off, adj);
        MS_K("vertex_neighbors", s); k_vertex_neighbors<false><<<ms_blocks(nv, 128), 128, 0, s>>>(faces, nv, off, adj, NULL, deg, NULL);
    }
    if (ms_exclusive_scan_i32(deg, nbr_off, nv, scan_ws, s)) return 2;
    if (nv > 0 && nbr) { MS_K("vertex_neighbors", s); k_vertex_neighbors<true><<<ms_blocks(nv, 128), 128, 0, s>>>(faces, nv, off, adj, nbr_off, NULL, nbr); }
    MS_K(NULL, s);
    MS_LAUNCH_CHECK();
    int32_t tot = 0;
    MS_CUDA(cudaMemcpyAsync(&tot, nbr_off + nv, sizeof(int32_t), cudaMemcpyDeviceToHost, s));
    MS_CUDA(cudaStreamSynchronize(s));
    *total_out = tot;
    return 0;
}

// TriMesh.compute_shells (trimesh.py:364-416): breadth-first hop count from the reference vertices.  A FIFO
// breadth-first search yields the graph distance to the nearest reference vertex, so the level-synchronous
// sweep below (all vertices of level L label their unvisited neighbours L+1; concurrent writers store the same
// value) gives identical shells whatever the visiting order.
__global__ void __launch_bounds__(256) k_shell_init(int64_t nv, int32_t *shells) {
    const int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (v < nv) shells[v] = -1;
}
__global__ void __launch_bounds__(256) k_shell_seed(const int32_t *__restrict__ ref, int64_t nref, int64_t nv, int32_t *shells, int32_t *bad) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nref) return;
    const int32_t r = ref[i];
    if (r < 0 || (int64_t)r >= nv) { atomicAdd(bad, 1); return; }
    shells[r] = 0;
}
__global__ void __launch_bounds__(256) k_shell_level(const int32_t *__restrict__ off, const int32_t *__restrict__ nbr, int64_t nv, int32_t level,
                                                     int32_t *shells, int32_t *changed) {
    const int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= nv || shells[v] != level) return;
    bool any = false;
    for (int j = off[v]; j < off[v + 1]; j++) {
        const int32_t u = nbr[j];
        if (((volatile int32_t*)shells)[u] == -1) { shells[u] = level + 1; any = true; }
    }
    if (any) *changed = 1;
}
__global__ void __launch_bounds__(256) k_shell_stats(int64_t nv, const int32_t *__restrict__ shells, int32_t *stats) {
    const int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= nv) return;
    const int32_t s = shells[v];
    if (s < 0) atomicAdd(stats, 1); else atomicMax(stats + 1, s);
}

#define SHELL_BATCH 16
extern "C" size_t ms_mesh_shells_workspace_bytes(int64_t nv) { (void)nv; return 4096; }

extern "C" int ms_mesh_shells(const int32_t *nbr_off, const int32_t *nbr, int64_t nv, const int32_t *ref, int64_t nref,
                              int32_t *shells, int64_t *n_unreached, int32_t *max_shell, void *ws, size_t ws_bytes, void *stream) {
    cudaStream_t s = (cudaStream_t)stream;
    MS_REQUIRE(nv > 0 && nref > 0 && nbr_off && ref && shells, "ms_mesh_shells: bad arguments");
    MS_REQUIRE(ws_bytes >= ms_mesh_shells_workspace_bytes(nv), "ms_mesh_shells: workspace too small");
    int32_t *flags = (int32_t*)ws;                 // [0] bad seeds, [1] unreached, [2] max shell, [8..8+SHELL_BATCH) changed per level
    MS_CUDA(cudaMemsetAsync(flags, 0, 4096, s));
    const int nb = ms_blocks(nv, 256);
    MS_K("shell_init", s); k_shell_init<<<nb, 256, 0, s>>>(nv, shells);
    MS_K("shell_seed", s); k_shell_seed<<<ms_blocks(nref, 256), 256, 0, s>>>(ref, nref, nv, shells, flags);
    int32_t h[8 + SHELL_BATCH];
    for (int64_t level = 0; level < nv; level += SHELL_BATCH) {
        MS_CUDA(cudaMemsetAsync(flags + 8, 0, sizeof(int32_t) * SHELL_BATCH, s));
        for (int k = 0; k < SHELL_BATCH; k++) {
            MS_K("shell_level", s); k_shell_level<<<nb, 256, 0, s>>>(nbr_off, nbr, nv, (int32_t)(level + k), shells, flags + 8 + k);
        }
        MS_K(NULL, s);
        MS_LAUNCH_CHECK();
        MS_CUDA(cudaMemcpyAsync(h, flags, sizeof(h), cudaMemcpyDeviceToHost, s));
        MS_CUDA(cudaStreamSynchronize(s));
        MS_REQUIRE(h[0] == 0, "ms_mesh_shells: %d reference vertices are out of range", h[0]);
        bool done = false;
        for (int k = 0; k < SHELL_BATCH; k++) done |= h[8 + k] == 0;
        if (done) break;
    }
    MS_K("shell_stats", s); k_shell_stats<<<nb, 256, 0, s>>>(nv, shells, flags + 1);
    MS_K(NULL, s);
    MS_LAUNCH_CHECK();
    MS_CUDA(cudaMemcpyAsync(h, flags, sizeof(int32_t) * 8, cudaMemcpyDeviceToHost, s));
    MS_CUDA(cudaStreamSynchronize(s));
    if (n_unreached) *n_unreached = h[1];
    if (max_shell) *max_shell = h[2];
    return 0;
}
