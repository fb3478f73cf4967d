// ms_geodesic.cu -- geodesic (type 1) KDE on B200.
//
// Replaces (LLNL/MemSurfer): TriMesh::compute_geodesics_fw memsurfer/src/TriMesh_kde.cpp:196-332 (all-pairs
// shortest paths along the mesh edges by Floyd-Warshall, O(V^3) sequential) and kde_2m :394-437 (a Gaussian of
// every graph distance), followed by the normalisation of TriMesh::kde :504-524.
//
// Design: the distance matrix D (Vp x Vp fp64, Vp = V rounded up to the 32-tile, caller-allocated) stays in
// HBM; Floyd-Warshall runs as the classic three-phase tiled sweep (pivot tile; its row and column; all other
// tiles), one 32x32 tile per CTA in shared memory, 3 launches per pivot.  min-plus is not a contraction the
// tensor cores do.  The reference's k,i,j loop and the tiled sweep perform the same relaxations on shortest
// paths; sums of the same edge lengths associate differently only where a path is relaxed through a different
// intermediate first, i.e. the results agree to fp64 rounding.  The density pass reads a column of D per
// target (coalesced across targets) and evaluates exp() in fp64.
#include "ms_common.cuh"
#include <float.h>

#define GT 32

__global__ void __launch_bounds__(256) k_geo_init(double *D, int64_t np) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= np * np) return;
    D[i] = (i / np == i % np) ? 0.0 : DBL_MAX;
}

// every face edge (a,b): D[a][b] = D[b][a] = sqrt(periodic-xy 3-D squared distance), DistanceKernels.hpp:120-135
__global__ void __launch_bounds__(256) k_geo_edges(const float *__restrict__ verts, const uint32_t *__restrict__ faces, int64_t nf, int64_t nv,
                                                   int periodic, double wx, double wy, double *D, int64_t np) {
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= 3 * nf) return;
    const int64_t f = e / 3; const int d = (int)(e - 3 * f);
    const int64_t a = faces[3*f + d], b = faces[3*f + (d == 2 ? 0 : d + 1)];
    if (a >= nv || b >= nv) return;
    double ax = verts[3*a], ay = verts[3*a + 1], bx = verts[3*b], by = verts[3*b + 1];
    if (periodic) {
        if (fabs(ax - bx) >= 0.5 * wx) { if (ax < bx) ax += wx; else bx += wx; }
        if (fabs(ay - by) >= 0.5 * wy) { if (ay < by) ay += wy; else by += wy; }
    }
    const double dz = (double)verts[3*a + 2] - (double)verts[3*b + 2];
    const double val = sqrt((ax - bx) * (ax - bx) + (ay - by) * (ay - by) + dz * dz);
    D[a * np + b] = val; D[b * np + a] = val;       // the edge of the neighbouring face stores the same value
}

__device__ __forceinline__ double geo_relax(double c, double a, double b) { const double s = a + b; return s < c ? s : c; }

// phase 1: the pivot tile (kb,kb) relaxed through its own 32 intermediates
__global__ void __launch_bounds__(GT * GT) k_geo_pivot(double *D, int64_t np, int kb) {
    __shared__ double P[GT][GT + 1];
    const int tx = threadIdx.x, ty = threadIdx.y;
    double *g = D + ((int64_t)kb * GT + ty) * np + (int64_t)kb * GT + tx;
    P[ty][tx] = *g;
    __syncthreads();
    for (int k = 0; k < GT; k++) {
        const double c = geo_relax(P[ty][tx], P[ty][k], P[k][tx]);
        __syncthreads();
        P[ty][tx] = c;
        __syncthreads();
    }
    *g = P[ty][tx];
}

// phase 2: blockIdx.y == 0: tiles (kb, j) of the pivot row; == 1: tiles (i, kb) of the pivot column
__global__ void __launch_bounds__(GT * GT) k_geo_cross(double *D, int64_t np, int kb) {
    const int t = blockIdx.x;
    if (t == kb) return;
    __shared__ double P[GT][GT + 1], C[GT][GT + 1];
    const int tx = threadIdx.x, ty = threadIdx.y;
    const bool row = blockIdx.y == 0;
    const int ti = row ? kb : t, tj = row ? t : kb;
    double *g = D + ((int64_t)ti * GT + ty) * np + (int64_t)tj * GT + tx;
    P[ty][tx] = D[((int64_t)kb * GT + ty) * np + (int64_t)kb * GT + tx];
    C[ty][tx] = *g;
    __syncthreads();
    for (int k = 0; k < GT; k++) {
        const double c = row ? geo_relax(C[ty][tx], P[ty][k], C[k][tx]) : geo_relax(C[ty][tx], C[ty][k], P[k][tx]);
        __syncthreads();
        C[ty][tx] = c;
        __syncthreads();
    }
    *g = C[ty][tx];
}

// phase 3: every other tile (i,j) through the finished column tile (i,kb) and row tile (kb,j)
__global__ void __launch_bounds__(GT * GT) k_geo_rest(double *D, int64_t np, int kb) {
    const int ti = blockIdx.y, tj = blockIdx.x;
    if (ti == kb || tj == kb) return;
    __shared__ double A[GT][GT + 1], B[GT][GT + 1];
    const int tx = threadIdx.x, ty = threadIdx.y;
    A[ty][tx] = D[((int64_t)ti * GT + ty) * np + (int64_t)kb * GT + tx];
    B[ty][tx] = D[((int64_t)kb * GT + ty) * np + (int64_t)tj * GT + tx];
    double *g = D + ((int64_t)ti * GT + ty) * np + (int64_t)tj * GT + tx;
    double c = *g;
    __syncthreads();
#pragma unroll 8
    for (int k = 0; k < GT; k++) c = geo_relax(c, A[ty][k], B[k][tx]);
    *g = c;
}

// kde_2m: all vertices: sum_i k(D[i][j]^2); selected ids: sum_i k(D[ids[i]][j]) (not squared, TriMesh_kde.cpp:432)
__global__ void __launch_bounds__(128) k_geo_kde(const double *__restrict__ D, int64_t nv, int64_t np, const int32_t *__restrict__ ids, int64_t nids,
                                                 double efactor, double sfactor, double *__restrict__ dens) {
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= nv) return;
    double acc = 0.0;
    if (nids == 0) for (int64_t i = 0; i < nv; i++) { const double d = D[i * np + j]; acc += sfactor * exp(d * d * efactor); }
    else for (int64_t i = 0; i < nids; i++) acc += sfactor * exp(D[(int64_t)ids[i] * np + j] * efactor);
    dens[j] = acc;
}

// deterministic normalisation: x 1/V, then (get_counts) x n / sum -- single block, fixed order
__global__ void __launch_bounds__(1024) k_geo_norm(const double *__restrict__ dens, int64_t nv, double npoints, int get_counts,
                                                   double *__restrict__ o64, float *__restrict__ o32) {
    __shared__ double sm[1024];
    const double n1 = 1.0 / (double)nv;
    double s = 0.0;
    for (int64_t i = threadIdx.x; i < nv; i += 1024) s += dens[i] * n1;
    sm[threadIdx.x] = s;
    __syncthreads();
    for (int o = 512; o > 0; o >>= 1) { if ((int)threadIdx.x < o) sm[threadIdx.x] += sm[threadIdx.x + o]; __syncthreads(); }
    const double scale = get_counts ? npoints / sm[0] : 1.0;
    for (int64_t i = threadIdx.x; i < nv; i += 1024) {
        const double d = dens[i] * n1 * scale;
        if (o64) o64[i] = d;
        if (o32) o32[i] = (float)d;
    }
}

extern "C" int64_t ms_geodesic_pitch(int64_t nv) { return (nv + GT - 1) / GT * GT; }
extern "C" size_t ms_kde_geodesic_workspace_bytes(int64_t nv) { return ms_align((size_t)nv * sizeof(double)) + 1024; }

// D: pitch x pitch fp64 (pitch = ms_geodesic_pitch(nv)), filled here; rows/columns >= nv are padding
extern "C" int ms_geodesics(const float *verts, int64_t nv, const uint32_t *faces, int64_t nf, int periodic, const float *box6,
                            double *D, void *stream) {
    cudaStream_t s = (cudaStream_t)stream;
    MS_REQUIRE(verts && faces && D && nv > 0 && nf >= 0, "ms_geodesics: bad arguments");
    MS_REQUIRE(nv <= 46000, "ms_geodesics: %lld vertices need a %.1f GB distance matrix", (long long)nv, (double)nv * nv * 8e-9);
    MS_REQUIRE(!periodic || box6, "ms_geodesics: periodic distances need the box");
    const int64_t np = ms_geodesic_pitch(nv);
    const int nb = (int)(np / GT);
    double wx = 0.0, wy = 0.0;
    if (periodic) { wx = (double)(float)(box6[3] - box6[0]); wy = (double)(float)(box6[4] - box6[1]); }
    MS_K("geo_init", s); k_geo_init<<<ms_blocks(np * np, 256), 256, 0, s>>>(D, np);
    if (nf > 0) { MS_K("geo_edges", s); k_geo_edges<<<ms_blocks(3 * nf, 256), 256, 0, s>>>(verts, faces, nf, nv, periodic, wx, wy, D, np); }
    const dim3 tile(GT, GT);
    for (int kb = 0; kb < nb; kb++) {
        MS_K("geo_pivot", s); k_geo_pivot<<<1, tile, 0, s>>>(D, np, kb);
        MS_K("geo_cross", s); k_geo_cross<<<dim3(nb, 2), tile, 0, s>>>(D, np, kb);
        MS_K("geo_rest", s);  k_geo_rest<<<dim3(nb, nb), tile, 0, s>>>(D, np, kb);
    }
    MS_K(NULL, s);
    MS_LAUNCH_CHECK();
    return 0;
}

extern "C" int ms_kde_geodesic(const double *D, int64_t nv, const int32_t *ids, int64_t nids, float sigma, int get_counts,
                               double *density_f64, float *density_f32, void *ws, size_t ws_bytes, void *stream) {
    cudaStream_t s = (cudaStream_t)stream;
    MS_REQUIRE(D && nv > 0 && nids >= 0 && (ids || nids == 0), "ms_kde_geodesic: bad arguments");
    MS_REQUIRE(sigma > 0.0f, "ms_kde_geodesic: sigma must be positive");
    MS_REQUIRE(ws_bytes >= ms_kde_geodesic_workspace_bytes(nv), "ms_kde_geodesic: workspace too small");
    // GaussianKernel2D (DensityKernels.hpp:50-66): both constants are stored as float
    const double sg = (double)sigma;
    const double efactor = (double)(float)(-1.0 / (2.0 * sg * sg));
    const double sfactor = (double)(float)(1.0 / (2.0 * 3.14159265358979323846 * sg * sg));
    double *dens = (double*)ws;
    MS_K("geo_kde", s); k_geo_kde<<<ms_blocks(nv, 128), 128, 0, s>>>(D, nv, ms_geodesic_pitch(nv), ids, nids, efactor, sfactor, dens);
    MS_K("geo_norm", s); k_geo_norm<<<1, 1024, 0, s>>>(dens, nv, (double)(nids > 0 ? nids : nv), get_counts, density_f64, density_f32);
    MS_K(NULL, s);
    MS_LAUNCH_CHECK();
    return 0;
}
