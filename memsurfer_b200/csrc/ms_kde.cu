// ms_kde.cu -- per-vertex Gaussian KDE densities as a tiled fp64 pair-sum on B200.
//
// Replaces (LLNL/MemSurfer): kde_2d / kde_3d memsurfer/src/TriMesh_kde.cpp:344-392 (the
// O(V*n_ids) double loop: two virtual calls + one libm exp per pair), the normalisations of
// TriMesh::kde TriMesh_kde.cpp:504-524, GaussianKernel2D/3D DensityKernels.hpp:44-84 and
// DistanceSquared / DistancePeriodicXYSquared DistanceKernels.hpp:75-136 (fused in).
//
// Design (not a contraction -> no tensor cores; bound = fp64 FMA pipe):
//   * coordinates are pre-scaled by sqrt(-efactor*log2(e)) once per field, so the pair term is
//     2^-(dx^2+dy^2[+dz^2]): no per-pair multiply by efactor, sfactor is applied once per target;
//   * single-wrap minimum image: the reference adds boxw to the smaller coordinate when
//     |dx| >= boxw/2; (|dx|-boxw)^2 == min(|dx|, boxw-|dx|)^2, which is branch-free;
//   * software exp2: round-to-nearest split with the 1.5*2^52 trick, degree-11 polynomial on
//     [-1/2,1/2] (truncation 6e-15 relative), scale 2^-n built from integer bits and folded into
//     the accumulating DFMA;
//   * source tiles are staged into shared memory by the TMA engine (cp.async.bulk + mbarrier,
//     double buffered); every thread keeps R targets in registers so one broadcast LDS.128
//     feeds R independent DFMA chains.
#include "ms_common.cuh"

#define KDE_TILE 1024           // source padding granule (entries)
#define KDE_TILE_BYTES 16384     // bytes per shared-memory stage: 1024 double2 / 512 double4
#define KDE_MAGIC 6755399441055744.0   // 1.5 * 2^52
#define KDE_CLAMP 1000.0
#define KDE_PAD 1e100

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    uint32_t ok = 0;
    while (!ok) {
        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                     : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    }
}

// 2^-r for r >= 0 (clamped), returned as (scale, poly) so the caller can fold scale*poly into an FMA
__device__ __forceinline__ void exp2neg_split(double r, double &scale, double &poly) {
    r = fmin(r, KDE_CLAMP);
    const double t = r + KDE_MAGIC;
    const int n = __double2loint(t);
    const double f = r - (t - KDE_MAGIC);             // in [-1/2, 1/2]
    // 2^-f = sum_k (-ln2)^k / k! f^k, k = 0..11 (truncation < 6.3e-15 on |f| <= 1/2)
    double p =            -4.44553827187081162e-10;
    p = fma(p, f,          7.05491162080112336e-09);
    p = fma(p, f,         -1.01780860092396999e-07);
    p = fma(p, f,          1.32154867901443095e-06);
    p = fma(p, f,         -1.52527338040598411e-05);
    p = fma(p, f,          1.54035303933816088e-04);
    p = fma(p, f,         -1.33335581464284433e-03);
    p = fma(p, f,          9.61812910762847688e-03);
    p = fma(p, f,         -5.55041086648215831e-02);
    p = fma(p, f,          2.40226506959100722e-01);
    p = fma(p, f,         -6.93147180559945286e-01);
    p = fma(p, f,          1.0);
    poly = p; scale = __hiloint2double((1023 - n) << 20, 0);
}

// ------------------------------------------------------------------ staging of scaled coordinates
template <int DIM> struct SrcT;
template <> struct SrcT<2> { typedef double2 type; };
template <> struct SrcT<3> { typedef double4 type; };

// tgt[j] = scaled vertex j (all vertices); src[i] = scaled vertex ids[i] (or i), padded to a
// multiple of KDE_TILE with far-away sentinels whose contribution underflows to ~2^-1000.
template <int DIM>
__global__ void __launch_bounds__(256) k_kde_stage(const float *__restrict__ verts, int64_t nv,
                                                   const int32_t *__restrict__ ids, int64_t ns, int64_t ns_pad,
                                                   double scale, typename SrcT<DIM>::type *tgt,
                                                   typename SrcT<DIM>::type *src) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    typename SrcT<DIM>::type o;
    if (i < nv) {
        o.x = ms_ld_f(verts + 3*i) * scale; o.y = ms_ld_f(verts + 3*i + 1) * scale;
        if (DIM == 3) { ((double*)&o)[2] = ms_ld_f(verts + 3*i + 2) * scale; ((double*)&o)[3] = 0.0; }
        tgt[i] = o;
    }
    if (i < ns_pad) {
        if (i < ns) {
            const int64_t s = ids ? (int64_t)__ldg(ids + i) : i;
            o.x = ms_ld_f(verts + 3*s) * scale; o.y = ms_ld_f(verts + 3*s + 1) * scale;
            if (DIM == 3) { ((double*)&o)[2] = ms_ld_f(verts + 3*s + 2) * scale; ((double*)&o)[3] = 0.0; }
        } else {
            o.x = KDE_PAD; o.y = KDE_PAD;
            if (DIM == 3) { ((double*)&o)[2] = 0.0; ((double*)&o)[3] = 0.0; }
        }
        src[i] = o;
    }
}

// ------------------------------------------------------------------ the pair-sum
// grid.x: target blocks (THREADS*R targets each); grid.y: source splits (tiles [y*tps,(y+1)*tps)).
// partial[y*nv + j] = sum over the split's sources of 2^-(scaled distance^2).
template <int DIM, bool PERIODIC, int THREADS, int R>
__global__ void __launch_bounds__(THREADS) k_kde_pairs(const typename SrcT<DIM>::type *__restrict__ tgt, int64_t nv,
                                                       const typename SrcT<DIM>::type *__restrict__ src, int ntiles, int tiles_per_split,
                                                       double wx, double wy, double *__restrict__ partial) {
    typedef typename SrcT<DIM>::type S;
    constexpr int TILE = KDE_TILE_BYTES / (int)sizeof(S);
    __shared__ __align__(128) S tile[2][TILE];
    __shared__ __align__(8) uint64_t bar[2];
    const int tid = threadIdx.x;
    const int t_begin = blockIdx.y * tiles_per_split;
    const int t_end = min(ntiles, t_begin + tiles_per_split);
    const int64_t j0 = (int64_t)blockIdx.x * (THREADS * R) + tid;

    double tx[R], ty[R], tz[R], acc[R];
#pragma unroll
    for (int r = 0; r < R; r++) {
        const int64_t j = j0 + (int64_t)r * THREADS;
        S q; q.x = KDE_PAD * 4; q.y = KDE_PAD * 4; if (DIM == 3) { ((double*)&q)[2] = 0.0; }
        if (j < nv) q = tgt[j];
        tx[r] = q.x; ty[r] = q.y; tz[r] = DIM == 3 ? ((double*)&q)[2] : 0.0; acc[r] = 0.0;
    }
    if (tid == 0) { mbar_init(&bar[0], 1); mbar_init(&bar[1], 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
    __syncthreads();
    const uint32_t tile_bytes = KDE_TILE_BYTES;
    if (tid == 0 && t_begin < t_end) { mbar_expect_tx(&bar[0], tile_bytes); bulk_g2s(tile[0], src + (size_t)t_begin * TILE, tile_bytes, &bar[0]); }
    uint32_t phase[2] = { 0u, 0u };
    for (int t = t_begin; t < t_end; t++) {
        const int s = (t - t_begin) & 1;
        if (tid == 0 && t + 1 < t_end) {   // buffer s^1 was released by the __syncthreads closing iteration t-1
            mbar_expect_tx(&bar[s ^ 1], tile_bytes);
            bulk_g2s(tile[s ^ 1], src + (size_t)(t + 1) * TILE, tile_bytes, &bar[s ^ 1]);
        }
        mbar_wait(&bar[s], phase[s]); phase[s] ^= 1u;
        const S *T = tile[s];
#pragma unroll 2
        for (int i = 0; i < TILE; i++) {
            const S sp = T[i];
#pragma unroll
            for (int r = 0; r < R; r++) {
                double dx = sp.x - tx[r], dy = sp.y - ty[r];
                if (PERIODIC) {
                    const double ax = fabs(dx), ay = fabs(dy);
                    dx = fmin(ax, wx - ax); dy = fmin(ay, wy - ay);
                }
                double r2 = fma(dy, dy, dx * dx);
                if (DIM == 3) { const double dz = ((const double*)&sp)[2] - tz[r]; r2 = fma(dz, dz, r2); }
                double sc, po; exp2neg_split(r2, sc, po);
                acc[r] = fma(sc, po, acc[r]);
            }
        }
        __syncthreads();
    }
#pragma unroll
    for (int r = 0; r < R; r++) {
        const int64_t j = j0 + (int64_t)r * THREADS;
        if (j < nv) partial[(size_t)blockIdx.y * nv + j] = acc[r];
    }
}

// ------------------------------------------------------------------ normalisation (TriMesh_kde.cpp:504-524)
#define RED_T 256
__global__ void __launch_bounds__(RED_T) k_kde_combine(const double *__restrict__ partial, int64_t nv, int nsplit,
                                                       double factor, double *__restrict__ dens, double *__restrict__ bsum) {
    __shared__ double sm[RED_T];
    const int64_t j = (int64_t)blockIdx.x * RED_T + threadIdx.x;
    double d = 0.0;
    if (j < nv) {
        for (int y = 0; y < nsplit; y++) d += partial[(size_t)y * nv + j];
        d *= factor;
        dens[j] = d;
    }
    sm[threadIdx.x] = d;
    __syncthreads();
    for (int o = RED_T / 2; o > 0; o >>= 1) { if (threadIdx.x < o) sm[threadIdx.x] += sm[threadIdx.x + o]; __syncthreads(); }
    if (threadIdx.x == 0) bsum[blockIdx.x] = sm[0];
}
__global__ void __launch_bounds__(RED_T) k_kde_total(const double *__restrict__ bsum, int64_t nb, double *total) {
    __shared__ double sm[RED_T];
    double s = 0.0;
    for (int64_t i = threadIdx.x; i < nb; i += RED_T) s += bsum[i];
    sm[threadIdx.x] = s;
    __syncthreads();
    for (int o = RED_T / 2; o > 0; o >>= 1) { if (threadIdx.x < o) sm[threadIdx.x] += sm[threadIdx.x + o]; __syncthreads(); }
    if (threadIdx.x == 0) *total = sm[0];
}
__global__ void __launch_bounds__(256) k_kde_emit(const double *__restrict__ dens, int64_t nv, const double *total, double npoints,
                                                  double *__restrict__ o64, float *__restrict__ o32) {
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= nv) return;
    double d = dens[j];
    if (total) d *= npoints / *total;
    if (o64) o64[j] = d;
    if (o32) o32[j] = (float)d;
}

struct KdePlan { int64_t ns, ns_pad; int threads, R, nsplit; int64_t blocks_x; };
static KdePlan kde_plan(int64_t nv, int64_t nids) {
    KdePlan p;
    p.ns = nids > 0 ? nids : nv;
    p.ns_pad = (p.ns + KDE_TILE - 1) / KDE_TILE * KDE_TILE;
    const bool big = nv >= (int64_t)MS_SM_COUNT * 128 * 4 * 2;
    p.threads = big ? 128 : 64; p.R = big ? 4 : 1;
    p.blocks_x = (nv + p.threads * p.R - 1) / (p.threads * p.R);
    const int64_t ntiles = p.ns_pad / 512;                      // smallest tile (3-D)
    int64_t want = (MS_SM_COUNT * 8 + p.blocks_x - 1) / (p.blocks_x > 0 ? p.blocks_x : 1);
    if (want < 1) want = 1; if (want > 64) want = 64; if (want > ntiles) want = ntiles > 0 ? ntiles : 1;
    p.nsplit = (int)want;
    return p;
}

extern "C" size_t ms_kde_workspace_bytes(int64_t nv, int64_t nids) {
    const KdePlan p = kde_plan(nv, nids);
    return ms_align((size_t)nv * 32) + ms_align((size_t)p.ns_pad * 32) + ms_align((size_t)p.nsplit * nv * 8)
         + ms_align((size_t)nv * 8) + ms_align(((size_t)nv / RED_T + 2) * 8) + 4096;
}

template <int DIM, bool PER>
static int kde_run(const float *verts, int64_t nv, const int32_t *ids, const KdePlan &p, double scale, double wx, double wy,
                   void *tgt_, void *src_, double *partial, cudaStream_t s) {
    typedef typename SrcT<DIM>::type S;
    S *tgt = (S*)tgt_, *src = (S*)src_;
    const int64_t nstage = nv > p.ns_pad ? nv : p.ns_pad;
    MS_K("kde_stage", s); k_kde_stage<DIM><<<ms_blocks(nstage, 256), 256, 0, s>>>(verts, nv, ids, p.ns, p.ns_pad, scale, tgt, src);
    const int tile = KDE_TILE_BYTES / (int)sizeof(S);
    const int ntiles = (int)(p.ns_pad / tile);
    const int tps = (ntiles + p.nsplit - 1) / p.nsplit;
    dim3 grid((unsigned)p.blocks_x, (unsigned)p.nsplit);
    if (p.R == 4) { MS_K("kde_pairs", s); k_kde_pairs<DIM, PER, 128, 4><<<grid, 128, 0, s>>>(tgt, nv, src, ntiles, tps, wx, wy, partial); }
    else          { MS_K("kde_pairs", s); k_kde_pairs<DIM, PER, 64, 1><<<grid, 64, 0, s>>>(tgt, nv, src, ntiles, tps, wx, wy, partial); }
    MS_K(NULL, s);
    MS_LAUNCH_CHECK();
    return 0;
}

extern "C" int ms_kde(const float *verts, int64_t nv, const int32_t *ids, int64_t nids,
                      int type, int periodic, const float *box6, float sigma, int get_counts,
                      double cutoff_sigmas, double *density_f64, float *density_f32,
                      void *ws, size_t ws_bytes, void *stream) {
    cudaStream_t s = (cudaStream_t)stream;
    MS_REQUIRE(type == 2 || type == 3, "ms_kde: invalid density type %d (should be 2 (2D) or 3 (3D); geodesic type 1 is not on the GPU path)", type);
    MS_REQUIRE(nv > 0 && nids >= 0 && sigma > 0.0f, "ms_kde: bad arguments nv=%lld nids=%lld sigma=%g", (long long)nv, (long long)nids, (double)sigma);
    MS_REQUIRE(!periodic || box6 != NULL, "ms_kde: periodic KDE needs the 6-float bbox");
    MS_REQUIRE(!(cutoff_sigmas > 0.0), "ms_kde: cut-off variant not built yet; pass cutoff_sigmas <= 0");
    MS_REQUIRE(ws_bytes >= ms_kde_workspace_bytes(nv, nids), "ms_kde: workspace too small");
    const KdePlan p = kde_plan(nv, nids);
    // constants rounded through float32 exactly as DensityKernels.hpp:47-48 stores them
    const double sg = (double)sigma;
    const float efactor = (float)(-1.0 / (2.0 * sg * sg));
    const float sfactor = type == 3 ? (float)(1.0 / pow(sqrt(2.0 * M_PI) * sg, 3.0)) : (float)(1.0 / (2.0 * M_PI * sg * sg));
    const double scale = sqrt(-(double)efactor * 1.4426950408889634074);
    double wx = 0.0, wy = 0.0;
    if (periodic) { wx = (double)(float)(box6[3] - box6[0]) * scale; wy = (double)(float)(box6[4] - box6[1]) * scale; }
    MsArena A(ws, ws_bytes);
    void *tgt = A.take<char>((size_t)nv * 32), *src = A.take<char>((size_t)p.ns_pad * 32);
    double *partial = A.take<double>((size_t)p.nsplit * nv), *dens = A.take<double>(nv);
    double *bsum = A.take<double>(nv / RED_T + 2);
    int rc;
    if (type == 2) rc = periodic ? kde_run<2, true>(verts, nv, nids > 0 ? ids : NULL, p, scale, wx, wy, tgt, src, partial, s)
                                 : kde_run<2, false>(verts, nv, nids > 0 ? ids : NULL, p, scale, wx, wy, tgt, src, partial, s);
    else           rc = periodic ? kde_run<3, true>(verts, nv, nids > 0 ? ids : NULL, p, scale, wx, wy, tgt, src, partial, s)
                                 : kde_run<3, false>(verts, nv, nids > 0 ? ids : NULL, p, scale, wx, wy, tgt, src, partial, s);
    if (rc) return rc;
    const int nb = ms_blocks(nv, RED_T);
    MS_K("kde_combine", s); k_kde_combine<<<nb, RED_T, 0, s>>>(partial, nv, p.nsplit, (double)sfactor * (1.0 / (double)nv), dens, bsum);
    double *total = bsum + nb;
    if (get_counts) { MS_K("kde_total", s); k_kde_total<<<1, RED_T, 0, s>>>(bsum, nb, total); }
    MS_K("kde_emit", s); k_kde_emit<<<ms_blocks(nv, 256), 256, 0, s>>>(dens, nv, get_counts ? total : NULL, (double)p.ns, density_f64, density_f32);
    MS_K(NULL, s);
    MS_LAUNCH_CHECK();
    return 0;
}
