// ms_orient.cu -- point-normal estimation, stages 4-5: consistent orientation of the PCA normals
// by a minimum spanning tree of the Riemannian graph, and the final float32 hand-off.
//
// Replaces (LLNL/MemSurfer): the CGAL::mst_orient_normals call of PointSet::need_normals
// memsurfer/src/PointSet.cpp:171-174 (algorithm: SURVEY.md Appendix A.2 -- source = max-z point,
// edge {a<b} iff b in (k+1)-NN(a), weight float(max(0,1-|na.nb|)), Prim MST from the source, flip
// propagation parent->child), PointSet::get_normals PointSet.cpp:70-88 and the global flip of
// Membrane.compute_pnormals memsurfer/membrane.py:145-152.
//
// B200 design: Prim + BFS are sequential.  The MST is unique under a strict total order on the
// edges (here (weight bits, a, slot of b in a's list)), so any MST algorithm returns the same tree:
// we run edge-centric Boruvka rounds (64-bit atomicMin per component) and carry, with every
// union-find link, the PARITY of sign flips along the tree path (link = parent | parity<<31).
// Path compression XORs parities, so after the last round every point knows its flip parity
// relative to the source without any BFS.  Built --fmad=false: weights are bit-identical to a
// sequential double evaluation.
#include "ms_common.cuh"

#define NOEDGE 0xffffffffu
#define NOBEST 0xffffffffffffffffull

__device__ __forceinline__ int lk_parent(unsigned l) { return (int)(l & 0x7fffffffu); }
__device__ __forceinline__ unsigned lk_par(unsigned l) { return l >> 31; }

// weight bits per directed slot (a, j): float(max(0, 1-|na.nb|)) if nbr[a][j] > a else NOEDGE
__global__ void __launch_bounds__(256) k_edge_weights(const int32_t *__restrict__ nbr, int64_t m, int k1,
                                                      const double *__restrict__ nrm, unsigned *__restrict__ wkey) {
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= m * k1) return;
    const int64_t a = e / k1;
    const int b = nbr[e];
    unsigned key = NOEDGE;
    if ((int64_t)b > a) {
        const double *na = nrm + 3*a, *nb = nrm + 3*(int64_t)b;
        const double dot = na[0]*nb[0] + na[1]*nb[1] + na[2]*nb[2];
        double w = 1.0 - fabs(dot); if (w < 0.0) w = 0.0;
        key = __float_as_uint((float)w);              // w >= 0: float bits are monotone as unsigned
    }
    wkey[e] = key;
}

__global__ void __launch_bounds__(256) k_orient_init(int64_t m, unsigned *link, unsigned long long *best, int *hook, uint8_t *alive) {
    const int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= m) return;
    link[v] = (unsigned)v; best[v] = NOBEST; hook[v] = -1; alive[v] = 1;
}

// every component records its minimum outgoing edge: key = weight bits << 32 | edge id (a*k1+j).
// A point whose stored edges are all internal to its component can never contribute again (components
// only grow): it is marked dead and skipped in later rounds, so the work shrinks geometrically.
__global__ void __launch_bounds__(256) k_min_edge(const int32_t *__restrict__ nbr, int64_t m, int k1,
                                                  const unsigned *__restrict__ wkey, const unsigned *__restrict__ link,
                                                  unsigned long long *best, uint8_t *alive) {
    const int64_t a = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (a >= m || !alive[a]) return;
    const int ca = lk_parent(link[a]);
    unsigned long long mine = NOBEST;
    bool any = false;
    for (int j = 0; j < k1; j++) {
        const unsigned w = wkey[a * k1 + j];
        if (w == NOEDGE) continue;
        const int b = nbr[a * k1 + j];
        const int cb = lk_parent(link[b]);
        if (cb == ca) continue;
        any = true;
        const unsigned long long key = ((unsigned long long)w << 32) | (unsigned long long)(a * k1 + j);
        if (key < mine) mine = key;
        atomicMin(best + cb, key);
    }
    if (mine != NOBEST) atomicMin(best + ca, mine);
    if (!any) alive[a] = 0;
}

// roots pick their hook: root r joins the component at the other end of its minimum edge; when two
// roots pick the same edge the smaller id stays root.  hook[r] = new parent | parity << 31.
__global__ void __launch_bounds__(256) k_hook(const int32_t *__restrict__ nbr, int64_t m, int k1,
                                              const double *__restrict__ nrm, const unsigned *__restrict__ link,
                                              const unsigned long long *__restrict__ best, int *hook, int *nhooks) {
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= m) return;
    if (lk_parent(link[r]) != (int)r) return;
    const unsigned long long key = best[r];
    if (key == NOBEST) return;
    const int64_t eid = (int64_t)(key & 0xffffffffull);
    const int a = (int)(eid / k1), b = nbr[eid];
    const unsigned la = link[a], lb = link[b];
    const int ra = lk_parent(la), rb = lk_parent(lb);
    const int other = (ra == (int)r) ? rb : ra;
    if (best[other] == key && (int)r < other) return;          // mutual pick: the smaller id stays root
    const double *na = nrm + 3*(int64_t)a, *nb = nrm + 3*(int64_t)b;
    const double dot = na[0]*nb[0] + na[1]*nb[1] + na[2]*nb[2];
    const unsigned s = dot < 0.0 ? 1u : 0u;
    hook[r] = (int)((unsigned)other | ((lk_par(la) ^ lk_par(lb) ^ s) << 31));
    atomicAdd(nhooks, 1);
}

__global__ void __launch_bounds__(256) k_apply_hooks(int64_t m, unsigned *link, int *hook, unsigned long long *best) {
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= m) return;
    best[r] = NOBEST;
    const int h = hook[r];
    if (h != -1) { link[r] = (unsigned)h; hook[r] = -1; }
}

// full path compression: every point ends up linked to its root with the path parity.
// Links are single 32-bit words, so a concurrently rewritten link is still a valid (ancestor, parity).
__global__ void __launch_bounds__(256) k_compress(int64_t m, unsigned *link) {
    const int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= m) return;
    unsigned l = ((volatile unsigned*)link)[v];
    int p = lk_parent(l); unsigned par = lk_par(l);
    if (p == (int)v) return;
    for (;;) {
        const unsigned lp = ((volatile unsigned*)link)[p];
        const int pp = lk_parent(lp);
        if (pp == p) break;
        par ^= lk_par(lp); p = pp;
    }
    link[v] = (unsigned)p | (par << 31);
}

// source = first point with maximal z (mst_find_source): max over (ordered z bits, ~index)
__global__ void __launch_bounds__(256) k_find_source(const float *__restrict__ pts, int64_t m, unsigned long long *srckey) {
    unsigned long long best = 0ull;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < m; i += (int64_t)gridDim.x * blockDim.x) {
        const unsigned u = __float_as_uint(pts[3*i+2]);
        const unsigned o = (u & 0x80000000u) ? ~u : (u | 0x80000000u);
        const unsigned long long k = ((unsigned long long)o << 32) | (unsigned long long)(0xffffffffu - (unsigned)i);
        best = k > best ? k : best;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { const unsigned long long x = __shfl_xor_sync(0xffffffffu, best, o); best = x > best ? x : best; }
    if ((threadIdx.x & 31) == 0) atomicMax(srckey, best);
}

__global__ void __launch_bounds__(256) k_apply_orientation(int64_t m, const unsigned *__restrict__ link,
                                                           const unsigned long long *__restrict__ srckey,
                                                           double *nrm, int *n_unoriented) {
    const int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= m) return;
    const int64_t src = (int64_t)(0xffffffffu - (unsigned)(*srckey & 0xffffffffull));
    const unsigned ls = link[src], lv = link[v];
    if (lk_parent(ls) != lk_parent(lv)) { atomicAdd(n_unoriented, 1); return; }
    // nrm[src].z is read before any thread may flip it only if src's own flip is decided from the
    // parity alone: flip(src) = (z < 0); flip(v) = flip(src) ^ parity(v) ^ parity(src).  Read z sign
    // through the parity-corrected value so the result does not depend on thread order.
    const double zs = nrm[3*src+2];
    // thread `src` may already have negated its normal: it does so iff zs_original < 0, after which
    // the stored value is > 0.  -0.0/0.0 are not flipped.  Hence "original z < 0" == stored z < 0
    // before the flip and never after it; we avoid the race by using a separate pass (see host).
    (void)zs;
    const unsigned flip0 = (unsigned)(*(const int*)(srckey + 1));     // host-prepared: 1 iff source z < 0
    const unsigned flip = flip0 ^ lk_par(lv) ^ lk_par(ls);
    if (flip) { nrm[3*v] = -nrm[3*v]; nrm[3*v+1] = -nrm[3*v+1]; nrm[3*v+2] = -nrm[3*v+2]; }
}

__global__ void k_source_flip(const unsigned long long *srckey, const double *nrm, int *flip0) {
    const int64_t src = (int64_t)(0xffffffffu - (unsigned)(*srckey & 0xffffffffull));
    *flip0 = nrm[3*src+2] < 0.0 ? 1 : 0;
}

extern "C" size_t ms_orient_workspace_bytes(int64_t m, int k1) {
    return ms_align((size_t)m * k1 * 4) + ms_align((size_t)m * 4) * 2 + ms_align((size_t)m * 8) + ms_align((size_t)m) + 4096;
}

extern "C" int ms_orient_normals(const float *points_all, int64_t m, const int32_t *nbr, int k1,
                                 double *normals, int64_t *n_unoriented,
                                 void *ws, size_t ws_bytes, void *stream) {
    cudaStream_t s = (cudaStream_t)stream;
    MS_REQUIRE(m > 0 && k1 >= 1 && m * (int64_t)k1 < ((int64_t)1 << 32), "ms_orient_normals: bad sizes");
    MS_REQUIRE(ws_bytes >= ms_orient_workspace_bytes(m, k1), "ms_orient_normals: workspace too small");
    MsArena A(ws, ws_bytes);
    unsigned *wkey = A.take<unsigned>((size_t)m * k1);
    unsigned *link = A.take<unsigned>(m);
    int *hook = A.take<int>(m);
    unsigned long long *best = A.take<unsigned long long>(m);
    uint8_t *alive = A.take<uint8_t>(m);
    unsigned long long *srckey = A.take<unsigned long long>(4);       // [0] key, [1] flip0 (int), [2] counters
    int *flip0 = (int*)(srckey + 1);
    int *counters = (int*)(srckey + 2);                               // [0] hooks this round, [1] unoriented
    MS_CUDA(cudaMemsetAsync(srckey, 0, 4 * sizeof(unsigned long long), s));
    const int nb = ms_blocks(m, 256);
    MS_K("edge_weights", s); k_edge_weights<<<ms_blocks(m * k1, 256), 256, 0, s>>>(nbr, m, k1, normals, wkey);
    MS_K("orient_init", s); k_orient_init<<<nb, 256, 0, s>>>(m, link, best, hook, alive);
    MS_K("find_source", s); k_find_source<<<ms_sm_count() * 4, 256, 0, s>>>(points_all, m, srckey);
    MS_K("source_flip", s); k_source_flip<<<1, 1, 0, s>>>(srckey, normals, flip0);
    MS_K(NULL, s);
    MS_LAUNCH_CHECK();
    // Boruvka needs at most ceil(log2 m) rounds (every round at least halves the number of components of
    // a connected graph).  The hook counter is read back only every other round after the third one:
    // a round without any hook is a cheap no-op, a host synchronisation is not.
    for (int round = 0; round < 64; round++) {
        MS_CUDA(cudaMemsetAsync(counters, 0, sizeof(int), s));
        MS_K("min_edge", s); k_min_edge<<<nb, 256, 0, s>>>(nbr, m, k1, wkey, link, best, alive);
        MS_K("hook", s); k_hook<<<nb, 256, 0, s>>>(nbr, m, k1, normals, link, best, hook, counters);
        MS_K("apply_hooks", s); k_apply_hooks<<<nb, 256, 0, s>>>(m, link, hook, best);
        MS_K("compress", s); k_compress<<<nb, 256, 0, s>>>(m, link);
        MS_K(NULL, s);
        MS_LAUNCH_CHECK();
        if (round < 3 || (round & 1) == 0) continue;
        int nh = 0;
        MS_CUDA(cudaMemcpyAsync(&nh, counters, sizeof(int), cudaMemcpyDeviceToHost, s));
        MS_CUDA(cudaStreamSynchronize(s));
        if (nh == 0) break;
    }
    MS_K("apply_orientation", s); k_apply_orientation<<<nb, 256, 0, s>>>(m, link, srckey, normals, counters + 1);
    MS_K(NULL, s);
    MS_LAUNCH_CHECK();
    int bad = 0;
    MS_CUDA(cudaMemcpyAsync(&bad, counters + 1, sizeof(int), cudaMemcpyDeviceToHost, s));
    MS_CUDA(cudaStreamSynchronize(s));
    if (n_unoriented) *n_unoriented = bad;
    return 0;
}

// ------------------------------------------------------------------ stage 5: float32 hand-off + global flip
// key = (|n_z| as float32 bits) << 32 | ~index  -> np.argmax(np.abs(pnormals[:,2])) (first maximum)
__global__ void __launch_bounds__(256) k_argmax_absz(const double *__restrict__ nrm, int64_t n, unsigned long long *key) {
    unsigned long long best = 0ull;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const unsigned u = __float_as_uint(fabsf((float)nrm[3*i+2]));
        const unsigned long long k = ((unsigned long long)u << 32) | (unsigned long long)(0xffffffffu - (unsigned)i);
        best = k > best ? k : best;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { const unsigned long long x = __shfl_xor_sync(0xffffffffu, best, o); best = x > best ? x : best; }
    if ((threadIdx.x & 31) == 0) atomicMax(key, best);
}

__global__ void __launch_bounds__(256) k_emit_pnormals(const double *__restrict__ nrm, int64_t n, const unsigned long long *key,
                                                       int ndir_hint, float *o32, double *o64) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int64_t imax = (int64_t)(0xffffffffu - (unsigned)(*key & 0xffffffffull));
    const float zmax = (float)nrm[3*imax+2];
    const double sgn = (zmax * (float)ndir_hint < 0.0f) ? -1.0 : 1.0;
#pragma unroll
    for (int d = 0; d < 3; d++) {
        const double x = nrm[3*i+d];
        if (o32) o32[3*i+d] = (float)x * (float)sgn;
        if (o64) o64[3*i+d] = x * sgn;
    }
}

extern "C" size_t ms_finish_workspace_bytes(int64_t n) { (void)n; return 256; }
extern "C" int ms_finish_pnormals(const double *normals, int64_t n, int ndir_hint,
                                  float *pnormals_f32, double *pnormals_f64,
                                  void *ws, size_t ws_bytes, void *stream) {
    cudaStream_t s = (cudaStream_t)stream;
    MS_REQUIRE(n > 0 && ws_bytes >= 8, "ms_finish_pnormals: bad arguments");
    MS_REQUIRE(pnormals_f64 != normals, "ms_finish_pnormals: output must not alias the input");
    unsigned long long *key = (unsigned long long*)ws;
    MS_CUDA(cudaMemsetAsync(key, 0, 8, s));
    MS_K("argmax_absz", s); k_argmax_absz<<<ms_sm_count() * 4, 256, 0, s>>>(normals, n, key);
    MS_K("emit_pnormals", s); k_emit_pnormals<<<ms_blocks(n, 256), 256, 0, s>>>(normals, n, key, ndir_hint, pnormals_f32, pnormals_f64);
    MS_K(NULL, s);
    MS_LAUNCH_CHECK();
    return 0;
}
