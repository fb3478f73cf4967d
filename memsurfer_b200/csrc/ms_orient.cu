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
// Boruvka rounds with 64-bit atomicMin per component, and every union-find link carries the PARITY of sign
// flips along the tree path (link = parent | parity << 31), so that after the last round every point knows
// its flip relative to the source without any BFS.
//
// The whole orientation is ONE cooperative kernel (one CTA of 1024 threads per SM, grid-wide barriers between
// phases): the number of rounds is decided on the device, the host neither reads anything back nor launches
// per-round kernels.  Work shrinks with the graph: round 0 walks the neighbour slots once and writes every edge as a
// (key, root a, root b) record; every later round re-roots the records it still has, drops the ones that became
// internal to a component and keeps the rest, so round r costs what is left of the graph, not m x k.  Every WARP
// owns a fixed segment of the record array (and of the root list) for the whole kernel and compacts it in place
// with a ballot: no append counters, no block barriers, the segment length lives in a register across the grid
// barriers.  Several records per lane are in flight at once (the first union-find hop of all of them is issued
// back to back).  A hooked root's link is compressed once, in the round after it hooked.  Built --fmad=false:
// weights are bit-identical to a sequential double evaluation.
#include "ms_common.cuh"
#include <cooperative_groups.h>
namespace cg = cooperative_groups;

#define NOBEST 0xffffffffffffffffull
#define OR_T 1024
#define OR_NONE 0xffffffffu
#define OR_U 2                         // slots per lane in flight (round 0)
#ifndef OR_UA
#define OR_UA 2                        // records per lane in flight (later rounds)
#endif
#define OR_MAX_WARPS 8192
#define OR_CHUNK 64                     // records per chunk of a warp's block-cyclic segment (= 32 * OR_U = 32 * OR_UA)

__device__ __forceinline__ int lk_parent(unsigned l) { return (int)(l & 0x7fffffffu); }
__device__ __forceinline__ unsigned lk_par(unsigned l) { return l >> 31; }

struct OrEdge { unsigned long long key; int ca, cb; };          // 16 bytes: (weight bits << 32 | edge id), roots of both ends

struct OrArgs {
    const float *pts; const int32_t *nbr; double *nrm; const int64_t *info_in; int64_t *info_out;
    int64_t m_cap; int k1;
    uint2 *lh;                         // [m_cap] .x = link (parent | parity << 31), .y = hook chosen in the current round (~0: none)
    unsigned long long *best[2];       // [m_cap] minimum outgoing edge of a root, double buffered by round parity
    OrEdge *E;                         // [m_cap * k1 + OR_CHUNK * warps] edge records, block-cyclic chunks per warp
    int *R;                            // [m_cap + warps] root list, one segment per warp
    unsigned *cnt;                     // [0..1] edges left after a round (by round parity), [4] unoriented
    unsigned long long *srckey;        // [0] (ordered z bits << 32 | ~index) of the source, [1] flip of the source
    unsigned long long *times;         // [64] globaltimer at the phase boundaries (diagnostic)
};

// one union-find hop: links are the compressed hooks of earlier rounds; a self-linked node may carry the hook it
// chose in the round that is being applied right now.  Every stored word is a valid (ancestor, parity).
__device__ __forceinline__ bool chase_hop(const uint2 v, int &x, unsigned &par) {
    const int p = lk_parent(v.x);
    if (p != x) { par ^= lk_par(v.x); x = p; return false; }
    if (v.y == OR_NONE) return true;
    par ^= lk_par(v.y); x = lk_parent(v.y);
    return false;
}
__device__ __forceinline__ int chase_root(const uint2 *lh, int x, unsigned &par) {
    for (;;) { const uint2 v = __ldcg(lh + x); if (chase_hop(v, x, par)) return x; }
}
// N chains at once: the loads of one hop of all unfinished chains are issued back to back
template <int N>
__device__ __forceinline__ void chase_roots(const uint2 *lh, int (&x)[N], const bool (&live)[N]) {
    bool busy[N];
#pragma unroll
    for (int c = 0; c < N; c++) busy[c] = live[c];
    for (;;) {
        uint2 v[N];
#pragma unroll
        for (int c = 0; c < N; c++) if (busy[c]) v[c] = __ldcg(lh + x[c]);
        bool any = false;
        unsigned par = 0u;
#pragma unroll
        for (int c = 0; c < N; c++) if (busy[c]) { busy[c] = !chase_hop(v[c], x[c], par); any |= busy[c]; }
        if (!any) break;
    }
}
// links only: the state BEFORE the hooks of the current round
__device__ __forceinline__ int chase_link(const uint2 *lh, int x, unsigned &par) {
    for (;;) {
        const unsigned l = __ldcg(&lh[x].x);
        const int p = lk_parent(l);
        if (p == x) return x;
        par ^= lk_par(l); x = p;
    }
}

__device__ __forceinline__ void hook_step(const OrArgs &A, int r, const unsigned long long *best) {
    const unsigned long long key = __ldcg(best + r);
    if (key == NOBEST) return;
    const unsigned eid = (unsigned)(key & 0xffffffffull);
    const int a = (int)(eid / (unsigned)A.k1), b = __ldcg(A.nbr + eid);
    unsigned pa = 0u, pb = 0u;
    const int ra = chase_link(A.lh, a, pa), rb = chase_link(A.lh, b, pb);
    const int other = (ra == r) ? rb : ra;
    if (__ldcg(best + other) == key && r < other) return;       // mutual pick: the smaller id stays root
    const double *na = A.nrm + 3 * (int64_t)a, *nb = A.nrm + 3 * (int64_t)b;
    const double dot = __ldcg(na)*__ldcg(nb) + __ldcg(na + 1)*__ldcg(nb + 1) + __ldcg(na + 2)*__ldcg(nb + 2);
    const unsigned s = dot < 0.0 ? 1u : 0u;
    A.lh[r].y = (unsigned)other | ((pa ^ pb ^ s) << 31);
}

// C part of a round: a root of the previous round either compresses its fresh hook into its link or, if it is
// still a root with something left to do, stays in the root list (and gets its `best` slot reset).
__device__ __forceinline__ bool settle_root(const OrArgs &A, int r, unsigned long long *best_prev) {
    const uint2 v = __ldcg(A.lh + r);
    if (v.y != OR_NONE) {
        unsigned par = 0u;
        const int root = chase_root(A.lh, r, par);
        A.lh[r].x = (unsigned)root | (par << 31);
        return false;
    }
    if (__ldcg(best_prev + r) == NOBEST) return false;           // no outgoing edge: this component is complete
    best_prev[r] = NOBEST;
    return true;
}

// atomicMin(best[c], key) for every lane with `on`, but ONE atomic per run of consecutive lanes with the same c
// (records stay sorted by their first end point, so runs are long): segmented min by shuffles.
#ifndef OR_PEEK
#define OR_PEEK 0
#endif
__device__ __forceinline__ void run_min(unsigned long long *best, int c, unsigned long long key, bool on, int lane) {
    if (!on) { c = -1 - lane; key = NOBEST; }                    // a run of its own, never issued
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned long long k2 = __shfl_down_sync(0xffffffffu, key, o);
        const int c2 = __shfl_down_sync(0xffffffffu, c, o);
        if (lane + o < 32 && c2 == c && k2 < key) key = k2;
    }
    const int cp = __shfl_up_sync(0xffffffffu, c, 1);
    if (on && (lane == 0 || cp != c)) {
        if (OR_PEEK && __ldcg(best + c) <= key) return;
        atomicMin(best + c, key);
    }
}
__device__ __forceinline__ void one_min(unsigned long long *best, int c, unsigned long long key) {
#ifdef OR_EXP_NOATOM
    if (key != 12345ull) return;
#endif
    if (OR_PEEK && __ldcg(best + c) <= key) return;              // `best` only ever decreases: a stale read is conservative
    atomicMin(best + c, key);
}

__device__ __forceinline__ unsigned long long or_now() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }

__global__ void __launch_bounds__(OR_T, 1) k_orient(OrArgs A) {
    cg::grid_group grid = cg::this_grid();
    const int64_t m = A.info_in ? min(A.info_in[MS_INFO_M], A.m_cap) : A.m_cap;
    const int64_t T = (int64_t)gridDim.x * OR_T, tid = (int64_t)blockIdx.x * OR_T + threadIdx.x;
    const int k1 = A.k1, lane = threadIdx.x & 31;
    const int64_t nw = T >> 5, gw = tid >> 5;                    // warps in the grid, this warp
    const unsigned lt = (1u << lane) - 1u;
    int tmark = 0;
#define OR_MARK() do { if (tid == 0 && tmark < 62) A.times[tmark] = or_now(); tmark++; } while (0)
    OR_MARK();

    // ---- init + source = first point with maximal z (mst_find_source): max over (ordered z bits, ~index)
    {
        unsigned long long bestk = 0ull;
        for (int64_t v = tid; v < m; v += T) {
            A.lh[v] = make_uint2((unsigned)v, OR_NONE);
            A.best[0][v] = NOBEST; A.best[1][v] = NOBEST;
            const unsigned u = __float_as_uint(A.pts[3*v+2]);
            const unsigned o = (u & 0x80000000u) ? ~u : (u | 0x80000000u);
            const unsigned long long k = ((unsigned long long)o << 32) | (unsigned long long)(0xffffffffu - (unsigned)v);
            bestk = k > bestk ? k : bestk;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) { const unsigned long long x = __shfl_xor_sync(0xffffffffu, bestk, o); bestk = x > bestk ? x : bestk; }
        if (lane == 0 && bestk) atomicMax(A.srckey, bestk);
    }
    grid.sync();
    OR_MARK();

    // ---- round 0, A part: every point is its own component, every stored edge crosses.  The warp walks its share
    // of the neighbour slots and writes the edges (b > a) compacted into its segment of E.
    // Ownership is BLOCK-CYCLIC: warp w owns the chunks w, w + nw, w + 2 nw, ... of OR_CHUNK slots / records, so that
    // at any moment the warps of the grid read and write neighbouring kilobytes (4736 private contiguous streams
    // thrash the DRAM row buffers: 2.3 TB/s) and, after compaction, the live records of all warps sit together at
    // the front of the array.  Logical record j of this warp lives at E[phys(j)].
    const int64_t ns = m * k1;
    OrEdge *const Ew = A.E + gw * OR_CHUNK;
    const int64_t cstride = nw * OR_CHUNK;
#define OR_PHYS(j) ((int64_t)((j) >> 6) * cstride + ((j) & 63))
    unsigned ecount = 0;                                         // records in this warp's segment
    {
        for (int64_t s0 = gw * OR_CHUNK; s0 < ns; s0 += cstride) {
            int b[OR_U], a[OR_U]; int64_t e[OR_U]; bool ok[OR_U];
#pragma unroll
            for (int u = 0; u < OR_U; u++) { e[u] = s0 + u * 32 + lane; ok[u] = e[u] < ns; b[u] = ok[u] ? A.nbr[e[u]] : -1; }
            double nb[OR_U][3], na[OR_U][3];
#pragma unroll
            for (int u = 0; u < OR_U; u++) {
                a[u] = ok[u] ? (int)((unsigned)e[u] / (unsigned)k1) : 0;
                ok[u] = ok[u] && b[u] > a[u];
                if (ok[u]) {
                    // random gathers go past L1 (measured: plain loads are 2.8x slower here)
                    const double *pa = A.nrm + 3 * (int64_t)a[u], *pb = A.nrm + 3 * (int64_t)b[u];
                    na[u][0] = pa[0]; na[u][1] = pa[1]; na[u][2] = pa[2];
                    nb[u][0] = __ldcg(pb); nb[u][1] = __ldcg(pb + 1); nb[u][2] = __ldcg(pb + 2);
                }
            }
#pragma unroll
            for (int u = 0; u < OR_U; u++) {
                OrEdge ed; ed.key = 0ull; ed.ca = a[u]; ed.cb = b[u];
                if (ok[u]) {
                    const double dot = na[u][0]*nb[u][0] + na[u][1]*nb[u][1] + na[u][2]*nb[u][2];
                    double wt = 1.0 - fabs(dot); if (wt < 0.0) wt = 0.0;
                    const unsigned w = __float_as_uint((float)wt);          // wt >= 0: float bits are monotone as unsigned
                    ed.key = ((unsigned long long)w << 32) | (unsigned long long)e[u];
                    one_min(A.best[0], b[u], ed.key);
                }
                run_min(A.best[0], a[u], ed.key, ok[u], lane);
                const unsigned mask = __ballot_sync(0xffffffffu, ok[u]);
                if (ok[u]) { const unsigned j = ecount + __popc(mask & lt); Ew[OR_PHYS(j)] = ed; }
                ecount += __popc(mask);
            }
        }
    }
    if (tid == 0) {
        const int64_t src = (int64_t)(0xffffffffu - (unsigned)(*A.srckey & 0xffffffffull));
        A.srckey[1] = (m > 0 && A.nrm[3*src+2] < 0.0) ? 1ull : 0ull;
    }
    grid.sync();
    OR_MARK();
    // ---- round 0, B part: every point picks its hook.  The warp's share of the points is also its root segment.
    const int64_t Sr = (m + nw - 1) / nw;
    int *Rseg = A.R + gw * Sr;
    unsigned rcount = 0;
    {
        const int64_t r_begin = gw * Sr, r_end = min(m, r_begin + Sr);
        for (int64_t r = r_begin + lane; r < r_end; r += 32) hook_step(A, (int)r, A.best[0]);
    }
    grid.sync();
    OR_MARK();

    unsigned rounds = 1;
    for (int round = 1; round < 64; round++, rounds++) {
        const int cur = round & 1, prev = cur ^ 1;
        // ---- C part: settle the roots of the previous round, keep the live ones (in place)
        {
            const int64_t r_begin = gw * Sr;
            const unsigned n_in = round == 1 ? (unsigned)max((int64_t)0, min(m, r_begin + Sr) - r_begin) : rcount;
            unsigned w = 0;
            for (unsigned i = 0; i < n_in; i += 32) {
                const bool valid = i + lane < n_in;
                int r = 0; bool keep = false;
                if (valid) { r = round == 1 ? (int)(r_begin + i + lane) : Rseg[i + lane]; keep = settle_root(A, r, A.best[prev]); }
                const unsigned mask = __ballot_sync(0xffffffffu, keep);
                if (keep) Rseg[w + __popc(mask & lt)] = r;
                w += __popc(mask);
            }
            rcount = w;
        }
        grid.sync();                                              // every link now points straight at a current root
        OR_MARK();
        // ---- A part: re-root the records (ONE load per end: every link is stable and points at a current root),
        // drop the internal ones, minimum per component of the rest.  Three-stage software pipeline per warp: the
        // records of iteration i+2 are streaming in and the link gathers of iteration i+1 are in flight while
        // iteration i is reduced and written back, so that neither latency sits on the loop-carried path.
        {
            const unsigned n_in = ecount;
            const unsigned *lk = (const unsigned*)A.lh;
            unsigned w = 0;
            OrEdge r0[OR_UA], r1[OR_UA]; unsigned la1[OR_UA], lb1[OR_UA]; bool v0[OR_UA], v1[OR_UA];
#define OR_LOAD(i_, r_, v_) _Pragma("unroll") for (int u = 0; u < OR_UA; u++) { const unsigned j = (i_) + u * 32 + lane; \
                                v_[u] = j < n_in; if (v_[u]) r_[u] = Ew[OR_PHYS(j)]; }
#define OR_GATHER(r_, v_, la_, lb_) _Pragma("unroll") for (int u = 0; u < OR_UA; u++) if (v_[u]) { \
                                la_[u] = __ldcg(lk + 2 * (size_t)r_[u].ca); lb_[u] = __ldcg(lk + 2 * (size_t)r_[u].cb); }
            OR_LOAD(0u, r1, v1)
            OR_GATHER(r1, v1, la1, lb1)
            OR_LOAD(32u * OR_UA, r0, v0)
            for (unsigned i = 0; i < n_in; i += 32 * OR_UA) {
                OrEdge ed[OR_UA]; unsigned la[OR_UA], lb[OR_UA]; bool valid[OR_UA];
#pragma unroll
                for (int u = 0; u < OR_UA; u++) { ed[u] = r1[u]; la[u] = la1[u]; lb[u] = lb1[u]; valid[u] = v1[u]; r1[u] = r0[u]; v1[u] = v0[u]; }
                OR_GATHER(r1, v1, la1, lb1)
                OR_LOAD(i + 64u * OR_UA, r0, v0)
#pragma unroll
                for (int u = 0; u < OR_UA; u++) {
                    const int ca = valid[u] ? lk_parent(la[u]) : 0, cb = valid[u] ? lk_parent(lb[u]) : 0;
                    const bool keep = valid[u] && ca != cb;
                    if (keep) { ed[u].ca = ca; ed[u].cb = cb; one_min(A.best[cur], cb, ed[u].key); }
                    run_min(A.best[cur], ca, keep ? ed[u].key : NOBEST, keep, lane);
                    const unsigned mask = __ballot_sync(0xffffffffu, keep);
                    if (keep) { const unsigned j = w + __popc(mask & lt); Ew[OR_PHYS(j)] = ed[u]; }     // j <= i + 32 u + lane: behind every read
                    w += __popc(mask);
                }
            }
            ecount = w;
            if (lane == 0 && w) atomicAdd(A.cnt + cur, w);
        }
        grid.sync();
        OR_MARK();
        const unsigned left = A.cnt[cur];                     // edges that still cross two components
        if (left == 0u) break;
        // ---- B part: the remaining roots pick their hooks; the counter of the finished round is recycled
        if (tid == 0) A.cnt[prev] = 0u;
        for (unsigned i = lane; i < rcount; i += 32) hook_step(A, Rseg[i], A.best[cur]);
        grid.sync();
        OR_MARK();
    }

    // ---- orientation: flip(v) = flip(source) ^ parity(v -> root) ^ parity(source -> root); points outside the
    // source's component are counted as unoriented and left as they are
    {
        const int64_t src = (int64_t)(0xffffffffu - (unsigned)(*A.srckey & 0xffffffffull));
        const unsigned flip0 = (unsigned)A.srckey[1];
        unsigned ps = 0u;
        const int rs = m > 0 ? chase_link(A.lh, (int)src, ps) : -1;
        unsigned bad = 0u;
        for (int64_t v = tid; v < m; v += T) {
            unsigned pv = 0u;
            const int rv = chase_link(A.lh, (int)v, pv);
            if (rv != rs) { bad++; continue; }
            if (flip0 ^ pv ^ ps) { A.nrm[3*v] = -A.nrm[3*v]; A.nrm[3*v+1] = -A.nrm[3*v+1]; A.nrm[3*v+2] = -A.nrm[3*v+2]; }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) bad += __shfl_xor_sync(0xffffffffu, bad, o);
        if (lane == 0 && bad) atomicAdd(A.cnt + 4, bad);
    }
    grid.sync();
    OR_MARK();
    if (tid == 0) {
        A.times[63] = (unsigned long long)tmark;
        if (A.info_out) { A.info_out[MS_INFO_UNORIENTED] = (int64_t)A.cnt[4]; A.info_out[MS_INFO_MST_ROUNDS] = (int64_t)rounds; }
    }
}

extern "C" size_t ms_orient_workspace_bytes(int64_t m, int k1) {
    return ms_align((size_t)m * 8) + 2 * ms_align((size_t)m * 8) + ms_align(((size_t)m * k1 + (size_t)OR_CHUNK * OR_MAX_WARPS) * sizeof(OrEdge))
         + ms_align(((size_t)m + OR_MAX_WARPS) * 4) + 4096;
}

static int orient_arena(OrArgs &A, int64_t m_cap, int k1, void *ws, size_t ws_bytes) {
    MsArena W(ws, ws_bytes);
    A.lh = W.take<uint2>(m_cap);
    A.best[0] = W.take<unsigned long long>(m_cap); A.best[1] = W.take<unsigned long long>(m_cap);
    A.E = W.take<OrEdge>((size_t)m_cap * k1 + (size_t)OR_CHUNK * OR_MAX_WARPS);
    A.R = W.take<int>((size_t)m_cap + OR_MAX_WARPS);
    A.cnt = W.take<unsigned>(16);
    A.srckey = W.take<unsigned long long>(4);
    A.times = W.take<unsigned long long>(64);
    return W.ok() ? 0 : 1;
}

extern "C" int ms_orient_normals(const float *points_all, int64_t m_cap, int64_t *info, const int32_t *nbr, int k1,
                                 double *normals, void *ws, size_t ws_bytes, void *stream) {
    cudaStream_t s = (cudaStream_t)stream;
    MS_REQUIRE(m_cap > 0 && k1 >= 1 && m_cap * (int64_t)k1 < ((int64_t)1 << 32), "ms_orient_normals: bad sizes");
    MS_REQUIRE(ws_bytes >= ms_orient_workspace_bytes(m_cap, k1), "ms_orient_normals: workspace too small");
    OrArgs A;
    A.pts = points_all; A.nbr = nbr; A.nrm = normals; A.info_in = info; A.info_out = info; A.m_cap = m_cap; A.k1 = k1;
    MS_REQUIRE(orient_arena(A, m_cap, k1, ws, ws_bytes) == 0, "ms_orient_normals: workspace too small");
    MS_CUDA(cudaMemsetAsync(A.cnt, 0, 16 * sizeof(unsigned), s));
    MS_CUDA(cudaMemsetAsync(A.srckey, 0, 4 * sizeof(unsigned long long), s));
    int dev = 0, coop = 0, per_sm = 0;
    MS_CUDA(cudaGetDevice(&dev));
    MS_CUDA(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev));
    MS_REQUIRE(coop, "ms_orient_normals: the device does not support cooperative launches");
    MS_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_orient, OR_T, 0));
    MS_REQUIRE(per_sm >= 1, "ms_orient_normals: k_orient does not fit an SM");
    const int sms = ms_sm_count();
    MS_REQUIRE((int64_t)sms * (OR_T / 32) <= OR_MAX_WARPS, "ms_orient_normals: more warps than the workspace was sized for");
    // Half the SMs per launch: a bilayer frame has two leaflets, each on its own stream, and two cooperative kernels
    // of sms/2 CTAs run side by side.  The kernel is bound by dependent gathers and grid barriers, not by issue
    // slots: alone it takes 1.6x longer on half the SMs (1.58 vs 0.98 ms, 499 849 points), two together finish in
    // 1.58 instead of 1.97 ms (frame 12.47 -> 12.24 ms; 100k-lipid trajectory 406 -> 436 frames/s).
#ifndef OR_SPLIT
#define OR_SPLIT 2
#endif
    dim3 grid((unsigned)(sms / OR_SPLIT)), block(OR_T);
    void *args[] = { (void*)&A };
    MS_K("orient", s);
    MS_CUDA(cudaLaunchCooperativeKernel((void*)k_orient, grid, block, args, 0, s));
    MS_K(NULL, s);
    MS_LAUNCH_CHECK();
    return 0;
}

// diagnostic: nanoseconds between the phase boundaries of the last ms_orient_normals call on this workspace
// (init, round 0 A, round 0 B, then A+C / B of every later round, orientation).  Syncs the stream.
extern "C" int ms_orient_phase_times(const void *ws, size_t ws_bytes, int64_t m_cap, int k1, double *ns_out, int cap, int *nphases, void *stream) {
    cudaStream_t s = (cudaStream_t)stream;
    MS_REQUIRE(ws && ns_out && nphases && ws_bytes >= ms_orient_workspace_bytes(m_cap, k1), "ms_orient_phase_times: bad arguments");
    OrArgs A;
    MS_REQUIRE(orient_arena(A, m_cap, k1, (void*)ws, ws_bytes) == 0, "ms_orient_phase_times: workspace too small");
    unsigned long long h[64];
    MS_CUDA(cudaMemcpyAsync(h, A.times, sizeof(h), cudaMemcpyDeviceToHost, s));
    MS_CUDA(cudaStreamSynchronize(s));
    int n = (int)h[63] - 1; if (n < 0) n = 0; if (n > 61) n = 61; if (n > cap) n = cap;
    for (int i = 0; i < n; i++) ns_out[i] = (double)(h[i + 1] - h[i]);
    *nphases = n;
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
