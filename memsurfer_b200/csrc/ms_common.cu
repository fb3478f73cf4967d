// ms_common.cu -- error string, version, exclusive scan, DFMA probe.
#include "ms_common.cuh"
#include <stdarg.h>
#include <string.h>
#include <atomic>
#include <vector>

static thread_local char g_err[512] = "";

void ms_set_error(const char *fmt, ...) {
    va_list ap; va_start(ap, fmt); vsnprintf(g_err, sizeof(g_err), fmt, ap); va_end(ap);
}

int ms_sm_count(void) {
    int dev = 0, sms = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms < 1) sms = 148;
    return sms;
}

extern "C" int ms_version(void) { return 200; }
extern "C" const char *ms_last_error(void) { return g_err; }

// ------------------------------------------------------------------ launch accounting / profiling
// Diagnostic state only (never read by a compute path): a process-wide launch counter and a
// per-thread list of (name, event) marks.
static std::atomic<long long> g_launches(0);
struct ProfMark { const char *name; cudaEvent_t ev; };
static thread_local bool g_prof_on = false;
static thread_local std::vector<ProfMark> *g_marks = nullptr;

void ms_prof_mark(const char *name, cudaStream_t s) {
    if (name) g_launches.fetch_add(1, std::memory_order_relaxed);
    if (!g_prof_on) return;
    if (!g_marks) g_marks = new std::vector<ProfMark>();
    if (!name && (g_marks->empty() || g_marks->back().name == nullptr)) return;
    ProfMark m; m.name = name;
    if (cudaEventCreate(&m.ev) != cudaSuccess) return;
    cudaEventRecord(m.ev, s);
    g_marks->push_back(m);
}

static void prof_clear() {
    if (!g_marks) return;
    for (auto &m : *g_marks) cudaEventDestroy(m.ev);
    g_marks->clear();
}

extern "C" int ms_profile_enable(int on) { prof_clear(); g_prof_on = on != 0; return 0; }
extern "C" int64_t ms_launch_count(int reset) {
    return reset ? g_launches.exchange(0) : g_launches.load();
}
extern "C" int ms_profile_report(char *buf, size_t cap) {
    MS_REQUIRE(buf && cap > 2, "ms_profile_report: need a buffer");
    struct Agg { const char *name; long long n; double ms; };
    std::vector<Agg> agg;
    if (g_marks && !g_marks->empty()) {
        MS_CUDA(cudaEventSynchronize(g_marks->back().ev));
        for (size_t i = 0; i + 1 < g_marks->size(); i++) {
            const ProfMark &a = (*g_marks)[i], &b = (*g_marks)[i + 1];
            if (!a.name) continue;
            float ms = 0.f;
            if (cudaEventElapsedTime(&ms, a.ev, b.ev) != cudaSuccess) { cudaGetLastError(); continue; }
            size_t k = 0;
            for (; k < agg.size(); k++) if (!strcmp(agg[k].name, a.name)) break;
            if (k == agg.size()) agg.push_back(Agg{ a.name, 0, 0.0 });
            agg[k].n++; agg[k].ms += ms;
        }
    }
    size_t o = 0;
    o += snprintf(buf + o, cap - o, "{");
    for (size_t k = 0; k < agg.size() && o < cap; k++)
        o += snprintf(buf + o, cap - o, "%s\"%s\": {\"launches\": %lld, \"ms\": %.6f}", k ? ", " : "", agg[k].name, agg[k].n, agg[k].ms);
    if (o < cap) o += snprintf(buf + o, cap - o, "}");
    MS_REQUIRE(o < cap, "ms_profile_report: buffer too small");
    prof_clear();
    return 0;
}

// ------------------------------------------------------------------ exclusive scan
#define SCAN_T 512
#define SCAN_I 4
#define SCAN_B (SCAN_T * SCAN_I)

__device__ __forceinline__ int block_exclusive_scan(int v, int *smem /* 32 */, int &total) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    int x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { int y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
    if (lane == 31) smem[w] = x;
    __syncthreads();
    if (w == 0) {
        int s = lane < (SCAN_T / 32) ? smem[lane] : 0;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { int y = __shfl_up_sync(0xffffffffu, s, o); if (lane >= o) s += y; }
        smem[lane] = s;
    }
    __syncthreads();
    const int woff = w > 0 ? smem[w - 1] : 0;
    total = smem[SCAN_T / 32 - 1];
    __syncthreads();
    return woff + x - v;
}

// Single-pass chained scan with decoupled look-back: one launch (+ one memset of the tile states) instead of
// the classic three.  Tile ids are handed out by an atomic counter in the order the CTAs start, so a tile only
// ever waits for tiles that are already running or done.  Tile state = one 64-bit word (flag << 62 | sum):
// a single store publishes flag and value together.
#define SCAN_FLAG_A (1ull << 62)       // aggregate of this tile only
#define SCAN_FLAG_P (2ull << 62)       // inclusive prefix up to and including this tile
__global__ void __launch_bounds__(SCAN_T) k_scan_lookback(const int32_t *__restrict__ in, int32_t *__restrict__ out, int64_t n,
                                                          unsigned long long *state /* [0] = tile counter, [1 + t] = tile t */) {
    __shared__ int sm[32];
    __shared__ int s_tile, s_prefix;
    if (threadIdx.x == 0) s_tile = (int)atomicAdd(state, 1ull);
    __syncthreads();
    const int t = s_tile;
    volatile unsigned long long *st = state + 1;
    const int64_t base = (int64_t)t * SCAN_B + threadIdx.x * SCAN_I;
    int v[SCAN_I]; int s = 0;
#pragma unroll
    for (int i = 0; i < SCAN_I; i++) { v[i] = base + i < n ? in[base + i] : 0; s += v[i]; }
    int total; int ex = block_exclusive_scan(s, sm, total);
    if (threadIdx.x == 0) {
        st[t] = (t == 0 ? SCAN_FLAG_P : SCAN_FLAG_A) | (unsigned long long)(unsigned)total;
        if (t == 0) s_prefix = 0;
    }
    if (t > 0 && threadIdx.x < 32) {
        const int lane = threadIdx.x;
        int prefix = 0;
        for (int j = t - 1; ; j -= 32) {
            unsigned long long w = SCAN_FLAG_P;              // tiles before tile 0: an empty inclusive prefix
            if (j - lane >= 0) { do { w = st[j - lane]; } while ((w >> 62) == 0ull); }
            const unsigned pmask = __ballot_sync(0xffffffffu, (w >> 62) == 2ull);
            const int first = pmask ? __ffs(pmask) - 1 : 32;  // nearest tile with a full prefix
            int val = lane <= first ? (int)(unsigned)(w & 0xffffffffull) : 0;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) val += __shfl_xor_sync(0xffffffffu, val, o);
            prefix += val;
            if (pmask) break;
        }
        if (lane == 0) { st[t] = SCAN_FLAG_P | (unsigned long long)(unsigned)(prefix + total); s_prefix = prefix; }
    }
    __syncthreads();
    ex += s_prefix;
#pragma unroll
    for (int i = 0; i < SCAN_I; i++) { if (base + i < n) out[base + i] = ex; ex += v[i]; }
    if (base <= n - 1 && n - 1 < base + SCAN_I) out[n] = ex;     // the thread that owns the last element (ex = grand total by now)
}

size_t ms_scan_ws_bytes(int64_t n) { return ms_align(((n + SCAN_B - 1) / SCAN_B + 2) * sizeof(unsigned long long)); }

int ms_exclusive_scan_i32(const int32_t *in, int32_t *out, int64_t n, void *ws, cudaStream_t s) {
    if (n <= 0) { MS_CUDA(cudaMemsetAsync(out, 0, sizeof(int32_t), s)); return 0; }
    const int64_t nb = (n + SCAN_B - 1) / SCAN_B;
    MS_CUDA(cudaMemsetAsync(ws, 0, (size_t)(nb + 1) * sizeof(unsigned long long), s));
    MS_K("scan", s); k_scan_lookback<<<(unsigned)nb, SCAN_T, 0, s>>>(in, out, n, (unsigned long long*)ws);
    MS_K(NULL, s);
    MS_LAUNCH_CHECK();
    return 0;
}

// ------------------------------------------------------------------ DFMA peak probe
// 8 independent dependent-chains per thread keep the fp64 pipe full; 2 FLOP per DFMA.
__global__ void __launch_bounds__(256) k_dfma_probe(double *out, int64_t iters) {
    double a0 = 1.0 + threadIdx.x * 1e-9, a1 = a0 + 1e-3, a2 = a0 + 2e-3, a3 = a0 + 3e-3;
    double a4 = a0 + 4e-3, a5 = a0 + 5e-3, a6 = a0 + 6e-3, a7 = a0 + 7e-3;
    const double m = 0.999999999, c = 1e-12;
    for (int64_t i = 0; i < iters; i++) {
        a0 = fma(a0, m, c); a1 = fma(a1, m, c); a2 = fma(a2, m, c); a3 = fma(a3, m, c);
        a4 = fma(a4, m, c); a5 = fma(a5, m, c); a6 = fma(a6, m, c); a7 = fma(a7, m, c);
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = ((a0 + a1) + (a2 + a3)) + ((a4 + a5) + (a6 + a7));
}

extern "C" int ms_dfma_probe(double *out, int64_t iters, double *flop, void *stream) {
    const int PROBE_BLOCKS = ms_sm_count() * 8;
    MS_REQUIRE(out != NULL && iters > 0, "ms_dfma_probe: out must hold %d doubles", PROBE_BLOCKS * 256);
    MS_K("dfma_probe", (cudaStream_t)stream); k_dfma_probe<<<PROBE_BLOCKS, 256, 0, (cudaStream_t)stream>>>(out, iters);
    MS_K(NULL, (cudaStream_t)stream);
    MS_LAUNCH_CHECK();
    if (flop) *flop = 2.0 * 8.0 * (double)iters * 256.0 * PROBE_BLOCKS;
    return 0;
}


// ------------------------------------------------------------------ per-frame field summary
// (sum, sum|x|, min, max) of an fp64 field in a fixed reduction order (deterministic): the record
// that ranks exchange when trajectory frames are sharded across GPUs.
#define SUM_T 256
#define SUM_BLOCKS 592            // partial sums per field: fixed, so that the reduction order (and the result) does not depend on the device
struct Sum4 { double s, a, mn, mx; };
__device__ __forceinline__ Sum4 sum4_merge(const Sum4 &x, const Sum4 &y) {
    Sum4 r; r.s = x.s + y.s; r.a = x.a + y.a; r.mn = fmin(x.mn, y.mn); r.mx = fmax(x.mx, y.mx); return r;
}
__device__ __forceinline__ Sum4 sum4_block(Sum4 v, Sum4 *sm) {
    sm[threadIdx.x] = v;
    __syncthreads();
    for (int o = SUM_T / 2; o > 0; o >>= 1) {
        if (threadIdx.x < o) sm[threadIdx.x] = sum4_merge(sm[threadIdx.x], sm[threadIdx.x + o]);
        __syncthreads();
    }
    return sm[0];
}
#define SUM_MAXF 64
struct SumFields { const double *p[SUM_MAXF]; };
// grid (blocks, nfields): field y of the batch
__global__ void __launch_bounds__(SUM_T) k_summary_partial(SumFields f, int64_t n, Sum4 *part) {
    __shared__ Sum4 sm[SUM_T];
    const double *__restrict__ x = f.p[blockIdx.y];
    Sum4 v; v.s = 0.0; v.a = 0.0; v.mn = INFINITY; v.mx = -INFINITY;
    for (int64_t i = (int64_t)blockIdx.x * SUM_T + threadIdx.x; i < n; i += (int64_t)gridDim.x * SUM_T) {
        const double d = x[i]; v.s += d; v.a += fabs(d); v.mn = fmin(v.mn, d); v.mx = fmax(v.mx, d);
    }
    v = sum4_block(v, sm);
    if (threadIdx.x == 0) part[(size_t)blockIdx.y * gridDim.x + blockIdx.x] = v;
}
__global__ void __launch_bounds__(SUM_T) k_summary_final(const Sum4 *__restrict__ part, int nb, double *out4) {
    __shared__ Sum4 sm[SUM_T];
    Sum4 v; v.s = 0.0; v.a = 0.0; v.mn = INFINITY; v.mx = -INFINITY;
    for (int i = threadIdx.x; i < nb; i += SUM_T) v = sum4_merge(v, part[(size_t)blockIdx.x * nb + i]);
    v = sum4_block(v, sm);
    if (threadIdx.x == 0) { double *o = out4 + 4 * (size_t)blockIdx.x; o[0] = v.s; o[1] = v.a; o[2] = v.mn; o[3] = v.mx; }
}
extern "C" size_t ms_field_summary_workspace_bytes(int nfields) {
    return ms_align((size_t)SUM_BLOCKS * (size_t)(nfields < 1 ? 1 : nfields) * sizeof(Sum4));
}
extern "C" int ms_field_summary(const double *const *fields, int nfields, int64_t n, double *out4,
                                void *ws, size_t ws_bytes, void *stream) {
    cudaStream_t s = (cudaStream_t)stream;
    MS_REQUIRE(fields && out4 && n > 0 && nfields >= 1 && nfields <= SUM_MAXF, "ms_field_summary: need 1..%d fields of n > 0 values", SUM_MAXF);
    MS_REQUIRE(ws_bytes >= ms_field_summary_workspace_bytes(nfields), "ms_field_summary: workspace too small");
    SumFields f;
    for (int i = 0; i < nfields; i++) { MS_REQUIRE(fields[i] != NULL, "ms_field_summary: field %d is NULL", i); f.p[i] = fields[i]; }
    int nb = ms_blocks(n, SUM_T); if (nb > SUM_BLOCKS) nb = SUM_BLOCKS;
    MS_K("summary_partial", s); k_summary_partial<<<dim3(nb, nfields), SUM_T, 0, s>>>(f, n, (Sum4*)ws);
    MS_K("summary_final", s); k_summary_final<<<nfields, SUM_T, 0, s>>>((const Sum4*)ws, nb, out4);
    MS_K(NULL, s);
    MS_LAUNCH_CHECK();
    return 0;
}
