// ms_common.cu -- error string, version, exclusive scan, DFMA probe.
#include "ms_common.cuh"
#include <stdarg.h>

static thread_local char g_err[512] = "";

void ms_set_error(const char *fmt, ...) {
    va_list ap; va_start(ap, fmt); vsnprintf(g_err, sizeof(g_err), fmt, ap); va_end(ap);
}

extern "C" int ms_version(void) { return 100; }
extern "C" const char *ms_last_error(void) { return g_err; }

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

__global__ void __launch_bounds__(SCAN_T) k_scan_blocksum(const int32_t *in, int64_t n, int32_t *bsum) {
    __shared__ int sm[32];
    const int64_t base = (int64_t)blockIdx.x * SCAN_B + threadIdx.x * SCAN_I;
    int s = 0;
#pragma unroll
    for (int i = 0; i < SCAN_I; i++) if (base + i < n) s += in[base + i];
    int total; block_exclusive_scan(s, sm, total);
    if (threadIdx.x == 0) bsum[blockIdx.x] = total;
}

__global__ void __launch_bounds__(SCAN_T) k_scan_top(int32_t *bsum, int64_t nb) {
    __shared__ int sm[32];
    int carry = 0;
    for (int64_t c = 0; c < nb; c += SCAN_T) {
        const int64_t i = c + threadIdx.x;
        const int v = i < nb ? bsum[i] : 0;
        int total; const int ex = block_exclusive_scan(v, sm, total);
        if (i < nb) bsum[i] = carry + ex;
        carry += total;
    }
    if (threadIdx.x == 0) bsum[nb] = carry;
}

__global__ void __launch_bounds__(SCAN_T) k_scan_final(const int32_t *in, int32_t *out, int64_t n, const int32_t *bsum) {
    __shared__ int sm[32];
    const int64_t base = (int64_t)blockIdx.x * SCAN_B + threadIdx.x * SCAN_I;
    int v[SCAN_I]; int s = 0;
#pragma unroll
    for (int i = 0; i < SCAN_I; i++) { v[i] = base + i < n ? in[base + i] : 0; s += v[i]; }
    int total; int ex = block_exclusive_scan(s, sm, total) + bsum[blockIdx.x];
#pragma unroll
    for (int i = 0; i < SCAN_I; i++) { if (base + i < n) out[base + i] = ex; ex += v[i]; }
    if (blockIdx.x == gridDim.x - 1 && threadIdx.x == 0) out[n] = bsum[gridDim.x];
}

size_t ms_scan_ws_bytes(int64_t n) { return ms_align(((n + SCAN_B - 1) / SCAN_B + 2) * sizeof(int32_t)); }

int ms_exclusive_scan_i32(const int32_t *in, int32_t *out, int64_t n, void *ws, cudaStream_t s) {
    int32_t *bsum = (int32_t*)ws;
    if (n <= 0) { MS_CUDA(cudaMemsetAsync(out, 0, sizeof(int32_t), s)); return 0; }
    const int64_t nb = (n + SCAN_B - 1) / SCAN_B;
    k_scan_blocksum<<<(unsigned)nb, SCAN_T, 0, s>>>(in, n, bsum);
    k_scan_top<<<1, SCAN_T, 0, s>>>(bsum, nb);
    k_scan_final<<<(unsigned)nb, SCAN_T, 0, s>>>(in, out, n, bsum);
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

#define PROBE_BLOCKS (MS_SM_COUNT * 8)
extern "C" int ms_dfma_probe(double *out, int64_t iters, double *flop, void *stream) {
    MS_REQUIRE(out != NULL && iters > 0, "ms_dfma_probe: out must hold %d doubles", PROBE_BLOCKS * 256);
    k_dfma_probe<<<PROBE_BLOCKS, 256, 0, (cudaStream_t)stream>>>(out, iters);
    MS_LAUNCH_CHECK();
    if (flop) *flop = 2.0 * 8.0 * (double)iters * 256.0 * PROBE_BLOCKS;
    return 0;
}
