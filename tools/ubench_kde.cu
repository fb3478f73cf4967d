// ubench_kde.cu -- issue-slot experiments for the KDE pair loop on B200 (not part of the library).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ubench_kde tools/ubench_kde.cu && ./ubench_kde
// Every warp runs the R=2 pair loop of k_kde_tiles over a shared-memory source array with no
// pipeline, barrier or global traffic.  Prints clk per warp-pair (ideal: 11 fp64 x 2 clk = 22).
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define MAGIC 6755399441055744.0
#define C1 (-0x1.62e42fefa39dap-9)
#define C2 (0x1.ebfbe3a9ac803p-19)
#define C3 (-0x1.c6b0902b2f6c8p-29)
#define NSRC 512

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

// MODE 0: full term.  1: no table (sc = 1).  2: no exponent arithmetic (table value used as is).
// 3: fp64 only: distance + split + polynomial, no integer work at all.  4: DFMA-only chains (11 per pair).
template <int MODE>
__device__ __forceinline__ void term(double r, uint32_t tbase, double &acc) {
    if (MODE == 4) {
        double p = r;
#pragma unroll
        for (int k = 0; k < 6; k++) p = fma(p, r, C1);
        acc = fma(p, r, acc);
        return;
    }
    const double t = r + MAGIC;
    const int n = __double2loint(t);
    const double f = r - (t - MAGIC);
    double p;
    if (MODE == 5) { p = fma(C2, f, C1); p = fma(p, f, 1.0); }
    else { p = fma(C3, f, C2); p = fma(p, f, C1); p = fma(p, f, 1.0); }
    if (MODE == 5) {
        int lo, hi;
        asm("{ .reg .b32 i, a; and.b32 i, %2, 511; mad.lo.u32 a, i, 128, %3; ld.shared.v2.b32 {%0, %1}, [a]; }"
            : "=r"(lo), "=r"(hi) : "r"(n), "r"(tbase));
        hi -= n << 11;
        acc = fma(__hiloint2double(hi, lo), p, acc);
        return;
    }
    if (MODE == 1 || MODE == 3) { acc = fma(p, MODE == 1 ? __hiloint2double(0x3ff00000 - (n << 12), 0) : 1.0, acc); return; }
    int lo, hi;
    asm("{ .reg .b32 i, a; and.b32 i, %2, 255; mad.lo.u32 a, i, 128, %3; ld.shared.v2.b32 {%0, %1}, [a]; }"
        : "=r"(lo), "=r"(hi) : "r"(n), "r"(tbase));
    if (MODE == 0) hi -= n << 12;
    acc = fma(__hiloint2double(hi, lo), p, acc);
}

template <int MODE, int U, int R>
__global__ void __launch_bounds__(1024 / (R > 2 ? 2 : 1)) k_loop(int reps, double *out, long long *clk) {
    extern __shared__ __align__(128) unsigned char sm[];
    double *tab = (double*)sm;                  // [256][16]
    double2 *src = (double2*)(sm + (MODE == 5 ? 65536 : 32768));      // [NSRC]
    for (int i = threadIdx.x; i < (MODE == 5 ? 8192 : 4096); i += blockDim.x) tab[i] = 1.0 / (1.0 + (i >> 4) / 256.0);
    for (int i = threadIdx.x; i < NSRC; i += blockDim.x) src[i] = make_double2(3.0 * (i % 23), 2.5 * (i % 17));
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const uint32_t tbase = smem_u32(tab) + 8u * (lane & 15);
    double tx[R], ty[R], acc[R];
#pragma unroll
    for (int r = 0; r < R; r++) { tx[r] = 0.37 * lane + 9.0 * r; ty[r] = 0.11 * lane + 5.0 * r; acc[r] = 0.0; }
    const long long c0 = clock64();
    for (int rep = 0; rep < reps; rep++) {
#pragma unroll U
        for (int i = 0; i < NSRC; i++) {
            const double2 sp = src[i];
#pragma unroll
            for (int r = 0; r < R; r++) {
                const double dx = sp.x - tx[r], dy = sp.y - ty[r];
                const double r2 = fma(dy, dy, dx * dx);
                term<MODE>(r2, tbase, acc[r]);
            }
        }
        tx[0] += 1e-3;
    }
    const long long c1 = clock64();
    double a = 0.0;
#pragma unroll
    for (int r = 0; r < R; r++) a += acc[r];
    out[blockIdx.x * blockDim.x + threadIdx.x] = a;
    if (threadIdx.x == 0) clk[blockIdx.x] = c1 - c0;
}

template <int MODE, int U, int R = 2>
static void run(const char *what, int warps_per_sm) {
    const int reps = 40, nblk = 148;
    double *out; long long *clk;
    cudaMalloc(&out, sizeof(double) * nblk * 1024); cudaMalloc(&clk, sizeof(long long) * nblk);
    const int smem = (MODE == 5 ? 65536 : 32768) + NSRC * 16;
    cudaFuncSetAttribute(k_loop<MODE, U, R>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    k_loop<MODE, U, R><<<nblk, warps_per_sm * 32, smem>>>(2, out, clk);
    k_loop<MODE, U, R><<<nblk, warps_per_sm * 32, smem>>>(reps, out, clk);
    cudaError_t e = cudaDeviceSynchronize();
    long long h[148]; cudaMemcpy(h, clk, sizeof(h), cudaMemcpyDeviceToHost);
    double avg = 0; for (int i = 0; i < nblk; i++) avg += (double)h[i]; avg /= nblk;
    // warp-pairs per scheduler = warps/4 * reps * NSRC * 2
    const double wp = (warps_per_sm / 4.0) * reps * NSRC * (double)R;
    printf("%-44s R=%d U=%d warps/sched %2d  clk per warp-pair %.2f  (%s)\n", what, R, U, warps_per_sm / 4, avg / wp, cudaGetErrorString(e));
    cudaFree(out); cudaFree(clk);
}

int main() {
    run<0, 2, 2>("full term", 16); run<0, 2, 2>("full term", 20); run<0, 2, 2>("full term", 24); run<0, 2, 2>("full term", 32);
    run<0, 1, 4>("full term", 8); run<0, 1, 4>("full term", 12); run<0, 1, 4>("full term", 16); run<0, 2, 4>("full term", 16);
    run<0, 1, 3>("full term", 16); run<0, 2, 3>("full term", 16); run<0, 2, 3>("full term", 20);
    run<5, 2, 2>("deg 2 / 512 table", 16); run<5, 2, 2>("deg 2 / 512 table", 24); run<5, 2, 2>("deg 2 / 512 table", 32);
    run<3, 1, 4>("fp64 only", 16); run<1, 1, 4>("no table", 16);
    return 0;
}
