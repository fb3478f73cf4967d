// ms_common.cuh -- shared helpers for libmemsurfer_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>
#include <stdio.h>
#include "../../include/memsurfer_b200.h"


void ms_set_error(const char *fmt, ...);

// SMs of the CURRENT device (queried on every call; 148 on a B200): grids of the grid-stride kernels are sized from it
int ms_sm_count(void);

// Launch accounting / optional per-kernel event timing (ms_profile_* in the C ABI).  Call
// ms_prof_mark("kernel name", stream) right before every kernel launch and
// ms_prof_mark(NULL, stream) after the last launch of an entry point.  With profiling off this
// only bumps the launch counter; with it on, a cudaEvent is recorded at every mark and the span
// up to the next mark on the same thread is attributed to the name.
void ms_prof_mark(const char *name, cudaStream_t s);
#define MS_K(name, s) ms_prof_mark(name, s)

#define MS_CUDA(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { \
    ms_set_error("%s:%d %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); return 2; } } while (0)
#define MS_LAUNCH_CHECK() MS_CUDA(cudaGetLastError())
#define MS_REQUIRE(cond, ...) do { if (!(cond)) { ms_set_error(__VA_ARGS__); return 1; } } while (0)

static inline size_t ms_align(size_t x) { return (x + 255) & ~(size_t)255; }
static inline int ms_blocks(int64_t n, int threads) { return (int)((n + threads - 1) / threads); }

// carve typed sub-buffers out of the caller's workspace
struct MsArena {
    char *base; size_t used, cap;
    MsArena(void *p, size_t c) : base((char*)p), used(0), cap(c) {}
    template <class T> T *take(size_t n) { size_t b = ms_align(n * sizeof(T)); T *r = (T*)(base + used); used += b; return r; }
    bool ok() const { return used <= cap; }
};

// exclusive scan of int32 (n+1 outputs: out[n] = total).  ws: ms_scan_ws_bytes(n)
size_t ms_scan_ws_bytes(int64_t n);
int ms_exclusive_scan_i32(const int32_t *in, int32_t *out, int64_t n, void *ws, cudaStream_t s);

__device__ __forceinline__ double ms_ld_f(const float *p) { return (double)__ldg(p); }
