// Shared helpers for libcnrf.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/cnrf.h"

#define CNRF_REQUIRE(cond, code) \
  do {                           \
    if (!(cond)) return (code);  \
  } while (0)

// Launch-error check that never synchronises: reports bad configurations, not asynchronous faults.
extern long long g_cnrf_launches;   // api.cu: kernels launched by this library (bench.py's gpu_launches claim)
#define CNRF_LAUNCH_CHECK()                      \
  do {                                           \
    ++g_cnrf_launches;                           \
    cudaError_t e__ = cudaGetLastError();        \
    if (e__ != cudaSuccess) return (int)e__;     \
  } while (0)

static inline cudaStream_t as_stream(cnrf_stream_t s) { return reinterpret_cast<cudaStream_t>(s); }

static inline unsigned int grid_for(int64_t work, int block) {
  int64_t g = (work + block - 1) / block;
  if (g < 1) g = 1;
  if (g > 0x7fffffffLL) g = 0x7fffffffLL;
  return (unsigned int)g;
}

constexpr unsigned FULL_MASK = 0xffffffffu;

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL_MASK, v, o);
  return v;
}

// inclusive warp scan (Hillis-Steele over shuffles)
__device__ __forceinline__ float warp_scan_inclusive(float v, int lane) {
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    float t = __shfl_up_sync(FULL_MASK, v, o);
    if (lane >= o) v += t;
  }
  return v;
}

// torch.nn.functional.softplus(x) with beta=1, threshold=20 (volumetric_render.py:32-33 calls it on raw-1).
__device__ __forceinline__ float softplus_t20(float x) { return x > 20.0f ? x : log1pf(expf(x)); }
__device__ __forceinline__ float sigmoidf_(float x) { return 1.0f / (1.0f + expf(-x)); }
