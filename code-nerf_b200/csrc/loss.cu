// The training loop's photometric loss (train.py:103-104): coarse_loss = mse(rgb_coarse[..., :3], target[..., :3]),
// fine_loss = mse(rgb_fine[..., :3], target[..., :3]), loss = coarse_loss + fine_loss -- both means and their sum in ONE
// launch, both gradients in one more (torch: 2 x (sub/pow kernel + mean reduction) + add forward, 2 x mse_backward +
// fills backward).  SURVEY section 8 row f2 (caller-side fusion, API-compatible: the caller still owns the target gather).
#include "common.cuh"
#include "internal.h"

namespace {
constexpr int LOSS_THREADS = 1024;

__device__ __forceinline__ float block_sum(float v, float* red) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, o);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  __syncthreads();                       // `red` may still be read from the previous reduction
  if (lane == 0) red[warp] = v;
  __syncthreads();
  float s = 0.f;
  if (warp == 0) {
    s = lane < (blockDim.x >> 5) ? red[lane] : 0.f;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xFFFFFFFFu, s, o);
  }
  return s;                              // valid in warp 0
}

// out[0] = mean((a - t)^2), out[1] = mean((b - t)^2), out[2] = out[0] + out[1] over N rows x 3 channels.
// a, b: rows of `ld_a` / `ld_b` floats, t: rows of `ld_t` floats (RGBA targets: ld_t = 4).
// One block: the sums are accumulated in a fixed order (deterministic), 12,288 elements per training chunk.
__global__ void __launch_bounds__(LOSS_THREADS) mse_pair_fwd_kernel(const float* __restrict__ a, int64_t ld_a,
                                                                    const float* __restrict__ b, int64_t ld_b,
                                                                    const float* __restrict__ t, int64_t ld_t, int64_t N,
                                                                    float* __restrict__ out) {
  __shared__ float red[32];
  float sa = 0.f, sb = 0.f;
  for (int64_t n = threadIdx.x; n < N; n += LOSS_THREADS) {
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const float tv = __ldg(t + n * ld_t + c);
      const float da = __ldg(a + n * ld_a + c) - tv, db = __ldg(b + n * ld_b + c) - tv;
      sa = fmaf(da, da, sa);
      sb = fmaf(db, db, sb);
    }
  }
  sa = block_sum(sa, red);
  sb = block_sum(sb, red);
  if (threadIdx.x == 0) {
    const float inv = 1.0f / (float)(N * 3);
    out[0] = sa * inv; out[1] = sb * inv; out[2] = sa * inv + sb * inv;
  }
}

// g_a = (g_out[0] + g_out[2]) * 2 (a - t) / (3N), g_b = (g_out[1] + g_out[2]) * 2 (b - t) / (3N); a NULL upstream
// gradient is zero.  g_a / g_b are dense [N,3].
__global__ void mse_pair_bwd_kernel(const float* __restrict__ a, int64_t ld_a, const float* __restrict__ b, int64_t ld_b,
                                    const float* __restrict__ t, int64_t ld_t, int64_t N, const float* __restrict__ g0,
                                    const float* __restrict__ g1, const float* __restrict__ g2, float* __restrict__ g_a,
                                    float* __restrict__ g_b) {
  const float gt = g2 ? __ldg(g2) : 0.f;
  const float scale = 2.0f / (float)(N * 3);
  const float ka = ((g0 ? __ldg(g0) : 0.f) + gt) * scale, kb = ((g1 ? __ldg(g1) : 0.f) + gt) * scale;
  const int64_t total = N * 3;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t n = i / 3;
    const int c = (int)(i - n * 3);
    const float tv = __ldg(t + n * ld_t + c);
    if (g_a) g_a[i] = ka * (__ldg(a + n * ld_a + c) - tv);
    if (g_b) g_b[i] = kb * (__ldg(b + n * ld_b + c) - tv);
  }
}
}  // namespace

extern "C" int cnrf_mse_pair_fwd(const float* rgb_a, int64_t ld_a, const float* rgb_b, int64_t ld_b, const float* target,
                                 int64_t ld_t, int64_t N, float* out3, cnrf_stream_t stream) {
  if (!rgb_a || !rgb_b || !target || !out3) return CNRF_E_NULL;
  if (N <= 0 || ld_a < 3 || ld_b < 3 || ld_t < 3) return CNRF_E_SHAPE;
  cudaStream_t st = (cudaStream_t)stream;
  mse_pair_fwd_kernel<<<1, LOSS_THREADS, 0, st>>>(rgb_a, ld_a, rgb_b, ld_b, target, ld_t, N, out3);
  CNRF_LAUNCH_CHECK();
  return CNRF_OK;
}

extern "C" int cnrf_mse_pair_bwd(const float* rgb_a, int64_t ld_a, const float* rgb_b, int64_t ld_b, const float* target,
                                 int64_t ld_t, int64_t N, const float* g_loss_a, const float* g_loss_b,
                                 const float* g_loss_sum, float* g_a, float* g_b, cnrf_stream_t stream) {
  if (!rgb_a || !rgb_b || !target) return CNRF_E_NULL;
  if (N <= 0 || ld_a < 3 || ld_b < 3 || ld_t < 3) return CNRF_E_SHAPE;
  cudaStream_t st = (cudaStream_t)stream;
  const int64_t total = N * 3;
  int grid = (int)((total + 255) / 256);
  if (grid > 148 * 8) grid = 148 * 8;
  mse_pair_bwd_kernel<<<grid, 256, 0, st>>>(rgb_a, ld_a, rgb_b, ld_b, target, ld_t, N, g_loss_a, g_loss_b, g_loss_sum,
                                            g_a, g_b);
  CNRF_LAUNCH_CHECK();
  return CNRF_OK;
}
