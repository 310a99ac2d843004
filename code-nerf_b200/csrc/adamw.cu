// Fused multi-tensor AdamW: the optimiser step of train.py:111-114 over the two MLPs and the two
// embedding tables in ONE launch.  torch's fused implementation hands every block a 65,536-element chunk (a few dozen
// blocks for these 2.2 M parameters: ~230 us per chunk step on B200, latency-bound); here a block owns 4,096 elements of
// one tensor, so the step is a ~60 MB streaming pass.  Same arithmetic as torch.optim.AdamW (decoupled weight decay,
// bias-corrected step, eps added after the square root).
#include "common.cuh"
#include "internal.h"

namespace {
constexpr int AW_MAX = 40;                 // tensors per launch (kernel-parameter space)
constexpr int AW_THREADS = 256;
constexpr int AW_PER_BLOCK = AW_THREADS * 16;

struct AdamWBatch {
  float* p[AW_MAX]; const float* g[AW_MAX]; float* m[AW_MAX]; float* v[AW_MAX];
  int64_t n[AW_MAX];
  float lr[AW_MAX], wd[AW_MAX], inv_bc1[AW_MAX], inv_bc2_sqrt[AW_MAX];
  int first_block[AW_MAX + 1];
  int count;
  float beta1, beta2, eps;
  const int64_t* step_dev;      // != NULL: the step count lives on the device (CUDA-graph replays: cnrf_adamw_step_dev)
};

__global__ void adamw_bump_kernel(int64_t* step) { *step += 1; }

__device__ __forceinline__ void adamw_one(float& p, float g, float& m, float& v, float lr, float wd, float b1, float b2,
                                          float eps, float inv_bc1, float inv_bc2_sqrt) {
  p -= lr * wd * p;
  m = m + (g - m) * (1.0f - b1);
  v = b2 * v + (1.0f - b2) * g * g;
  const float denom = sqrtf(v) * inv_bc2_sqrt + eps;
  p -= (lr * inv_bc1) * m / denom;
}

__global__ void __launch_bounds__(AW_THREADS) adamw_kernel(const __grid_constant__ AdamWBatch b) {
  int t = 0;
  while (t + 1 < b.count && (int)blockIdx.x >= b.first_block[t + 1]) ++t;
  const int64_t base = (int64_t)(blockIdx.x - b.first_block[t]) * AW_PER_BLOCK;
  const int64_t n = b.n[t];
  float* __restrict__ p = b.p[t];
  const float* __restrict__ g = b.g[t];
  float* __restrict__ m = b.m[t];
  float* __restrict__ v = b.v[t];
  const float lr = b.lr[t], wd = b.wd[t];
  float i1 = b.inv_bc1[t], i2 = b.inv_bc2_sqrt[t];
  if (b.step_dev) {             // the same double-precision bias corrections the host path computes, once per block
    __shared__ float s_bc[2];
    if (threadIdx.x == 0) {
      const double step = (double)*b.step_dev;
      s_bc[0] = (float)(1.0 / (1.0 - pow((double)b.beta1, step)));
      s_bc[1] = (float)(1.0 / sqrt(1.0 - pow((double)b.beta2, step)));
    }
    __syncthreads();
    i1 = s_bc[0]; i2 = s_bc[1];
  }
  const bool vec = ((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) | reinterpret_cast<uintptr_t>(m) |
                     reinterpret_cast<uintptr_t>(v)) & 15) == 0;
#pragma unroll
  for (int u = 0; u < 4; ++u) {
    const int64_t i = base + ((int64_t)u * AW_THREADS + threadIdx.x) * 4;
    if (i >= n) break;
    if (vec && i + 3 < n) {
      float4 pp = *reinterpret_cast<float4*>(p + i), mm = *reinterpret_cast<float4*>(m + i), vv = *reinterpret_cast<float4*>(v + i);
      const float4 gg = __ldg(reinterpret_cast<const float4*>(g + i));
      adamw_one(pp.x, gg.x, mm.x, vv.x, lr, wd, b.beta1, b.beta2, b.eps, i1, i2);
      adamw_one(pp.y, gg.y, mm.y, vv.y, lr, wd, b.beta1, b.beta2, b.eps, i1, i2);
      adamw_one(pp.z, gg.z, mm.z, vv.z, lr, wd, b.beta1, b.beta2, b.eps, i1, i2);
      adamw_one(pp.w, gg.w, mm.w, vv.w, lr, wd, b.beta1, b.beta2, b.eps, i1, i2);
      *reinterpret_cast<float4*>(p + i) = pp; *reinterpret_cast<float4*>(m + i) = mm; *reinterpret_cast<float4*>(v + i) = vv;
    } else {
      for (int j = 0; j < 4 && i + j < n; ++j) adamw_one(p[i + j], g[i + j], m[i + j], v[i + j], lr, wd, b.beta1, b.beta2, b.eps, i1, i2);
    }
  }
}
}  // namespace

static int adamw_step_impl(const cnrf_adamw_tensor* tensors, int32_t count, float beta1, float beta2, float eps,
                           int64_t* step_dev, cnrf_stream_t stream) {
  if (count < 0 || (count > 0 && !tensors)) return CNRF_E_NULL;
  cudaStream_t st = as_stream(stream);
  if (step_dev && count > 0) {
    adamw_bump_kernel<<<1, 1, 0, st>>>(step_dev);
    CNRF_LAUNCH_CHECK();
  }
  // validate everything before the first launch: a bad entry must not leave earlier tensors already stepped
  for (int i = 0; i < count; ++i) {
    const cnrf_adamw_tensor& t = tensors[i];
    if (t.numel <= 0) continue;
    if (!t.param || !t.grad || !t.exp_avg || !t.exp_avg_sq) return CNRF_E_NULL;
    if (!step_dev && t.step < 1.0f) return CNRF_E_SHAPE;
  }
  // `next` is the first table row the batch did NOT consume (empty tensors are skipped without taking a batch entry, so a
  // batch may cover more than AW_MAX rows)
  for (int at = 0, next = 0; at < count; at = next) {
    AdamWBatch b;
    b.count = 0; b.beta1 = beta1; b.beta2 = beta2; b.eps = eps; b.step_dev = step_dev;
    int blocks = 0;
    for (next = at; next < count && b.count < AW_MAX; ++next) {
      const cnrf_adamw_tensor& t = tensors[next];
      if (t.numel <= 0) continue;
      const int k = b.count++;
      b.p[k] = t.param; b.g[k] = t.grad; b.m[k] = t.exp_avg; b.v[k] = t.exp_avg_sq; b.n[k] = t.numel;
      b.lr[k] = t.lr; b.wd[k] = t.weight_decay;
      const double stp = step_dev ? 1.0 : (double)t.step;      // (device-step mode: the kernel computes the corrections)
      const double bc1 = 1.0 - pow((double)beta1, stp), bc2 = 1.0 - pow((double)beta2, stp);
      b.inv_bc1[k] = (float)(1.0 / bc1);
      b.inv_bc2_sqrt[k] = (float)(1.0 / sqrt(bc2));
      b.first_block[k] = blocks;
      blocks += (int)((t.numel + AW_PER_BLOCK - 1) / AW_PER_BLOCK);
    }
    b.first_block[b.count] = blocks;
    if (blocks == 0) continue;
    adamw_kernel<<<blocks, AW_THREADS, 0, st>>>(b);
    CNRF_LAUNCH_CHECK();
  }
  return CNRF_OK;
}

extern "C" int cnrf_adamw_step(const cnrf_adamw_tensor* tensors, int32_t count, float beta1, float beta2, float eps,
                               cnrf_stream_t stream) {
  return adamw_step_impl(tensors, count, beta1, beta2, eps, nullptr, stream);
}

extern "C" int cnrf_adamw_step_dev(const cnrf_adamw_tensor* tensors, int32_t count, float beta1, float beta2, float eps,
                                   int64_t* step_dev, cnrf_stream_t stream) {
  if (!step_dev) return CNRF_E_NULL;
  return adamw_step_impl(tensors, count, beta1, beta2, eps, step_dev, stream);
}
