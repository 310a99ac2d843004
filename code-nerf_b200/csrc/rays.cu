// Ray generation: RaySampler.get_bundle / .sample  (view_synthesis/nerf/ray_sampler.py:31-51, 66-80, 84-99).
// One thread per ray: the camera-frame direction is recomputed from the pixel index with the same IEEE operations
// the reference uses to build its (H,W,3) table, rotated by the pose, and written straight to the packed (B*n,3)
// output -- the full-image bundle and the advanced-index gather of the reference never exist in HBM.
#include "common.cuh"
#include "internal.h"

__device__ __forceinline__ void camera_dir(int64_t flat, int W, float focal, float cx, float cy, float d[3]) {
  int h = (int)(flat / W), w = (int)(flat % W);        // flat = h*W + w (flatten(1,2), ray_sampler.py:69)
  d[0] = __fdiv_rn(__fsub_rn((float)w, cx), focal);      // (ii - cx) / f            ray_sampler.py:46
  d[1] = __fdiv_rn(-__fsub_rn((float)h, cy), focal);     // -(jj - cy) / f           ray_sampler.py:47
  d[2] = -1.0f;                                          // -ones                    ray_sampler.py:48
}

// A block owns 256 consecutive rays.  Full blocks stage their 768 + 768 floats in shared memory and write them as float4
// (whole 128-byte lines per warp instruction instead of 4-byte stores at a 12-byte stride); the last block, or
// pointers that are not 16-byte aligned, take the scalar path.
__global__ void __launch_bounds__(256) ray_bundle_fwd_kernel(const float* __restrict__ pose, int64_t B, int H, int W, float focal,
                                                             float cx, float cy, const int64_t* __restrict__ select, int64_t n,
                                                             float* __restrict__ ro, float* __restrict__ rd, int vec_ok) {
  __shared__ __align__(16) float s_rd[768];
  __shared__ __align__(16) float s_ro[768];
  const int64_t total = B * n;
  // grid-stride over tiles of 256 rays: a few thousand resident blocks write the whole bundle (a block per tile spent
  // a measurable share of a 50 us, write-only kernel on block scheduling)
  for (int64_t i0 = (int64_t)blockIdx.x * 256; i0 < total; i0 += (int64_t)gridDim.x * 256) {
  const int64_t i = i0 + threadIdx.x;
  const bool full = vec_ok && i0 + 256 <= total;
  float v[3] = {0.f, 0.f, 0.f}, o[3] = {0.f, 0.f, 0.f};
  if (i < total) {
    int64_t b, flat;
    if (total < (1ll << 31)) { const uint32_t bb = (uint32_t)i / (uint32_t)n; b = bb; flat = select ? select[i] : (int64_t)((uint32_t)i - bb * (uint32_t)n); }
    else { b = i / n; flat = select ? select[i] : (i - b * n); }
    const float* P = pose + b * 16;
    float d[3];
    camera_dir(flat, W, focal, cx, cy, d);
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      // einsum 'hwij,bji->bhwj': rd_j = sum_i d_i * R[j][i]   (ray_sampler.py:97)
      float a = __fmul_rn(d[0], __ldg(P + j * 4 + 0));
      a = __fmaf_rn(d[1], __ldg(P + j * 4 + 1), a);
      a = __fmaf_rn(d[2], __ldg(P + j * 4 + 2), a);
      v[j] = a;
      o[j] = __ldg(P + j * 4 + 3);                          // ro = pose[:3, -1]  (ray_sampler.py:98)
    }
  }
  if (!full) {
    if (i < total) {
#pragma unroll
      for (int j = 0; j < 3; ++j) {
        rd[i * 3 + j] = v[j];
        if (ro) ro[i * 3 + j] = o[j];
      }
    }
    continue;                                              // block-uniform: `full` depends on i0 only
  }
#pragma unroll
  for (int j = 0; j < 3; ++j) { s_rd[threadIdx.x * 3 + j] = v[j]; s_ro[threadIdx.x * 3 + j] = o[j]; }
  __syncthreads();
  if (threadIdx.x < 192) {
    __stcs(reinterpret_cast<float4*>(rd + i0 * 3) + threadIdx.x, reinterpret_cast<const float4*>(s_rd)[threadIdx.x]);
    if (ro) __stcs(reinterpret_cast<float4*>(ro + i0 * 3) + threadIdx.x, reinterpret_cast<const float4*>(s_ro)[threadIdx.x]);
  }
  __syncthreads();                                         // the staging rows are rewritten by the next tile
  }
}

// One block per image: d pose = sum over the image's rays; block-level reduction of 12 numbers.
__global__ void ray_bundle_bwd_kernel(const float* __restrict__ g_ro, const float* __restrict__ g_rd, int H, int W,
                                      float focal, float cx, float cy, const int64_t* __restrict__ select, int64_t n,
                                      float* __restrict__ g_pose) {
  int64_t b = blockIdx.x;
  float acc[12];
#pragma unroll
  for (int k = 0; k < 12; ++k) acc[k] = 0.f;
  for (int64_t r = threadIdx.x; r < n; r += blockDim.x) {
    int64_t i = b * n + r;
    int64_t flat = select ? select[i] : r;
    float d[3];
    camera_dir(flat, W, focal, cx, cy, d);
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      float g = g_rd[i * 3 + j];
      acc[j * 4 + 0] += g * d[0];
      acc[j * 4 + 1] += g * d[1];
      acc[j * 4 + 2] += g * d[2];
      if (g_ro) acc[j * 4 + 3] += g_ro[i * 3 + j];
    }
  }
  __shared__ float red[12][32];
  int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < 12; ++k) {
    float v = warp_sum(acc[k]);
    if (lane == 0) red[k][wid] = v;
  }
  __syncthreads();
  if (threadIdx.x < 16) {
    float v = 0.f;
    if (threadIdx.x < 12) {
      int nw = blockDim.x >> 5;
      for (int w = 0; w < nw; ++w) v += red[threadIdx.x][w];
    }
    g_pose[b * 16 + threadIdx.x] = v;                     // row 3 of the 4x4 gets zeros
  }
}

extern "C" int cnrf_ray_bundle_fwd(const float* pose, int64_t B, int32_t H, int32_t W, float focal, float cx, float cy,
                                   const int64_t* select, int64_t n, float* ro, float* rd, cnrf_stream_t stream) {
  CNRF_REQUIRE(pose && rd, CNRF_E_NULL);
  CNRF_REQUIRE(B > 0 && H > 0 && W > 0 && n > 0 && n <= (int64_t)H * W, CNRF_E_SHAPE);
  CNRF_REQUIRE(select || n == (int64_t)H * W, CNRF_E_SHAPE);
  const int vec_ok = ((reinterpret_cast<uintptr_t>(rd) | reinterpret_cast<uintptr_t>(ro)) & 15) == 0;
  const unsigned tiles = grid_for(B * n, 256), cap = (unsigned)cnrf_sm_count() * 8u;      // 8 resident blocks of 256 per SM
  ray_bundle_fwd_kernel<<<tiles < cap ? tiles : cap, 256, 0, as_stream(stream)>>>(pose, B, H, W, focal, cx, cy, select, n,
                                                                                  ro, rd, vec_ok);
  CNRF_LAUNCH_CHECK();
  return CNRF_OK;
}

extern "C" int cnrf_ray_bundle_bwd(const float* g_ro, const float* g_rd, int64_t B, int32_t H, int32_t W, float focal,
                                   float cx, float cy, const int64_t* select, int64_t n, float* g_pose,
                                   cnrf_stream_t stream) {
  CNRF_REQUIRE(g_rd && g_pose, CNRF_E_NULL);
  CNRF_REQUIRE(B > 0 && H > 0 && W > 0 && n > 0 && n <= (int64_t)H * W, CNRF_E_SHAPE);
  CNRF_REQUIRE(select || n == (int64_t)H * W, CNRF_E_SHAPE);
  ray_bundle_bwd_kernel<<<(unsigned)B, 256, 0, as_stream(stream)>>>(g_ro, g_rd, H, W, focal, cx, cy, select, n, g_pose);
  CNRF_LAUNCH_CHECK();
  return CNRF_OK;
}
