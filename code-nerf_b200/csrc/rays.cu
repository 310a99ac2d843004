// Ray generation: RaySampler.get_bundle / .sample  (view_synthesis/nerf/ray_sampler.py:31-51, 66-80, 84-99).
// One thread per ray: the camera-frame direction is recomputed from the pixel index with the same IEEE operations
// the reference uses to build its (H,W,3) table, rotated by the pose, and written straight to the packed (B*n,3)
// output -- the full-image bundle and the advanced-index gather of the reference never exist in HBM.
#include "common.cuh"
#include "internal.h"

__device__ __forceinline__ void camera_dir(int64_t flat, int W, float focal, float cx, float cy, float d[3]) {
  // flat = h*W + w (flatten(1,2), ray_sampler.py:69); a pixel index fits 32 bits (H, W are ints): one 32-bit division instead
  // of the emulated 64-bit division + modulo that made this write-only kernel issue-bound
  const uint32_t f = (uint32_t)flat;
  const int h = (int)(f / (uint32_t)W), w = (int)(f - (uint32_t)h * (uint32_t)W);
  d[0] = __fdiv_rn(__fsub_rn((float)w, cx), focal);      // (ii - cx) / f            ray_sampler.py:46
  d[1] = __fdiv_rn(-__fsub_rn((float)h, cy), focal);     // -(jj - cy) / f           ray_sampler.py:47
  d[2] = -1.0f;                                          // -ones                    ray_sampler.py:48
}

// q = x / d, r = x - q d for x < 2^31, d < 2^15 through one multiply-high (m = floor(2^32 / d) + 1) and one correction:
// this write-only kernel was ISSUE-bound (~100 instructions per ray: two emulated divisions, two IEEE divides, twelve
// scalar pose loads), not bandwidth-bound -- a memset of the same buffers runs at 0.94 of the copy peak, it ran at 0.54.
__device__ __forceinline__ void divmod_u32(uint32_t x, uint32_t d, uint32_t m, uint32_t& q, uint32_t& r) {
  q = __umulhi(x, m);
  r = x - q * d;
  if (r >= d) { q -= 1; r += d; }          // m over-estimates 1/d: q can only be one too large (r wrapped around)
}

// A block owns tiles of 256 consecutive rays (grid-stride).  The camera-frame direction components depend on the pixel's
// column (x) or row (y) only: the block computes the W + H IEEE quotients of ray_sampler.py:46-47 ONCE into shared memory
// (bit-identical values, two table reads per ray instead of two divides).  Full tiles stage their 768 + 768 floats in
// shared memory and write them as float4 (whole 128-byte lines per warp instruction instead of 4-byte stores at a
// 12-byte stride); the last tile, or pointers that are not 16-byte aligned, take the scalar path.
constexpr int RAY_TABLE_MAX = 1024;          // H, W up to 1024 use the tables (8 KB); larger images compute per ray
__global__ void __launch_bounds__(256) ray_bundle_fwd_kernel(const float* __restrict__ pose, int64_t B, int H, int W, float focal,
                                                             float cx, float cy, const int64_t* __restrict__ select, int64_t n,
                                                             float* __restrict__ ro, float* __restrict__ rd, int vec_ok,
                                                             uint32_t magic_n, uint32_t magic_w) {
  __shared__ __align__(16) float s_rd_buf[2][768];       // double-buffered: one barrier per tile
  __shared__ __align__(16) float s_ro_buf[2][768];
  __shared__ float s_dx[RAY_TABLE_MAX], s_dy[RAY_TABLE_MAX];
  const bool tables = H <= RAY_TABLE_MAX && W <= RAY_TABLE_MAX;
  if (tables) {
    for (int k = threadIdx.x; k < W; k += 256) s_dx[k] = __fdiv_rn(__fsub_rn((float)k, cx), focal);     // (ii - cx) / f   :46
    for (int k = threadIdx.x; k < H; k += 256) s_dy[k] = __fdiv_rn(-__fsub_rn((float)k, cy), focal);    // -(jj - cy) / f  :47
    __syncthreads();
  }
  int phase = 0;
  const int64_t total = B * n;
  const bool small = total < (1ll << 31) && n < (1ll << 15) * (1ll << 15) && W < (1 << 15);
  const bool pose_vec = (reinterpret_cast<uintptr_t>(pose) & 15) == 0;
  for (int64_t i0 = (int64_t)blockIdx.x * 256; i0 < total; i0 += (int64_t)gridDim.x * 256) {
  const int64_t i = i0 + threadIdx.x;
  const bool full = vec_ok && i0 + 256 <= total;
  float v[3] = {0.f, 0.f, 0.f}, o[3] = {0.f, 0.f, 0.f};
  if (i < total) {
    int64_t b, flat;
    if (small && magic_n) {
      uint32_t bb, rem;
      divmod_u32((uint32_t)i, (uint32_t)n, magic_n, bb, rem);
      b = bb; flat = select ? __ldg(select + i) : (int64_t)rem;
    } else { b = i / n; flat = select ? __ldg(select + i) : (i - b * n); }
    float d[3];
    if (tables && small && magic_w) {
      uint32_t h, w;
      divmod_u32((uint32_t)flat, (uint32_t)W, magic_w, h, w);        // flat = h*W + w (flatten(1,2), ray_sampler.py:69)
      d[0] = s_dx[w]; d[1] = s_dy[h]; d[2] = -1.0f;                  // -ones   ray_sampler.py:48
    } else {
      camera_dir(flat, W, focal, cx, cy, d);
    }
    const float* P = pose + b * 16;
    float R[3][4];
    if (pose_vec) {
#pragma unroll
      for (int j = 0; j < 3; ++j) { const float4 r4 = __ldg(reinterpret_cast<const float4*>(P) + j); R[j][0] = r4.x; R[j][1] = r4.y; R[j][2] = r4.z; R[j][3] = r4.w; }
    } else {
#pragma unroll
      for (int j = 0; j < 3; ++j)
#pragma unroll
        for (int k = 0; k < 4; ++k) R[j][k] = __ldg(P + j * 4 + k);
    }
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      // einsum 'hwij,bji->bhwj': rd_j = sum_i d_i * R[j][i]   (ray_sampler.py:97)
      float a = __fmul_rn(d[0], R[j][0]);
      a = __fmaf_rn(d[1], R[j][1], a);
      a = __fmaf_rn(d[2], R[j][2], a);
      v[j] = a;
      o[j] = R[j][3];                                       // ro = pose[:3, -1]  (ray_sampler.py:98)
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
  float* s_rd = s_rd_buf[phase];
  float* s_ro = s_ro_buf[phase];
  phase ^= 1;
#pragma unroll
  for (int j = 0; j < 3; ++j) { s_rd[threadIdx.x * 3 + j] = v[j]; s_ro[threadIdx.x * 3 + j] = o[j]; }
  __syncthreads();        // (a buffer is rewritten two tiles later: every thread has passed the next tile's barrier by then)
  if (threadIdx.x < 192) {
    __stcs(reinterpret_cast<float4*>(rd + i0 * 3) + threadIdx.x, reinterpret_cast<const float4*>(s_rd)[threadIdx.x]);
    if (ro) __stcs(reinterpret_cast<float4*>(ro + i0 * 3) + threadIdx.x, reinterpret_cast<const float4*>(s_ro)[threadIdx.x]);
  }
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
  // multiply-high reciprocals of n and W (0 = not representable: the kernel divides)
  const uint32_t magic_n = n < (1ll << 30) ? (uint32_t)((1ull << 32) / (uint64_t)n) + 1u : 0u;
  const uint32_t magic_w = (uint32_t)((1ull << 32) / (uint64_t)W) + 1u;
  ray_bundle_fwd_kernel<<<tiles < cap ? tiles : cap, 256, 0, as_stream(stream)>>>(pose, B, H, W, focal, cx, cy, select, n,
                                                                                  ro, rd, vec_ok, n > 1 ? magic_n : 0u,
                                                                                  W > 1 ? magic_w : 0u);
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
