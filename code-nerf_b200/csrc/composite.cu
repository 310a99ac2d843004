// Alpha compositing forward and backward: volume_render (view_synthesis/nerf/volumetric_render.py:36-66).
// One warp per ray, one pass over HBM: lane l of chunk c owns sample 32c+l, (r,g,b,sigma) arrives as one float4 per
// lane (512 B per warp request), transmittance is an exclusive warp scan of sigma*delta with a carry between chunks.
// The backward keeps every per-sample intermediate of the ray in registers (NCH chunks, compile-time), so raw/z are
// read exactly once and g_raw written exactly once; the suffix sum sum_{j>i} gw_j w_j runs as a reverse warp scan.
#include "common.cuh"

constexpr int CW = 4;  // warps (rays) per block

struct Sample {
  float sr, sg, sb;   // sigmoid(raw rgb)
  float sigma, delta, dist, z;
  float T, E;         // transmittance before the sample, exp(-sigma*delta)
  float dsig;         // d softplus(x-1) / dx
  bool valid;
};

__device__ __forceinline__ Sample load_sample(const float4* __restrict__ raw4, const float* __restrict__ z, int i, int S,
                                              float norm) {
  Sample s;
  s.valid = i < S;
  float4 r = s.valid ? raw4[i] : make_float4(0.f, 0.f, 0.f, 0.f);
  s.z = s.valid ? z[i] : 0.f;
  float zn = (i + 1 < S) ? z[i + 1] : 0.f;
  s.dist = (i + 1 < S) ? __fsub_rn(zn, s.z) : 1e10f;        // last distance forced to 1e10   volumetric_render.py:41-43
  s.delta = __fmul_rn(s.dist, norm);                         // dists * ||rd||                 :45
  float x = __fsub_rn(r.w, 1.0f);
  s.sigma = s.valid ? softplus_t20(x) : 0.f;                 // shifted_softplus               :32-33,48
  s.dsig = x > 20.0f ? 1.0f : sigmoidf_(x);
  s.sr = sigmoidf_(r.x); s.sg = sigmoidf_(r.y); s.sb = sigmoidf_(r.z);
  s.E = expf(-__fmul_rn(s.sigma, s.delta));
  s.T = 1.f;
  return s;
}
__device__ __forceinline__ float widen(float sg) { return __fsub_rn(__fmul_rn(sg, 1.002f), 0.001f); }  // :28-29

// Forward over one ray.  Leaves T in every Sample, returns the five reductions.
template <int NCH>
__device__ __forceinline__ void ray_forward(const float4* __restrict__ raw4, const float* __restrict__ z, int S,
                                            float norm, int lane, Sample (&sm)[NCH], float& R, float& G, float& B,
                                            float& D, float& A) {
  float carry = 0.f;
  R = G = B = D = A = 0.f;
#pragma unroll
  for (int c = 0; c < NCH; ++c) {
    Sample s = load_sample(raw4, z, c * 32 + lane, S, norm);
    float sd = s.valid ? __fmul_rn(s.sigma, s.delta) : 0.f;
    float inc = warp_scan_inclusive(sd, lane);
    float prev = __shfl_up_sync(FULL_MASK, inc, 1);
    float excl = carry + (lane > 0 ? prev : 0.f);            // exclusive cumsum             :52-55
    carry += __shfl_sync(FULL_MASK, inc, 31);
    s.T = expf(-excl);
    float w = s.valid ? __fmul_rn(__fsub_rn(1.0f, s.E), s.T) : 0.f;  // alpha * transmittance :56-57
    R += w * widen(s.sr); G += w * widen(s.sg); B += w * widen(s.sb);
    D += w * s.z; A += w;
    sm[c] = s;
  }
  R = warp_sum(R); G = warp_sum(G); B = warp_sum(B); D = warp_sum(D); A = warp_sum(A);
}

template <int NCH>
__global__ void __launch_bounds__(CW * 32)
composite_fwd_kernel(const float* __restrict__ raw, const float* __restrict__ z, const float* __restrict__ rd,
                     int64_t N, int S, float* __restrict__ rgb, float* __restrict__ disp, float* __restrict__ acc,
                     float* __restrict__ weights, float* __restrict__ depth) {
  int lane = threadIdx.x & 31;
  int64_t n = (int64_t)blockIdx.x * CW + (threadIdx.x >> 5);
  if (n >= N) return;
  float dx = rd[n * 3], dy = rd[n * 3 + 1], dz = rd[n * 3 + 2];
  float norm = sqrtf(dx * dx + dy * dy + dz * dz);
  const float4* raw4 = reinterpret_cast<const float4*>(raw) + n * (int64_t)S;
  const float* zr = z + n * (int64_t)S;
  Sample sm[NCH];
  float R, G, B, D, A;
  ray_forward<NCH>(raw4, zr, S, norm, lane, sm, R, G, B, D, A);
  if (weights) {
#pragma unroll
    for (int c = 0; c < NCH; ++c) {
      int i = c * 32 + lane;
      if (i < S) weights[n * (int64_t)S + i] = __fmul_rn(__fsub_rn(1.0f, sm[c].E), sm[c].T);
    }
  }
  if (lane == 0) {
    if (rgb) { rgb[n * 3] = R; rgb[n * 3 + 1] = G; rgb[n * 3 + 2] = B; }
    if (depth) depth[n] = D;
    if (acc) acc[n] = A;
    if (disp) disp[n] = __fdiv_rn(1.0f, fmaxf(1e-10f, __fdiv_rn(D, A)));   // :64
  }
}

template <int NCH>
__global__ void __launch_bounds__(CW * 32)
composite_bwd_kernel(const float* __restrict__ raw, const float* __restrict__ z, const float* __restrict__ rd,
                     int64_t N, int S, const float* __restrict__ g_rgb, const float* __restrict__ g_disp,
                     const float* __restrict__ g_acc, const float* __restrict__ g_weights,
                     const float* __restrict__ g_depth, float* __restrict__ g_raw, float* __restrict__ g_z,
                     float* __restrict__ g_rd) {
  int lane = threadIdx.x & 31;
  int64_t n = (int64_t)blockIdx.x * CW + (threadIdx.x >> 5);
  if (n >= N) return;
  float dx = rd[n * 3], dy = rd[n * 3 + 1], dz = rd[n * 3 + 2];
  float norm = sqrtf(dx * dx + dy * dy + dz * dz);
  const float4* raw4 = reinterpret_cast<const float4*>(raw) + n * (int64_t)S;
  const float* zr = z + n * (int64_t)S;
  Sample sm[NCH];
  float R, G, B, D, A;
  ray_forward<NCH>(raw4, zr, S, norm, lane, sm, R, G, B, D, A);

  float gR = g_rgb ? g_rgb[n * 3] : 0.f, gG = g_rgb ? g_rgb[n * 3 + 1] : 0.f, gB = g_rgb ? g_rgb[n * 3 + 2] : 0.f;
  float gD = g_depth ? g_depth[n] : 0.f, gA = g_acc ? g_acc[n] : 0.f;
  if (g_disp) {                                               // disp = 1 / max(1e-10, D/A)
    float q = D / A;
    if (q > 1e-10f) {
      float gq = -g_disp[n] / (q * q);
      gD += gq / A;
      gA -= gq * D / (A * A);
    }
  }
  // reverse pass: suffix_i = sum_{j>i} gw_j w_j
  float carry = 0.f, g_norm = 0.f;
  float ddist[NCH];       // d L / d dists_i, kept for the g_z sweep below
#pragma unroll
  for (int c = NCH - 1; c >= 0; --c) {
    const Sample& s = sm[c];
    int i = c * 32 + lane;
    float w = s.valid ? (1.0f - s.E) * s.T : 0.f;
    float cr = widen(s.sr), cg = widen(s.sg), cb = widen(s.sb);
    float gw = gR * cr + gG * cg + gB * cb + gD * s.z + gA + ((g_weights && s.valid) ? g_weights[n * (int64_t)S + i] : 0.f);
    float v = s.valid ? gw * w : 0.f;
    // inclusive suffix scan within the warp
    float sfx = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      float t = __shfl_down_sync(FULL_MASK, sfx, o);
      if (lane + o < 32) sfx += t;
    }
    float excl_sfx = sfx - v + carry;
    carry += __shfl_sync(FULL_MASK, sfx, 0);
    float ds = gw * s.T * s.E - excl_sfx;                     // dL / d(sigma*delta)
    float dsigma = ds * s.delta, ddelta = ds * s.sigma;
    ddist[c] = (s.valid && i + 1 < S) ? ddelta * norm : 0.f;
    if (s.valid) {
      g_norm += ddelta * s.dist;
      float4 o4;
      o4.x = gR * w * 1.002f * s.sr * (1.0f - s.sr);
      o4.y = gG * w * 1.002f * s.sg * (1.0f - s.sg);
      o4.z = gB * w * 1.002f * s.sb * (1.0f - s.sb);
      o4.w = dsigma * s.dsig;
      reinterpret_cast<float4*>(g_raw)[n * (int64_t)S + i] = o4;
    }
  }
  if (g_z) {   // dists_i = z_{i+1} - z_i  ->  g_z[i] = gD*w_i - ddist_i + ddist_{i-1}
    float prev_last = 0.f;
#pragma unroll
    for (int c = 0; c < NCH; ++c) {
      int i = c * 32 + lane;
      float up = __shfl_up_sync(FULL_MASK, ddist[c], 1);
      float below = lane > 0 ? up : prev_last;
      prev_last = __shfl_sync(FULL_MASK, ddist[c], 31);
      if (i < S) g_z[n * (int64_t)S + i] = gD * (1.0f - sm[c].E) * sm[c].T - ddist[c] + below;
    }
  }
  if (g_rd) {
    g_norm = warp_sum(g_norm);
    if (lane == 0) {
      float k = norm > 0.f ? g_norm / norm : 0.f;
      g_rd[n * 3] = k * dx; g_rd[n * 3 + 1] = k * dy; g_rd[n * 3 + 2] = k * dz;
    }
  }
}

#define DISPATCH_NCH(S, CALL)                                  \
  switch (((S) + 31) / 32) {                                   \
    case 1: { constexpr int NCH = 1; CALL; } break;            \
    case 2: { constexpr int NCH = 2; CALL; } break;            \
    case 3: { constexpr int NCH = 3; CALL; } break;            \
    case 4: { constexpr int NCH = 4; CALL; } break;            \
    case 5: { constexpr int NCH = 5; CALL; } break;            \
    case 6: { constexpr int NCH = 6; CALL; } break;            \
    case 7: case 8: { constexpr int NCH = 8; CALL; } break;    \
    default: return CNRF_E_SHAPE;                              \
  }

extern "C" int cnrf_composite_fwd(const float* raw, const float* z, const float* rd, int64_t N, int32_t S, float* rgb,
                                  float* disp, float* acc, float* weights, float* depth, cnrf_stream_t stream) {
  CNRF_REQUIRE(raw && z && rd, CNRF_E_NULL);
  CNRF_REQUIRE(N > 0 && S > 0 && S <= 256, CNRF_E_SHAPE);
  DISPATCH_NCH(S, (composite_fwd_kernel<NCH><<<grid_for(N, CW), CW * 32, 0, as_stream(stream)>>>(
                      raw, z, rd, N, S, rgb, disp, acc, weights, depth)));
  CNRF_LAUNCH_CHECK();
  return CNRF_OK;
}

extern "C" int cnrf_composite_bwd(const float* raw, const float* z, const float* rd, int64_t N, int32_t S,
                                  const float* g_rgb, const float* g_disp, const float* g_acc, const float* g_weights,
                                  const float* g_depth, float* g_raw, float* g_z, float* g_rd, cnrf_stream_t stream) {
  CNRF_REQUIRE(raw && z && rd && g_raw, CNRF_E_NULL);
  CNRF_REQUIRE(N > 0 && S > 0 && S <= 256, CNRF_E_SHAPE);
  DISPATCH_NCH(S, (composite_bwd_kernel<NCH><<<grid_for(N, CW), CW * 32, 0, as_stream(stream)>>>(
                      raw, z, rd, N, S, g_rgb, g_disp, g_acc, g_weights, g_depth, g_raw, g_z, g_rd)));
  CNRF_LAUNCH_CHECK();
  return CNRF_OK;
}
