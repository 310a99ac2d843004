// Alpha compositing forward and backward: volume_render (view_synthesis/nerf/volumetric_render.py:36-66).
// One warp per ray (R rays per warp when a ray is a single 32-sample chunk), one pass over HBM: lane l of chunk c owns
// sample 32c+l, (r,g,b,sigma) arrives as one float4 per lane (512 B per warp request).  EVERY load of the warp's rays is
// issued before the first dependent instruction (the kernel is a stream: bytes in flight per SM are what set its
// speed), the next-sample depth comes from a shuffle instead of a second load, transmittance is an exclusive warp scan
// of sigma*delta with a carry between chunks.  The backward keeps the per-sample intermediates of the ray in registers
// (NCH chunks, compile-time), so raw/z are read exactly once and g_raw written exactly once; the suffix sum
// sum_{j>i} gw_j w_j runs as a reverse warp scan.
//
// Transcendentals: the first version used expf / log1pf / IEEE division everywhere and was ISSUE-bound (~250 executed
// instructions per 32 samples, 0.3-0.5 of the HBM peak).  exp(-x) of sigmoid, alpha and transmittance now go through
// ex2.approx (__expf: |error| <= 2 ulp + 1.2e-7*|x| relative, results <= 1 so <= 1e-7 absolute), reciprocals through
// rcp.approx (1 ulp), softplus through lg2.approx above e^x = 1/32 and a 4-term log1p series below it (relative error
// < 2e-7; exp(raw-1) itself stays the accurate expf because the LAST sample multiplies sigma by 1e10, :43).  Parity with
// the reference stays within 1e-6 absolute (tests/test_gpu_parity.py::test_volume_render_*: 2e-6 abs on every output).
#include "common.cuh"
#include "internal.h"

constexpr int CWARPS = 8;  // warps per block

// target pixel channel of ray n for the loss epilogue (internal.h: LossTarget); the uint8 form is the loader's own
// (float)((double)v / 255.0) (datasets/dataset.py:72,90)
__device__ __forceinline__ float loss_target(const LossTarget& t, int64_t n, int c) {
  if (t.views) return (float)((double)__ldg(t.views + __ldg(t.pixel + n) * t.channels + c) / 255.0);
  return __ldg(t.rows + n * t.ld + c);
}

// single-instruction SFU forms (flush-to-zero: a subnormal exp(-x) or alpha is 0 to every tolerance used here)
__device__ __forceinline__ float ex2_ftz(float x) { float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float lg2_ftz(float x) { float y; asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float rcp_fast(float x) { float y; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float exp_neg(float x) { return ex2_ftz(x * -1.4426950408889634f); }       // exp(-x)
__device__ __forceinline__ float sigmoid_fast(float x) { return rcp_fast(1.0f + exp_neg(x)); }
__device__ __forceinline__ float widen(float sg) { return __fsub_rn(__fmul_rn(sg, 1.002f), 0.001f); }  // :28-29

// Sum K (power of two <= 32) per-lane values over the warp with K-1 + log2(32/K) shuffles instead of 5K: at every
// halving step a lane keeps one value of each pair and sends the other to its partner.  Lane l returns the total of
// value index  bit0 = l&16, bit1 = l&8, bit2 = l&4, bit3 = l&2, bit4 = l&1  (as far as K reaches).
template <int K>
__device__ __forceinline__ float warp_reduce_multi(float (&v)[K], int lane) {
  int off = 16;
#pragma unroll
  for (int k = K; k > 1; k >>= 1, off >>= 1) {
    const bool hi = (lane & off) != 0;
#pragma unroll
    for (int i = 0; i < k / 2; ++i) {
      const float keep = hi ? v[2 * i + 1] : v[2 * i], send = hi ? v[2 * i] : v[2 * i + 1];
      v[i] = keep + __shfl_xor_sync(FULL_MASK, send, off);
    }
  }
  float c = v[0];
#pragma unroll
  for (; off > 0; off >>= 1) c += __shfl_xor_sync(FULL_MASK, c, off);
  return c;
}

// sigma = softplus(x) with torch's threshold 20 (:32-33,48); dsig = d sigma / dx
template <bool WANT_D>
__device__ __forceinline__ void softplus_fast(float x, float& sigma, float& dsig) {
  const float e = expf(x);
  const float u = 1.0f + e;
  const float series = e * fmaf(e, fmaf(e, fmaf(e, -0.25f, 0.33333334f), -0.5f), 1.0f);   // log1p(e), e < 1/32
  const float sp = e < 0.03125f ? series : lg2_ftz(u) * 0.6931471805599453f;
  sigma = x > 20.0f ? x : sp;
  if (WANT_D) dsig = x > 20.0f ? 1.0f : e * rcp_fast(u);
}

// The registers of one ray: NCH chunks of 32 samples.
template <int NCH>
struct RayRegs {
  float4 raw[NCH];
  float z[NCH];
};

template <int NCH>
__device__ __forceinline__ void ray_load(RayRegs<NCH>& q, const float* __restrict__ raw, const float* __restrict__ z,
                                         int64_t n, int S, int lane, bool live) {
  const float4* raw4 = reinterpret_cast<const float4*>(raw) + n * (int64_t)S;
  const float* zr = z + n * (int64_t)S;
#pragma unroll
  for (int c = 0; c < NCH; ++c) {
    const int i = c * 32 + lane;
    const bool ok = live && i < S;
    q.raw[c] = ok ? __ldg(raw4 + i) : make_float4(0.f, 0.f, 0.f, 0.f);
    q.z[c] = ok ? __ldg(zr + i) : 0.f;
  }
}

// dists (:41-43): z[i+1] - z[i], the last one forced to 1e10.  z[i+1] of lane 31 is lane 0 of the next chunk.
template <int NCH>
__device__ __forceinline__ float next_dist(const RayRegs<NCH>& q, int c, int S, int lane) {
  float zn = __shfl_down_sync(FULL_MASK, q.z[c], 1);
  if (c + 1 < NCH) {
    const float first = __shfl_sync(FULL_MASK, q.z[c + 1], 0);
    zn = lane == 31 ? first : zn;
  }
  const int i = c * 32 + lane;
  return (i + 1 < S) ? __fsub_rn(zn, q.z[c]) : 1e10f;
}

struct Sums { float R, G, B, D, A; };

// ---------------------------------------------------------------------------------------------------------------
// LOSS: the photometric loss rides in the epilogue (CompositeLossFwd); a separate instantiation, so that the plain kernels
// stay exactly the code they were (the loss arguments cost the non-loss launches 3-16 % when they shared one body)
template <int NCH, int R, bool LOSS>
__global__ void __launch_bounds__(CWARPS * 32)
composite_fwd_kernel(const float* __restrict__ raw, const float* __restrict__ z, const float* __restrict__ rd,
                     int64_t N, int S, float* __restrict__ rgb, float* __restrict__ disp, float* __restrict__ acc,
                     float* __restrict__ weights, float* __restrict__ depth, const CompositeLossFwd lf) {
  const int lane = threadIdx.x & 31;
  const int64_t n0 = ((int64_t)blockIdx.x * CWARPS + (threadIdx.x >> 5)) * R;
  if (n0 >= N) return;
  RayRegs<NCH> q[R];
  float nrm[R];
#pragma unroll
  for (int r = 0; r < R; ++r) ray_load<NCH>(q[r], raw, z, n0 + r, S, lane, n0 + r < N);
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const int64_t n = min(n0 + r, N - 1);
    const float dx = __ldg(rd + n * 3), dy = __ldg(rd + n * 3 + 1), dz = __ldg(rd + n * 3 + 2);
    nrm[r] = sqrtf(dx * dx + dy * dy + dz * dz);
  }
  float sums[R * 8];     // per ray: R, G, B, depth, acc, 0, 0, 0
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const int64_t n = n0 + r;
    float carry = 0.f;
    Sums s = {0.f, 0.f, 0.f, 0.f, 0.f};
    {   // rays past N compute on zero-filled registers (no divergent region around the shuffles); only stores are guarded
#pragma unroll
      for (int c = 0; c < NCH; ++c) {
        const int i = c * 32 + lane;
        const bool valid = i < S;
        const float4 rw = q[r].raw[c];
        const float delta = __fmul_rn(next_dist<NCH>(q[r], c, S, lane), nrm[r]);      // dists * ||rd||      :45
        float sigma, unused;
        softplus_fast<false>(__fsub_rn(rw.w, 1.0f), sigma, unused);
        const float sd = valid ? __fmul_rn(sigma, delta) : 0.f;
        const float E = exp_neg(sd);
        const float inc = warp_scan_inclusive(sd, lane);
        const float prev = __shfl_up_sync(FULL_MASK, inc, 1);
        const float excl = carry + (lane > 0 ? prev : 0.f);                            // exclusive cumsum    :52-55
        carry += __shfl_sync(FULL_MASK, inc, 31);
        const float T = exp_neg(excl);
        const float w = valid ? __fmul_rn(__fsub_rn(1.0f, E), T) : 0.f;                // alpha * T           :56-57
        s.R = fmaf(w, widen(sigmoid_fast(rw.x)), s.R);
        s.G = fmaf(w, widen(sigmoid_fast(rw.y)), s.G);
        s.B = fmaf(w, widen(sigmoid_fast(rw.z)), s.B);
        s.D = fmaf(w, q[r].z[c], s.D);
        s.A += w;
        if (weights && valid && n < N) weights[n * (int64_t)S + i] = w;
      }
    }
    sums[r * 8 + 0] = s.R; sums[r * 8 + 1] = s.G; sums[r * 8 + 2] = s.B; sums[r * 8 + 3] = s.D; sums[r * 8 + 4] = s.A;
    sums[r * 8 + 5] = 0.f; sums[r * 8 + 6] = 0.f; sums[r * 8 + 7] = 0.f;
  }
  // one joint reduction: lane l ends with quantity qn = (l>>4 & 1) | (l>>3 & 1) << 1 | (l>>2 & 1) << 2 of ray
  // r = (l>>1 & 1) | (l & 1) << 1 (for R = 4; fewer ray bits for smaller R); the other lanes hold copies.
  const float tot = warp_reduce_multi<R * 8>(sums, lane);
  const int qn = ((lane >> 4) & 1) | (((lane >> 3) & 1) << 1) | (((lane >> 2) & 1) << 2);
  const int rr = R == 1 ? 0 : (R == 2 ? ((lane >> 1) & 1) : (((lane >> 1) & 1) | ((lane & 1) << 1)));
  const bool writer = R == 1 ? (lane & 3) == 0 : (R == 2 ? (lane & 1) == 0 : true);
  const float other = __shfl_xor_sync(FULL_MASK, tot, 28);          // depth lane (qn 3) <-> acc lane (qn 4)
  const int64_t n = n0 + rr;
  if (writer && n < N) {
    if (qn < 3) { if (rgb) rgb[n * 3 + qn] = tot; }
    else if (qn == 3) {
      if (depth) depth[n] = tot;
      if (disp) disp[n] = rcp_fast(fmaxf(1e-10f, tot * rcp_fast(other)));             // 1 / max(1e-10, depth / acc)  :64
    } else if (qn == 4) { if (acc) acc[n] = tot; }
  }
  if (LOSS) {           // photometric loss folded into the epilogue (train.py:103-104): sum over rays and channels of (rgb - target)^2
    float e = 0.f;
    if (writer && n < N && qn < 3) { const float d = tot - loss_target(lf.t, n, qn); e = d * d; }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) e += __shfl_xor_sync(FULL_MASK, e, o);
    if (lane == 0) {
      atomicAdd(lf.sq_sum, e);
      if (lf.out3) {    // the last warp of the fine pass turns the two sums into (coarse_loss, fine_loss, loss)
        __threadfence();
        if (atomicAdd(lf.done, 1u) == lf.num_warps - 1u) {
          __threadfence();
          const float a = *(volatile float*)lf.other_sq * lf.inv, b = *(volatile float*)lf.sq_sum * lf.inv;
          lf.out3[0] = a; lf.out3[1] = b; lf.out3[2] = a + b;
        }
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------------------------
// GEO: the caller wants g_z and / or g_rd (camera-pose optimisation); training needs neither.
template <int NCH, int R, bool GEO, bool LOSS>
__global__ void __launch_bounds__(CWARPS * 32)
composite_bwd_kernel(const float* __restrict__ raw, const float* __restrict__ z, const float* __restrict__ rd,
                     int64_t N, int S, const float* __restrict__ g_rgb, const float* __restrict__ g_disp,
                     const float* __restrict__ g_acc, const float* __restrict__ g_weights,
                     const float* __restrict__ g_depth, float* __restrict__ g_raw, float* __restrict__ g_z,
                     float* __restrict__ g_rd, const CompositeLossBwd lb) {
  const int lane = threadIdx.x & 31;
  const int64_t n0 = ((int64_t)blockIdx.x * CWARPS + (threadIdx.x >> 5)) * R;
  if (n0 >= N) return;
  RayRegs<NCH> q[R];
  float gwt[R][NCH];
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const int64_t n = n0 + r;
    ray_load<NCH>(q[r], raw, z, n, S, lane, n < N);
#pragma unroll
    for (int c = 0; c < NCH; ++c) {
      const int i = c * 32 + lane;
      gwt[r][c] = (g_weights && n < N && i < S) ? __ldg(g_weights + n * (int64_t)S + i) : 0.f;
    }
  }
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const int64_t n = n0 + r;
    if (n >= N) break;
    const float dx = __ldg(rd + n * 3), dy = __ldg(rd + n * 3 + 1), dz = __ldg(rd + n * 3 + 2);
    const float norm = sqrtf(dx * dx + dy * dy + dz * dz);
    float gR = g_rgb ? __ldg(g_rgb + n * 3) : 0.f, gG = g_rgb ? __ldg(g_rgb + n * 3 + 1) : 0.f,
          gB = g_rgb ? __ldg(g_rgb + n * 3 + 2) : 0.f;
    if (LOSS) {         // the loss lives in the forward epilogue: d loss / d rgb = (g_term + g_sum) * 2 (rgb - target) / (3 N)
      const float k = ((lb.g_term ? __ldg(lb.g_term) : 0.f) + (lb.g_sum ? __ldg(lb.g_sum) : 0.f)) * lb.scale;
      gR = k * (__ldg(lb.rgb_out + n * 3) - loss_target(lb.t, n, 0));
      gG = k * (__ldg(lb.rgb_out + n * 3 + 1) - loss_target(lb.t, n, 1));
      gB = k * (__ldg(lb.rgb_out + n * 3 + 2) - loss_target(lb.t, n, 2));
    }
    float gD = g_depth ? __ldg(g_depth + n) : 0.f, gA = g_acc ? __ldg(g_acc + n) : 0.f;

    // ---- forward sweep: per-sample state stays in registers
    float sr[NCH], sg[NCH], sb[NCH], sigma[NCH], dsig[NCH], delta[NCH], T[NCH], E[NCH];
    float carry = 0.f, D = 0.f, A = 0.f;
#pragma unroll
    for (int c = 0; c < NCH; ++c) {
      const int i = c * 32 + lane;
      const bool valid = i < S;
      const float4 rw = q[r].raw[c];
      delta[c] = __fmul_rn(next_dist<NCH>(q[r], c, S, lane), norm);
      softplus_fast<true>(__fsub_rn(rw.w, 1.0f), sigma[c], dsig[c]);
      const float sd = valid ? __fmul_rn(sigma[c], delta[c]) : 0.f;
      E[c] = exp_neg(sd);
      const float inc = warp_scan_inclusive(sd, lane);
      const float prev = __shfl_up_sync(FULL_MASK, inc, 1);
      const float excl = carry + (lane > 0 ? prev : 0.f);
      carry += __shfl_sync(FULL_MASK, inc, 31);
      T[c] = exp_neg(excl);
      sr[c] = sigmoid_fast(rw.x); sg[c] = sigmoid_fast(rw.y); sb[c] = sigmoid_fast(rw.z);
      if (g_disp) {
        const float w = valid ? (1.0f - E[c]) * T[c] : 0.f;
        D = fmaf(w, q[r].z[c], D); A += w;
      }
    }
    if (g_disp) {                                               // disp = 1 / max(1e-10, D/A)
      D = warp_sum(D); A = warp_sum(A);
      const float qq = D / A;
      if (qq > 1e-10f) {
        const float gq = -__ldg(g_disp + n) / (qq * qq);
        gD += gq / A;
        gA -= gq * D / (A * A);
      }
    }
    // ---- reverse sweep: suffix_i = sum_{j>i} gw_j w_j
    float rcarry = 0.f, g_norm = 0.f;
    float ddist[GEO ? NCH : 1];       // d L / d dists_i, kept for the g_z sweep below
#pragma unroll
    for (int c = NCH - 1; c >= 0; --c) {
      const int i = c * 32 + lane;
      const bool valid = i < S;
      const float w = valid ? (1.0f - E[c]) * T[c] : 0.f;
      const float gw = gR * widen(sr[c]) + gG * widen(sg[c]) + gB * widen(sb[c]) + gD * q[r].z[c] + gA + gwt[r][c];
      const float v = gw * w;
      float sfx = v;                                            // inclusive suffix scan within the warp
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const float t = __shfl_down_sync(FULL_MASK, sfx, o);
        if (lane + o < 32) sfx += t;
      }
      const float excl_sfx = sfx - v + rcarry;
      rcarry += __shfl_sync(FULL_MASK, sfx, 0);
      const float ds = gw * T[c] * E[c] - excl_sfx;             // dL / d(sigma*delta)
      const float ddelta = ds * sigma[c];
      if constexpr (GEO) {
        ddist[c] = (valid && i + 1 < S) ? ddelta * norm : 0.f;
        if (valid) g_norm += ddelta * (delta[c] / norm);        // * dists_i
      }
      if (valid) {
        float4 o4;
        o4.x = gR * w * 1.002f * sr[c] * (1.0f - sr[c]);
        o4.y = gG * w * 1.002f * sg[c] * (1.0f - sg[c]);
        o4.z = gB * w * 1.002f * sb[c] * (1.0f - sb[c]);
        o4.w = ds * delta[c] * dsig[c];
        reinterpret_cast<float4*>(g_raw)[n * (int64_t)S + i] = o4;
      }
    }
    if constexpr (GEO) {
      if (g_z) {   // dists_i = z_{i+1} - z_i  ->  g_z[i] = gD*w_i - ddist_i + ddist_{i-1}
        float prev_last = 0.f;
#pragma unroll
        for (int c = 0; c < NCH; ++c) {
          const int i = c * 32 + lane;
          const float up = __shfl_up_sync(FULL_MASK, ddist[c], 1);
          const float below = lane > 0 ? up : prev_last;
          prev_last = __shfl_sync(FULL_MASK, ddist[c], 31);
          if (i < S) g_z[n * (int64_t)S + i] = gD * (1.0f - E[c]) * T[c] - ddist[c] + below;
        }
      }
      if (g_rd) {
        g_norm = warp_sum(g_norm);
        if (lane == 0) {
          const float k = norm > 0.f ? g_norm / norm : 0.f;
          g_rd[n * 3] = k * dx; g_rd[n * 3 + 1] = k * dy; g_rd[n * 3 + 2] = k * dz;
        }
      }
    }
  }
}

template <int NCH, int R>
static void launch_fwd(const float* raw, const float* z, const float* rd, int64_t N, int S, float* rgb, float* disp,
                       float* acc, float* weights, float* depth, CompositeLossFwd lf, cudaStream_t st) {
  lf.num_warps = (unsigned)((N + R - 1) / R);
  if (lf.sq_sum)
    composite_fwd_kernel<NCH, R, true><<<grid_for(N, CWARPS * R), CWARPS * 32, 0, st>>>(raw, z, rd, N, S, rgb, disp, acc, weights,
                                                                                         depth, lf);
  else
    composite_fwd_kernel<NCH, R, false><<<grid_for(N, CWARPS * R), CWARPS * 32, 0, st>>>(raw, z, rd, N, S, rgb, disp, acc, weights,
                                                                                          depth, lf);
}
template <int NCH, int R>
static void launch_bwd(const float* raw, const float* z, const float* rd, int64_t N, int S, const float* g_rgb,
                       const float* g_disp, const float* g_acc, const float* g_weights, const float* g_depth,
                       float* g_raw, float* g_z, float* g_rd, const CompositeLossBwd& lb, cudaStream_t st) {
  const unsigned grid = grid_for(N, CWARPS * R);
  if (lb.rgb_out)      // (the loss variant serves the fused training entry: rays carry no gradient there)
    composite_bwd_kernel<NCH, R, false, true><<<grid, CWARPS * 32, 0, st>>>(raw, z, rd, N, S, g_rgb, g_disp, g_acc, g_weights,
                                                                             g_depth, g_raw, nullptr, nullptr, lb);
  else if (g_z || g_rd)
    composite_bwd_kernel<NCH, R, true, false><<<grid, CWARPS * 32, 0, st>>>(raw, z, rd, N, S, g_rgb, g_disp, g_acc, g_weights,
                                                                             g_depth, g_raw, g_z, g_rd, lb);
  else
    composite_bwd_kernel<NCH, R, false, false><<<grid, CWARPS * 32, 0, st>>>(raw, z, rd, N, S, g_rgb, g_disp, g_acc, g_weights,
                                                                              g_depth, g_raw, g_z, g_rd, lb);
}

// chunks of 32 samples per ray -> (NCH, rays per warp forward, rays per warp backward)
#define DISPATCH_NCH(S, FWD_OR_BWD, ...)                                  \
  switch (((S) + 31) / 32) {                                              \
    case 1: FWD_OR_BWD(1, 4, 2, __VA_ARGS__); break;                      \
    case 2: FWD_OR_BWD(2, 2, 1, __VA_ARGS__); break;                      \
    case 3: FWD_OR_BWD(3, 1, 1, __VA_ARGS__); break;                      \
    case 4: FWD_OR_BWD(4, 1, 1, __VA_ARGS__); break;                      \
    case 5: FWD_OR_BWD(5, 1, 1, __VA_ARGS__); break;                      \
    case 6: FWD_OR_BWD(6, 1, 1, __VA_ARGS__); break;                      \
    case 7: case 8: FWD_OR_BWD(8, 1, 1, __VA_ARGS__); break;              \
    default: return CNRF_E_SHAPE;                                         \
  }
#define CALL_FWD(NCH, RF, RB, ...) launch_fwd<NCH, RF>(__VA_ARGS__)
#define CALL_BWD(NCH, RF, RB, ...) launch_bwd<NCH, RB>(__VA_ARGS__)

int cnrf_composite_fwd_loss(const float* raw, const float* z, const float* rd, int64_t N, int32_t S, float* rgb, float* disp,
                            float* acc, float* weights, float* depth, const CompositeLossFwd& lf, cudaStream_t st) {
  CNRF_REQUIRE(raw && z && rd, CNRF_E_NULL);
  CNRF_REQUIRE(N > 0 && S > 0 && S <= 256, CNRF_E_SHAPE);
  DISPATCH_NCH(S, CALL_FWD, raw, z, rd, N, S, rgb, disp, acc, weights, depth, lf, st);
  CNRF_LAUNCH_CHECK();
  return CNRF_OK;
}
extern "C" int cnrf_composite_fwd(const float* raw, const float* z, const float* rd, int64_t N, int32_t S, float* rgb,
                                  float* disp, float* acc, float* weights, float* depth, cnrf_stream_t stream) {
  return cnrf_composite_fwd_loss(raw, z, rd, N, S, rgb, disp, acc, weights, depth, CompositeLossFwd(), as_stream(stream));
}

int cnrf_composite_bwd_loss(const float* raw, const float* z, const float* rd, int64_t N, int32_t S, const float* g_rgb,
                            const float* g_disp, const float* g_acc, const float* g_weights, const float* g_depth, float* g_raw,
                            float* g_z, float* g_rd, const CompositeLossBwd& lb, cudaStream_t st) {
  CNRF_REQUIRE(raw && z && rd && g_raw, CNRF_E_NULL);
  CNRF_REQUIRE(!(lb.rgb_out && (g_z || g_rd)), CNRF_E_UNSUPPORTED);
  CNRF_REQUIRE(N > 0 && S > 0 && S <= 256, CNRF_E_SHAPE);
  DISPATCH_NCH(S, CALL_BWD, raw, z, rd, N, S, g_rgb, g_disp, g_acc, g_weights, g_depth, g_raw, g_z, g_rd, lb, st);
  CNRF_LAUNCH_CHECK();
  return CNRF_OK;
}
extern "C" int cnrf_composite_bwd(const float* raw, const float* z, const float* rd, int64_t N, int32_t S,
                                  const float* g_rgb, const float* g_disp, const float* g_acc, const float* g_weights,
                                  const float* g_depth, float* g_raw, float* g_z, float* g_rd, cnrf_stream_t stream) {
  return cnrf_composite_bwd_loss(raw, z, rd, N, S, g_rgb, g_disp, g_acc, g_weights, g_depth, g_raw, g_z, g_rd, CompositeLossBwd(),
                                 as_stream(stream));
}
