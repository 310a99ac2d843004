// Stratified + hierarchical sampling: PointSampler.sample_uniform / .sample_pdf
// (view_synthesis/nerf/point_sampler.py:49-120).  One warp per ray.
//   sample_uniform : z = lower + (upper-lower)*t  and pts = ro + rd*z, float4 stores of the packed (S,3) row.
//   sample_pdf     : warp-scan CDF in registers -> staged in shared memory, per-lane binary search for
//                    searchsorted(right=True), lerp, in-warp bitonic sort of [z_coarse | samples], float4 stores.
// Multiplies and adds are kept un-fused (__fmul_rn/__fadd_rn) wherever the reference's eager ops are, so that given
// an identical CDF the samples, sorted depths and points are bit-identical to the reference.
#include "common.cuh"
#include <math_constants.h>
#include <stdlib.h>

constexpr int WARPS_PER_BLOCK = 4;

// pts row writer: the ray's S*3 floats are contiguous; element e -> (s = e/3, c = e%3).  One float4 covers floats
// 4q..4q+3 = samples s = 4(q/3) + q%3 and s+1, components rotated by q%3: no per-float division.
__device__ __forceinline__ void write_points(float* __restrict__ dst, const float* zsm, int S, float ro0, float ro1,
                                             float ro2, float rd0, float rd1, float rd2, int lane) {
  const int total = S * 3;
  if ((total & 3) == 0 && ((reinterpret_cast<uintptr_t>(dst) & 15) == 0)) {
    float4* d4 = reinterpret_cast<float4*>(dst);
    for (int q = lane; q < total / 4; q += 32) {
      const int m = q / 3, rem = q - 3 * m, s = 4 * m + rem;
      const float za = zsm[s], zb = zsm[s + 1];
      const float oA = rem == 0 ? ro0 : (rem == 1 ? ro1 : ro2), dA = rem == 0 ? rd0 : (rem == 1 ? rd1 : rd2);
      const float oB = rem == 0 ? ro1 : (rem == 1 ? ro2 : ro0), dB = rem == 0 ? rd1 : (rem == 1 ? rd2 : rd0);
      const float oC = rem == 0 ? ro2 : (rem == 1 ? ro0 : ro1), dC = rem == 0 ? rd2 : (rem == 1 ? rd0 : rd1);
      const float z2 = rem == 2 ? zb : za, z3 = rem == 0 ? za : zb;
      d4[q] = make_float4(__fadd_rn(oA, __fmul_rn(dA, za)), __fadd_rn(oB, __fmul_rn(dB, z2)),      // ro + rd * z
                          __fadd_rn(oC, __fmul_rn(dC, z3)), __fadd_rn(oA, __fmul_rn(dA, zb)));      // (point_sampler.py:70,118)
    }
  } else {
    for (int e = lane; e < total; e += 32) {
      const int s = e / 3, c = e - 3 * s;
      const float o = c == 0 ? ro0 : (c == 1 ? ro1 : ro2);
      const float d = c == 0 ? rd0 : (c == 1 ? rd1 : rd2);
      dst[e] = __fadd_rn(o, __fmul_rn(d, zsm[s]));
    }
  }
}

// Four consecutive samples = 12 floats = three float4 stores from one lane (one LDS.128 for the depths, no selects).
__device__ __forceinline__ void write_point_group(float* __restrict__ dst /*16-byte aligned, 12 floats*/, float4 z, float o0,
                                                  float o1, float o2, float d0, float d1, float d2) {
  float4* d4 = reinterpret_cast<float4*>(dst);
  d4[0] = make_float4(__fadd_rn(o0, __fmul_rn(d0, z.x)), __fadd_rn(o1, __fmul_rn(d1, z.x)), __fadd_rn(o2, __fmul_rn(d2, z.x)),
                      __fadd_rn(o0, __fmul_rn(d0, z.y)));
  d4[1] = make_float4(__fadd_rn(o1, __fmul_rn(d1, z.y)), __fadd_rn(o2, __fmul_rn(d2, z.y)), __fadd_rn(o0, __fmul_rn(d0, z.z)),
                      __fadd_rn(o1, __fmul_rn(d1, z.z)));
  d4[2] = make_float4(__fadd_rn(o2, __fmul_rn(d2, z.z)), __fadd_rn(o0, __fmul_rn(d0, z.w)), __fadd_rn(o1, __fmul_rn(d1, z.w)),
                      __fadd_rn(o2, __fmul_rn(d2, z.w)));
}

// Fast path of sample_uniform for S <= 32*NCH: R rays per warp, every load issued before the first store.
constexpr int SU_WARPS = 8;
template <int NCH, int R>
__global__ void __launch_bounds__(SU_WARPS * 32)
sample_uniform_fast_kernel(const float* __restrict__ ro, const float* __restrict__ rd, int64_t N, int S,
                           const float* __restrict__ z_table, const float* __restrict__ lower,
                           const float* __restrict__ upper, const float* __restrict__ t_rand, float* __restrict__ pts,
                           float* __restrict__ z_out) {
  __shared__ __align__(16) float smem[SU_WARPS * R * NCH * 32];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int64_t n0 = ((int64_t)blockIdx.x * SU_WARPS + wid) * R;
  if (n0 >= N) return;
  const bool grouped = S == NCH * 32 && (reinterpret_cast<uintptr_t>(pts) & 15) == 0;   // one group of 4 samples per lane
  float t[R][NCH], o[R][3] = {}, d[R][3] = {}, lo[NCH], up[NCH];
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const int64_t n = min(n0 + r, N - 1);
#pragma unroll
    for (int c = 0; c < NCH; ++c) {
      const int s = c * 32 + lane;
      t[r][c] = (t_rand && s < S) ? __ldg(t_rand + n * S + s) : 0.f;
    }
    if (!grouped) {
#pragma unroll
      for (int k = 0; k < 3; ++k) { o[r][k] = __ldg(ro + n * 3 + k); d[r][k] = __ldg(rd + n * 3 + k); }
    }
  }
#pragma unroll
  for (int c = 0; c < NCH; ++c) {
    const int s = c * 32 + lane;
    if (t_rand) { lo[c] = s < S ? __ldg(lower + s) : 0.f; up[c] = s < S ? __ldg(upper + s) : 0.f; }
    else { lo[c] = s < S ? __ldg(z_table + s) : 0.f; up[c] = lo[c]; }
  }
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const int64_t n = n0 + r;
    if (n >= N) break;
    float* zsm = smem + (size_t)(wid * R + r) * NCH * 32;
#pragma unroll
    for (int c = 0; c < NCH; ++c) {
      const int s = c * 32 + lane;
      const float z = t_rand ? __fadd_rn(lo[c], __fmul_rn(__fsub_rn(up[c], lo[c]), t[r][c])) : lo[c];   // :65 / :67
      zsm[s] = z;
      if (s < S) z_out[n * S + s] = z;
    }
  }
  __syncwarp();
  if (grouped) {
    // R rays x NCH*8 groups of four samples = exactly 32 groups: one group per lane, ray and origin re-read per lane
    constexpr int GPR = NCH * 8;
    const int r = lane / GPR, g = lane % GPR;
    const int64_t n = n0 + r;
    if (n < N) {
      const float4 zq = *reinterpret_cast<const float4*>(smem + (size_t)(wid * R + r) * NCH * 32 + 4 * g);
      write_point_group(pts + (n * (int64_t)S + 4 * g) * 3, zq, __ldg(ro + n * 3), __ldg(ro + n * 3 + 1), __ldg(ro + n * 3 + 2),
                        __ldg(rd + n * 3), __ldg(rd + n * 3 + 1), __ldg(rd + n * 3 + 2));
    }
    return;
  }
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const int64_t n = n0 + r;
    if (n >= N) break;
    write_points(pts + n * (int64_t)S * 3, smem + (size_t)(wid * R + r) * NCH * 32, S, o[r][0], o[r][1], o[r][2], d[r][0],
                 d[r][1], d[r][2], lane);
  }
}

__global__ void __launch_bounds__(WARPS_PER_BLOCK * 32)
sample_uniform_kernel(const float* __restrict__ ro, const float* __restrict__ rd, int64_t N, int S,
                      const float* __restrict__ z_table, const float* __restrict__ lower,
                      const float* __restrict__ upper, const float* __restrict__ t_rand, float* __restrict__ pts,
                      float* __restrict__ z_out) {
  extern __shared__ float smem[];
  int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  float* zsm = smem + (size_t)wid * S;
  int64_t n = (int64_t)blockIdx.x * WARPS_PER_BLOCK + wid;
  if (n >= N) return;
  for (int s = lane; s < S; s += 32) {
    float z;
    if (t_rand) {
      float lo = lower[s], up = upper[s];
      z = __fadd_rn(lo, __fmul_rn(__fsub_rn(up, lo), t_rand[n * S + s]));   // point_sampler.py:65
    } else {
      z = z_table[s];                                                        // point_sampler.py:67
    }
    zsm[s] = z;
    z_out[n * S + s] = z;
  }
  __syncwarp();
  write_points(pts + n * (int64_t)S * 3, zsm, S, ro[n * 3], ro[n * 3 + 1], ro[n * 3 + 2], rd[n * 3], rd[n * 3 + 1],
               rd[n * 3 + 2], lane);
}

// d ro = sum_s g_pts ; d rd = sum_s g_pts * z.  One warp per ray.
__global__ void __launch_bounds__(WARPS_PER_BLOCK * 32)
points_bwd_kernel(const float* __restrict__ g_pts, const float* __restrict__ z, int64_t N, int S,
                  float* __restrict__ g_ro, float* __restrict__ g_rd) {
  int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  int64_t n = (int64_t)blockIdx.x * WARPS_PER_BLOCK + wid;
  if (n >= N) return;
  float a[3] = {0.f, 0.f, 0.f}, b[3] = {0.f, 0.f, 0.f};
  const float* g = g_pts + n * (int64_t)S * 3;
  for (int s = lane; s < S; s += 32) {
    float zz = z[n * S + s];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      float v = g[s * 3 + c];
      a[c] += v;
      b[c] += v * zz;
    }
  }
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    a[c] = warp_sum(a[c]);
    b[c] = warp_sum(b[c]);
  }
  if (lane == 0) {
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      g_ro[n * 3 + c] = a[c];
      g_rd[n * 3 + c] = b[c];
    }
  }
}

// searchsorted(cdf, u, right=True): number of entries <= u in a non-decreasing row (point_sampler.py:101).
// Branch-free descent with a warp-uniform trip count (the row length), so the lanes of a warp never diverge.
__device__ __forceinline__ int upper_bound_smem(const float* cdf, int C, float u) {
  int pos = 0;
  for (int step = 1 << (31 - __clz(C)); step > 0; step >>= 1) {
    const int np = pos + step;
    if (np <= C && cdf[np - 1] <= u) pos = np;
  }
  return pos;
}

// inverse-CDF sample for one u given the row's cdf and bins in shared memory (point_sampler.py:101-113)
__device__ __forceinline__ float invert_one(const float* cdf, const float* bins, int C, float u, int* ind_out) {
  int ind = upper_bound_smem(cdf, C, u);
  int below = max(0, ind - 1), above = min(C - 1, ind);
  float cb = cdf[below], ca = cdf[above], bb = bins[below], ba = bins[above];
  float denom = __fsub_rn(ca, cb);
  denom = denom < 1e-5f ? 1.0f : denom;
  float t = __fdiv_rn(__fsub_rn(u, cb), denom);
  *ind_out = ind;
  return __fadd_rn(bb, __fmul_rn(t, __fsub_rn(ba, bb)));
}

// in-warp bitonic sort of P (power of two) floats in shared memory, ascending
__device__ __forceinline__ void bitonic_sort_smem(float* a, int P, int lane) {
  for (int k = 2; k <= P; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int i = lane; i < P; i += 32) {
        int l = i ^ j;
        if (l > i) {
          float x = a[i], y = a[l];
          bool up = (i & k) == 0;
          if ((x > y) == up) { a[i] = y; a[l] = x; }
        }
      }
      __syncwarp();
    }
  }
}

__global__ void __launch_bounds__(WARPS_PER_BLOCK * 32)
sample_pdf_kernel(const float* __restrict__ ro, const float* __restrict__ rd, const float* __restrict__ weights,
                  int64_t w_stride, const float* __restrict__ z_coarse, int64_t N, int Sc, int Nf, int P,
                  const float* __restrict__ u_in, float* __restrict__ pts, float* __restrict__ z_out,
                  float* __restrict__ cdf_out, int64_t* __restrict__ inds_out) {
  extern __shared__ float smem[];
  const int C = Sc - 1;                                     // entries in cdf and bins
  int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  float* cdf = smem + (size_t)wid * (2 * C + P);
  float* bins = cdf + C;
  float* zs = bins + C;                                     // P floats: [z_coarse | samples | +inf pad]
  int64_t n = (int64_t)blockIdx.x * WARPS_PER_BLOCK + wid;
  if (n >= N) return;
  const float* w = weights + n * w_stride;
  const float* zc = z_coarse + n * (int64_t)Sc;
  const int S = Sc + Nf;

  // pdf normaliser: sum(w + 1e-5)   (point_sampler.py:86-87)
  float part = 0.f;
  for (int j = lane; j < C - 1; j += 32) part += __fadd_rn(w[j], 1e-5f);
  float total = warp_sum(part);
  // cdf = [0, cumsum(pdf)]  (point_sampler.py:88-89): chunked warp scan with carry
  if (lane == 0) cdf[0] = 0.f;
  float carry = 0.f;
  for (int base = 0; base < C - 1; base += 32) {
    int j = base + lane;
    float p = j < C - 1 ? __fdiv_rn(__fadd_rn(w[j], 1e-5f), total) : 0.f;
    float inc = warp_scan_inclusive(p, lane) + carry;
    if (j < C - 1) cdf[j + 1] = inc;
    carry = __shfl_sync(FULL_MASK, inc, 31);
  }
  for (int j = lane; j < C; j += 32) bins[j] = __fmul_rn(0.5f, __fadd_rn(zc[j + 1], zc[j]));  // :85
  for (int j = lane; j < Sc; j += 32) zs[j] = zc[j];
  for (int j = S + lane; j < P; j += 32) zs[j] = CUDART_INF_F;
  __syncwarp();
  if (cdf_out) for (int j = lane; j < C; j += 32) cdf_out[n * C + j] = cdf[j];

  for (int k = lane; k < Nf; k += 32) {
    // u: the caller's rand draw (:93) or linspace(0,1,Nf) (:95): torch's linspace is start + step*k below the midpoint
    // and end - step*(Nf-1-k) above it.
    float u;
    if (u_in) {
      u = u_in[n * (int64_t)Nf + k];
    } else {
      float step = Nf > 1 ? __fdiv_rn(1.0f, (float)(Nf - 1)) : 0.f;
      u = (k < Nf / 2) ? __fmul_rn(step, (float)k) : __fsub_rn(1.0f, __fmul_rn(step, (float)(Nf - 1 - k)));
    }
    int ind;
    zs[Sc + k] = invert_one(cdf, bins, C, u, &ind);
    if (inds_out) inds_out[n * (int64_t)Nf + k] = ind;
  }
  __syncwarp();
  bitonic_sort_smem(zs, P, lane);                           // sort(cat(z_vals, z_samples))  (:116)
  for (int j = lane; j < S; j += 32) z_out[n * (int64_t)S + j] = zs[j];
  write_points(pts + n * (int64_t)S * 3, zs, S, ro[n * 3], ro[n * 3 + 1], ro[n * 3 + 2], rd[n * 3], rd[n * 3 + 1],
               rd[n * 3 + 2], lane);
}

__global__ void __launch_bounds__(WARPS_PER_BLOCK * 32)
invert_cdf_kernel(const float* __restrict__ cdf_g, const float* __restrict__ bins_g, const float* __restrict__ u,
                  int64_t N, int C, int K, float* __restrict__ samples, int64_t* __restrict__ inds) {
  extern __shared__ float smem[];
  int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  float* cdf = smem + (size_t)wid * 2 * C;
  float* bins = cdf + C;
  int64_t n = (int64_t)blockIdx.x * WARPS_PER_BLOCK + wid;
  if (n >= N) return;
  for (int j = lane; j < C; j += 32) { cdf[j] = cdf_g[n * C + j]; bins[j] = bins_g[n * C + j]; }
  __syncwarp();
  for (int k = lane; k < K; k += 32) {
    int ind;
    float s = invert_one(cdf, bins, C, u[n * (int64_t)K + k], &ind);
    samples[n * (int64_t)K + k] = s;
    inds[n * (int64_t)K + k] = ind;
  }
}

// ---- fast hierarchical sampling (Nf <= 256): no full-width sort ---------------------------------------------------------
// The first version bitonic-sorted all next_pow2(Sc+Nf) depths in shared memory: 36 stages x 8 elements per lane of
// LDS/STS + __syncwarp = ~4,000 instructions per ray, 0.08 of the HBM peak.  z_coarse is already ascending
// (stratified bins), so only the Nf new samples are sorted -- EPL per lane in REGISTERS (element i = lane*EPL + e:
// exchange distances below EPL stay inside the lane, the others are one shfl_xor) -- and the two ascending lists are
// merged by rank: output slot of sample p = p + #{coarse <= v}, of coarse j = j + #{samples < c} (binary searches in
// shared memory).  Same arithmetic per sample as above, so given an identical CDF the depths are bit-identical.
constexpr int SP_WARPS = 8;

template <int EPL>
__device__ __forceinline__ void bitonic_sort_regs(float (&a)[EPL], int lane) {
  constexpr int P = EPL * 32;
#pragma unroll
  for (int k = 2; k <= P; k <<= 1) {
#pragma unroll
    for (int j = k >> 1; j > 0; j >>= 1) {
      if (j >= EPL) {
        const int lj = j / EPL;
        const bool lower = (lane & lj) == 0;
#pragma unroll
        for (int e = 0; e < EPL; ++e) {
          const float y = __shfl_xor_sync(FULL_MASK, a[e], lj);
          const bool up = (((lane * EPL + e) & k) == 0);
          a[e] = (lower == up) ? fminf(a[e], y) : fmaxf(a[e], y);
        }
      } else {
#pragma unroll
        for (int e = 0; e < EPL; ++e) {
          const int l = e ^ j;
          if (l > e) {
            const bool up = (((lane * EPL + e) & k) == 0);
            const float x = a[e], y = a[l];
            const float mn = fminf(x, y), mx = fmaxf(x, y);
            a[e] = up ? mn : mx;
            a[l] = up ? mx : mn;
          }
        }
      }
    }
  }
}

// number of entries < v (strict) / <= v (!strict) in an ascending shared-memory row of n floats
template <bool STRICT>
__device__ __forceinline__ int count_below(const float* a, int n, float v) {
  int pos = 0;
  for (int step = 1 << (31 - __clz(n)); step > 0; step >>= 1) {
    const int np = pos + step;
    if (np <= n && (STRICT ? (a[np - 1] < v) : (a[np - 1] <= v))) pos = np;
  }
  return pos;
}

template <int EPL>
__global__ void __launch_bounds__(SP_WARPS * 32)
sample_pdf_fast_kernel(const float* __restrict__ ro, const float* __restrict__ rd, const float* __restrict__ weights,
                       int64_t w_stride, const float* __restrict__ z_coarse, int64_t N, int Sc, int Nf,
                       const float* __restrict__ u_in, float* __restrict__ pts, float* __restrict__ z_out,
                       float* __restrict__ cdf_out, int64_t* __restrict__ inds_out) {
  extern __shared__ float smem[];
  constexpr int P = EPL * 32;
  const int C = Sc - 1, S = Sc + Nf;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  float* cdf = smem + (size_t)wid * (2 * C + Sc + P + S + 1);
  float* bins = cdf + C;
  float* zcs = bins + C;        // Sc coarse depths
  float* ss = zcs + Sc;         // P sorted new samples (+inf padded)
  float* zs = ss + P;           // S merged depths
  const int64_t n = (int64_t)blockIdx.x * SP_WARPS + wid;
  if (n >= N) return;
  const float* w = weights + n * w_stride;
  const float* zc = z_coarse + n * (int64_t)Sc;

  // every global load of the ray first: the uniform draws, the coarse depths, the weights, the ray
  float u[EPL];
  const bool vec4 = EPL == 4 && u_in && Nf == P && ((reinterpret_cast<uintptr_t>(u_in) & 15) == 0);
  if (vec4) {
    const float4 t = __ldg(reinterpret_cast<const float4*>(u_in + n * (int64_t)Nf) + lane);
    u[0] = t.x; u[1 % EPL] = t.y; u[2 % EPL] = t.z; u[3 % EPL] = t.w;
  } else {
#pragma unroll
    for (int e = 0; e < EPL; ++e) {
      const int k = lane * EPL + e;
      if (u_in) {
        u[e] = k < Nf ? __ldg(u_in + n * (int64_t)Nf + k) : 0.f;
      } else {   // linspace(0,1,Nf) (:95): torch computes start + step*k below the midpoint, end - step*(Nf-1-k) above it
        const float step = Nf > 1 ? __fdiv_rn(1.0f, (float)(Nf - 1)) : 0.f;
        u[e] = (k < Nf / 2) ? __fmul_rn(step, (float)k) : __fsub_rn(1.0f, __fmul_rn(step, (float)(Nf - 1 - k)));
      }
    }
  }
  const float o0 = __ldg(ro + n * 3), o1 = __ldg(ro + n * 3 + 1), o2 = __ldg(ro + n * 3 + 2);
  const float d0 = __ldg(rd + n * 3), d1 = __ldg(rd + n * 3 + 1), d2 = __ldg(rd + n * 3 + 2);
  for (int j = lane; j < Sc; j += 32) zcs[j] = __ldg(zc + j);

  // pdf normaliser sum(w + 1e-5) (:86-87), cdf = [0, cumsum(pdf)] (:88-89): chunked warp scan with carry
  float part = 0.f;
  for (int j = lane; j < C - 1; j += 32) part += __fadd_rn(__ldg(w + j), 1e-5f);
  const float total = warp_sum(part);
  if (lane == 0) cdf[0] = 0.f;
  float carry = 0.f;
  for (int base = 0; base < C - 1; base += 32) {
    const int j = base + lane;
    const float p = j < C - 1 ? __fdiv_rn(__fadd_rn(__ldg(w + j), 1e-5f), total) : 0.f;
    const float inc = warp_scan_inclusive(p, lane) + carry;
    if (j < C - 1) cdf[j + 1] = inc;
    carry = __shfl_sync(FULL_MASK, inc, 31);
  }
  __syncwarp();
  for (int j = lane; j < C; j += 32) bins[j] = __fmul_rn(0.5f, __fadd_rn(zcs[j + 1], zcs[j]));  // :85
  __syncwarp();
  if (cdf_out) for (int j = lane; j < C; j += 32) cdf_out[n * C + j] = cdf[j];

  float a[EPL];
#pragma unroll
  for (int e = 0; e < EPL; ++e) {
    const int k = lane * EPL + e;
    int ind = 0;
    const float v = invert_one(cdf, bins, C, u[e], &ind);
    a[e] = k < Nf ? v : CUDART_INF_F;
    if (inds_out && k < Nf) inds_out[n * (int64_t)Nf + k] = ind;
  }
  bitonic_sort_regs<EPL>(a, lane);
#pragma unroll
  for (int e = 0; e < EPL; ++e) ss[lane * EPL + e] = a[e];
  __syncwarp();
  // merge by rank (ties: coarse first -- only the values are observable, :116 keeps the sorted values alone)
#pragma unroll
  for (int e = 0; e < EPL; ++e) {
    const int k = lane * EPL + e;
    if (k < Nf) zs[k + count_below<false>(zcs, Sc, a[e])] = a[e];
  }
  for (int j = lane; j < Sc; j += 32) {
    const float c = zcs[j];
    zs[j + count_below<true>(ss, Nf, c)] = c;
  }
  __syncwarp();
  for (int j = lane; j < S; j += 32) z_out[n * (int64_t)S + j] = zs[j];
  write_points(pts + n * (int64_t)S * 3, zs, S, o0, o1, o2, d0, d1, d2, lane);
}

// ---- compile-time shapes (the live configs: 32+128, 64+64, 64+128): every loop unrolled, searches padded ------------
// ncu on the runtime-shape kernel above: 2,070 executed instructions per ray at IPC 3.4 -- issue-bound, a quarter of the
// HBM peak; the searches (runtime trip counts, bounds checks), the generic loop prologues and the select-heavy
// compare-exchange were most of it.  Here the cdf / coarse-depth rows are padded to a power of two with +inf, so a
// search is log2 steps of (LDS, FSETP, predicated add) with immediate offsets.
template <int SC, int NF>
struct __align__(16) SpdfSmem {
  float zs[SC + NF];    // merged depths (first: 16-byte aligned for LDS.128)
  float cdf[SC];        // C = SC-1 entries, then +inf
  float bins[SC];
  float ss[NF];         // sorted new samples
  float zc[SC];
  float4 tab[SC];       // per searchsorted index: (cdf[below], denominator, bins[below], bins[above] - bins[below])
  int last[NF];         // last[c] = 1 + index of the last coarse depth that has exactly c new samples below it (else 0)
};

// Direction-free bitonic network: every merge of width k opens with the MIRROR exchange i <-> i ^ (k-1) and continues with
// half-cleaners i <-> i ^ j (j = k/4 .. 1); every comparator is ascending, so the only predicates are "am I the lower
// lane for lane-distance 1/2/4/8/16" (five, computed once) -- the direction-per-element form spent as many instructions
// on predicates as on FMNMX + SHFL.
template <int EPL>
__device__ __forceinline__ void bitonic_sort_regs_v2(float (&a)[EPL], int lane) {
  constexpr int P = EPL * 32;
  bool low[5];
#pragma unroll
  for (int b = 0; b < 5; ++b) low[b] = (lane & (1 << b)) == 0;
#pragma unroll
  for (int k = 2; k <= P; k <<= 1) {
    // ---- mirror step
    if (k <= EPL) {
#pragma unroll
      for (int e = 0; e < EPL; ++e) {
        const int l = e ^ (k - 1);
        if (l > e) { const float x = a[e], y = a[l]; a[e] = fminf(x, y); a[l] = fmaxf(x, y); }
      }
    } else {
      const int lm = k / EPL - 1;                  // lane xor mask; registers mirror e -> EPL-1-e
      int hb = 0;                                   // log2 of the top flipped lane bit: lower side <=> that bit is 0
#pragma unroll
      for (int t = 1; t < 5; ++t) if ((k / EPL / 2) >> t) hb = t;
      float y[EPL];
#pragma unroll
      for (int e = 0; e < EPL; ++e) y[e] = __shfl_xor_sync(FULL_MASK, a[EPL - 1 - e], lm);
#pragma unroll
      for (int e = 0; e < EPL; ++e) a[e] = low[hb] ? fminf(a[e], y[e]) : fmaxf(a[e], y[e]);
    }
    // ---- half-cleaners
#pragma unroll
    for (int j = k >> 2; j > 0; j >>= 1) {
      if (j >= EPL) {
        int hb = 0;
#pragma unroll
        for (int t = 1; t < 5; ++t) if ((j / EPL) >> t) hb = t;
#pragma unroll
        for (int e = 0; e < EPL; ++e) {
          const float y = __shfl_xor_sync(FULL_MASK, a[e], j / EPL);
          a[e] = low[hb] ? fminf(a[e], y) : fmaxf(a[e], y);
        }
      } else {
#pragma unroll
        for (int e = 0; e < EPL; ++e) {
          const int l = e ^ j;
          if (l > e) { const float x = a[e], y = a[l]; a[e] = fminf(x, y); a[l] = fmaxf(x, y); }
        }
      }
    }
  }
}

// Shared-memory loads by 32-bit address with an immediate offset: a search step is LDS + FSETP + predicated add.
template <int IMM>
__device__ __forceinline__ float lds_imm(uint32_t addr) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1+%2];" : "=f"(v) : "r"(addr), "n"(IMM) : "memory");
  return v;
}
template <int STEP, bool STRICT>
__device__ __forceinline__ void search_steps(uint32_t& addr, float v) {
  if constexpr (STEP > 0) {
    const float t = lds_imm<(STEP - 1) * 4>(addr);
    if (STRICT ? (t < v) : (t <= v)) addr += STEP * 4;
    search_steps<STEP / 2, STRICT>(addr, v);
  }
}
// byte offset 4 * #entries <= v (or < v) of an ascending LEN-entry row (LEN a power of two); FINAL adds the last probe
// that lets the count reach LEN itself (rows without a +inf pad)
template <int LEN, bool STRICT, bool FINAL>
__device__ __forceinline__ uint32_t search_bytes(uint32_t base, float v) {
  uint32_t addr = base;
  search_steps<LEN / 2, STRICT>(addr, v);
  if (FINAL) {
    const float t = lds_imm<0>(addr);
    if (STRICT ? (t < v) : (t <= v)) addr += 4;
  }
  return addr - base;
}

// #entries <= v (or < v) among the first LEN (power of two) floats of an ascending row; rows are +inf padded
template <int LEN, bool STRICT>
__device__ __forceinline__ int count_pow2(const float* a, float v) {
  int pos = 0;
#pragma unroll
  for (int step = LEN / 2; step > 0; step >>= 1) {
    const float t = a[pos + step - 1];
    if (STRICT ? (t < v) : (t <= v)) pos += step;
  }
  const float t = a[pos];
  pos += (STRICT ? (t < v) : (t <= v)) ? 1 : 0;
  return pos;
}

template <int SC, int NF, int VAR>
__global__ void __launch_bounds__(SP_WARPS * 32)
sample_pdf_static_kernel(const float* __restrict__ ro, const float* __restrict__ rd, const float* __restrict__ weights,
                         int64_t w_stride, const float* __restrict__ z_coarse, int64_t N, const float* __restrict__ u_in,
                         float* __restrict__ pts, float* __restrict__ z_out, float* __restrict__ cdf_out,
                         int64_t* __restrict__ inds_out) {
  static_assert(SC % 32 == 0 && NF % 32 == 0 && (NF & (NF - 1)) == 0 && (SC & (SC - 1)) == 0, "shape");
  constexpr int C = SC - 1, S = SC + NF, EPL = NF / 32, NCC = SC / 32;
  __shared__ SpdfSmem<SC, NF> sm_all[SP_WARPS];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  SpdfSmem<SC, NF>& sm = sm_all[wid];
  const int64_t n = (int64_t)blockIdx.x * SP_WARPS + wid;
  if (n >= N) return;
  const float* w = weights + n * w_stride;
  const float* zc = z_coarse + n * (int64_t)SC;

  // ---- every global load of the ray first
  float u[EPL];
  if (u_in) {
    const float* ur = u_in + n * (int64_t)NF;
    if (EPL % 4 == 0 && (reinterpret_cast<uintptr_t>(u_in) & 15) == 0) {
#pragma unroll
      for (int e4 = 0; e4 < EPL / 4; ++e4) {
        const float4 t = __ldg(reinterpret_cast<const float4*>(ur) + lane * (EPL / 4) + e4);
        u[e4 * 4] = t.x; u[(e4 * 4 + 1) % EPL] = t.y; u[(e4 * 4 + 2) % EPL] = t.z; u[(e4 * 4 + 3) % EPL] = t.w;
      }
    } else {
#pragma unroll
      for (int e = 0; e < EPL; ++e) u[e] = __ldg(ur + lane * EPL + e);
    }
  } else {   // linspace(0,1,NF) (:95): torch computes start + step*k below the midpoint, end - step*(NF-1-k) above it
    const float step = __fdiv_rn(1.0f, (float)(NF - 1));
#pragma unroll
    for (int e = 0; e < EPL; ++e) {
      const int k = lane * EPL + e;
      u[e] = (k < NF / 2) ? __fmul_rn(step, (float)k) : __fsub_rn(1.0f, __fmul_rn(step, (float)(NF - 1 - k)));
    }
  }
  float zcr[NCC], wr[NCC];
#pragma unroll
  for (int c = 0; c < NCC; ++c) {
    const int j = c * 32 + lane;
    zcr[c] = __ldg(zc + j);
    wr[c] = j < C - 1 ? __ldg(w + j) : 0.f;
  }
  const float o0 = __ldg(ro + n * 3), o1 = __ldg(ro + n * 3 + 1), o2 = __ldg(ro + n * 3 + 2);
  const float d0 = __ldg(rd + n * 3), d1 = __ldg(rd + n * 3 + 1), d2 = __ldg(rd + n * 3 + 2);

  // ---- pdf normaliser sum(w + 1e-5) (:86-87); cdf = [0, cumsum(pdf)] (:88-89); bins = mids of z_coarse (:85)
  float part = 0.f;
#pragma unroll
  for (int c = 0; c < NCC; ++c) if (c * 32 + lane < C - 1) part += __fadd_rn(wr[c], 1e-5f);
  const float total = warp_sum(part);
  if (lane == 0) sm.cdf[0] = 0.f;
  float carry = 0.f;
#pragma unroll
  for (int c = 0; c < NCC; ++c) {
    const int j = c * 32 + lane;
    const float p = j < C - 1 ? __fdiv_rn(__fadd_rn(wr[c], 1e-5f), total) : 0.f;
    const float inc = warp_scan_inclusive(p, lane) + carry;
    if (j < C - 1) sm.cdf[j + 1] = inc;
    carry = __shfl_sync(FULL_MASK, inc, 31);
    if constexpr ((VAR & 2) == 0) sm.zc[j] = zcr[c];
    float zn = __shfl_down_sync(FULL_MASK, zcr[c], 1);
    if (c + 1 < NCC) { const float first = __shfl_sync(FULL_MASK, zcr[c + 1 < NCC ? c + 1 : c], 0); zn = lane == 31 ? first : zn; }
    if (j < C) sm.bins[j] = __fmul_rn(0.5f, __fadd_rn(zn, zcr[c]));
  }
  if (lane == 0) { sm.cdf[C] = CUDART_INF_F; sm.bins[C] = 0.f; }
  __syncwarp();
  if (cdf_out) {
#pragma unroll
    for (int c = 0; c < NCC; ++c) if (c * 32 + lane < C) cdf_out[n * C + c * 32 + lane] = sm.cdf[c * 32 + lane];
  }

  constexpr bool TAB = (VAR & 1) != 0, PMAX = (VAR & 2) != 0, ZFOLD = (VAR & 4) != 0;
  // ---- per-index lerp table: what :104-111 gather and subtract for searchsorted index i is the same for every u in it
  if constexpr (TAB) {
#pragma unroll
    for (int c = 0; c < NCC; ++c) {
      const int i = c * 32 + lane;
      const int below = max(0, i - 1), above = min(C - 1, i);
      const float cb = sm.cdf[below], ca = sm.cdf[above], bb = sm.bins[below], ba = sm.bins[above];
      float denom = __fsub_rn(ca, cb);
      denom = denom < 1e-5f ? 1.0f : denom;
      sm.tab[i] = make_float4(cb, denom, bb, __fsub_rn(ba, bb));
    }
  }
  if constexpr (PMAX) {
#pragma unroll
    for (int e = 0; e < EPL; ++e) sm.last[lane * EPL + e] = 0;
  }
  if constexpr (TAB) __syncwarp();
  // ---- inverse CDF (:101-113)
  const uint32_t cdf_s = (uint32_t)__cvta_generic_to_shared(sm.cdf), tab_s = (uint32_t)__cvta_generic_to_shared(sm.tab);
  const uint32_t ss_s = (uint32_t)__cvta_generic_to_shared(sm.ss), zs_s = (uint32_t)__cvta_generic_to_shared(sm.zs);
  const uint32_t zc_s = (uint32_t)__cvta_generic_to_shared(sm.zc);
  float a[EPL];
#pragma unroll
  for (int e = 0; e < EPL; ++e) {
    // searchsorted(cdf, u, right=True): entries <= u; the +inf pad keeps the descent inside the C real entries
    const uint32_t ib = search_bytes<SC, false, false>(cdf_s, u[e]);
    if constexpr (TAB) {
      float4 q;
      asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(q.x), "=f"(q.y), "=f"(q.z), "=f"(q.w) : "r"(tab_s + ib * 4) : "memory");
      const float t = __fdiv_rn(__fsub_rn(u[e], q.x), q.y);
      a[e] = __fadd_rn(q.z, __fmul_rn(t, q.w));
    } else {
      const uint32_t below = cdf_s + (ib >= 4 ? ib - 4 : 0), above = cdf_s + min(ib, (uint32_t)(C - 1) * 4);
      const float cb = lds_imm<0>(below), ca = lds_imm<0>(above);
      const float bb = lds_imm<SC * 4>(below), ba = lds_imm<SC * 4>(above);      // bins follow cdf in SpdfSmem
      float denom = __fsub_rn(ca, cb);
      denom = denom < 1e-5f ? 1.0f : denom;
      const float t = __fdiv_rn(__fsub_rn(u[e], cb), denom);
      a[e] = __fadd_rn(bb, __fmul_rn(t, __fsub_rn(ba, bb)));
    }
    if (inds_out) inds_out[n * (int64_t)NF + lane * EPL + e] = ib >> 2;
  }
  bitonic_sort_regs_v2<EPL>(a, lane);
#pragma unroll
  for (int e = 0; e < EPL; ++e) sm.ss[lane * EPL + e] = a[e];
  __syncwarp();
  // ---- merge by rank (ties: coarse first -- only the values are observable, :116 keeps the sorted values alone)
  // coarse depth j lands at j + cnt_j, cnt_j = #{new samples < c_j} (one search per coarse depth).  New sample p lands at
  // p + #{j : c_j <= v_p} = p + #{j : cnt_j <= p} = p + max{ j+1 : cnt_j <= p } (cnt is non-decreasing): the last lane of
  // every run of equal cnt publishes j+1 at last[cnt], and a prefix MAX over last[] replaces the per-sample searches.
  int cnt[NCC];
#pragma unroll
  for (int c = 0; c < NCC; ++c) {
    const uint32_t cb = search_bytes<NF, true, true>(ss_s, zcr[c]);
    cnt[c] = (int)(cb >> 2);
    asm volatile("st.shared.f32 [%0], %1;" ::"r"(zs_s + (c * 32 + lane) * 4 + cb), "f"(zcr[c]) : "memory");
  }
  if constexpr (PMAX) {
#pragma unroll
    for (int c = 0; c < NCC; ++c) {
      int nxt = __shfl_down_sync(FULL_MASK, cnt[c], 1);
      if (c + 1 < NCC) { const int first = __shfl_sync(FULL_MASK, cnt[c + 1 < NCC ? c + 1 : c], 0); nxt = lane == 31 ? first : nxt; }
      else if (lane == 31) nxt = -1;
      if (nxt != cnt[c] && cnt[c] < NF) sm.last[cnt[c]] = c * 32 + lane + 1;
    }
    __syncwarp();
    int pm[EPL];
#pragma unroll
    for (int e = 0; e < EPL; ++e) { pm[e] = sm.last[lane * EPL + e]; if (e > 0) pm[e] = max(pm[e], pm[e - 1]); }
    int run = pm[EPL - 1];
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(FULL_MASK, run, o); if (lane >= o) run = max(run, t); }
    int before = __shfl_up_sync(FULL_MASK, run, 1);
    if (lane == 0) before = 0;
#pragma unroll
    for (int e = 0; e < EPL; ++e)
      asm volatile("st.shared.f32 [%0], %1;" ::"r"(zs_s + (lane * EPL + e + max(before, pm[e])) * 4), "f"(a[e]) : "memory");
  } else {
#pragma unroll
    for (int e = 0; e < EPL; ++e) {
      const uint32_t dst = zs_s + (lane * EPL + e) * 4 + search_bytes<SC, false, true>(zc_s, a[e]);
      asm volatile("st.shared.f32 [%0], %1;" ::"r"(dst), "f"(a[e]) : "memory");
    }
  }
  __syncwarp();
  float* zo = z_out + n * (int64_t)S;
  if (((reinterpret_cast<uintptr_t>(pts) | reinterpret_cast<uintptr_t>(z_out)) & 15) == 0) {
    if constexpr (!ZFOLD) {
#pragma unroll
      for (int c = 0; c < S / 32; ++c) zo[c * 32 + lane] = sm.zs[c * 32 + lane];
    }
#pragma unroll
    for (int g0 = 0; g0 < S / 4; g0 += 32) {
      const int g = g0 + lane;
      if (g < S / 4) {
        const float4 zq = *reinterpret_cast<const float4*>(sm.zs + 4 * g);
        if constexpr (ZFOLD) *reinterpret_cast<float4*>(zo + 4 * g) = zq;
        write_point_group(pts + (n * (int64_t)S + 4 * g) * 3, zq, o0, o1, o2, d0, d1, d2);
      }
    }
  } else {
#pragma unroll
    for (int c = 0; c < S / 32; ++c) zo[c * 32 + lane] = sm.zs[c * 32 + lane];
    write_points(pts + n * (int64_t)S * 3, sm.zs, S, o0, o1, o2, d0, d1, d2, lane);
  }
}

template <int EPL>
static int launch_sample_pdf_fast(const float* ro, const float* rd, const float* weights, int64_t w_stride,
                                  const float* z_coarse, int64_t N, int Sc, int Nf, const float* u, float* pts, float* z,
                                  float* cdf_out, int64_t* inds_out, cudaStream_t st) {
  const size_t smem = (size_t)SP_WARPS * (2 * (Sc - 1) + Sc + EPL * 32 + Sc + Nf + 1) * sizeof(float);
  if (smem > 48 * 1024) return CNRF_E_UNSUPPORTED;
  sample_pdf_fast_kernel<EPL><<<grid_for(N, SP_WARPS), SP_WARPS * 32, smem, st>>>(ro, rd, weights, w_stride, z_coarse, N, Sc,
                                                                                  Nf, u, pts, z, cdf_out, inds_out);
  return CNRF_OK;
}

static int next_pow2(int v) { int p = 1; while (p < v) p <<= 1; return p; }

extern "C" int cnrf_sample_uniform_fwd(const float* ro, const float* rd, int64_t N, int32_t S, const float* z_table,
                                       const float* lower, const float* upper, const float* t_rand, float* pts,
                                       float* z, cnrf_stream_t stream) {
  CNRF_REQUIRE(ro && rd && pts && z, CNRF_E_NULL);
  CNRF_REQUIRE(t_rand ? (lower && upper) : (z_table != nullptr), CNRF_E_NULL);
  CNRF_REQUIRE(N > 0 && S > 0 && S <= 4096, CNRF_E_SHAPE);
  if (S <= 32) {
    sample_uniform_fast_kernel<1, 4><<<grid_for(N, SU_WARPS * 4), SU_WARPS * 32, 0, as_stream(stream)>>>(
        ro, rd, N, S, z_table, lower, upper, t_rand, pts, z);
    CNRF_LAUNCH_CHECK();
    return CNRF_OK;
  }
  if (S <= 64) {
    sample_uniform_fast_kernel<2, 2><<<grid_for(N, SU_WARPS * 2), SU_WARPS * 32, 0, as_stream(stream)>>>(
        ro, rd, N, S, z_table, lower, upper, t_rand, pts, z);
    CNRF_LAUNCH_CHECK();
    return CNRF_OK;
  }
  size_t smem = (size_t)WARPS_PER_BLOCK * S * sizeof(float);
  sample_uniform_kernel<<<grid_for(N, WARPS_PER_BLOCK), WARPS_PER_BLOCK * 32, smem, as_stream(stream)>>>(
      ro, rd, N, S, z_table, lower, upper, t_rand, pts, z);
  CNRF_LAUNCH_CHECK();
  return CNRF_OK;
}

extern "C" int cnrf_points_bwd(const float* g_pts, const float* z, int64_t N, int32_t S, float* g_ro, float* g_rd,
                               cnrf_stream_t stream) {
  CNRF_REQUIRE(g_pts && z && g_ro && g_rd, CNRF_E_NULL);
  CNRF_REQUIRE(N > 0 && S > 0, CNRF_E_SHAPE);
  points_bwd_kernel<<<grid_for(N, WARPS_PER_BLOCK), WARPS_PER_BLOCK * 32, 0, as_stream(stream)>>>(g_pts, z, N, S, g_ro,
                                                                                                 g_rd);
  CNRF_LAUNCH_CHECK();
  return CNRF_OK;
}

extern "C" int cnrf_sample_pdf_fwd(const float* ro, const float* rd, const float* weights, int64_t w_stride,
                                   const float* z_coarse, int64_t N, int32_t Sc, int32_t Nf, const float* u, float* pts,
                                   float* z, float* cdf_out, int64_t* inds_out, cnrf_stream_t stream) {
  CNRF_REQUIRE(ro && rd && weights && z_coarse && pts && z, CNRF_E_NULL);
  CNRF_REQUIRE(N > 0 && Sc >= 3 && Nf > 0 && Sc + Nf <= 2048, CNRF_E_SHAPE);
  // VAR 6 = prefix-max merge + depths stored from the point writer, no lerp table: the measured best of the eight
  // combinations on B200 at 32+128 (453 us per 524,288 rays = 0.60 of the HBM peak; the per-index lerp table costs 30 us).
#define SP_STATIC(SC_, NF_)                                                                                          \
  if (Sc == SC_ && Nf == NF_) {                                                                                      \
    sample_pdf_static_kernel<SC_, NF_, 6><<<grid_for(N, SP_WARPS), SP_WARPS * 32, 0, as_stream(stream)>>>(          \
        ro, rd, weights, w_stride, z_coarse, N, u, pts, z, cdf_out, inds_out);                                      \
    CNRF_LAUNCH_CHECK();                                                                                             \
    return CNRF_OK;                                                                                                  \
  }
  SP_STATIC(32, 128) SP_STATIC(64, 64) SP_STATIC(64, 128)
#undef SP_STATIC
  if (Nf <= 256) {
    const int epl = next_pow2((Nf + 31) / 32);
    int rc = CNRF_E_UNSUPPORTED;
#define SP_ARGS ro, rd, weights, w_stride, z_coarse, N, Sc, Nf, u, pts, z, cdf_out, inds_out, as_stream(stream)
    if (epl == 1) rc = launch_sample_pdf_fast<1>(SP_ARGS);
    else if (epl == 2) rc = launch_sample_pdf_fast<2>(SP_ARGS);
    else if (epl == 4) rc = launch_sample_pdf_fast<4>(SP_ARGS);
    else if (epl == 8) rc = launch_sample_pdf_fast<8>(SP_ARGS);
#undef SP_ARGS
    if (rc == CNRF_OK) { CNRF_LAUNCH_CHECK(); return CNRF_OK; }
  }
  int P = next_pow2(Sc + Nf);
  size_t smem = (size_t)WARPS_PER_BLOCK * (2 * (Sc - 1) + P) * sizeof(float);
  CNRF_REQUIRE(smem <= 48 * 1024, CNRF_E_SHAPE);
  sample_pdf_kernel<<<grid_for(N, WARPS_PER_BLOCK), WARPS_PER_BLOCK * 32, smem, as_stream(stream)>>>(
      ro, rd, weights, w_stride, z_coarse, N, Sc, Nf, P, u, pts, z, cdf_out, inds_out);
  CNRF_LAUNCH_CHECK();
  return CNRF_OK;
}

extern "C" int cnrf_invert_cdf(const float* cdf, const float* bins, const float* u, int64_t N, int32_t C, int32_t K,
                               float* samples, int64_t* inds, cnrf_stream_t stream) {
  CNRF_REQUIRE(cdf && bins && u && samples && inds, CNRF_E_NULL);
  CNRF_REQUIRE(N > 0 && C > 0 && K > 0 && C <= 1024, CNRF_E_SHAPE);
  size_t smem = (size_t)WARPS_PER_BLOCK * 2 * C * sizeof(float);
  invert_cdf_kernel<<<grid_for(N, WARPS_PER_BLOCK), WARPS_PER_BLOCK * 32, smem, as_stream(stream)>>>(cdf, bins, u, N, C,
                                                                                                    K, samples, inds);
  CNRF_LAUNCH_CHECK();
  return CNRF_OK;
}
