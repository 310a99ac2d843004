// fp32 SIMT GEMM toolkit used by the CNRF_PRECISION_FP32 field path and by the per-ray (latent-code) layers of both
// precisions.  Plain CUDA-core FMA: this is the parity mode (1e-3 gate), not the throughput mode -- the per-sample
// contraction of the throughput mode lives in mlp_tc.cu (tcgen05).  64x64x16 tiles, 256 threads, 4x4 per thread,
// arbitrary operand strides so that X*W^T, dY*W and dY^T*X all map onto the same kernel without transposed copies.
#include "common.cuh"
#include "internal.h"

constexpr int BM = 64, BN = 64, BK = 16, GT = 256;

template <bool A_KCONTIG, bool B_NCONTIG>
__global__ void __launch_bounds__(GT) sgemm_kernel(GemmArgs g) {
  __shared__ float As[BK][BM + 4];
  __shared__ float Bs[BK][BN + 4];
  const int t = threadIdx.x;
  const int64_t m0 = (int64_t)blockIdx.x * BM;
  const int n0 = blockIdx.y * BN;
  // K range of this split
  int64_t kchunk = (g.K + g.split_k - 1) / g.split_k;
  kchunk = (kchunk + BK - 1) / BK * BK;
  const int64_t kbeg = (int64_t)blockIdx.z * kchunk;
  const int64_t kend = min(g.K, kbeg + kchunk);
  const int ty = t / 16, tx = t % 16;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  for (int64_t k0 = kbeg; k0 < kend; k0 += BK) {
    // ---- stage A tile (BM x BK) and B tile (BK x BN); 4 elements per thread each
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      int mm, kk;
      if (A_KCONTIG) { int idx = t * 4 + e; mm = idx / BK; kk = idx % BK; }     // consecutive threads walk k
      else           { int idx = e * GT + t; kk = idx / BM; mm = idx % BM; }     // consecutive threads walk m
      int64_t m = m0 + mm, k = k0 + kk;
      As[kk][mm] = (m < g.M && k < kend) ? g.A[m * g.a_rs + k * g.a_cs] : 0.f;
    }
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      int kk, nn;
      if (B_NCONTIG) { int idx = e * GT + t; kk = idx / BN; nn = idx % BN; }     // consecutive threads walk n
      else           { int idx = t * 4 + e; nn = idx / BK; kk = idx % BK; }      // consecutive threads walk k
      int64_t k = k0 + kk; int n = n0 + nn;
      Bs[kk][nn] = (k < kend && n < g.N) ? g.B[k * g.b_rs + (int64_t)n * g.b_cs] : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < BK; ++kk) {
      float4 a4 = *reinterpret_cast<const float4*>(&As[kk][ty * 4]);
      float4 b4 = *reinterpret_cast<const float4*>(&Bs[kk][tx * 4]);
      float a[4] = {a4.x, a4.y, a4.z, a4.w}, b[4] = {b4.x, b4.y, b4.z, b4.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    int64_t m = m0 + ty * 4 + i;
    if (m >= g.M) continue;
    const float* brow = g.bias ? g.bias + ((m / g.bias_div) % g.bias_mod) * g.ld_bias : nullptr;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int n = n0 + tx * 4 + j;
      if (n >= g.N) continue;
      float v = acc[i][j];
      float* c = g.C + m * g.ldc + n;
      if (g.split_k > 1) { atomicAdd(c, v); continue; }
      if (g.accumulate) { *c += v; continue; }
      if (brow) v += brow[n];
      if (g.relu) v = fmaxf(v, 0.f);
      if (g.mask) v = g.mask[m * g.ld_mask + n] > 0.f ? v : 0.f;
      *c = v;
    }
  }
}

int cnrf_sgemm(const GemmArgs& a, cudaStream_t st) {
  if (!a.A || !a.B || !a.C) return CNRF_E_NULL;
  if (a.M <= 0 || a.N <= 0 || a.K <= 0 || a.split_k < 1) return CNRF_E_SHAPE;
  dim3 grid((unsigned)((a.M + BM - 1) / BM), (unsigned)((a.N + BN - 1) / BN), (unsigned)a.split_k);
  const bool ak = a.a_cs == 1, bn = a.b_cs == 1;
  if (ak && bn) sgemm_kernel<true, true><<<grid, GT, 0, st>>>(a);
  else if (ak && !bn) sgemm_kernel<true, false><<<grid, GT, 0, st>>>(a);
  else if (!ak && bn) sgemm_kernel<false, true><<<grid, GT, 0, st>>>(a);
  else sgemm_kernel<false, false><<<grid, GT, 0, st>>>(a);
  CNRF_LAUNCH_CHECK();
  return CNRF_OK;
}

__global__ void strided_rowsum_kernel(const float* __restrict__ X, int64_t ld_x, int64_t G, int64_t g_stride, int64_t Q,
                                      int64_t q_stride, int N, float* __restrict__ out, int64_t ld_out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= G * N) return;
  int64_t gi = i / N;
  int n = (int)(i - gi * N);
  const float* p = X + (gi * g_stride) * ld_x + n;
  float s = 0.f;
  for (int64_t q = 0; q < Q; ++q) s += p[q * q_stride * ld_x];
  out[gi * ld_out + n] = s;
}

int cnrf_strided_rowsum(const float* X, int64_t ld_x, int64_t G, int64_t g_stride, int64_t Q, int64_t q_stride, int N,
                        float* out, int64_t ld_out, cudaStream_t st) {
  strided_rowsum_kernel<<<grid_for(G * N, 256), 256, 0, st>>>(X, ld_x, G, g_stride, Q, q_stride, N, out, ld_out);
  CNRF_LAUNCH_CHECK();
  return CNRF_OK;
}

constexpr int CS_ROWS = 32;
__global__ void colsum_kernel(const float* __restrict__ X, int64_t ld_x, int64_t M, int N, float* __restrict__ out) {
  int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  int64_t m0 = (int64_t)blockIdx.y * CS_ROWS, m1 = min(M, m0 + CS_ROWS);
  float s = 0.f;
  for (int64_t m = m0; m < m1; ++m) s += X[m * ld_x + n];
  atomicAdd(out + n, s);
}

int cnrf_fill_zero(void* p, size_t bytes, cudaStream_t st) {
  cudaError_t e = cudaMemsetAsync(p, 0, bytes, st);
  return e == cudaSuccess ? CNRF_OK : (int)e;
}

int cnrf_colsum(const float* X, int64_t ld_x, int64_t M, int N, float* out, cudaStream_t st) {
  int rc = cnrf_fill_zero(out, (size_t)N * sizeof(float), st);
  if (rc) return rc;
  dim3 grid((N + 127) / 128, (unsigned)((M + CS_ROWS - 1) / CS_ROWS));
  colsum_kernel<<<grid, 128, 0, st>>>(X, ld_x, M, N, out);
  CNRF_LAUNCH_CHECK();
  return CNRF_OK;
}

__global__ void fill_batch_kernel(const __grid_constant__ FillBatch b) {
  const int i = blockIdx.y;
  float* d = b.dst[i];
  const float v = b.value[i];
  for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < b.n[i]; j += (int64_t)gridDim.x * blockDim.x) d[j] = v;
}
int cnrf_fill_batch(const FillBatch& b, cudaStream_t st) {
  if (b.count <= 0) return CNRF_OK;
  if (b.count > CNRF_FILL_BATCH_MAX) return CNRF_E_SHAPE;
  fill_batch_kernel<<<dim3(64, b.count), 256, 0, st>>>(b);
  CNRF_LAUNCH_CHECK();
  return CNRF_OK;
}
