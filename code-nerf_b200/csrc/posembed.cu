// Positional encoding as a standalone operator: PositionalEmbedder.embed (view_synthesis/nerf/position_embed.py:35-53).
// The fused tensor-core path never calls this (the encoding is produced inside the first MLP layer's operand
// staging); it backs the public `embed()` method and the fp32 parity path.  Accurate sinf/cosf only: arguments reach
// 2^9 * |x|, so no fast-math intrinsics.
#include "common.cuh"

// One thread per (row m, band k, channel d): ONE sincosf serves the sin and the cos column of the band (k == 0 of an
// include_input encoding copies x instead).  Row-major thread order, so a warp writes a contiguous run of each row.
__global__ void pos_embed_fwd_kernel(const float* __restrict__ x, int64_t M, int D, const float* __restrict__ freqs,
                                     int L, int inc, int64_t ld_out, float* __restrict__ out) {
  const int per_row = D * (L + inc);
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= M * per_row) return;
  int64_t m;
  int j;
  if (M * per_row < (int64_t)1 << 32) {                    // 32-bit division when the index fits
    const uint32_t mi = (uint32_t)i / (uint32_t)per_row;
    m = mi; j = (int)((uint32_t)i - mi * (uint32_t)per_row);
  } else {
    m = i / per_row; j = (int)(i - m * per_row);
  }
  const int k = j / D, d = j - k * D;
  const float v = __ldg(x + m * D + d);
  float* o = out + m * ld_out;
  if (inc && k == 0) { o[d] = v; return; }
  const int b = k - inc;
  float sn, cs;
  sincosf(__fmul_rn(v, __ldg(freqs + b)), &sn, &cs);        // tensor * freq       position_embed.py:48
  o[(inc + 2 * b) * D + d] = sn;                            // [sin, cos] per band position_embed.py:47
  o[(inc + 2 * b + 1) * D + d] = cs;
}

__global__ void pos_embed_bwd_kernel(const float* __restrict__ x, const float* __restrict__ g_out, int64_t M, int D,
                                     const float* __restrict__ freqs, int L, int inc, int64_t ld_g,
                                     float* __restrict__ g_x) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= M * D) return;
  int64_t m = i / D;
  int d = (int)(i - m * D);
  const float* g = g_out + m * ld_g;
  float v = x[i];
  float acc = inc ? g[d] : 0.f;
  for (int k = 0; k < L; ++k) {
    float f = freqs[k];
    float s, c;
    sincosf(__fmul_rn(v, f), &s, &c);
    float gs = g[(inc + 2 * k) * D + d], gc = g[(inc + 2 * k + 1) * D + d];
    acc += f * (gs * c - gc * s);
  }
  g_x[i] = acc;
}

// internal entry points with a row stride (used by the fp32 field path to write into padded buffers)
int cnrf_pos_embed_fwd_ld(const float* x, int64_t M, int D, const float* freqs, int L, int inc, int64_t ld_out,
                          float* out, cudaStream_t st) {
  pos_embed_fwd_kernel<<<grid_for(M * D * (inc + L), 256), 256, 0, st>>>(x, M, D, freqs, L, inc, ld_out, out);
  CNRF_LAUNCH_CHECK();
  return CNRF_OK;
}
int cnrf_pos_embed_bwd_ld(const float* x, const float* g_out, int64_t M, int D, const float* freqs, int L, int inc,
                          int64_t ld_g, float* g_x, cudaStream_t st) {
  pos_embed_bwd_kernel<<<grid_for(M * D, 256), 256, 0, st>>>(x, g_out, M, D, freqs, L, inc, ld_g, g_x);
  CNRF_LAUNCH_CHECK();
  return CNRF_OK;
}

extern "C" int cnrf_pos_embed_fwd(const float* x, int64_t M, int32_t D, const float* freqs, int32_t L,
                                  int32_t include_input, float* out, cnrf_stream_t stream) {
  CNRF_REQUIRE(x && freqs && out, CNRF_E_NULL);
  CNRF_REQUIRE(M > 0 && D > 0 && L > 0, CNRF_E_SHAPE);
  int inc = include_input ? 1 : 0;
  return cnrf_pos_embed_fwd_ld(x, M, D, freqs, L, inc, (int64_t)D * (inc + 2 * L), out, as_stream(stream));
}

extern "C" int cnrf_pos_embed_bwd(const float* x, const float* g_out, int64_t M, int32_t D, const float* freqs,
                                  int32_t L, int32_t include_input, float* g_x, cnrf_stream_t stream) {
  CNRF_REQUIRE(x && g_out && freqs && g_x, CNRF_E_NULL);
  CNRF_REQUIRE(M > 0 && D > 0 && L > 0, CNRF_E_SHAPE);
  int inc = include_input ? 1 : 0;
  return cnrf_pos_embed_bwd_ld(x, g_out, M, D, freqs, L, inc, (int64_t)D * (inc + 2 * L), g_x, as_stream(stream));
}
