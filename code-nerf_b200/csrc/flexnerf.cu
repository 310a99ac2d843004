// Vanilla NeRF field: FlexibleNeRFModel.forward with view directions (view_synthesis/models/model.py:51-74), forward and
// backward, behind cnrf_flexnerf_* (SURVEY section 8 row a15 / "next" row f1; config/lego.yml = 8 x 256, skip [4]).
//
// Layer by layer through the GEMM toolkit (sgemm.cu fp32 / gemm_tc.cu split-bf16 tcgen05): each layer is one launch with
// bias + ReLU (forward) or the ReLU mask (backward) in the epilogue.  Two things are folded the way the CodeNeRF path
// folds its codes: the view-direction half of layers_dir[0] depends on the RAY only, so it is one [N, H/2] GEMM whose
// rows enter the per-sample layer as bias rows (pairing as nerf/__init__.py:127); and the activation feeding a
// skip layer is written straight into the left part of a [M, H + dim_xyz] buffer whose right part holds gamma(x), so
// cat(h, xyz) (models/model.py:63) never runs as a copy of h.
#include <algorithm>
#include <stdlib.h>
#include "common.cuh"
#include "internal.h"

namespace {

struct Arena {
  char* base; size_t cap; size_t off = 0; bool ok = true;
  Arena(void* p, size_t c) : base(static_cast<char*>(p)), cap(c) {}
  float* take(size_t nfloat) {
    size_t bytes = (nfloat * sizeof(float) + 255) & ~size_t(255);
    if (off + bytes > cap) { ok = false; return nullptr; }
    float* r = reinterpret_cast<float*>(base + off);
    off += bytes;
    return r;
  }
};
inline size_t al(size_t nfloat) { return (nfloat * sizeof(float) + 255) & ~size_t(255); }

#define RUN(x) do { int rc__ = (x); if (rc__) return rc__; } while (0)

struct FDims {
  int L, H, dx, dd, S, pairing;   // L = number of layers_xyz (num_layers - 1)
  uint32_t skip;
  int64_t N, M;
  bool tc;
  bool viewdirs() const { return dd > 0; }      // dim_dir == 0: the use_viewdirs=False network (fc_out head, model.py:47,76)
  bool is_skip(int i) const { return (skip >> i) & 1u; }
  // activation a (0 = output of layer1, i+1 = output of layers_xyz[i]) feeds layers_xyz[a] (or fc_feat when a == L)
  int64_t ld_act(int a) const { return (a < L && is_skip(a)) ? H + dx : H; }
};

inline int dims_of(const cnrf_flex_shape* s, FDims& d) {
  if (!s) return CNRF_E_NULL;
  if (s->num_layers < 1 || s->num_layers - 1 > CNRF_FLEX_MAX_XYZ_LAYERS || s->hidden < 2 || (s->hidden & 1) ||
      s->dim_xyz <= 0 || s->dim_dir < 0 || s->num_rays <= 0 || s->num_samples <= 0)
    return CNRF_E_SHAPE;
  if (s->pairing != CNRF_PAIRING_REFERENCE && s->pairing != CNRF_PAIRING_CORRECT) return CNRF_E_SHAPE;
  if (s->precision != CNRF_PRECISION_FP32 && s->precision != CNRF_PRECISION_BF16 && s->precision != CNRF_PRECISION_SPLIT) return CNRF_E_UNSUPPORTED;
  d.L = s->num_layers - 1; d.H = s->hidden; d.dx = s->dim_xyz; d.dd = s->dim_dir; d.S = s->num_samples;
  d.pairing = s->pairing; d.skip = s->skip_mask; d.N = s->num_rays; d.M = s->num_rays * (int64_t)s->num_samples;
  d.tc = s->precision != CNRF_PRECISION_FP32;      // BF16 and SPLIT: per-layer split-bf16 tensor-core GEMMs
  return CNRF_OK;
}

int run_gemm(bool tc, const GemmArgs& g, cudaStream_t st) {
  if (tc) {
    int rc = cnrf_gemm_tc(g, st);
    if (rc != CNRF_E_UNSUPPORTED) return rc;
  }
  return cnrf_sgemm(g, st);
}
// Y[rows,Nout] = act(X[rows,K] W[Nout,K]^T + bias[(m/div)%mod])
int lin(bool tc, const float* X, int64_t ldx, int64_t rows, int K, const float* W, int64_t ldw, int Nout, float* Y, int64_t ldy,
        const float* bias, int64_t ld_bias, int64_t div, int64_t mod, int relu, cudaStream_t st) {
  GemmArgs g;
  g.A = X; g.a_rs = ldx; g.a_cs = 1; g.B = W; g.b_rs = 1; g.b_cs = ldw; g.C = Y; g.ldc = ldy;
  g.M = rows; g.N = Nout; g.K = K; g.bias = bias; g.ld_bias = ld_bias; g.bias_div = div; g.bias_mod = mod; g.relu = relu;
  return run_gemm(tc, g, st);
}
// dX[rows,K] (+)= (dY[rows,Nout] W[Nout,K]) * (mask > 0)
int dgrad(bool tc, const float* dY, int64_t ldy, int64_t rows, int Nout, const float* W, int64_t ldw, int K, float* dX,
          int64_t ldx, const float* mask, int64_t ldm, int accumulate, cudaStream_t st) {
  GemmArgs g;
  g.A = dY; g.a_rs = ldy; g.a_cs = 1; g.B = W; g.b_rs = ldw; g.b_cs = 1; g.C = dX; g.ldc = ldx;
  g.M = rows; g.N = K; g.K = Nout; g.mask = mask; g.ld_mask = ldm; g.accumulate = accumulate;
  return run_gemm(tc, g, st);
}
// dW[Nout,K] += dY[rows,Nout]^T X[rows,K]   (dW zero-filled by the caller: split-K partials are added atomically)
int wgrad(bool tc, const float* dY, int64_t ldy, const float* X, int64_t ldx, int64_t rows, int Nout, int K, float* dW,
          int64_t ldw, cudaStream_t st) {
  GemmArgs g;
  g.A = dY; g.a_rs = 1; g.a_cs = ldy; g.B = X; g.b_rs = ldx; g.b_cs = 1; g.C = dW; g.ldc = ldw;
  g.M = Nout; g.N = K; g.K = rows;
  int64_t tiles = ((Nout + 63) / 64) * (int64_t)((K + 63) / 64);
  int64_t want = (1184 + tiles - 1) / tiles, cap = (rows + 511) / 512;
  g.split_k = (int)std::max<int64_t>(1, std::min<int64_t>(want, cap));
  if (g.split_k == 1) g.accumulate = 1;
  return run_gemm(tc, g, st);
}

inline void pair_index(const FDims& d, int64_t& div, int64_t& mod) {
  if (d.pairing == CNRF_PAIRING_REFERENCE) { div = 1; mod = d.N; } else { div = d.S; mod = INT64_MAX; }
}

// dst[m, 0..w) = src[m, 0..w)  (strided rows): gamma(x) into the right part of a skip buffer, g_raw columns apart
int copy_cols(const float* src, int64_t ld_src, float* dst, int64_t ld_dst, int64_t rows, int w, cudaStream_t st) {
  cudaError_t e = cudaMemcpy2DAsync(dst, (size_t)ld_dst * sizeof(float), src, (size_t)ld_src * sizeof(float),
                                    (size_t)w * sizeof(float), (size_t)rows, cudaMemcpyDeviceToDevice, st);
  return e == cudaSuccess ? CNRF_OK : (int)e;
}

// out[m,n] = (in[m,n] + gs[m*ld_gs] * w[n]) * (act[m,n] > 0):  d feat = (d v Wdir_feat + g_sigma fc_alpha) masked by feat
__global__ void add_outer_mask_kernel(float* __restrict__ io, const float* __restrict__ gs, int64_t ld_gs,
                                      const float* __restrict__ w, const float* __restrict__ act, int64_t M, int H) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= M * H) return;
  const int64_t m = i / H;
  const int n = (int)(i - m * H);
  const float v = io[i] + gs[m * ld_gs] * __ldg(w + n);
  io[i] = act[i] > 0.f ? v : 0.f;
}

// saved activations: act[0..L] (ld per FDims::ld_act), feat [M,H], v [M,H/2], betad [N,H/2]
struct Saved { float* act[CNRF_FLEX_MAX_XYZ_LAYERS + 1]; float *feat, *v; };
inline size_t saved_bytes_of(const FDims& d) {
  size_t b = 0;
  for (int a = 0; a <= d.L; ++a) b += al(d.M * d.ld_act(a));
  return b + al(d.M * d.H) + al(d.M * (d.H / 2));
}
inline bool carve_saved(const FDims& d, void* p, size_t bytes, Saved& s) {
  Arena a(p, bytes);
  for (int i = 0; i <= d.L; ++i) s.act[i] = a.take(d.M * d.ld_act(i));
  s.feat = a.take(d.M * d.H);
  s.v = a.take(d.M * (d.H / 2));
  return a.ok;
}

}  // namespace

extern "C" size_t cnrf_flexnerf_saved_bytes(const cnrf_flex_shape* s) {
  FDims d;
  if (dims_of(s, d) != CNRF_OK || !s->keep) return 0;
  return saved_bytes_of(d);
}

extern "C" size_t cnrf_flexnerf_workspace_bytes(const cnrf_flex_shape* s, int32_t backward) {
  FDims d;
  if (dims_of(s, d) != CNRF_OK) return 0;
  const size_t betad = al(d.N * (d.H / 2));
  // inference: two ping-pong activation buffers wide enough for a skip input, + feat + v live in the workspace
  if (!backward && cnrf_flex_fused_eligible(s))            // fused chain: betad + the packed bf16 weight image only
    return betad + 1024 + cnrf_flex_fused_packed_bytes(d.L);
  if (!backward) return betad + (s->keep ? 0 : 2 * al(d.M * (d.H + d.dx)) + al(d.M * d.H) + al(d.M * (d.H / 2)));
  // backward: two ping-pong [M,H] gradient buffers, d v [M,H/2], d betad [N,H/2], ones [M]
  return 2 * al(d.M * d.H) + al(d.M * (d.H / 2)) + betad + al(d.M);
}

namespace {
// inference: two ping-pong activation buffers wide enough for a skip input, + feat + v
inline bool carve_scratch(const FDims& d, Arena& a, Saved& s) {
  float* b0 = a.take(d.M * (d.H + d.dx));
  float* b1 = a.take(d.M * (d.H + d.dx));
  for (int i = 0; i <= d.L; ++i) s.act[i] = (i & 1) ? b1 : b0;
  s.feat = a.take(d.M * d.H);
  s.v = a.take(d.M * (d.H / 2));
  return a.ok;
}
}  // namespace

extern "C" int cnrf_flexnerf_fwd(const cnrf_flex_shape* s, const cnrf_flex_params* p, const float* x_xyz, int64_t ld_xyz,
                                 const float* x_dir, int64_t ld_dir, float* raw, void* saved, size_t saved_bytes,
                                 void* workspace, size_t workspace_bytes, cnrf_stream_t stream) {
  FDims d;
  RUN(dims_of(s, d));
  CNRF_REQUIRE(p && x_xyz && (x_dir || !d.viewdirs()) && raw && workspace, CNRF_E_NULL);
  CNRF_REQUIRE(ld_xyz >= d.dx && ld_dir >= d.dd, CNRF_E_SHAPE);
  cudaStream_t st = as_stream(stream);
  const int H = d.H, H2 = d.H / 2;
  Arena ws(workspace, workspace_bytes);
  float* betad = ws.take(d.N * H2);
  static const bool no_fused = getenv("CNRF_FLEX_NO_FUSED") != nullptr;      // A/B switch: per-layer toolkit GEMMs
  if (cnrf_flex_fused_eligible(s) && !no_fused) {
    // inference in bf16: the whole per-sample chain is ONE persistent tcgen05 kernel (flex_fwd_tc.cu)
    float* pk = ws.take((cnrf_flex_fused_packed_bytes(d.L) + 1024) / sizeof(float));
    CNRF_REQUIRE(ws.ok, CNRF_E_WORKSPACE);
    uint8_t* packed = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(pk) + 1023) & ~uintptr_t(1023));
    return cnrf_flex_fused_fwd(s, p, x_xyz, ld_xyz, x_dir, ld_dir, raw, packed, st);
  }
  Saved sv;
  if (s->keep) {
    CNRF_REQUIRE(saved, CNRF_E_NULL);
    CNRF_REQUIRE(carve_saved(d, saved, saved_bytes, sv), CNRF_E_WORKSPACE);
  } else {
    CNRF_REQUIRE(carve_scratch(d, ws, sv), CNRF_E_WORKSPACE);
  }
  CNRF_REQUIRE(ws.ok, CNRF_E_WORKSPACE);

  // per ray: betad = Wdir[:, H:] gamma(d) + bdir                                      models/model.py:69-71
  if (d.viewdirs()) RUN(lin(d.tc, x_dir, ld_dir, d.N, d.dd, p->w_dir + H, H + d.dd, H2, betad, H2, p->b_dir, 0, 1, INT64_MAX, 0, st));
  // h = relu(layer1 xyz)                                                              :59
  RUN(lin(d.tc, x_xyz, ld_xyz, d.M, d.dx, p->w_layer1, d.dx, H, sv.act[0], d.ld_act(0), p->b_layer1, 0, 1, INT64_MAX, 1, st));
  for (int i = 0; i < d.L; ++i) {                                                      // :60-64
    const int K = d.is_skip(i) ? H + d.dx : H;
    if (d.is_skip(i)) RUN(copy_cols(x_xyz, ld_xyz, sv.act[i] + H, H + d.dx, d.M, d.dx, st));
    RUN(lin(d.tc, sv.act[i], d.ld_act(i), d.M, K, p->w_xyz[i], K, H, sv.act[i + 1], d.ld_act(i + 1), p->b_xyz[i], 0, 1,
            INT64_MAX, 1, st));
  }
  if (!d.viewdirs())   // raw = fc_out h_L  (fc_out [4,H] rides in w_rgb / b_rgb)                                  :76
    return lin(d.tc, sv.act[d.L], d.ld_act(d.L), d.M, H, p->w_rgb, H, 4, raw, 4, p->b_rgb, 0, 1, INT64_MAX, 0, st);
  RUN(lin(d.tc, sv.act[d.L], H, d.M, H, p->w_feat, H, H, sv.feat, H, p->b_feat, 0, 1, INT64_MAX, 1, st));      // :67
  RUN(lin(d.tc, sv.feat, H, d.M, H, p->w_alpha, H, 1, raw + 3, 4, p->b_alpha, 0, 1, INT64_MAX, 0, st));         // :68
  int64_t div, mod;
  pair_index(d, div, mod);
  RUN(lin(d.tc, sv.feat, H, d.M, H, p->w_dir, H + d.dd, H2, sv.v, H2, betad, H2, div, mod, 1, st));             // :69-71
  RUN(lin(d.tc, sv.v, H2, d.M, H2, p->w_rgb, H2, 3, raw, 4, p->b_rgb, 0, 1, INT64_MAX, 0, st));                 // :72-73
  return CNRF_OK;
}

extern "C" int32_t cnrf_flexnerf_fused_inference(const cnrf_flex_shape* s) {
  FDims d;
  static const bool no_fused = getenv("CNRF_FLEX_NO_FUSED") != nullptr;
  return dims_of(s, d) == CNRF_OK && cnrf_flex_fused_eligible(s) && !no_fused ? 1 : 0;
}

extern "C" int cnrf_flexnerf_fwd_pts(const cnrf_flex_shape* s, const cnrf_flex_params* p, const float* pts, int64_t ld_pts,
                                     const float* freqs_xyz, int32_t num_freqs, int32_t include_input, const float* x_dir,
                                     int64_t ld_dir, float* raw, void* workspace, size_t workspace_bytes,
                                     cnrf_stream_t stream) {
  FDims d;
  RUN(dims_of(s, d));
  CNRF_REQUIRE(p && pts && freqs_xyz && x_dir && raw && workspace, CNRF_E_NULL);
  CNRF_REQUIRE(ld_pts >= 3 && ld_dir >= d.dd, CNRF_E_SHAPE);
  if (!cnrf_flexnerf_fused_inference(s)) return CNRF_E_UNSUPPORTED;
  cudaStream_t st = as_stream(stream);
  const int H = d.H, H2 = d.H / 2;
  Arena ws(workspace, workspace_bytes);
  float* betad = ws.take(d.N * H2);
  float* pk = ws.take((cnrf_flex_fused_packed_bytes(d.L) + 1024) / sizeof(float));
  CNRF_REQUIRE(ws.ok, CNRF_E_WORKSPACE);
  uint8_t* packed = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(pk) + 1023) & ~uintptr_t(1023));
  const FlexPts gp = {pts, ld_pts, freqs_xyz, num_freqs, include_input ? 1 : 0};
  return cnrf_flex_fused_fwd(s, p, nullptr, 0, x_dir, ld_dir, raw, packed, st, &gp);
}

extern "C" int cnrf_flexnerf_bwd(const cnrf_flex_shape* s, const cnrf_flex_params* p, const float* x_xyz, int64_t ld_xyz,
                                 const float* x_dir, int64_t ld_dir, const float* g_raw, const void* saved,
                                 size_t saved_bytes, const cnrf_flex_grads* g, float* g_x_xyz, float* g_x_dir,
                                 void* workspace, size_t workspace_bytes, cnrf_stream_t stream) {
  FDims d;
  RUN(dims_of(s, d));
  CNRF_REQUIRE(p && x_xyz && (x_dir || !d.viewdirs()) && g_raw && saved && g && workspace, CNRF_E_NULL);
  cudaStream_t st = as_stream(stream);
  const int H = d.H, H2 = d.H / 2;
  Saved sv;
  CNRF_REQUIRE(carve_saved(d, const_cast<void*>(saved), saved_bytes, sv), CNRF_E_WORKSPACE);
  Arena ws(workspace, workspace_bytes);
  float* ga = ws.take(d.M * H);
  float* gb = ws.take(d.M * H);
  float* dv = ws.take(d.M * H2);
  float* dbetad = ws.take(d.N * H2);
  (void)ws.take(d.M);                          // (was the ones vector of the bias-gradient GEMMs; the workspace layout stays)
  CNRF_REQUIRE(ws.ok, CNRF_E_WORKSPACE);

  // zero-fill every parameter gradient (split-K partial sums are added atomically)
  {
    FillBatch f;
    auto flush_if_full = [&]() -> int { if (f.count == CNRF_FILL_BATCH_MAX) { int rc = cnrf_fill_batch(f, st); f.count = 0; return rc; } return CNRF_OK; };
    auto add = [&](float* ptr, int64_t n, float v = 0.f) -> int { f.add(ptr, n, v); return flush_if_full(); };
    RUN(add(g->w_layer1, (int64_t)H * d.dx)); RUN(add(g->b_layer1, H));
    for (int i = 0; i < d.L; ++i) { RUN(add(g->w_xyz[i], (int64_t)H * (d.is_skip(i) ? H + d.dx : H))); RUN(add(g->b_xyz[i], H)); }
    if (d.viewdirs()) {
      RUN(add(g->w_feat, (int64_t)H * H)); RUN(add(g->b_feat, H));
      RUN(add(g->w_dir, (int64_t)H2 * (H + d.dd))); RUN(add(g->b_dir, H2));
      RUN(add(g->w_alpha, H)); RUN(add(g->b_alpha, 1));
      RUN(add(g->w_rgb, (int64_t)3 * H2)); RUN(add(g->b_rgb, 3));
    } else {
      RUN(add(g->w_rgb, (int64_t)4 * H)); RUN(add(g->b_rgb, 4));
    }
    if (g_x_xyz) RUN(add(g_x_xyz, d.M * d.dx));
    RUN(cnrf_fill_batch(f, st));
  }
  // db = column sums of dY, exact fp32: as a GEMM against a ones vector it was a 128 x 256 tensor-core tile with one live row
  // that re-read dY at a fraction of the HBM rate (a tenth of this backward's launches)
  auto bias_grad = [&](const float* dY, int64_t ldy, int64_t rows, int Nout, float* db) {
    return cnrf_colsum(dY, ldy, rows, Nout, db, st);
  };
  float *cur = gb, *nxt = ga;   // cur receives d h_L
  if (!d.viewdirs()) {
    // raw = fc_out h_L                                                                  models/model.py:76
    RUN(wgrad(d.tc, g_raw, 4, sv.act[d.L], d.ld_act(d.L), d.M, 4, H, g->w_rgb, H, st));
    RUN(bias_grad(g_raw, 4, d.M, 4, g->b_rgb));
    RUN(dgrad(d.tc, g_raw, 4, d.M, 4, p->w_rgb, H, H, cur, H, sv.act[d.L], d.ld_act(d.L), 0, st));
  } else {
  // rgb = fc_rgb v                                                                      models/model.py:72
  RUN(wgrad(d.tc, g_raw, 4, sv.v, H2, d.M, 3, H2, g->w_rgb, H2, st));
  RUN(bias_grad(g_raw, 4, d.M, 3, g->b_rgb));
  RUN(dgrad(d.tc, g_raw, 4, d.M, 3, p->w_rgb, H2, H2, dv, H2, sv.v, H2, 0, st));       // d v (masked by v > 0)
  // v = relu(Wdir_feat feat + betad[pair])                                              :69-71
  RUN(wgrad(d.tc, dv, H2, sv.feat, H, d.M, H2, H, g->w_dir, H + d.dd, st));
  if (d.pairing == CNRF_PAIRING_REFERENCE) RUN(cnrf_strided_rowsum(dv, H2, d.N, 1, d.S, d.N, H2, dbetad, H2, st));
  else RUN(cnrf_strided_rowsum(dv, H2, d.N, d.S, d.S, 1, H2, dbetad, H2, st));
  RUN(wgrad(d.tc, dbetad, H2, x_dir, ld_dir, d.N, H2, d.dd, g->w_dir + H, H + d.dd, st));
  RUN(bias_grad(dbetad, H2, d.N, H2, g->b_dir));
  if (g_x_dir) RUN(dgrad(d.tc, dbetad, H2, d.N, H2, p->w_dir + H, H + d.dd, d.dd, g_x_dir, d.dd, nullptr, 0, 0, st));
  // sigma = fc_alpha feat                                                               :68
  RUN(wgrad(d.tc, g_raw + 3, 4, sv.feat, H, d.M, 1, H, g->w_alpha, H, st));
  RUN(bias_grad(g_raw + 3, 4, d.M, 1, g->b_alpha));
  // d feat = (d v Wdir_feat + g_sigma fc_alpha) * (feat > 0)
  RUN(dgrad(d.tc, dv, H2, d.M, H2, p->w_dir, H + d.dd, H, ga, H, nullptr, 0, 0, st));
  add_outer_mask_kernel<<<grid_for(d.M * H, 256), 256, 0, st>>>(ga, g_raw + 3, 4, p->w_alpha, sv.feat, d.M, H);
  CNRF_LAUNCH_CHECK();
  // feat = relu(fc_feat h_L)                                                            :67
  RUN(wgrad(d.tc, ga, H, sv.act[d.L], d.ld_act(d.L), d.M, H, H, g->w_feat, H, st));
  RUN(bias_grad(ga, H, d.M, H, g->b_feat));
  RUN(dgrad(d.tc, ga, H, d.M, H, p->w_feat, H, H, cur, H, sv.act[d.L], d.ld_act(d.L), 0, st));
  }
  for (int i = d.L - 1; i >= 0; --i) {                                                   // :60-64
    const int K = d.is_skip(i) ? H + d.dx : H;
    RUN(wgrad(d.tc, cur, H, sv.act[i], d.ld_act(i), d.M, H, K, g->w_xyz[i], K, st));
    RUN(bias_grad(cur, H, d.M, H, g->b_xyz[i]));
    if (d.is_skip(i) && g_x_xyz) RUN(dgrad(d.tc, cur, H, d.M, H, p->w_xyz[i] + H, K, d.dx, g_x_xyz, d.dx, nullptr, 0, 1, st));
    RUN(dgrad(d.tc, cur, H, d.M, H, p->w_xyz[i], K, H, nxt, H, sv.act[i], d.ld_act(i), 0, st));
    std::swap(cur, nxt);
  }
  // h_0 = relu(layer1 xyz)                                                              :59
  RUN(wgrad(d.tc, cur, H, x_xyz, ld_xyz, d.M, H, d.dx, g->w_layer1, d.dx, st));
  RUN(bias_grad(cur, H, d.M, H, g->b_layer1));
  if (g_x_xyz) RUN(dgrad(d.tc, cur, H, d.M, H, p->w_layer1, d.dx, d.dx, g_x_xyz, d.dx, nullptr, 0, 1, st));
  return CNRF_OK;
}
