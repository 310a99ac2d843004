// CodeNeRF radiance field: CodeNeRFModel.forward (view_synthesis/models/model.py:160-194) and forward_pass
// (view_synthesis/nerf/__init__.py:94-134), forward and backward, behind cnrf_codenerf_* / cnrf_field_*.
//
// Code folding: the three latent-code layers and the code halves of layer_xyz2 / fc_out / fc_rgb, and the
// view-direction half of layer_dir1, depend on the RAY only.  They are evaluated once per ray (fp32, sgemm.cu) into
// bias tables
//     beta2 = W2c relu(Ws1 z_s + bs1) + b2     betao = Woc relu(Ws2 z_s + bs2) + bo
//     betar = Wrt relu(Wt  z_t + bt ) + br     betad = Wdv gamma(d^) + bd1
// and the per-SAMPLE chain is  h1 = relu(W1 gamma(x) + b1), h2 = relu(W2h h1 + beta2[ray]), o = Woh h2 + betao[ray],
// v1 = relu(Wdf o[1:] + betad[pair]), v2 = relu(Wd2 v1 + bd2), rgb = Wrv v2 + betar[ray].   `pair` reproduces the
// reference's view-direction tiling (row k -> ray k mod N, nerf/__init__.py:127) unless CNRF_PAIRING_CORRECT.
// The chain runs either as fp32 SIMT GEMMs (this file; parity mode) or as the fused tcgen05 kernels of mlp_tc.cu.
//
// fc_out fold (tensor-core mode only).  models/model.py:182-189 apply no activation between fc_out and layer_dir1, so
//     v1 = relu(Wdf (Wout' [h2 | as2] + bo') + Wdv gamma(d) + bd1) = relu(Wfh h2 + betaf[ray] + Wdv gamma(d) + bd1),
//     Wfh = Wdf Wout'h  [256 x 256],   betaf = Wdf betao',   betao' = Wout'c as2 + bo'   (Wout' = fc_out rows 1.., bo' likewise)
// and the per-sample chain has four 256-wide layers instead of five.  Wfh is rebuilt from the fp32 parameters at every
// call as four K-slices written to separate buffers (four one-round-trip problems in the first per-ray launch) that the
// weight packer adds in a fixed order -- the forward stays bit-reproducible.  The backward returns d Wfh (weight-gradient
// kernel) and d betaf (its per-ray reducers); they are unfolded here:
//     d Wdf = d Wfh Wout'h^T + d betaf^T betao',   d Wout'h = Wdf^T d Wfh,   d betao' = d betaf Wdf
// after which the code halves of fc_out take the same path as before (per_ray_bwd).
#include <algorithm>
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

struct Dims {
  int H, C, dx, dd, S, pairing;
  int64_t N, M;
  bool tc;   // per-ray GEMMs on the tensor cores (gemm_tc.cu) instead of the fp32 SIMT toolkit
  bool split = false;   // CNRF_PRECISION_SPLIT: the per-SAMPLE GEMMs of the fp32-layout chain as split-bf16 tensor-core GEMMs
  const int32_t* rows_dev = nullptr;   // shared-code mode: device-resident number of code runs (<= N), see code_runs.cu
};
inline Dims dims_of(const cnrf_field_shape* s) {
  Dims d;
  d.H = s->hidden; d.C = s->code; d.dx = s->dim_xyz; d.dd = s->dim_dir; d.S = s->num_samples; d.pairing = s->pairing;
  d.N = s->num_rays; d.M = s->num_rays * (int64_t)s->num_samples;
  d.tc = s->precision == CNRF_PRECISION_BF16;
  d.split = s->precision == CNRF_PRECISION_SPLIT;
  return d;
}
// CNRF_PRECISION_SPLIT runs the fp32 path (same saved layout, per-layer GEMMs with fp32 activations) with the per-sample
// contractions on the tensor cores as split-bf16 (gemm_tc.cu; SIMT where the toolkit declines a shape); the per-ray code
// layers (1 / S of the work) stay fp32 FMA
inline int layout_of(const cnrf_field_shape* s) { return s->precision == CNRF_PRECISION_BF16 ? CNRF_PRECISION_BF16 : CNRF_PRECISION_FP32; }
inline bool fp32_layout(const cnrf_field_shape* s) { return s->precision != CNRF_PRECISION_BF16; }
inline int check_shape(const cnrf_field_shape* s) {
  if (!s) return CNRF_E_NULL;
  if (s->hidden <= 0 || s->code <= 0 || s->dim_xyz <= 0 || s->dim_dir <= 0 || s->num_rays <= 0 || s->num_samples <= 0)
    return CNRF_E_SHAPE;
  if (s->pairing != CNRF_PAIRING_REFERENCE && s->pairing != CNRF_PAIRING_CORRECT) return CNRF_E_SHAPE;
  if (s->precision != CNRF_PRECISION_FP32 && s->precision != CNRF_PRECISION_BF16 && s->precision != CNRF_PRECISION_SPLIT) return CNRF_E_UNSUPPORTED;
  return CNRF_OK;
}

// saved-for-backward layout (fp32 precision)
struct SavedF32 { float *h1, *h2, *o, *v1, *v2, *as1, *as2, *at; };
inline size_t saved_f32_bytes(const Dims& d) {
  return 4 * al(d.M * d.H) + al(d.M * (d.C + 1)) + 3 * al(d.N * d.C);
}
inline bool carve_saved(const Dims& d, void* p, size_t bytes, SavedF32& s) {
  Arena a(p, bytes);
  s.h1 = a.take(d.M * d.H); s.h2 = a.take(d.M * d.H); s.o = a.take(d.M * (d.C + 1));
  s.v1 = a.take(d.M * d.H); s.v2 = a.take(d.M * d.H);
  s.as1 = a.take(d.N * d.C); s.as2 = a.take(d.N * d.C); s.at = a.take(d.N * d.C);
  return a.ok;
}
// per-ray bias tables (forward workspace) and their gradients (backward workspace)
// bof/bos: fc_out bias table split (aligned feat / sigma) for the tc path; bf: bias table of the folded layer_dir1 (forward)
// or its gradient (backward); wf [4][H x H]: the four K-slices of this call's fold Wfh; dwfh [H x H]: gradient of Wfh
struct Tables { float *b2, *bo, *br, *bd, *bof, *bos, *bf, *wf, *dwfh; };
inline size_t tables_bytes(const Dims& d) {
  return 2 * al(d.N * d.H) + al(d.N * (d.C + 1)) + al(d.N * 3) + al(d.N * d.C) + al(d.N) + al(d.N * d.H) + al((size_t)4 * d.H * d.H) +
         al((size_t)d.H * d.H);
}
inline void carve_tables(const Dims& d, Arena& a, Tables& t) {
  t.b2 = a.take(d.N * d.H); t.bo = a.take(d.N * (d.C + 1)); t.br = a.take(d.N * 3); t.bd = a.take(d.N * d.H);
  t.bof = a.take(d.N * d.C); t.bos = a.take(d.N);
  t.bf = a.take(d.N * d.H); t.wf = a.take((size_t)4 * d.H * d.H); t.dwfh = a.take((size_t)d.H * d.H);
}

#define RUN(x) do { int rc__ = (x); if (rc__) return rc__; } while (0)

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
// dW[Nout,K] = dY[rows,Nout]^T X[rows,K]   (dW must be zero-filled: split-K partials are added atomically)
int wgrad(bool tc, const float* dY, int64_t ldy, const float* X, int64_t ldx, int64_t rows, int Nout, int K, float* dW,
          int64_t ldw, cudaStream_t st) {
  GemmArgs g;
  g.A = dY; g.a_rs = 1; g.a_cs = ldy; g.B = X; g.b_rs = ldx; g.b_cs = 1; g.C = dW; g.ldc = ldw;
  g.M = Nout; g.N = K; g.K = rows;
  int64_t tiles = ((Nout + 63) / 64) * (int64_t)((K + 63) / 64);
  int64_t want = (1184 + tiles - 1) / tiles, cap = (rows + 511) / 512;
  g.split_k = (int)std::max<int64_t>(1, std::min<int64_t>(want, cap));
  if (g.split_k == 1) g.accumulate = 1;   // still += into the zero-filled buffer
  return run_gemm(tc, g, st);
}

__global__ void copy_col_kernel(const float* __restrict__ src, int64_t ld_src, float* __restrict__ dst, int64_t ld_dst,
                                int64_t M) {
  int64_t m = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (m < M) dst[m * ld_dst] = src[m * ld_src];
}
int copy_col(const float* src, int64_t ld_src, float* dst, int64_t ld_dst, int64_t M, cudaStream_t st) {
  copy_col_kernel<<<grid_for(M, 256), 256, 0, st>>>(src, ld_src, dst, ld_dst, M);
  CNRF_LAUNCH_CHECK();
  return CNRF_OK;
}

inline void pair_index(const Dims& d, int64_t& div, int64_t& mod) {
  if (d.pairing == CNRF_PAIRING_REFERENCE) { div = 1; mod = d.N; } else { div = d.S; mod = INT64_MAX; }
}

// ---- per-ray layers ------------------------------------------------------------------------------------------
// Mutually independent GEMMs are collected and issued as ONE tcgen05 launch (gemm_tc.cu) in the bf16 mode; the fp32
// mode runs them one by one through the SIMT kernel.
GemmArgs mk_lin(const float* X, int64_t ldx, int64_t rows, int K, const float* W, int64_t ldw, int Nout, float* Y,
                int64_t ldy, const float* bias, int relu) {
  GemmArgs g;
  g.A = X; g.a_rs = ldx; g.a_cs = 1; g.B = W; g.b_rs = 1; g.b_cs = ldw; g.C = Y; g.ldc = ldy;
  g.M = rows; g.N = Nout; g.K = K; g.bias = bias; g.ld_bias = 0; g.relu = relu;
  g.rows_dim = 1;
  return g;
}
GemmArgs mk_dgrad(const float* dY, int64_t ldy, int64_t rows, int Nout, const float* W, int64_t ldw, int K, float* dX,
                  int64_t ldx, const float* mask, int64_t ldm, int accumulate) {
  GemmArgs g;
  g.A = dY; g.a_rs = ldy; g.a_cs = 1; g.B = W; g.b_rs = ldw; g.b_cs = 1; g.C = dX; g.ldc = ldx;
  g.M = rows; g.N = K; g.K = Nout; g.mask = mask; g.ld_mask = ldm; g.accumulate = accumulate;
  g.rows_dim = 1;
  return g;
}
GemmArgs mk_wgrad(const float* dY, int64_t ldy, const float* X, int64_t ldx, int64_t rows, int Nout, int K, float* dW,
                  int64_t ldw) {
  GemmArgs g;
  g.A = dY; g.a_rs = 1; g.a_cs = ldy; g.B = X; g.b_rs = ldx; g.b_cs = 1; g.C = dW; g.ldc = ldw;
  g.M = Nout; g.N = K; g.K = rows;
  int64_t tiles = ((Nout + 63) / 64) * (int64_t)((K + 63) / 64);
  int64_t want = (1184 + tiles - 1) / tiles, cap = (rows + 511) / 512;
  g.split_k = (int)std::max<int64_t>(1, std::min<int64_t>(want, cap));
  if (g.split_k == 1) g.accumulate = 1;   // += into the zero-filled buffer
  g.rows_dim = 2;
  return g;
}
// out[0, n] += sum_m X[m, n] as ones^T X (out zero-filled by the caller)
GemmArgs mk_colsum(const float* ones, const float* X, int64_t ldx, int64_t rows, int N, float* out) {
  return mk_wgrad(ones, 1, X, ldx, rows, 1, N, out, N);
}
struct Batch {
  GemmArgs g[CNRF_GEMM_BATCH_MAX];
  int n = 0;
  const int32_t* rows_dev = nullptr;       // every problem of the batch runs over this many rows (device-resident)
  explicit Batch(const int32_t* rd = nullptr) : rows_dev(rd) {}
  void add(const GemmArgs& a) { g[n] = a; g[n].rows_dev = rows_dev; ++n; }
  void add_fixed(const GemmArgs& a) { g[n] = a; g[n].rows_dev = nullptr; g[n].rows_dim = 0; ++n; }   // shape known on the host
  int flush(bool tc, cudaStream_t st) {
    int rc = tc ? cnrf_gemm_tc_batch(g, n, st) : CNRF_E_UNSUPPORTED;
    if (rc == CNRF_E_UNSUPPORTED && rows_dev) return rc;      // only the tensor-core kernel reads a device-side row count
    if (rc == CNRF_E_UNSUPPORTED) {
      rc = CNRF_OK;
      for (int i = 0; i < n && rc == CNRF_OK; ++i) rc = cnrf_sgemm(g[i], st);
    }
    n = 0;
    return rc;
  }
};

// plain strided GEMM C[M,N] += A B without epilogue (the fold / unfold products); C zero-filled by the caller
GemmArgs mk_mm(const float* A, int64_t a_rs, int64_t a_cs, const float* B, int64_t b_rs, int64_t b_cs, float* Cm, int64_t ldc,
               int64_t M, int N, int64_t K) {
  GemmArgs g;
  g.A = A; g.a_rs = a_rs; g.a_cs = a_cs; g.B = B; g.b_rs = b_rs; g.b_cs = b_cs; g.C = Cm; g.ldc = ldc; g.M = M; g.N = N; g.K = K;
  g.accumulate = 1; g.k_slice = 64;      // += onto a zero-filled C, one 64-wide K slice per CTA (see GemmArgs::k_slice)
  return g;
}

// The per-ray layers as dependency LEVELS of mutually independent GEMMs: a caller adds one level of one model (or of both
// models of a render call: the pair path below) to a Batch and flushes it as one launch.
// bof (folded mode): the fc_out feature bias table betao' [N,C]; betaf = betao' Wdf^T follows from it, and the backward's
// d Wdf needs it again
constexpr int PER_RAY_FWD_LEVELS = 3;
int per_ray_fwd_add(int level, const Dims& d, const cnrf_codenerf_params* p, const float* x_dir, int64_t ld_dir, const float* z_s,
                    int64_t ld_zs, const float* z_t, int64_t ld_zt, float* as1, float* as2, float* at, const Tables& t,
                    bool split_bo, float* bof, Batch& b) {
  const int H = d.H, C = d.C;
  if (level == 0) {
    if (split_bo) {   // the fold Wfh = Wdf Wout'h as four K-slices (Wdf = layer_dir1[:, :C], Wout'h = fc_out[1:, :H]): slice j goes to
      if (!bof || C % 4) return CNRF_E_NULL;      // its own [H x H] buffer (plain stores; pack_weights_kernel adds them in order)
      for (int j = 0; j < 4; ++j) {
        GemmArgs g = mk_mm(p->weight[CNRF_L_DIR1] + j * (C / 4), C + d.dd, 1, p->weight[CNRF_L_FC_OUT] + (int64_t)(1 + j * (C / 4)) * (H + C),
                           H + C, 1, t.wf + (size_t)j * H * H, H, H, H, C / 4);
        g.accumulate = 0; g.k_slice = 0;
        b.add_fixed(g);
      }
    }
    b.add(mk_lin(z_s, ld_zs, d.N, C, p->weight[CNRF_L_SHAPE1], C, C, as1, C, p->bias[CNRF_L_SHAPE1], 1));     // model.py:174
    b.add(mk_lin(z_s, ld_zs, d.N, C, p->weight[CNRF_L_SHAPE2], C, C, as2, C, p->bias[CNRF_L_SHAPE2], 1));     // :175
    b.add(mk_lin(z_t, ld_zt, d.N, C, p->weight[CNRF_L_TEXTURE1], C, C, at, C, p->bias[CNRF_L_TEXTURE1], 1));  // :177
    // the tensor-core chain applies the view-direction half of layer_dir1 as 32 extra input columns instead of a bias table
    if (!d.tc) b.add(mk_lin(x_dir, ld_dir, d.N, d.dd, p->weight[CNRF_L_DIR1] + C, C + d.dd, H, t.bd, H, p->bias[CNRF_L_DIR1], 0));
  } else if (level == 1) {
    b.add(mk_lin(as1, C, d.N, C, p->weight[CNRF_L_XYZ2] + H, H + C, H, t.b2, H, p->bias[CNRF_L_XYZ2], 0));
    if (split_bo) {   // tensor-core chain: betao' and the sigma column here, the folded layer_dir1 bias rows one level later
      b.add(mk_lin(as2, C, d.N, C, p->weight[CNRF_L_FC_OUT] + (H + C) + H, H + C, C, bof, C, p->bias[CNRF_L_FC_OUT] + 1, 0));
      b.add(mk_lin(as2, C, d.N, C, p->weight[CNRF_L_FC_OUT] + H, H + C, 1, t.bos, 1, p->bias[CNRF_L_FC_OUT], 0));
    } else {
      b.add(mk_lin(as2, C, d.N, C, p->weight[CNRF_L_FC_OUT] + H, H + C, C + 1, t.bo, C + 1, p->bias[CNRF_L_FC_OUT], 0));
    }
    b.add(mk_lin(at, C, d.N, C, p->weight[CNRF_L_FC_RGB] + H, H + C, 3, t.br, 3, p->bias[CNRF_L_FC_RGB], 0));
  } else if (split_bo) {   // betaf = betao' Wdf^T
    b.add(mk_lin(bof, C, d.N, C, p->weight[CNRF_L_DIR1], C + d.dd, H, t.bf, H, nullptr, 0));
  }
  return CNRF_OK;
}
int per_ray_fwd(const Dims& d, const cnrf_codenerf_params* p, const float* x_dir, int64_t ld_dir, const float* z_s,
                int64_t ld_zs, const float* z_t, int64_t ld_zt, float* as1, float* as2, float* at, const Tables& t,
                bool split_bo, cudaStream_t st, float* bof = nullptr) {
  Batch b(d.rows_dev);
  for (int level = 0; level < PER_RAY_FWD_LEVELS; ++level) {
    RUN(per_ray_fwd_add(level, d, p, x_dir, ld_dir, z_s, ld_zs, z_t, ld_zt, as1, as2, at, t, split_bo, bof, b));
    RUN(b.flush(d.tc, st));
  }
  return CNRF_OK;
}

// gradients of the bias tables -> code-layer weights, code halves of the split layers, z_s / z_t / x_dir.
// tmp: 3 x [N,C] scratch + [N] (a ones vector for the bias-gradient column sums of the tensor-core mode).
// bof (tensor-core mode): the forward's betao' table [N,C]; dt.bf / dt.dwfh then hold d betaf and d Wfh (see the fold
// note at the top of the file) and dt.bo only its sigma column.
// Levels: 0 (tensor-core mode only) the unfold products; 1 the bias tables' layers; 2 the three latent-code layers.
// The tensor-core mode adds both shape branches into g_z_s atomically: the caller zero-fills g_z_s and sets the ones
// vector (per_ray_bwd_fills) before level 0; the fp32 mode finishes with column-sum launches (per_ray_bwd_colsums).
// acc_codes: g_z_s / g_z_t already hold another model's code gradients (pair path): add to them
constexpr int PER_RAY_BWD_LEVELS = 3;
inline float* per_ray_ones(const Dims& d, float* tmp) { return tmp + 3 * (al(d.N * d.C) / sizeof(float)); }
int per_ray_bwd_add(int level, const Dims& d, const cnrf_codenerf_params* p, const cnrf_codenerf_grads* g, const float* x_dir,
                    int64_t ld_dir, const float* z_s, int64_t ld_zs, const float* z_t, int64_t ld_zt, const float* as1,
                    const float* as2, const float* at, const Tables& dt, float* tmp, float* g_z_s, float* g_z_t,
                    float* g_x_dir, const float* bof, bool acc_codes, Batch& b) {
  const int H = d.H, C = d.C;
  const size_t nc = al(d.N * C) / sizeof(float);
  float *tmp_t = tmp, *tmp_s2 = tmp + nc, *tmp_s1 = tmp + 2 * nc, *ones = tmp + 3 * nc;
  if (d.tc && !g) return CNRF_E_NULL;
  if (level == 0) {
    if (d.tc) {     // unfold:  d Wdf = d Wfh Wout'h^T (+ d betaf^T betao' below),  d Wout'h = Wdf^T d Wfh,  d betao' = d betaf Wdf
      if (!bof) return CNRF_E_NULL;
      b.add_fixed(mk_mm(dt.dwfh, H, 1, p->weight[CNRF_L_FC_OUT] + (H + C), 1, H + C, g->weight[CNRF_L_DIR1], C + d.dd, H, C, H));
      b.add_fixed(mk_mm(p->weight[CNRF_L_DIR1], 1, C + d.dd, dt.dwfh, H, 1, g->weight[CNRF_L_FC_OUT] + (H + C), H + C, C, H, H));
      b.add(mk_dgrad(dt.bf, H, d.N, H, p->weight[CNRF_L_DIR1], C + d.dd, C, dt.bo + 1, C + 1, nullptr, 0, 0));
    }
  } else if (level == 1) {
    if (d.tc) b.add(mk_wgrad(dt.bf, H, bof, C, d.N, H, C, g->weight[CNRF_L_DIR1], C + d.dd));     // += onto the product above
    // texture branch: betar = Wrt at + br;   shape branch 2: betao = Woc as2 + bo;   shape branch 1: beta2 = W2c as1 + b2;
    // view-direction half of layer_dir1: betad = Wdv gamma(d) + bd1
    if (g) b.add(mk_wgrad(dt.br, 3, at, C, d.N, 3, C, g->weight[CNRF_L_FC_RGB] + H, H + C));
    b.add(mk_dgrad(dt.br, 3, d.N, 3, p->weight[CNRF_L_FC_RGB] + H, H + C, C, tmp_t, C, at, C, 0));
    if (g) b.add(mk_wgrad(dt.bo, C + 1, as2, C, d.N, C + 1, C, g->weight[CNRF_L_FC_OUT] + H, H + C));
    b.add(mk_dgrad(dt.bo, C + 1, d.N, C + 1, p->weight[CNRF_L_FC_OUT] + H, H + C, C, tmp_s2, C, as2, C, 0));
    if (g) b.add(mk_wgrad(dt.b2, H, as1, C, d.N, H, C, g->weight[CNRF_L_XYZ2] + H, H + C));
    b.add(mk_dgrad(dt.b2, H, d.N, H, p->weight[CNRF_L_XYZ2] + H, H + C, C, tmp_s1, C, as1, C, 0));
    if (d.rows_dev && g_x_dir) return CNRF_E_UNSUPPORTED;   // shared-code rays carry no pose gradient
    if (!d.tc && g) b.add(mk_wgrad(dt.bd, H, x_dir, ld_dir, d.N, H, d.dd, g->weight[CNRF_L_DIR1] + C, C + d.dd));
    if (g_x_dir) b.add(mk_dgrad(dt.bd, H, d.N, H, p->weight[CNRF_L_DIR1] + C, C + d.dd, d.dd, g_x_dir, d.dd, nullptr, 0, 0));
    if (d.tc) {
      b.add(mk_colsum(ones, dt.br, 3, d.N, 3, g->bias[CNRF_L_FC_RGB]));
      b.add(mk_colsum(ones, dt.bo, C + 1, d.N, C + 1, g->bias[CNRF_L_FC_OUT]));
      b.add(mk_colsum(ones, dt.b2, H, d.N, H, g->bias[CNRF_L_XYZ2]));
    }
  } else {
    // the three latent-code layers
    if (g) b.add(mk_wgrad(tmp_t, C, z_t, ld_zt, d.N, C, C, g->weight[CNRF_L_TEXTURE1], C));
    b.add(mk_dgrad(tmp_t, C, d.N, C, p->weight[CNRF_L_TEXTURE1], C, C, g_z_t, C, nullptr, 0, acc_codes ? 1 : 0));
    if (g) b.add(mk_wgrad(tmp_s2, C, z_s, ld_zs, d.N, C, C, g->weight[CNRF_L_SHAPE2], C));
    if (g) b.add(mk_wgrad(tmp_s1, C, z_s, ld_zs, d.N, C, C, g->weight[CNRF_L_SHAPE1], C));
    // both shape branches land in g_z_s: added atomically onto the zero-filled buffer when they share a launch
    b.add(mk_dgrad(tmp_s2, C, d.N, C, p->weight[CNRF_L_SHAPE2], C, C, g_z_s, C, nullptr, 0, d.tc ? 1 : 0));
    b.add(mk_dgrad(tmp_s1, C, d.N, C, p->weight[CNRF_L_SHAPE1], C, C, g_z_s, C, nullptr, 0, 1));
    if (d.tc) {
      b.add(mk_colsum(ones, tmp_t, C, d.N, C, g->bias[CNRF_L_TEXTURE1]));
      b.add(mk_colsum(ones, tmp_s2, C, d.N, C, g->bias[CNRF_L_SHAPE2]));
      b.add(mk_colsum(ones, tmp_s1, C, d.N, C, g->bias[CNRF_L_SHAPE1]));
    }
  }
  return CNRF_OK;
}
int per_ray_bwd_colsums(int level, const Dims& d, const cnrf_codenerf_grads* g, const Tables& dt, const float* tmp, cudaStream_t st) {
  const int H = d.H, C = d.C;
  if (!g) return CNRF_OK;
  const size_t nc = al(d.N * C) / sizeof(float);
  if (level == 1) {
    RUN(cnrf_colsum(dt.br, 3, d.N, 3, g->bias[CNRF_L_FC_RGB], st));
    RUN(cnrf_colsum(dt.bo, C + 1, d.N, C + 1, g->bias[CNRF_L_FC_OUT], st));
    RUN(cnrf_colsum(dt.b2, H, d.N, H, g->bias[CNRF_L_XYZ2], st));
    RUN(cnrf_colsum(dt.bd, H, d.N, H, g->bias[CNRF_L_DIR1], st));
  } else if (level == 2) {
    RUN(cnrf_colsum(tmp, C, d.N, C, g->bias[CNRF_L_TEXTURE1], st));
    RUN(cnrf_colsum(tmp + nc, C, d.N, C, g->bias[CNRF_L_SHAPE2], st));
    RUN(cnrf_colsum(tmp + 2 * nc, C, d.N, C, g->bias[CNRF_L_SHAPE1], st));
  }
  return CNRF_OK;
}
int per_ray_bwd(const Dims& d, const cnrf_codenerf_params* p, const cnrf_codenerf_grads* g, const float* x_dir,
                int64_t ld_dir, const float* z_s, int64_t ld_zs, const float* z_t, int64_t ld_zt, const float* as1,
                const float* as2, const float* at, const Tables& dt, float* tmp, float* g_z_s, float* g_z_t,
                float* g_x_dir, cudaStream_t st, const float* bof = nullptr) {
  if (d.tc) {       // bias gradients were zero-filled with the weights (zero_weight_grads)
    FillBatch f;
    f.add(g_z_s, d.N * d.C);
    f.add(per_ray_ones(d, tmp), d.N, 1.0f);
    RUN(cnrf_fill_batch(f, st));
  }
  Batch b(d.rows_dev);
  for (int level = 0; level < PER_RAY_BWD_LEVELS; ++level) {
    RUN(per_ray_bwd_add(level, d, p, g, x_dir, ld_dir, z_s, ld_zs, z_t, ld_zt, as1, as2, at, dt, tmp, g_z_s, g_z_t, g_x_dir, bof,
                        false, b));
    RUN(b.flush(d.tc, st));
    if (!d.tc) RUN(per_ray_bwd_colsums(level, d, g, dt, tmp, st));
  }
  return CNRF_OK;
}

// ---- per-sample chain, fp32 SIMT ------------------------------------------------------------------------------
int chain_fwd_f32(const Dims& d, const cnrf_codenerf_params* p, const float* x_xyz, int64_t ld_xyz, const Tables& t,
                  const SavedF32& s, float* raw, cudaStream_t st) {
  const int H = d.H, C = d.C;
  const int64_t INF = INT64_MAX;
  int64_t pdiv, pmod;
  pair_index(d, pdiv, pmod);
  RUN(lin(d.split, x_xyz, ld_xyz, d.M, d.dx, p->weight[CNRF_L_XYZ1], d.dx, H, s.h1, H, p->bias[CNRF_L_XYZ1], 0, 1, INF, 1, st)); // model.py:179
  RUN(lin(d.split, s.h1, H, d.M, H, p->weight[CNRF_L_XYZ2], H + C, H, s.h2, H, t.b2, H, d.S, INF, 1, st));                        // :180-181
  RUN(lin(d.split, s.h2, H, d.M, H, p->weight[CNRF_L_FC_OUT], H + C, C + 1, s.o, C + 1, t.bo, C + 1, d.S, INF, 0, st));           // :182-184
  RUN(lin(d.split, s.o + 1, C + 1, d.M, C, p->weight[CNRF_L_DIR1], C + d.dd, H, s.v1, H, t.bd, H, pdiv, pmod, 1, st));            // :186-189
  RUN(lin(d.split, s.v1, H, d.M, H, p->weight[CNRF_L_DIR2], H, H, s.v2, H, p->bias[CNRF_L_DIR2], 0, 1, INF, 1, st));              // :190
  RUN(lin(d.split, s.v2, H, d.M, H, p->weight[CNRF_L_FC_RGB], H + C, 3, raw, 4, t.br, 3, d.S, INF, 0, st));                       // :191-192
  RUN(copy_col(s.o, C + 1, raw + 3, 4, d.M, st));                                                                       // :194  [rgb | sigma]
  return CNRF_OK;
}

// ga, gb: [M, max(H, C+1)] scratch.  Fills the hidden halves of the weight grads, b1 / bd2 grads, the bias-table
// gradients dt, and (optionally) g_x_xyz.
int chain_bwd_f32(const Dims& d, const cnrf_codenerf_params* p, const cnrf_codenerf_grads* g, const float* x_xyz,
                  int64_t ld_xyz, const float* g_raw, const SavedF32& s, float* ga, float* gb, const Tables& dt,
                  float* g_x_xyz, cudaStream_t st) {
  const int H = d.H, C = d.C;
  // fc_rgb
  if (g) RUN(wgrad(d.split, g_raw, 4, s.v2, H, d.M, 3, H, g->weight[CNRF_L_FC_RGB], H + C, st));
  RUN(cnrf_strided_rowsum(g_raw, 4, d.N, d.S, d.S, 1, 3, dt.br, 3, st));
  RUN(dgrad(d.split, g_raw, 4, d.M, 3, p->weight[CNRF_L_FC_RGB], H + C, H, ga, H, s.v2, H, 0, st));
  // layer_dir2
  if (g) RUN(wgrad(d.split, ga, H, s.v1, H, d.M, H, H, g->weight[CNRF_L_DIR2], H, st));
  if (g) RUN(cnrf_colsum(ga, H, d.M, H, g->bias[CNRF_L_DIR2], st));
  RUN(dgrad(d.split, ga, H, d.M, H, p->weight[CNRF_L_DIR2], H, H, gb, H, s.v1, H, 0, st));
  // layer_dir1 (feature half)
  if (g) RUN(wgrad(d.split, gb, H, s.o + 1, C + 1, d.M, H, C, g->weight[CNRF_L_DIR1], C + d.dd, st));
  if (d.pairing == CNRF_PAIRING_REFERENCE) RUN(cnrf_strided_rowsum(gb, H, d.N, 1, d.S, d.N, H, dt.bd, H, st));
  else RUN(cnrf_strided_rowsum(gb, H, d.N, d.S, d.S, 1, H, dt.bd, H, st));
  RUN(dgrad(d.split, gb, H, d.M, H, p->weight[CNRF_L_DIR1], C + d.dd, C, ga + 1, C + 1, nullptr, 0, 0, st));
  RUN(copy_col(g_raw + 3, 4, ga, C + 1, d.M, st));                       // d sigma_raw -> column 0 of d o
  // fc_out
  if (g) RUN(wgrad(d.split, ga, C + 1, s.h2, H, d.M, C + 1, H, g->weight[CNRF_L_FC_OUT], H + C, st));
  RUN(cnrf_strided_rowsum(ga, C + 1, d.N, d.S, d.S, 1, C + 1, dt.bo, C + 1, st));
  RUN(dgrad(d.split, ga, C + 1, d.M, C + 1, p->weight[CNRF_L_FC_OUT], H + C, H, gb, H, s.h2, H, 0, st));
  // layer_xyz2
  if (g) RUN(wgrad(d.split, gb, H, s.h1, H, d.M, H, H, g->weight[CNRF_L_XYZ2], H + C, st));
  RUN(cnrf_strided_rowsum(gb, H, d.N, d.S, d.S, 1, H, dt.b2, H, st));
  RUN(dgrad(d.split, gb, H, d.M, H, p->weight[CNRF_L_XYZ2], H + C, H, ga, H, s.h1, H, 0, st));
  // layer_xyz1
  if (g) RUN(wgrad(d.split, ga, H, x_xyz, ld_xyz, d.M, H, d.dx, g->weight[CNRF_L_XYZ1], d.dx, st));
  if (g) RUN(cnrf_colsum(ga, H, d.M, H, g->bias[CNRF_L_XYZ1], st));
  if (g_x_xyz) RUN(dgrad(d.split, ga, H, d.M, H, p->weight[CNRF_L_XYZ1], d.dx, d.dx, g_x_xyz, d.dx, nullptr, 0, 0, st));
  return CNRF_OK;
}

int add_weight_grad_fills(const Dims& d, const cnrf_codenerf_grads* g, FillBatch& f) {
  const int H = d.H, C = d.C;
  const size_t n[CNRF_NUM_LAYERS] = {(size_t)H * d.dx, (size_t)H * (H + C), (size_t)(C + 1) * (H + C), (size_t)C * C,
                                     (size_t)C * C, (size_t)C * C, (size_t)H * (C + d.dd), (size_t)H * H,
                                     (size_t)3 * (H + C)};
  const size_t nb[CNRF_NUM_LAYERS] = {(size_t)H, (size_t)H, (size_t)C + 1, (size_t)C, (size_t)C, (size_t)C, (size_t)H,
                                      (size_t)H, 3};
  for (int l = 0; l < CNRF_NUM_LAYERS; ++l) {
    if (!g->weight[l] || !g->bias[l]) return CNRF_E_NULL;
    f.add(g->weight[l], (int64_t)n[l]);
    f.add(g->bias[l], (int64_t)nb[l]);
  }
  return CNRF_OK;
}
int zero_weight_grads(const Dims& d, const cnrf_codenerf_grads* g, cudaStream_t st) {
  FillBatch f;
  RUN(add_weight_grad_fills(d, g, f));
  return cnrf_fill_batch(f, st);
}

inline size_t max_sz(size_t a, size_t b) { return a > b ? a : b; }

}  // namespace

// tensor-core chain (mlp_tc.cu)
size_t cnrf_tc_saved_bytes(int64_t M);
size_t cnrf_tc_workspace_bytes(int64_t N, int64_t M, int backward);

extern "C" size_t cnrf_codenerf_saved_bytes(const cnrf_field_shape* s) {
  if (check_shape(s) != CNRF_OK || !s->keep) return 0;
  Dims d = dims_of(s);
  if (fp32_layout(s)) return saved_f32_bytes(d);
  // (+ shared-code mode: the code runs and one code row per run, found once in the forward and reused by the backward)
  return 4 * al(d.N * d.C) + cnrf_code_runs_bytes(d.N) + 256 + 2 * al(d.N * d.C) + cnrf_tc_saved_bytes(d.M) + 2048;
}

extern "C" size_t cnrf_codenerf_workspace_bytes(const cnrf_field_shape* s, int32_t backward) {
  if (check_shape(s) != CNRF_OK) return 0;
  Dims d = dims_of(s);
  size_t b = tables_bytes(d);
  if (fp32_layout(s)) {
    if (!backward) return b + (s->keep ? 0 : saved_f32_bytes(d));     // inference still needs the activation buffers
    return b + 2 * al(d.M * max_sz(d.H, d.C + 1)) + 3 * al(d.N * d.C) + al(d.N);
  }
  return b + 3 * al(d.N * d.C) + al(d.N) + (s->keep || backward ? 0 : 4 * al(d.N * d.C)) + cnrf_tc_workspace_bytes(d.N, d.M, backward) +
         cnrf_code_runs_bytes(d.N) + 256 + 4 * al(d.N * d.C);      // shared-code mode: runs, one code row and gradient row per run
}

// tensor-core per-sample chain (mlp_tc.cu).  x_xyz == nullptr means "encode pts in-kernel" (field mode).
struct TcChainArgs {
  int64_t N, M; int S, pairing, dx, dd;
  const cnrf_codenerf_params* p;
  const float* x_xyz; int64_t ld_xyz;         // pre-embedded input, or
  const float* pts; const float* freqs_xyz; int L_xyz; int inc_xyz;   // points to encode
  const float *b2, *bf, *bos, *br;            // per-ray bias tables [N,256], [N,256] (folded layer_dir1), [N] (sigma), [N,3]
  const float* wf;                            // the four K-slices [4][256 x 256] of this call's fold Wfh = Wdf Wout'h
  const float* x_dir; int64_t ld_dir;         // gamma(d^) per ray [N, dd]
  const int32_t* gid;                         // shared-code mode: row of the bias tables that ray r uses (NULL: r)
  uint8_t* packed = nullptr;                  // != NULL: the weight images were packed there already (cnrf_tc_pack_pair)
  bool prezeroed = false;                     // backward: the caller zero-filled every accumulation target
};
uint8_t* cnrf_tc_packed_in_saved(void* saved, int64_t M);
size_t cnrf_tc_packed_bytes();
int cnrf_tc_pack_pair(const cnrf_codenerf_params* p0, const float* wf0, uint8_t* at0, const cnrf_codenerf_params* p1,
                      const float* wf1, uint8_t* at1, int dx, int dd, cudaStream_t st);
int cnrf_tc_chain_fwd(const TcChainArgs& a, float* raw, void* saved, size_t saved_bytes, void* ws, size_t ws_bytes,
                      cudaStream_t st);
int cnrf_tc_chain_bwd(const TcChainArgs& a, const float* g_raw, const void* saved, size_t saved_bytes,
                      const cnrf_codenerf_grads* g, float* d_b2, float* d_bo, float* d_br, float* d_bd, float* d_bf,
                      float* dwfh, float* g_x_xyz /*[M,dx] or NULL*/, void* ws, size_t ws_bytes, cudaStream_t st);

namespace {

struct FieldIn {   // everything the two entry-point families share
  const float* x_xyz; int64_t ld_xyz; const float* x_dir; int64_t ld_dir;
  const float* z_s; int64_t ld_zs; const float* z_t; int64_t ld_zt;
  const float* pts; const float* freqs_xyz; int L_xyz; int inc_xyz;     // only for the fused bf16 field path
  // shared-code mode (bf16 only): z_s / z_t are TABLES [num_codes, C] indexed by ids[ray]; g_z_s / g_z_t of the backward
  // are table-shaped dense gradients
  const int64_t* ids = nullptr; int64_t num_codes = 0;
};

// shared-code mode: find the runs of equal ids and gather one code row per run
struct SharedCodes { CodeRuns runs; float *zc_s, *zc_t; };
// `compute` = false: the arena already holds them (the backward, over the forward's saved buffer)
int setup_shared_codes(Dims& d, const FieldIn& in, Arena& w, SharedCodes& sc, bool compute, cudaStream_t st) {
  size_t rb = cnrf_code_runs_bytes(d.N);
  float* rbase = w.take(rb / sizeof(float));
  sc.zc_s = w.take(d.N * d.C); sc.zc_t = w.take(d.N * d.C);
  if (!w.ok) return CNRF_E_WORKSPACE;
  sc.runs = cnrf_code_runs_carve(rbase, d.N);
  if (compute) {
    RUN(cnrf_code_runs(in.ids, d.N, in.num_codes, sc.runs, st));
    RUN(cnrf_gather_codes(in.ids, sc.runs, d.N, d.C, in.z_s, in.ld_zs, sc.zc_s, st));
    RUN(cnrf_gather_codes(in.ids, sc.runs, d.N, d.C, in.z_t, in.ld_zt, sc.zc_t, st));
  }
  d.rows_dev = sc.runs.count;
  return CNRF_OK;
}

int codenerf_fwd_impl(const cnrf_field_shape* s, const cnrf_codenerf_params* p, const FieldIn& in, float* raw,
                      void* saved, size_t saved_bytes, void* ws, size_t ws_bytes, cudaStream_t st) {
  Dims d = dims_of(s);
  Arena w(ws, ws_bytes);
  Tables t;
  carve_tables(d, w, t);
  if (!w.ok) return CNRF_E_WORKSPACE;
  if (in.ids && s->precision != CNRF_PRECISION_BF16) return CNRF_E_UNSUPPORTED;
  if (fp32_layout(s)) {
    SavedF32 sv;
    if (s->keep) {
      if (!saved || !carve_saved(d, saved, saved_bytes, sv)) return CNRF_E_WORKSPACE;
    } else {
      size_t need = saved_f32_bytes(d);
      if (w.off + need > w.cap) return CNRF_E_WORKSPACE;
      carve_saved(d, w.base + w.off, need, sv);
    }
    RUN(per_ray_fwd(d, p, in.x_dir, in.ld_dir, in.z_s, in.ld_zs, in.z_t, in.ld_zt, sv.as1, sv.as2, sv.at, t, false, st));
    return chain_fwd_f32(d, p, in.x_xyz, in.ld_xyz, t, sv, raw, st);
  }
  // bf16 tensor-core chain: needs H = C = 256
  if (d.H != 256 || d.C != 256 || d.dx > 63 || d.dd > 31) return CNRF_E_UNSUPPORTED;
  float *as1, *as2, *at;
  char* tc_saved = nullptr; size_t tc_saved_bytes = 0;
  if (s->keep && !saved) return CNRF_E_WORKSPACE;
  Arena sa(s->keep ? saved : nullptr, s->keep ? saved_bytes : 0);
  Arena& act = s->keep ? sa : w;            // what the backward needs again lives in the saved buffer
  as1 = act.take(d.N * d.C); as2 = act.take(d.N * d.C); at = act.take(d.N * d.C);
  float* bof = act.take(d.N * d.C);         // betao' of the folded fc_out: only the backward reads it
  if (!act.ok) return CNRF_E_WORKSPACE;
  const float *zs = in.z_s, *zt = in.z_t;
  int64_t ld_zs = in.ld_zs, ld_zt = in.ld_zt;
  SharedCodes sc;
  if (in.ids) {
    RUN(setup_shared_codes(d, in, act, sc, true, st));
    zs = sc.zc_s; zt = sc.zc_t; ld_zs = ld_zt = d.C;
  }
  if (s->keep) { tc_saved = sa.base + sa.off; tc_saved_bytes = sa.cap - sa.off; }
  RUN(per_ray_fwd(d, p, in.x_dir, in.ld_dir, zs, ld_zs, zt, ld_zt, as1, as2, at, t, true, st, bof));
  TcChainArgs a{d.N, d.M, d.S, d.pairing, d.dx, d.dd, p, in.x_xyz, in.ld_xyz, in.pts, in.freqs_xyz, in.L_xyz,
                in.inc_xyz, t.b2, t.bf, t.bos, t.br, t.wf, in.x_dir, in.ld_dir, in.ids ? sc.runs.gid : nullptr};
  return cnrf_tc_chain_fwd(a, raw, tc_saved, tc_saved_bytes, w.base + w.off, w.cap - w.off, st);
}

// accumulate_tables (shared-code mode only): g_z_s / g_z_t already hold another model's table gradients -- add to them
int codenerf_bwd_impl(const cnrf_field_shape* s, const cnrf_codenerf_params* p, const FieldIn& in, const float* g_raw,
                      const void* saved, size_t saved_bytes, const cnrf_codenerf_grads* g, float* g_z_s, float* g_z_t,
                      float* g_x_xyz, float* g_x_dir, void* ws, size_t ws_bytes, cudaStream_t st, bool accumulate_tables = false) {
  Dims d = dims_of(s);
  if (!s->keep || !saved) return CNRF_E_WORKSPACE;
  Arena w(ws, ws_bytes);
  Tables dt;
  carve_tables(d, w, dt);
  float* tmp = w.take(d.N * d.C);          // 3 x [N,C] + [N], contiguous (per_ray_bwd)
  w.take(d.N * d.C); w.take(d.N * d.C); w.take(d.N);
  if (!w.ok) return CNRF_E_WORKSPACE;
  if (in.ids && s->precision != CNRF_PRECISION_BF16) return CNRF_E_UNSUPPORTED;
  if (!g && !fp32_layout(s)) return CNRF_E_NULL;       // only the fp32-layout chains skip the weight gradients of a frozen network
  if (g) RUN(zero_weight_grads(d, g, st));
  if (fp32_layout(s)) {
    SavedF32 sv;
    if (!carve_saved(d, const_cast<void*>(saved), saved_bytes, sv)) return CNRF_E_WORKSPACE;
    size_t wide = max_sz(d.H, d.C + 1);
    float* ga = w.take(d.M * wide);
    float* gb = w.take(d.M * wide);
    if (!w.ok) return CNRF_E_WORKSPACE;
    RUN(chain_bwd_f32(d, p, g, in.x_xyz, in.ld_xyz, g_raw, sv, ga, gb, dt, g_x_xyz, st));
    return per_ray_bwd(d, p, g, in.x_dir, in.ld_dir, in.z_s, in.ld_zs, in.z_t, in.ld_zt, sv.as1, sv.as2, sv.at, dt,
                       tmp, g_z_s, g_z_t, g_x_dir, st);
  }
  if (d.H != 256 || d.C != 256 || d.dx > 63 || d.dd > 31) return CNRF_E_UNSUPPORTED;
  Arena sa(const_cast<void*>(saved), saved_bytes);
  float* as1 = sa.take(d.N * d.C); float* as2 = sa.take(d.N * d.C); float* at = sa.take(d.N * d.C);
  const float* bof = sa.take(d.N * d.C);
  if (!sa.ok) return CNRF_E_WORKSPACE;
  const float *zs = in.z_s, *zt = in.z_t;
  int64_t ld_zs = in.ld_zs, ld_zt = in.ld_zt;
  float *gz_s = g_z_s, *gz_t = g_z_t;
  SharedCodes sc;
  if (in.ids) {          // runs and per-run code rows: the forward left them in the saved buffer
    RUN(setup_shared_codes(d, in, sa, sc, false, st));
    zs = sc.zc_s; zt = sc.zc_t; ld_zs = ld_zt = d.C;
    gz_s = w.take(d.N * d.C); gz_t = w.take(d.N * d.C);
    if (!w.ok) return CNRF_E_WORKSPACE;
  }
  TcChainArgs a{d.N, d.M, d.S, d.pairing, d.dx, d.dd, p, in.x_xyz, in.ld_xyz, in.pts, in.freqs_xyz, in.L_xyz,
                in.inc_xyz, nullptr, nullptr, nullptr, nullptr, nullptr, in.x_dir, in.ld_dir,
                in.ids ? sc.runs.gid : nullptr};
  RUN(cnrf_tc_chain_bwd(a, g_raw, sa.base + sa.off, sa.cap - sa.off, g, dt.b2, dt.bo, dt.br, g_x_dir ? dt.bd : nullptr, dt.bf,
                        dt.dwfh, g_x_xyz, w.base + w.off, w.cap - w.off, st));
  RUN(per_ray_bwd(d, p, g, in.x_dir, in.ld_dir, zs, ld_zs, zt, ld_zt, as1, as2, at, dt, tmp, gz_s, gz_t, g_x_dir, st, bof));
  if (in.ids) {          // one row per run -> dense table gradients (what torch.nn.Embedding's backward produces)
    if (!accumulate_tables) {
      FillBatch f;
      f.add(g_z_s, in.num_codes * d.C);
      f.add(g_z_t, in.num_codes * d.C);
      RUN(cnrf_fill_batch(f, st));
    }
    RUN(cnrf_scatter_code_grads(in.ids, sc.runs, d.N, d.C, gz_s, g_z_s, d.C, st));
    RUN(cnrf_scatter_code_grads(in.ids, sc.runs, d.N, d.C, gz_t, g_z_t, d.C, st));
  }
  return CNRF_OK;
}

int check_common(const cnrf_field_shape* s, const cnrf_codenerf_params* p) {
  int rc = check_shape(s);
  if (rc) return rc;
  if (!p) return CNRF_E_NULL;
  for (int l = 0; l < CNRF_NUM_LAYERS; ++l)
    if (!p->weight[l] || !p->bias[l]) return CNRF_E_NULL;
  return CNRF_OK;
}

}  // namespace

extern "C" int cnrf_codenerf_fwd(const cnrf_field_shape* s, const cnrf_codenerf_params* p, const float* x_xyz,
                                 int64_t ld_xyz, const float* x_dir, int64_t ld_dir, const float* z_s, int64_t ld_zs,
                                 const float* z_t, int64_t ld_zt, float* raw, void* saved, size_t saved_bytes,
                                 void* workspace, size_t workspace_bytes, cnrf_stream_t stream) {
  RUN(check_common(s, p));
  CNRF_REQUIRE(x_xyz && x_dir && z_s && z_t && raw && workspace, CNRF_E_NULL);
  FieldIn in{x_xyz, ld_xyz, x_dir, ld_dir, z_s, ld_zs, z_t, ld_zt, nullptr, nullptr, 0, 0};
  return codenerf_fwd_impl(s, p, in, raw, saved, saved_bytes, workspace, workspace_bytes, as_stream(stream));
}

extern "C" int cnrf_codenerf_bwd(const cnrf_field_shape* s, const cnrf_codenerf_params* p, const float* x_xyz,
                                 int64_t ld_xyz, const float* x_dir, int64_t ld_dir, const float* z_s, int64_t ld_zs,
                                 const float* z_t, int64_t ld_zt, const float* g_raw, const void* saved,
                                 size_t saved_bytes, const cnrf_codenerf_grads* g, float* g_z_s, float* g_z_t,
                                 float* g_x_xyz, float* g_x_dir, void* workspace, size_t workspace_bytes,
                                 cnrf_stream_t stream) {
  RUN(check_common(s, p));
  CNRF_REQUIRE(x_xyz && x_dir && z_s && z_t && g_raw && g_z_s && g_z_t && workspace, CNRF_E_NULL);
  FieldIn in{x_xyz, ld_xyz, x_dir, ld_dir, z_s, ld_zs, z_t, ld_zt, nullptr, nullptr, 0, 0};
  return codenerf_bwd_impl(s, p, in, g_raw, saved, saved_bytes, g, g_z_s, g_z_t, g_x_xyz, g_x_dir, workspace,
                           workspace_bytes, as_stream(stream));
}

// ---- forward_pass: encode + pair + field (view_synthesis/nerf/__init__.py:116-134) ------------------------------
namespace {

// viewdirs = rd / ||rd||   (nerf/__init__.py:126)
__global__ void normalize_dirs_kernel(const float* __restrict__ rd, int64_t N, float* __restrict__ out) {
  int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  float x = rd[n * 3], y = rd[n * 3 + 1], z = rd[n * 3 + 2];
  float nr = sqrtf(x * x + y * y + z * z);
  out[n * 3] = __fdiv_rn(x, nr); out[n * 3 + 1] = __fdiv_rn(y, nr); out[n * 3 + 2] = __fdiv_rn(z, nr);
}
// g_rd = (g - d^ (d^ . g)) / ||rd||
__global__ void normalize_dirs_bwd_kernel(const float* __restrict__ rd, const float* __restrict__ g, int64_t N,
                                          float* __restrict__ g_rd) {
  int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  float x = rd[n * 3], y = rd[n * 3 + 1], z = rd[n * 3 + 2];
  float nr = sqrtf(x * x + y * y + z * z), inv = 1.0f / nr;
  float ux = x * inv, uy = y * inv, uz = z * inv;
  float gx = g[n * 3], gy = g[n * 3 + 1], gz = g[n * 3 + 2];
  float dot = ux * gx + uy * gy + uz * gz;
  g_rd[n * 3] = (gx - ux * dot) * inv; g_rd[n * 3 + 1] = (gy - uy * dot) * inv; g_rd[n * 3 + 2] = (gz - uz * dot) * inv;
}

struct FieldScratch { float *dirn, *demb, *xemb, *g_dirn, *g_demb, *g_xemb; };
inline size_t field_scratch_bytes(const Dims& d, int precision, int backward) {
  size_t b = al(d.N * 3) + al(d.N * d.dd);
  if (precision == CNRF_PRECISION_FP32) b += al(d.M * d.dx);
  if (backward) b += al(d.N * 3) + al(d.N * d.dd) + al(d.M * d.dx);
  return b;
}
inline void carve_field(const Dims& d, int precision, int backward, Arena& a, FieldScratch& f) {
  f.dirn = a.take(d.N * 3); f.demb = a.take(d.N * d.dd);
  f.xemb = precision == CNRF_PRECISION_FP32 ? a.take(d.M * d.dx) : nullptr;
  f.g_dirn = f.g_demb = f.g_xemb = nullptr;
  if (backward) { f.g_dirn = a.take(d.N * 3); f.g_demb = a.take(d.N * d.dd); f.g_xemb = a.take(d.M * d.dx); }
}

int field_encode(const Dims& d, int precision, const float* pts, const float* rd, const float* fx, int Lx, int incx,
                 const float* fd, int Ld, int incd, const FieldScratch& f, cudaStream_t st) {
  normalize_dirs_kernel<<<grid_for(d.N, 256), 256, 0, st>>>(rd, d.N, f.dirn);
  CNRF_LAUNCH_CHECK();
  RUN(cnrf_pos_embed_fwd_ld(f.dirn, d.N, 3, fd, Ld, incd, d.dd, f.demb, st));
  if (precision == CNRF_PRECISION_FP32) RUN(cnrf_pos_embed_fwd_ld(pts, d.M, 3, fx, Lx, incx, d.dx, f.xemb, st));
  return CNRF_OK;
}

int check_field(const cnrf_field_shape* s, int Lx, int Ld, int& incx, int& incd) {
  if (Lx <= 0 || Ld <= 0) return CNRF_E_SHAPE;
  incx = s->dim_xyz - 6 * Lx; incd = s->dim_dir - 6 * Ld;        // dim = 3*(include_input + 2L)
  if ((incx != 0 && incx != 3) || (incd != 0 && incd != 3)) return CNRF_E_SHAPE;
  incx /= 3; incd /= 3;
  return CNRF_OK;
}

}  // namespace

extern "C" size_t cnrf_field_workspace_bytes(const cnrf_field_shape* s, int32_t backward) {
  if (check_shape(s) != CNRF_OK) return 0;
  return field_scratch_bytes(dims_of(s), layout_of(s), backward) + cnrf_codenerf_workspace_bytes(s, backward);
}

int cnrf_field_fwd_impl(const cnrf_field_shape* s, const cnrf_codenerf_params* p, const float* pts, const float* rd,
                   const float* freqs_xyz, int32_t L_xyz, const float* freqs_dir, int32_t L_dir, const int64_t* ids,
                   int64_t num_codes, const float* z_s, int64_t ld_zs, const float* z_t, int64_t ld_zt, float* raw,
                   void* saved, size_t saved_bytes, void* workspace, size_t workspace_bytes, cudaStream_t st) {
  RUN(check_common(s, p));
  CNRF_REQUIRE(pts && rd && freqs_xyz && freqs_dir && z_s && z_t && raw && workspace, CNRF_E_NULL);
  int incx, incd;
  RUN(check_field(s, L_xyz, L_dir, incx, incd));
  Dims d = dims_of(s);
  Arena w(workspace, workspace_bytes);
  FieldScratch f;
  carve_field(d, layout_of(s), 0, w, f);
  if (!w.ok) return CNRF_E_WORKSPACE;
  RUN(field_encode(d, layout_of(s), pts, rd, freqs_xyz, L_xyz, incx, freqs_dir, L_dir, incd, f, st));
  FieldIn in{f.xemb, d.dx, f.demb, d.dd, z_s, ld_zs, z_t, ld_zt, pts, freqs_xyz, L_xyz, incx, ids, num_codes};
  return codenerf_fwd_impl(s, p, in, raw, saved, saved_bytes, w.base + w.off, w.cap - w.off, st);
}

int cnrf_field_bwd_impl(const cnrf_field_shape* s, const cnrf_codenerf_params* p, const float* pts, const float* rd,
                   const float* freqs_xyz, int32_t L_xyz, const float* freqs_dir, int32_t L_dir, const int64_t* ids,
                   int64_t num_codes, const float* z_s, int64_t ld_zs, const float* z_t, int64_t ld_zt,
                   const float* g_raw, const void* saved, size_t saved_bytes, const cnrf_codenerf_grads* g, float* g_z_s,
                   float* g_z_t, float* g_pts, float* g_rd, void* workspace, size_t workspace_bytes, cudaStream_t st,
                   bool accumulate_tables) {
  RUN(check_common(s, p));
  // the bf16 chain saved its own encoded inputs: the points are only read again in the fp32 mode or for point gradients
  CNRF_REQUIRE((pts || (s->precision == CNRF_PRECISION_BF16 && !g_pts)) && rd && freqs_xyz && freqs_dir && z_s && z_t && g_raw &&
               g_z_s && g_z_t && workspace, CNRF_E_NULL);
  int incx, incd;
  RUN(check_field(s, L_xyz, L_dir, incx, incd));
  Dims d = dims_of(s);
  Arena w(workspace, workspace_bytes);
  FieldScratch f;
  carve_field(d, layout_of(s), 1, w, f);
  if (!w.ok) return CNRF_E_WORKSPACE;
  RUN(field_encode(d, layout_of(s), pts, rd, freqs_xyz, L_xyz, incx, freqs_dir, L_dir, incd, f, st));
  FieldIn in{f.xemb, d.dx, f.demb, d.dd, z_s, ld_zs, z_t, ld_zt, pts, freqs_xyz, L_xyz, incx, ids, num_codes};
  RUN(codenerf_bwd_impl(s, p, in, g_raw, saved, saved_bytes, g, g_z_s, g_z_t, g_pts ? f.g_xemb : nullptr,
                        g_rd ? f.g_demb : nullptr, w.base + w.off, w.cap - w.off, st, accumulate_tables));
  if (g_pts) RUN(cnrf_pos_embed_bwd_ld(pts, f.g_xemb, d.M, 3, freqs_xyz, L_xyz, incx, d.dx, g_pts, st));
  if (g_rd) {
    RUN(cnrf_pos_embed_bwd_ld(f.dirn, f.g_demb, d.N, 3, freqs_dir, L_dir, incd, d.dd, f.g_dirn, st));
    normalize_dirs_bwd_kernel<<<grid_for(d.N, 256), 256, 0, st>>>(rd, f.g_dirn, d.N, g_rd);
    CNRF_LAUNCH_CHECK();
  }
  return CNRF_OK;
}

extern "C" int cnrf_field_fwd(const cnrf_field_shape* s, const cnrf_codenerf_params* p, const float* pts,
                              const float* rd, const float* freqs_xyz, int32_t L_xyz, const float* freqs_dir,
                              int32_t L_dir, const float* z_s, int64_t ld_zs, const float* z_t, int64_t ld_zt,
                              float* raw, void* saved, size_t saved_bytes, void* workspace, size_t workspace_bytes,
                              cnrf_stream_t stream) {
  return cnrf_field_fwd_impl(s, p, pts, rd, freqs_xyz, L_xyz, freqs_dir, L_dir, nullptr, 0, z_s, ld_zs, z_t, ld_zt, raw, saved,
                        saved_bytes, workspace, workspace_bytes, as_stream(stream));
}

extern "C" int cnrf_field_bwd(const cnrf_field_shape* s, const cnrf_codenerf_params* p, const float* pts,
                              const float* rd, const float* freqs_xyz, int32_t L_xyz, const float* freqs_dir,
                              int32_t L_dir, const float* z_s, int64_t ld_zs, const float* z_t, int64_t ld_zt,
                              const float* g_raw, const void* saved, size_t saved_bytes, const cnrf_codenerf_grads* g,
                              float* g_z_s, float* g_z_t, float* g_pts, float* g_rd, void* workspace,
                              size_t workspace_bytes, cnrf_stream_t stream) {
  return cnrf_field_bwd_impl(s, p, pts, rd, freqs_xyz, L_xyz, freqs_dir, L_dir, nullptr, 0, z_s, ld_zs, z_t, ld_zt, g_raw, saved,
                             saved_bytes, g, g_z_s, g_z_t, g_pts, g_rd, workspace, workspace_bytes, as_stream(stream), false);
}

// Shared-code variants: the embedding lookup of train.py:96 (ShapeTextureEmbedding.forward, models/model.py:100-103)
// fused into the field.  ids [N] int64 on the device; tab_s / tab_t [num_codes, C] are the two embedding tables.
extern "C" int cnrf_field_fwd_ids(const cnrf_field_shape* s, const cnrf_codenerf_params* p, const float* pts,
                                  const float* rd, const float* freqs_xyz, int32_t L_xyz, const float* freqs_dir,
                                  int32_t L_dir, const int64_t* ids, int64_t num_codes, const float* tab_s, int64_t ld_s,
                                  const float* tab_t, int64_t ld_t, float* raw, void* saved, size_t saved_bytes,
                                  void* workspace, size_t workspace_bytes, cnrf_stream_t stream) {
  CNRF_REQUIRE(ids, CNRF_E_NULL);
  CNRF_REQUIRE(num_codes > 0, CNRF_E_SHAPE);
  return cnrf_field_fwd_impl(s, p, pts, rd, freqs_xyz, L_xyz, freqs_dir, L_dir, ids, num_codes, tab_s, ld_s, tab_t, ld_t, raw,
                        saved, saved_bytes, workspace, workspace_bytes, as_stream(stream));
}

extern "C" int cnrf_field_bwd_ids(const cnrf_field_shape* s, const cnrf_codenerf_params* p, const float* pts,
                                  const float* rd, const float* freqs_xyz, int32_t L_xyz, const float* freqs_dir,
                                  int32_t L_dir, const int64_t* ids, int64_t num_codes, const float* tab_s, int64_t ld_s,
                                  const float* tab_t, int64_t ld_t, const float* g_raw, const void* saved,
                                  size_t saved_bytes, const cnrf_codenerf_grads* g, float* g_tab_s, float* g_tab_t,
                                  void* workspace, size_t workspace_bytes, cnrf_stream_t stream) {
  CNRF_REQUIRE(ids, CNRF_E_NULL);
  CNRF_REQUIRE(num_codes > 0, CNRF_E_SHAPE);
  return cnrf_field_bwd_impl(s, p, pts, rd, freqs_xyz, L_xyz, freqs_dir, L_dir, ids, num_codes, tab_s, ld_s, tab_t, ld_t, g_raw,
                             saved, saved_bytes, g, g_tab_s, g_tab_t, nullptr, nullptr, workspace, workspace_bytes,
                             as_stream(stream), false);
}


// ---- both models of one render call (tensor-core mode): see internal.h --------------------------------------------------
namespace {

struct PairModel { float *as1, *as2, *at, *bof; char* tc_saved; size_t tc_saved_bytes; };
struct PairCommon { CodeRuns runs; float *zc_s, *zc_t; };

// the layout codenerf_fwd_impl gives a model's saved buffer (tensor-core mode): as1, as2, at, betao', [runs, one code row
// per run x 2], then the chain's own saved images
bool carve_pair_acts(const Dims& d, bool ids, Arena& a, PairModel& m, PairCommon* c) {
  m.as1 = a.take(d.N * d.C); m.as2 = a.take(d.N * d.C); m.at = a.take(d.N * d.C); m.bof = a.take(d.N * d.C);
  if (ids) {
    float* rbase = a.take(cnrf_code_runs_bytes(d.N) / sizeof(float));
    float* zs = a.take(d.N * d.C); float* zt = a.take(d.N * d.C);
    if (a.ok && c) { c->runs = cnrf_code_runs_carve(rbase, d.N); c->zc_s = zs; c->zc_t = zt; }
  }
  m.tc_saved = a.base + a.off; m.tc_saved_bytes = a.cap - a.off;
  return a.ok;
}

struct PairFwd {           // everything cnrf_pair_fwd_* derive from (in, side)
  Dims d[2]; Tables t[2]; PairModel m[2]; PairCommon c; float *dirn, *demb; uint8_t* packed[2]; bool ids, keep;
  bool bcast;      // every ray carries the same code row (stride-0 expand, eval.py:149): treated as ONE run, no ids needed
  bool shared() const { return ids || bcast; }
};
inline size_t up1k_sz(size_t b) { return (b + 1023) & ~size_t(1023); }
size_t pair_fwd_side_bytes(const Dims& d, bool keep) {
  size_t b = al(d.N * 3) + al(d.N * d.dd) + 2 * tables_bytes(d);
  if (!keep) b += 2 * 4 * al(d.N * d.C) + cnrf_code_runs_bytes(d.N) + 256 + 2 * al(d.N * d.C) + 2 * up1k_sz(cnrf_tc_packed_bytes()) + 2048;
  return b + 1024;
}
int carve_pair_fwd(const CnrfPairIn& in, void* side, size_t side_bytes, PairFwd& f) {
  f.ids = in.ids != nullptr; f.keep = in.shape[0]->keep != 0;
  f.bcast = !f.ids && in.ld_zs == 0 && in.ld_zt == 0;
  for (int i = 0; i < 2; ++i) f.d[i] = dims_of(in.shape[i]);
  Arena a(side, side_bytes);
  f.dirn = a.take(f.d[0].N * 3); f.demb = a.take(f.d[0].N * f.d[0].dd);
  for (int i = 0; i < 2; ++i) carve_tables(f.d[i], a, f.t[i]);
  if (!a.ok) return CNRF_E_WORKSPACE;
  if (f.keep) {
    for (int i = 0; i < 2; ++i) {
      if (!in.saved[i]) return CNRF_E_WORKSPACE;
      Arena sa(in.saved[i], in.saved_bytes[i]);
      if (!carve_pair_acts(f.d[i], f.shared(), sa, f.m[i], i == 0 ? &f.c : nullptr)) return CNRF_E_WORKSPACE;   // the runs live with the coarse model
      f.packed[i] = cnrf_tc_packed_in_saved(f.m[i].tc_saved, f.d[i].M);
    }
  } else {
    for (int i = 0; i < 2; ++i) {
      f.m[i].as1 = a.take(f.d[i].N * f.d[i].C); f.m[i].as2 = a.take(f.d[i].N * f.d[i].C);
      f.m[i].at = a.take(f.d[i].N * f.d[i].C); f.m[i].bof = a.take(f.d[i].N * f.d[i].C);
      f.m[i].tc_saved = nullptr; f.m[i].tc_saved_bytes = 0;
    }
    if (f.shared()) {
      float* rbase = a.take(cnrf_code_runs_bytes(f.d[0].N) / sizeof(float));
      f.c.zc_s = a.take(f.d[0].N * f.d[0].C); f.c.zc_t = a.take(f.d[0].N * f.d[0].C);
      if (!a.ok) return CNRF_E_WORKSPACE;
      f.c.runs = cnrf_code_runs_carve(rbase, f.d[0].N);
    }
    for (int i = 0; i < 2; ++i) {
      const size_t at = up1k_sz(reinterpret_cast<size_t>(a.base + a.off));
      f.packed[i] = reinterpret_cast<uint8_t*>(at);
      a.off = (at - reinterpret_cast<size_t>(a.base)) + up1k_sz(cnrf_tc_packed_bytes());
      if (a.off > a.cap) return CNRF_E_WORKSPACE;
    }
  }
  if (!a.ok) return CNRF_E_WORKSPACE;
  if (f.shared()) f.d[0].rows_dev = f.d[1].rows_dev = f.c.runs.count;
  return CNRF_OK;
}

struct PairBwd { Dims d[2]; Tables dt[2]; float* tmp[2]; PairModel m[2]; PairCommon c; float *gz_s, *gz_t; bool ids, bcast;
                 bool shared() const { return ids || bcast; } };
size_t pair_bwd_side_bytes(const Dims& d) {
  return 2 * (tables_bytes(d) + 3 * al(d.N * d.C) + al(d.N)) + 2 * al(d.N * d.C) + 1024;
}
int carve_pair_bwd(const CnrfPairIn& in, void* side, size_t side_bytes, PairBwd& f) {
  f.ids = in.ids != nullptr;
  f.bcast = !f.ids && in.ld_zs == 0 && in.ld_zt == 0;
  Arena a(side, side_bytes);
  for (int i = 0; i < 2; ++i) {
    f.d[i] = dims_of(in.shape[i]);
    carve_tables(f.d[i], a, f.dt[i]);
    f.tmp[i] = a.take(f.d[i].N * f.d[i].C); a.take(f.d[i].N * f.d[i].C); a.take(f.d[i].N * f.d[i].C); a.take(f.d[i].N);
    if (!in.saved[i]) return CNRF_E_WORKSPACE;
    Arena sa(in.saved[i], in.saved_bytes[i]);
    if (!carve_pair_acts(f.d[i], f.shared(), sa, f.m[i], i == 0 ? &f.c : nullptr)) return CNRF_E_WORKSPACE;
  }
  f.gz_s = a.take(f.d[0].N * f.d[0].C); f.gz_t = a.take(f.d[0].N * f.d[0].C);
  if (!a.ok) return CNRF_E_WORKSPACE;
  if (f.shared()) f.d[0].rows_dev = f.d[1].rows_dev = f.c.runs.count;
  return CNRF_OK;
}

int check_pair(const CnrfPairIn& in, int& incx, int& incd) {
  for (int i = 0; i < 2; ++i) RUN(check_common(in.shape[i], in.params[i]));
  if (!cnrf_pair_eligible(in.shape[0], in.shape[1])) return CNRF_E_UNSUPPORTED;
  CNRF_REQUIRE(in.rd && in.freqs_xyz && in.freqs_dir && in.z_s && in.z_t, CNRF_E_NULL);
  if (in.ids && in.num_codes <= 0) return CNRF_E_SHAPE;
  return check_field(in.shape[0], in.L_xyz, in.L_dir, incx, incd);
}

}  // namespace

bool cnrf_pair_eligible(const cnrf_field_shape* c, const cnrf_field_shape* f) {
  if (check_shape(c) != CNRF_OK || check_shape(f) != CNRF_OK) return false;
  if (c->precision != CNRF_PRECISION_BF16 || f->precision != CNRF_PRECISION_BF16) return false;
  if (c->hidden != 256 || c->code != 256 || c->dim_xyz > 63 || c->dim_dir > 31) return false;
  return c->hidden == f->hidden && c->code == f->code && c->dim_xyz == f->dim_xyz && c->dim_dir == f->dim_dir &&
         c->num_rays == f->num_rays && c->pairing == f->pairing && c->keep == f->keep;
}

size_t cnrf_pair_side_bytes(const cnrf_field_shape* coarse, int32_t backward) {
  if (check_shape(coarse) != CNRF_OK) return 0;
  const Dims d = dims_of(coarse);
  return backward ? pair_bwd_side_bytes(d) : pair_fwd_side_bytes(d, coarse->keep != 0);
}

int cnrf_pair_fwd_prepare(const CnrfPairIn& in, void* side, size_t side_bytes, cudaStream_t st) {
  int incx, incd;
  RUN(check_pair(in, incx, incd));
  PairFwd f;
  RUN(carve_pair_fwd(in, side, side_bytes, f));
  const Dims& d = f.d[0];
  // viewdirs = rd / ||rd|| and their encoding: the same rays serve both passes                nerf/__init__.py:126-129
  normalize_dirs_kernel<<<grid_for(d.N, 256), 256, 0, st>>>(in.rd, d.N, f.dirn);
  CNRF_LAUNCH_CHECK();
  RUN(cnrf_pos_embed_fwd_ld(f.dirn, d.N, 3, in.freqs_dir, in.L_dir, incd, d.dd, f.demb, st));
  const float *zs = in.z_s, *zt = in.z_t;
  int64_t ld_zs = in.ld_zs, ld_zt = in.ld_zt;
  if (f.ids) {     // both models read the same two embedding tables through the same ids: one run scan, one row per run
    RUN(cnrf_code_runs(in.ids, d.N, in.num_codes, f.c.runs, st));
    RUN(cnrf_gather_codes(in.ids, f.c.runs, d.N, d.C, in.z_s, in.ld_zs, f.c.zc_s, st));
    RUN(cnrf_gather_codes(in.ids, f.c.runs, d.N, d.C, in.z_t, in.ld_zt, f.c.zc_t, st));
    zs = f.c.zc_s; zt = f.c.zc_t; ld_zs = ld_zt = d.C;
  } else if (f.bcast) {   // one code row for every ray: one run, its row is row 0 of the caller's tensors
    RUN(cnrf_code_runs_broadcast(d.N, f.c.runs, st));
    ld_zs = ld_zt = d.C;
  }
  Batch b(d.rows_dev);
  for (int level = 0; level < PER_RAY_FWD_LEVELS; ++level) {
    for (int i = 0; i < 2; ++i)
      RUN(per_ray_fwd_add(level, f.d[i], in.params[i], f.demb, d.dd, zs, ld_zs, zt, ld_zt, f.m[i].as1, f.m[i].as2, f.m[i].at, f.t[i],
                          true, f.m[i].bof, b));
    RUN(b.flush(true, st));
  }
  return cnrf_tc_pack_pair(in.params[0], f.t[0].wf, f.packed[0], in.params[1], f.t[1].wf, f.packed[1], d.dx, d.dd, st);
}

int cnrf_pair_fwd_chain(const CnrfPairIn& in, int which, const float* pts, float* raw, void* side, size_t side_bytes, void* ws,
                        size_t ws_bytes, cudaStream_t st) {
  int incx, incd;
  RUN(check_pair(in, incx, incd));
  CNRF_REQUIRE(pts && raw && ws && (which == 0 || which == 1), CNRF_E_NULL);
  PairFwd f;
  RUN(carve_pair_fwd(in, side, side_bytes, f));
  const Dims& d = f.d[which];
  const Tables& t = f.t[which];
  TcChainArgs a{d.N, d.M, d.S, d.pairing, d.dx, d.dd, in.params[which], nullptr, d.dx, pts, in.freqs_xyz, in.L_xyz, incx,
                t.b2, t.bf, t.bos, t.br, t.wf, f.demb, d.dd, f.shared() ? f.c.runs.gid : nullptr};
  a.packed = f.packed[which];
  return cnrf_tc_chain_fwd(a, raw, f.m[which].tc_saved, f.m[which].tc_saved_bytes, ws, ws_bytes, st);
}

int cnrf_pair_bwd_prepare(const CnrfPairIn& in, const cnrf_codenerf_grads* const g[2], void* side, size_t side_bytes, cudaStream_t st) {
  int incx, incd;
  RUN(check_pair(in, incx, incd));
  CNRF_REQUIRE(g && g[0] && g[1], CNRF_E_NULL);
  PairBwd f;
  RUN(carve_pair_bwd(in, side, side_bytes, f));
  // every accumulation target of the pass, both models, one launch: parameter gradients (the weight-gradient kernel and the
  // split-K per-ray GEMMs add into them), the gradient tables the weight-gradient kernel's reducers add into, the ones vectors
  FillBatch fb;
  for (int i = 0; i < 2; ++i) {
    const Dims& d = f.d[i];
    RUN(add_weight_grad_fills(d, g[i], fb));
    fb.add(f.dt[i].b2, d.N * d.H); fb.add(f.dt[i].bo, d.N * (d.C + 1)); fb.add(f.dt[i].br, d.N * 3);
    fb.add(f.dt[i].bf, d.N * d.H); fb.add(f.dt[i].dwfh, (int64_t)d.H * d.H);
    fb.add(per_ray_ones(d, f.tmp[i]), d.N, 1.0f);
  }
  fb.add(f.gz_s, f.d[0].N * f.d[0].C); fb.add(f.gz_t, f.d[0].N * f.d[0].C);
  return cnrf_fill_batch(fb, st);
}

int cnrf_pair_bwd_chain(const CnrfPairIn& in, int which, const float* g_raw, const cnrf_codenerf_grads* g, void* side,
                        size_t side_bytes, void* ws, size_t ws_bytes, cudaStream_t st) {
  int incx, incd;
  RUN(check_pair(in, incx, incd));
  CNRF_REQUIRE(g_raw && g && ws && (which == 0 || which == 1), CNRF_E_NULL);
  PairBwd f;
  RUN(carve_pair_bwd(in, side, side_bytes, f));
  const Dims& d = f.d[which];
  const Tables& dt = f.dt[which];
  TcChainArgs a{d.N, d.M, d.S, d.pairing, d.dx, d.dd, in.params[which], nullptr, d.dx, nullptr, in.freqs_xyz, in.L_xyz, incx,
                nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, d.dd, f.shared() ? f.c.runs.gid : nullptr};
  a.prezeroed = true;
  return cnrf_tc_chain_bwd(a, g_raw, f.m[which].tc_saved, f.m[which].tc_saved_bytes, g, dt.b2, dt.bo, dt.br, nullptr, dt.bf, dt.dwfh,
                           nullptr, ws, ws_bytes, st);
}

int cnrf_pair_bwd_finish(const CnrfPairIn& in, const cnrf_codenerf_grads* const g[2], float* g_z_s, float* g_z_t, void* side,
                         size_t side_bytes, cudaStream_t st) {
  int incx, incd;
  RUN(check_pair(in, incx, incd));
  CNRF_REQUIRE(g && g[0] && g[1] && g_z_s && g_z_t, CNRF_E_NULL);
  PairBwd f;
  RUN(carve_pair_bwd(in, side, side_bytes, f));
  const Dims& d = f.d[0];
  const float *zs = in.z_s, *zt = in.z_t;
  int64_t ld_zs = in.ld_zs, ld_zt = in.ld_zt;
  float *gz_s = g_z_s, *gz_t = g_z_t;      // per-ray codes: both models add straight into the caller's [N, C] gradients
  FillBatch fb;
  if (f.ids) {     // one gradient row per run (zero-filled by the prepare step), scattered into the dense tables at the end
    zs = f.c.zc_s; zt = f.c.zc_t; ld_zs = ld_zt = d.C;
    gz_s = f.gz_s; gz_t = f.gz_t;
    fb.add(g_z_s, in.num_codes * d.C); fb.add(g_z_t, in.num_codes * d.C);
  } else {                // (broadcast codes: the single run's gradient lands in row 0, the other rows stay zero -- the sum over
    fb.add(g_z_s, d.N * d.C); fb.add(g_z_t, d.N * d.C);      //  rows that the caller's expand() backward takes is the same)
    if (f.bcast) ld_zs = ld_zt = d.C;
  }
  RUN(cnrf_fill_batch(fb, st));
  Batch b(d.rows_dev);
  for (int level = 0; level < PER_RAY_BWD_LEVELS; ++level) {
    for (int i = 0; i < 2; ++i)
      RUN(per_ray_bwd_add(level, f.d[i], in.params[i], g[i], nullptr, d.dd, zs, ld_zs, zt, ld_zt, f.m[i].as1, f.m[i].as2, f.m[i].at,
                          f.dt[i], f.tmp[i], gz_s, gz_t, nullptr, f.m[i].bof, true, b));
    RUN(b.flush(true, st));
  }
  if (f.ids) {
    RUN(cnrf_scatter_code_grads(in.ids, f.c.runs, d.N, d.C, gz_s, g_z_s, d.C, st));
    RUN(cnrf_scatter_code_grads(in.ids, f.c.runs, d.N, d.C, gz_t, g_z_t, d.C, st));
  }
  return CNRF_OK;
}
