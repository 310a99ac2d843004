// Internal C++ interfaces between the translation units of libcnrf.so (not part of the C ABI).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

// ---- per-device host-side caches -----------------------------------------------------------------------------------
// A process may drive several GPUs (eval on cuda:1 after cuda:0, a test looping over devices): function attributes
// (dynamic shared-memory opt-in), __device__ symbol addresses and SM counts are per device, so every cache of them is an
// array indexed by the current device ordinal.  Ordinals beyond the table are simply not cached (-1).
constexpr int CNRF_MAX_DEVICES = 64;
inline int cnrf_device_slot() {
  int d = 0;
  if (cudaGetDevice(&d) != cudaSuccess || d < 0 || d >= CNRF_MAX_DEVICES) return -1;
  return d;
}
inline int cnrf_sm_count() {
  static int n[CNRF_MAX_DEVICES] = {};
  const int slot = cnrf_device_slot();
  if (slot >= 0 && n[slot] > 0) return n[slot];
  int dev = 0, v = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev);
  if (v <= 0) v = 148;
  if (slot >= 0) n[slot] = v;
  return v;
}
// `done` is a function-local `static bool done[CNRF_MAX_DEVICES]`; runs `set` (-> cudaError_t) once per device
template <class F>
inline cudaError_t cnrf_once_per_device(bool* done, F set) {
  const int slot = cnrf_device_slot();
  if (slot >= 0 && done[slot]) return cudaSuccess;
  const cudaError_t e = set();
  if (e == cudaSuccess && slot >= 0) done[slot] = true;
  return e;
}
// address of a __device__ symbol on the current device
template <class T>
inline T* cnrf_symbol_ptr(T** cache, const T& symbol) {
  const int slot = cnrf_device_slot();
  if (slot >= 0 && cache[slot]) return cache[slot];
  T* p = nullptr;
  cudaGetSymbolAddress(reinterpret_cast<void**>(&p), symbol);
  if (slot >= 0) cache[slot] = p;
  return p;
}

// ---- sgemm.cu : fp32 SIMT GEMM toolkit -------------------------------------------------------------------------
// C(m,n) = epi( sum_k A(m,k) * B(k,n) ),  A(m,k) = A[m*a_rs + k*a_cs],  B(k,n) = B[k*b_rs + n*b_cs], C row-major (ldc).
struct GemmArgs {
  const float* A = nullptr; int64_t a_rs = 0, a_cs = 0;
  const float* B = nullptr; int64_t b_rs = 0, b_cs = 0;
  float* C = nullptr; int64_t ldc = 0;
  int64_t M = 0; int N = 0; int64_t K = 0;
  // epilogue: + bias[((m / bias_div) % bias_mod) * ld_bias + n], relu, * (mask(m,n) > 0)
  const float* bias = nullptr; int64_t ld_bias = 0; int64_t bias_div = 1; int64_t bias_mod = INT64_MAX;
  int relu = 0;
  const float* mask = nullptr; int64_t ld_mask = 0;
  int accumulate = 0;   // C += result (no bias/relu/mask)
  int split_k = 1;      // > 1: partial sums are atomically added into C, which the caller zero-fills
  // tensor-core kernel only: > 0 (a multiple of 64) = contraction slice per CTA; the slices' partial sums are added
  // atomically, so C must be zero-filled (or hold the value to add to) and `accumulate` set.  Short-and-wide problems
  // (the fc_out fold, K = 256) are latency-bound: one load round trip per CTA instead of K / 64 serial ones.
  int64_t k_slice = 0;
  // tensor-core kernel only: a device-resident row count that shrinks M (rows_dim = 1) or K (rows_dim = 2) at run time;
  // M / K above are then the worst case the launch is sized for (code_runs.cu)
  const int32_t* rows_dev = nullptr; int rows_dim = 0;
};
int cnrf_sgemm(const GemmArgs& a, cudaStream_t st);
// gemm_tc.cu: the same contract on the tensor cores (bf16 operands, fp32 accumulate); CNRF_E_UNSUPPORTED when neither
// stride of an operand is 1 or K > 320 without split-K permission -- the caller then uses cnrf_sgemm.
int cnrf_gemm_tc(const GemmArgs& a, cudaStream_t st);
// up to CNRF_GEMM_BATCH_MAX mutually independent problems in one launch
constexpr int CNRF_GEMM_BATCH_MAX = 24;     // (a 4.9 KB kernel parameter: CUDA >= 12.1 passes up to 32 KB)
int cnrf_gemm_tc_batch(const GemmArgs* list, int count, cudaStream_t st);
// gemm_stream.cu: persistent, pipelined split-bf16 GEMM for ONE big per-sample problem (M >= 296 x 128 rows, 64 <= N <= 256,
// K <= 512, A row-major, B a weight matrix); CNRF_E_UNSUPPORTED otherwise.  cnrf_gemm_tc tries it first.
int cnrf_gemm_stream(const GemmArgs& a, cudaStream_t st);
int cnrf_gemm_stream_error();
// one launch: dst[i][0..n[i]) = value[i]  (zero-filling gradient buffers, ones vectors)
constexpr int CNRF_FILL_BATCH_MAX = 64;
struct FillBatch { float* dst[CNRF_FILL_BATCH_MAX]; int64_t n[CNRF_FILL_BATCH_MAX]; float value[CNRF_FILL_BATCH_MAX]; int count = 0;
  void add(float* p, int64_t cnt, float v = 0.f) { dst[count] = p; n[count] = cnt; value[count] = v; ++count; } };
int cnrf_fill_batch(const FillBatch& b, cudaStream_t st);
// out[g*ld_out + n] = sum_{q<Q} X[(g*g_stride + q*q_stride)*ld_x + n],  g < G, n < N
int cnrf_strided_rowsum(const float* X, int64_t ld_x, int64_t G, int64_t g_stride, int64_t Q, int64_t q_stride, int N,
                        float* out, int64_t ld_out, cudaStream_t st);
// out[n] = sum_m X[m*ld_x + n]  (out is zero-filled by the callee)
int cnrf_colsum(const float* X, int64_t ld_x, int64_t M, int N, float* out, cudaStream_t st);
int cnrf_fill_zero(void* p, size_t bytes, cudaStream_t st);

// ---- codenerf.cu: forward_pass with or without the embedding lookup folded in (ids != NULL: z_s / z_t are the tables) ------
int cnrf_field_fwd_impl(const cnrf_field_shape* s, const cnrf_codenerf_params* p, const float* pts, const float* rd,
                        const float* freqs_xyz, int32_t L_xyz, const float* freqs_dir, int32_t L_dir, const int64_t* ids,
                        int64_t num_codes, const float* z_s, int64_t ld_zs, const float* z_t, int64_t ld_zt, float* raw,
                        void* saved, size_t saved_bytes, void* workspace, size_t workspace_bytes, cudaStream_t st);
// accumulate_tables (ids mode): g_z_s / g_z_t already hold another model's dense table gradients; add instead of overwrite
int cnrf_field_bwd_impl(const cnrf_field_shape* s, const cnrf_codenerf_params* p, const float* pts, const float* rd,
                        const float* freqs_xyz, int32_t L_xyz, const float* freqs_dir, int32_t L_dir, const int64_t* ids,
                        int64_t num_codes, const float* z_s, int64_t ld_zs, const float* z_t, int64_t ld_zt,
                        const float* g_raw, const void* saved, size_t saved_bytes, const cnrf_codenerf_grads* g, float* g_z_s,
                        float* g_z_t, float* g_pts, float* g_rd, void* workspace, size_t workspace_bytes, cudaStream_t st,
                        bool accumulate_tables);

// ---- codenerf.cu: both models of one render call (tensor-core mode) -- the fused entry of render.cu ------------------------
// The per-ray work of the coarse and the fine model (code runs, code rows, the three levels of code-layer GEMMs, the fc_out
// folds, weight packing, direction encoding; in the backward the zero fills, the unfold / code-layer gradient GEMMs and
// the scatter into the embedding tables) does not depend on the samples, so one set of launches serves both models:
//   forward   cnrf_pair_fwd_prepare -> cnrf_pair_fwd_chain(0) ... cnrf_pair_fwd_chain(1)
//   backward  cnrf_pair_bwd_prepare -> cnrf_pair_bwd_chain(1) -> cnrf_pair_bwd_chain(0) -> cnrf_pair_bwd_finish
// `side` is caller scratch of cnrf_pair_side_bytes() that must stay untouched between the calls of one direction.
struct CnrfPairIn {
  const cnrf_field_shape* shape[2];            // [0] coarse, [1] fine: same rays, codes, widths; num_samples differ
  const cnrf_codenerf_params* params[2];
  const float* rd; const float* freqs_xyz; int32_t L_xyz; const float* freqs_dir; int32_t L_dir;
  const int64_t* ids; int64_t num_codes;       // ids != NULL: z_s / z_t are the embedding tables (shared-code mode)
  const float* z_s; int64_t ld_zs; const float* z_t; int64_t ld_zt;
  void* saved[2]; size_t saved_bytes[2];       // cnrf_codenerf_saved_bytes(shape[i]) each when shape[i]->keep
};
bool cnrf_pair_eligible(const cnrf_field_shape* coarse, const cnrf_field_shape* fine);
size_t cnrf_pair_side_bytes(const cnrf_field_shape* coarse, int32_t backward);
int cnrf_pair_fwd_prepare(const CnrfPairIn& in, void* side, size_t side_bytes, cudaStream_t st);
int cnrf_pair_fwd_chain(const CnrfPairIn& in, int which, const float* pts, float* raw, void* side, size_t side_bytes, void* ws,
                        size_t ws_bytes, cudaStream_t st);
int cnrf_pair_bwd_prepare(const CnrfPairIn& in, const cnrf_codenerf_grads* const g[2], void* side, size_t side_bytes, cudaStream_t st);
int cnrf_pair_bwd_chain(const CnrfPairIn& in, int which, const float* g_raw, const cnrf_codenerf_grads* g, void* side,
                        size_t side_bytes, void* ws, size_t ws_bytes, cudaStream_t st);
// g_z_s / g_z_t: dense table gradients [num_codes, C] (ids mode, overwritten) or per-ray code gradients [N, C] (overwritten)
int cnrf_pair_bwd_finish(const CnrfPairIn& in, const cnrf_codenerf_grads* const g[2], float* g_z_s, float* g_z_t, void* side,
                         size_t side_bytes, cudaStream_t st);

// ---- composite.cu: the photometric loss of train.py:103-104 folded into the compositing epilogues --------
// target of ray n, channel c: rows[n * ld + c] (already gathered float rows), or -- the gather of train.py:77-80 fused as
// well -- views[pixel[n] * channels + c] / 255.0 from uint8 views resident in HBM (data.cu: the loader's own arithmetic)
struct LossTarget { const float* rows = nullptr; int64_t ld = 0; const uint8_t* views = nullptr; const int64_t* pixel = nullptr; int channels = 0; };
struct CompositeLossFwd {
  LossTarget t;
  float* sq_sum = nullptr;          // += sum (rgb - target)^2 of this pass (zero-filled by the caller); NULL: no loss
  float* out3 = nullptr;            // fine pass only: (coarse_loss, fine_loss, loss) written by the last warp to finish
  const float* other_sq = nullptr;  // ... the coarse pass's sum
  unsigned* done = nullptr;         // ... warps finished (zero-filled by the caller)
  unsigned num_warps = 0;           // set by the launcher
  float inv = 0.f;                  // 1 / (3 N)
};
struct CompositeLossBwd {
  LossTarget t;
  const float* rgb_out = nullptr;   // the forward's rgb [N,3]; NULL: the upstream gradient is g_rgb
  const float* g_term = nullptr;    // device scalars (NULL = 0): d / d (this pass's loss term), d / d (the sum)
  const float* g_sum = nullptr;
  float scale = 0.f;                // 2 / (3 N)
};
int cnrf_composite_fwd_loss(const float* raw, const float* z, const float* rd, int64_t N, int32_t S, float* rgb, float* disp,
                            float* acc, float* weights, float* depth, const CompositeLossFwd& lf, cudaStream_t st);
int cnrf_composite_bwd_loss(const float* raw, const float* z, const float* rd, int64_t N, int32_t S, const float* g_rgb,
                            const float* g_disp, const float* g_acc, const float* g_weights, const float* g_depth, float* g_raw,
                            float* g_z, float* g_rd, const CompositeLossBwd& lb, cudaStream_t st);

// ---- posembed.cu -------------------------------------------------------------------------------------------------
int cnrf_pos_embed_fwd_ld(const float* x, int64_t M, int D, const float* freqs, int L, int inc, int64_t ld_out,
                          float* out, cudaStream_t st);
int cnrf_pos_embed_bwd_ld(const float* x, const float* g_out, int64_t M, int D, const float* freqs, int L, int inc,
                          int64_t ld_g, float* g_x, cudaStream_t st);

// ---- ktimer.cu: CUDA-event timers around the tensor-core kernels (debug / bench) ----------------------------------------
enum { CNRF_KT_FWD = 0, CNRF_KT_BWD = 1, CNRF_KT_DW = 2, CNRF_KT_COUNT = 3 };
bool cnrf_ktimer_begin(int which, int64_t rows, cudaStream_t st);   // true: timing, call cnrf_ktimer_end after the launch
void cnrf_ktimer_end(int which, cudaStream_t st);

// ---- code_runs.cu: runs of rays that share one latent-code row (device-resident run count) ------------------------------
struct CodeRuns { int32_t* gid = nullptr; int32_t* rep = nullptr; int32_t* count = nullptr; };
size_t cnrf_code_runs_bytes(int64_t N);
CodeRuns cnrf_code_runs_carve(void* base /*256-byte aligned*/, int64_t N);
int cnrf_code_runs(const int64_t* ids, int64_t N, int64_t num_codes, const CodeRuns& r, cudaStream_t st);
int cnrf_code_runs_broadcast(int64_t N, const CodeRuns& r, cudaStream_t st);     // all rays share one code row: a single run
int cnrf_gather_codes(const int64_t* ids, const CodeRuns& r, int64_t N, int C, const float* tab, int64_t ld, float* out,
                      cudaStream_t st);
int cnrf_scatter_code_grads(const int64_t* ids, const CodeRuns& r, int64_t N, int C, const float* g_codes, float* g_tab,
                            int64_t ld, cudaStream_t st);

// ---- flex_fwd_tc.cu: fused tcgen05 forward chain of the vanilla NeRF field (inference, bf16, hidden 256) -----------------
bool cnrf_flex_fused_eligible(const cnrf_flex_shape* s);
size_t cnrf_flex_fused_packed_bytes(int num_xyz_layers);
// gp != NULL: the kernel encodes the sample positions itself (x_xyz unused)
struct FlexPts { const float* pts; int64_t ld; const float* freqs; int nfreq; int inc; };
int cnrf_flex_fused_fwd(const cnrf_flex_shape* s, const cnrf_flex_params* p, const float* x_xyz, int64_t ld_xyz,
                        const float* x_dir /*[N,dim_dir]*/, int64_t ld_dir, float* raw, uint8_t* packed /*1024-byte aligned*/,
                        cudaStream_t st,
                        const FlexPts* gp = nullptr);
int cnrf_flex_fused_error();
