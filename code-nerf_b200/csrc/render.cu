// cnrf_render_fwd / cnrf_render_bwd: predict_radiance_and_render (view_synthesis/nerf/__init__.py:74-91) as ONE entry
// point per direction -- stratified sampling -> coarse field -> compositing -> inverse-CDF sampling -> fine field ->
// compositing, and the matching backward -- so the host makes two library calls per chunk instead of driving ~58 launches
// through Python.  The kernels are the ones behind the per-stage entry points (sampling.cu, codenerf.cu + the tensor-core
// chain, composite.cu); this file only sequences them on the caller's stream over caller-owned `saved` / `workspace`
// buffers.  Rays carry no gradient here (training and rendering); camera-pose optimisation (eval.py:129-168) keeps the
// per-stage path, whose autograd graph reaches ro / rd.
#include "common.cuh"
#include "internal.h"

namespace {

inline size_t al256(size_t bytes) { return (bytes + 255) & ~size_t(255); }

struct Cut {     // carve 256-byte aligned pieces out of a caller buffer
  char* base; size_t cap; size_t off = 0; bool ok = true;
  Cut(void* p, size_t c) : base(static_cast<char*>(p)), cap(c) {
    const size_t mis = reinterpret_cast<uintptr_t>(p) & 255;
    if (mis) off = 256 - mis;
  }
  float* take(size_t nfloat) {
    const size_t bytes = al256(nfloat * sizeof(float));
    if (!base || off + bytes > cap) { ok = false; return nullptr; }
    float* r = reinterpret_cast<float*>(base + off);
    off += bytes;
    return r;
  }
  void* rest() { return base ? base + off : nullptr; }
  size_t rest_bytes() const { return cap > off ? cap - off : 0; }
};

struct Dims { int64_t N; int Sc, Sf; bool f32; };
inline Dims dims_of(const cnrf_render_shape* s) { return {s->num_rays, s->num_coarse, s->num_coarse + s->num_fine, s->precision != CNRF_PRECISION_BF16}; }

inline cnrf_field_shape field_shape(const cnrf_render_shape* s, int S, int keep) {
  cnrf_field_shape f;
  f.hidden = s->hidden; f.code = s->code; f.dim_xyz = s->dim_xyz; f.dim_dir = s->dim_dir; f.num_rays = s->num_rays;
  f.num_samples = S; f.pairing = s->pairing; f.precision = s->precision; f.keep = keep;
  return f;
}

inline int check(const cnrf_render_shape* s) {
  if (!s) return CNRF_E_NULL;
  if (s->num_rays <= 0 || s->num_coarse < 3 || s->num_fine <= 0) return CNRF_E_SHAPE;     // sample_pdf needs Sc - 2 >= 1 weights
  return CNRF_OK;
}

// what the backward needs from the forward, apart from the two fields' own saved buffers
struct Kept { float *z_c, *raw_c, *z_f, *raw_f, *pts_c, *pts_f; };
inline size_t kept_bytes(const Dims& d, bool with_pts) {
  size_t b = al256(d.N * d.Sc * 4) + al256(d.N * d.Sc * 16) + al256(d.N * d.Sf * 4) + al256(d.N * d.Sf * 16);
  if (with_pts) b += al256(d.N * d.Sc * 12) + al256(d.N * d.Sf * 12);
  return b + 256;
}
inline void carve_kept(const Dims& d, Cut& c, bool with_pts, Kept& k) {
  k.z_c = c.take(d.N * d.Sc); k.raw_c = c.take(d.N * d.Sc * 4); k.z_f = c.take(d.N * d.Sf); k.raw_f = c.take(d.N * d.Sf * 4);
  k.pts_c = with_pts ? c.take(d.N * d.Sc * 3) : nullptr;
  k.pts_f = with_pts ? c.take(d.N * d.Sf * 3) : nullptr;
}

__global__ void add_inplace_kernel(float* __restrict__ dst, const float* __restrict__ src, int64_t n4) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n4) return;
  float4 a = reinterpret_cast<float4*>(dst)[i];
  const float4 b = reinterpret_cast<const float4*>(src)[i];
  a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
  reinterpret_cast<float4*>(dst)[i] = a;
}

#define RUN(x) do { int rc__ = (x); if (rc__) return rc__; } while (0)
}  // namespace

extern "C" size_t cnrf_render_saved_bytes(const cnrf_render_shape* s) {
  if (check(s) != CNRF_OK || !s->keep) return 0;
  const Dims d = dims_of(s);
  const cnrf_field_shape fc = field_shape(s, d.Sc, 1), ff = field_shape(s, d.Sf, 1);
  // the fp32 chain re-embeds the points in its backward; the bf16 chain saved its own encoded inputs
  return kept_bytes(d, d.f32) + al256(cnrf_codenerf_saved_bytes(&fc)) + al256(cnrf_codenerf_saved_bytes(&ff)) + 512;
}

extern "C" size_t cnrf_render_workspace_bytes(const cnrf_render_shape* s, int32_t backward) {
  if (check(s) != CNRF_OK) return 0;
  const Dims d = dims_of(s);
  const cnrf_field_shape fc = field_shape(s, d.Sc, s->keep), ff = field_shape(s, d.Sf, s->keep);
  const size_t fw = cnrf_field_workspace_bytes(&fc, backward), fwf = cnrf_field_workspace_bytes(&ff, backward);
  size_t field = al256(fw > fwf ? fw : fwf) + 256;
  if (cnrf_pair_eligible(&fc, &ff)) field += al256(cnrf_pair_side_bytes(&fc, backward)) + 256;     // the two-model per-ray state
  if (!backward) {
    size_t b = al256(d.N * d.Sc * 4) /*weights*/ + 256 /*loss sums*/ + field + 512;
    if (!s->keep) b += kept_bytes(d, true);                       // inference: nothing is kept, everything is scratch
    else if (!d.f32) b += al256(d.N * d.Sc * 12) + al256(d.N * d.Sf * 12);     // bf16 training: points are scratch
    return b;
  }
  return al256(d.N * d.Sf * 16) /*g_raw*/ + 2 * al256(d.N * (size_t)s->code * 4) /*coarse per-ray code gradients*/ + field + 512;
}

namespace {
inline int loss_target_of(const cnrf_loss_target* t, LossTarget& lt) {
  if (!t) return CNRF_E_NULL;
  if (t->views) {
    CNRF_REQUIRE(t->pixel && t->channels >= 3 && t->channels <= 4, CNRF_E_SHAPE);
    lt.views = t->views; lt.pixel = t->pixel; lt.channels = t->channels;
  } else {
    CNRF_REQUIRE(t->rows && t->ld_rows >= 3, CNRF_E_SHAPE);
    lt.rows = t->rows; lt.ld = t->ld_rows;
  }
  return CNRF_OK;
}

// tgt != NULL: the photometric loss of train.py:103-104 rides in the two compositing epilogues and lands in loss3
int render_fwd_impl(const cnrf_render_shape* s, const cnrf_codenerf_params* coarse, const cnrf_codenerf_params* fine,
                    const cnrf_render_inputs* in, const cnrf_loss_target* tgt, float* loss3, float* rgb_coarse, float* rgb_fine,
                    float* extras_fine, void* saved, size_t saved_bytes, void* workspace, size_t workspace_bytes,
                    cnrf_stream_t stream) {
  RUN(check(s));
  CNRF_REQUIRE(coarse && fine && in && rgb_coarse && rgb_fine && workspace, CNRF_E_NULL);
  CNRF_REQUIRE(in->ro && in->rd && in->z_table && in->lower && in->upper && in->freqs_xyz && in->freqs_dir && in->z_s && in->z_t,
               CNRF_E_NULL);
  if (s->keep && !saved) return CNRF_E_WORKSPACE;
  const Dims d = dims_of(s);
  cudaStream_t st = as_stream(stream);
  Cut sv(s->keep ? saved : nullptr, s->keep ? saved_bytes : 0), ws(workspace, workspace_bytes);
  Kept k;
  if (s->keep) {
    carve_kept(d, sv, d.f32, k);
    if (!d.f32) { k.pts_c = ws.take(d.N * d.Sc * 3); k.pts_f = ws.take(d.N * d.Sf * 3); }
  } else {
    carve_kept(d, ws, true, k);
  }
  float* weights = ws.take(d.N * d.Sc);
  float* loss_acc = ws.take(64);           // [0] coarse sum of squares, [1] fine, [2] warps of the fine pass that finished
  if (!ws.ok || (s->keep && !sv.ok)) return CNRF_E_WORKSPACE;
  CompositeLossFwd lc, lf;
  if (tgt) {
    CNRF_REQUIRE(loss3, CNRF_E_NULL);
    RUN(loss_target_of(tgt, lc.t));
    lf.t = lc.t;
    lc.sq_sum = loss_acc; lf.sq_sum = loss_acc + 1; lf.other_sq = loss_acc; lf.done = reinterpret_cast<unsigned*>(loss_acc + 2);
    lf.out3 = loss3; lf.inv = 1.0f / (3.0f * (float)d.N);
    RUN(cnrf_fill_zero(loss_acc, 3 * sizeof(float), as_stream(stream)));
  }
  const cnrf_field_shape fc = field_shape(s, d.Sc, s->keep), ff = field_shape(s, d.Sf, s->keep);
  const size_t sb_c = s->keep ? al256(cnrf_codenerf_saved_bytes(&fc)) : 0, sb_f = s->keep ? al256(cnrf_codenerf_saved_bytes(&ff)) : 0;
  char* saved_c = s->keep ? static_cast<char*>(sv.rest()) : nullptr;
  char* saved_f = s->keep ? saved_c + sb_c : nullptr;
  if (s->keep && sv.rest_bytes() < sb_c + sb_f) return CNRF_E_WORKSPACE;

  // tensor-core mode: the per-ray work of BOTH models (code runs, code layers, folds, weight packing, direction encoding)
  // in one set of launches ahead of the two per-sample chains (codenerf.cu, "both models of one render call")
  const bool pair = cnrf_pair_eligible(&fc, &ff);
  CnrfPairIn pin{{&fc, &ff}, {coarse, fine}, in->rd, in->freqs_xyz, in->L_xyz, in->freqs_dir, in->L_dir, in->ids, in->num_codes,
                 in->z_s, in->ld_zs, in->z_t, in->ld_zt, {saved_c, saved_f}, {sb_c, sb_f}};
  void* side = nullptr; size_t side_bytes = 0;
  if (pair) {
    side_bytes = al256(cnrf_pair_side_bytes(&fc, 0));
    side = ws.take(side_bytes / sizeof(float));
    if (!ws.ok) return CNRF_E_WORKSPACE;
    RUN(cnrf_pair_fwd_prepare(pin, side, side_bytes, st));
  }
  // coarse pass                                                                          nerf/__init__.py:82-84
  RUN(cnrf_sample_uniform_fwd(in->ro, in->rd, d.N, d.Sc, in->z_table, in->lower, in->upper, in->t_rand, k.pts_c, k.z_c, stream));
  if (pair) RUN(cnrf_pair_fwd_chain(pin, 0, k.pts_c, k.raw_c, side, side_bytes, ws.rest(), ws.rest_bytes(), st));
  else RUN(cnrf_field_fwd_impl(&fc, coarse, k.pts_c, in->rd, in->freqs_xyz, in->L_xyz, in->freqs_dir, in->L_dir, in->ids, in->num_codes,
                               in->z_s, in->ld_zs, in->z_t, in->ld_zt, k.raw_c, saved_c, sb_c, ws.rest(), ws.rest_bytes(), st));
  RUN(cnrf_composite_fwd_loss(k.raw_c, k.z_c, in->rd, d.N, d.Sc, rgb_coarse, nullptr, nullptr, weights, nullptr, lc, st));
  // fine pass: importance samples from weights[..., 1:-1]                                nerf/__init__.py:87-89
  RUN(cnrf_sample_pdf_fwd(in->ro, in->rd, weights + 1, d.Sc, k.z_c, d.N, d.Sc, s->num_fine, in->u, k.pts_f, k.z_f, nullptr, nullptr,
                          stream));
  if (pair) RUN(cnrf_pair_fwd_chain(pin, 1, k.pts_f, k.raw_f, side, side_bytes, ws.rest(), ws.rest_bytes(), st));
  else RUN(cnrf_field_fwd_impl(&ff, fine, k.pts_f, in->rd, in->freqs_xyz, in->L_xyz, in->freqs_dir, in->L_dir, in->ids, in->num_codes,
                               in->z_s, in->ld_zs, in->z_t, in->ld_zt, k.raw_f, saved_f, sb_f, ws.rest(), ws.rest_bytes(), st));
  float* disp = extras_fine, *acc = extras_fine ? extras_fine + d.N : nullptr, *depth = extras_fine ? extras_fine + 2 * d.N : nullptr;
  return cnrf_composite_fwd_loss(k.raw_f, k.z_f, in->rd, d.N, d.Sf, rgb_fine, disp, acc, nullptr, depth, lf, st);
}
}  // namespace

extern "C" int cnrf_render_fwd(const cnrf_render_shape* s, const cnrf_codenerf_params* coarse, const cnrf_codenerf_params* fine,
                               const cnrf_render_inputs* in, float* rgb_coarse, float* rgb_fine, float* extras_fine,
                               void* saved, size_t saved_bytes, void* workspace, size_t workspace_bytes, cnrf_stream_t stream) {
  return render_fwd_impl(s, coarse, fine, in, nullptr, nullptr, rgb_coarse, rgb_fine, extras_fine, saved, saved_bytes, workspace,
                         workspace_bytes, stream);
}

extern "C" int cnrf_render_loss_fwd(const cnrf_render_shape* s, const cnrf_codenerf_params* coarse, const cnrf_codenerf_params* fine,
                                    const cnrf_render_inputs* in, const cnrf_loss_target* target, float* loss3, float* rgb_coarse,
                                    float* rgb_fine, void* saved, size_t saved_bytes, void* workspace, size_t workspace_bytes,
                                    cnrf_stream_t stream) {
  CNRF_REQUIRE(target && loss3, CNRF_E_NULL);
  return render_fwd_impl(s, coarse, fine, in, target, loss3, rgb_coarse, rgb_fine, nullptr, saved, saved_bytes, workspace,
                         workspace_bytes, stream);
}

namespace {
// lbc / lbf: the upstream gradients of the two compositing passes come from the loss folded into the forward epilogues
int render_bwd_impl(const cnrf_render_shape* s, const cnrf_codenerf_params* coarse, const cnrf_codenerf_params* fine,
                    const cnrf_render_inputs* in, const float* g_rgb_coarse, const float* g_rgb_fine, const CompositeLossBwd& lbc,
                    const CompositeLossBwd& lbf, const void* saved, size_t saved_bytes, const cnrf_codenerf_grads* g_coarse,
                    const cnrf_codenerf_grads* g_fine, float* g_z_s, float* g_z_t, void* workspace, size_t workspace_bytes,
                    cnrf_stream_t stream) {
  RUN(check(s));
  CNRF_REQUIRE(coarse && fine && in && saved && g_coarse && g_fine && g_z_s && g_z_t && workspace, CNRF_E_NULL);
  CNRF_REQUIRE(g_rgb_coarse || g_rgb_fine || lbc.rgb_out || lbf.rgb_out, CNRF_E_NULL);
  CNRF_REQUIRE(in->rd && in->freqs_xyz && in->freqs_dir && in->z_s && in->z_t, CNRF_E_NULL);
  if (!s->keep) return CNRF_E_WORKSPACE;
  const Dims d = dims_of(s);
  cudaStream_t st = as_stream(stream);
  Cut sv(const_cast<void*>(saved), saved_bytes), ws(workspace, workspace_bytes);
  Kept k;
  carve_kept(d, sv, d.f32, k);
  float* g_raw = ws.take(d.N * d.Sf * 4);
  float* gc_s = ws.take(d.N * s->code);
  float* gc_t = ws.take(d.N * s->code);
  if (!ws.ok || !sv.ok) return CNRF_E_WORKSPACE;
  const cnrf_field_shape fc = field_shape(s, d.Sc, 1), ff = field_shape(s, d.Sf, 1);
  const size_t sb_c = al256(cnrf_codenerf_saved_bytes(&fc)), sb_f = al256(cnrf_codenerf_saved_bytes(&ff));
  const char* saved_c = static_cast<const char*>(sv.rest());
  const char* saved_f = saved_c + sb_c;
  if (sv.rest_bytes() < sb_c + sb_f) return CNRF_E_WORKSPACE;
  const bool tables = in->ids != nullptr;
  // A missing upstream gradient is a zero gradient: the pass still runs so that every parameter gradient is written
  // (the caller's buffers are uninitialised) -- compositing treats a NULL g_rgb as zero.
  // fine pass first: its dW kernel is the long one, and a caller that overlaps a gradient exchange wants it early
  if (cnrf_pair_eligible(&fc, &ff)) {     // tensor-core mode: zero fills, per-ray gradient GEMMs and table scatters once for both models
    CnrfPairIn pin{{&fc, &ff}, {coarse, fine}, in->rd, in->freqs_xyz, in->L_xyz, in->freqs_dir, in->L_dir, in->ids, in->num_codes,
                   in->z_s, in->ld_zs, in->z_t, in->ld_zt, {const_cast<char*>(saved_c), const_cast<char*>(saved_f)}, {sb_c, sb_f}};
    const cnrf_codenerf_grads* const gs[2] = {g_coarse, g_fine};
    const size_t side_bytes = al256(cnrf_pair_side_bytes(&fc, 1));
    void* side = ws.take(side_bytes / sizeof(float));
    if (!ws.ok) return CNRF_E_WORKSPACE;
    RUN(cnrf_pair_bwd_prepare(pin, gs, side, side_bytes, st));
    RUN(cnrf_composite_bwd_loss(k.raw_f, k.z_f, in->rd, d.N, d.Sf, g_rgb_fine, nullptr, nullptr, nullptr, nullptr, g_raw, nullptr, nullptr, lbf, st));
    RUN(cnrf_pair_bwd_chain(pin, 1, g_raw, g_fine, side, side_bytes, ws.rest(), ws.rest_bytes(), st));
    RUN(cnrf_composite_bwd_loss(k.raw_c, k.z_c, in->rd, d.N, d.Sc, g_rgb_coarse, nullptr, nullptr, nullptr, nullptr, g_raw, nullptr, nullptr, lbc, st));
    RUN(cnrf_pair_bwd_chain(pin, 0, g_raw, g_coarse, side, side_bytes, ws.rest(), ws.rest_bytes(), st));
    return cnrf_pair_bwd_finish(pin, gs, g_z_s, g_z_t, side, side_bytes, st);
  }
  RUN(cnrf_composite_bwd_loss(k.raw_f, k.z_f, in->rd, d.N, d.Sf, g_rgb_fine, nullptr, nullptr, nullptr, nullptr, g_raw, nullptr, nullptr, lbf, st));
  RUN(cnrf_field_bwd_impl(&ff, fine, k.pts_f, in->rd, in->freqs_xyz, in->L_xyz, in->freqs_dir, in->L_dir, in->ids, in->num_codes,
                          in->z_s, in->ld_zs, in->z_t, in->ld_zt, g_raw, saved_f, sb_f, g_fine, g_z_s, g_z_t, nullptr, nullptr,
                          ws.rest(), ws.rest_bytes(), st, false));
  RUN(cnrf_composite_bwd_loss(k.raw_c, k.z_c, in->rd, d.N, d.Sc, g_rgb_coarse, nullptr, nullptr, nullptr, nullptr, g_raw, nullptr, nullptr, lbc, st));
  // coarse model: table gradients are scatter-added onto the fine model's (shared-code mode); per-ray code gradients go
  // through a scratch pair and one add
  RUN(cnrf_field_bwd_impl(&fc, coarse, k.pts_c, in->rd, in->freqs_xyz, in->L_xyz, in->freqs_dir, in->L_dir, in->ids, in->num_codes,
                          in->z_s, in->ld_zs, in->z_t, in->ld_zt, g_raw, saved_c, sb_c, g_coarse, tables ? g_z_s : gc_s,
                          tables ? g_z_t : gc_t, nullptr, nullptr, ws.rest(), ws.rest_bytes(), st, tables));
  if (!tables) {
    const int64_t n = d.N * s->code;
    if (n % 4) return CNRF_E_SHAPE;
    add_inplace_kernel<<<grid_for(n / 4, 256), 256, 0, st>>>(g_z_s, gc_s, n / 4);
    CNRF_LAUNCH_CHECK();
    add_inplace_kernel<<<grid_for(n / 4, 256), 256, 0, st>>>(g_z_t, gc_t, n / 4);
    CNRF_LAUNCH_CHECK();
  }
  return CNRF_OK;
}
}  // namespace

extern "C" int cnrf_render_bwd(const cnrf_render_shape* s, const cnrf_codenerf_params* coarse, const cnrf_codenerf_params* fine,
                               const cnrf_render_inputs* in, const float* g_rgb_coarse, const float* g_rgb_fine, const void* saved,
                               size_t saved_bytes, const cnrf_codenerf_grads* g_coarse, const cnrf_codenerf_grads* g_fine,
                               float* g_z_s, float* g_z_t, void* workspace, size_t workspace_bytes, cnrf_stream_t stream) {
  return render_bwd_impl(s, coarse, fine, in, g_rgb_coarse, g_rgb_fine, CompositeLossBwd(), CompositeLossBwd(), saved, saved_bytes,
                         g_coarse, g_fine, g_z_s, g_z_t, workspace, workspace_bytes, stream);
}

extern "C" int cnrf_render_loss_bwd(const cnrf_render_shape* s, const cnrf_codenerf_params* coarse, const cnrf_codenerf_params* fine,
                                    const cnrf_render_inputs* in, const cnrf_loss_target* target, const float* rgb_coarse,
                                    const float* rgb_fine, const float* g_loss_coarse, const float* g_loss_fine,
                                    const float* g_loss_sum, const void* saved, size_t saved_bytes,
                                    const cnrf_codenerf_grads* g_coarse, const cnrf_codenerf_grads* g_fine, float* g_z_s,
                                    float* g_z_t, void* workspace, size_t workspace_bytes, cnrf_stream_t stream) {
  CNRF_REQUIRE(s && target && rgb_coarse && rgb_fine, CNRF_E_NULL);
  CNRF_REQUIRE(g_loss_coarse || g_loss_fine || g_loss_sum, CNRF_E_NULL);
  CompositeLossBwd lbc, lbf;
  RUN(loss_target_of(target, lbc.t));
  lbf.t = lbc.t;
  lbc.rgb_out = rgb_coarse; lbc.g_term = g_loss_coarse; lbc.g_sum = g_loss_sum;
  lbf.rgb_out = rgb_fine; lbf.g_term = g_loss_fine; lbf.g_sum = g_loss_sum;
  lbc.scale = lbf.scale = 2.0f / (3.0f * (float)s->num_rays);
  return render_bwd_impl(s, coarse, fine, in, nullptr, nullptr, lbc, lbf, saved, saved_bytes, g_coarse, g_fine, g_z_s, g_z_t,
                         workspace, workspace_bytes, stream);
}
