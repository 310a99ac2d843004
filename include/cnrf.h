/*
 * cnrf.h -- C ABI of libcnrf.so, the sm_100a implementation of code-nerf's volumetric-rendering hot path.
 *
 * The reference (akashsharma02/code-nerf) is pure Python/PyTorch: it has no FFI.  Its "plugin boundary" for
 * this path is the Python call surface of view_synthesis.nerf / view_synthesis.models.
 * Each entry point below states the reference function (file:line, relative to the reference tree) whose
 * arithmetic it replaces; the drop-in Python package code-nerf_b200/view_synthesis binds them with ctypes
 * (INTEGRATION.md shows the stub).
 *
 * Conventions
 *   - extern "C", plain pointers and sizes; no torch / C++ types cross this boundary.
 *   - every pointer is a DEVICE pointer unless named host_*; the caller owns every buffer.
 *   - the library never allocates device memory, never synchronises, never throws; work is enqueued on
 *     `stream` (a cudaStream_t passed as void*).  Scratch space is passed in as `workspace`; its size comes
 *     from the matching *_workspace_bytes query.
 *   - return value: 0 = ok, < 0 = argument error (CNRF_E_*), > 0 = a cudaError_t from launch/enqueue.
 *   - all floating-point tensors are fp32, row-major, densely packed unless a stride argument says otherwise.
 *   - indices are int64 ("long") exactly as torch / numpy hand them over.
 */
#ifndef CNRF_H_
#define CNRF_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CNRF_ABI_VERSION 1

#define CNRF_OK 0
#define CNRF_E_NULL (-1)     /* a required pointer is NULL            */
#define CNRF_E_SHAPE (-2)    /* a size is out of the supported range  */
#define CNRF_E_WORKSPACE (-3)/* workspace / saved buffer too small    */
#define CNRF_E_UNSUPPORTED (-4)
#define CNRF_E_DEVICE (-5)   /* not an sm_100 device                  */
#define CNRF_E_SCRATCH_LIMIT (-6) /* cnrf_field_fwd_noalloc: the shape needs more than CNRF_NOALLOC_SCRATCH_BYTES */

typedef void* cnrf_stream_t; /* cudaStream_t */

int cnrf_abi_version(void);
const char* cnrf_error_string(int code);
/* 0 when the current device can run the tcgen05 kernels (compute capability 10.x), else CNRF_E_DEVICE. */
int cnrf_device_check(void);
/* Number of kernels this library has launched in the calling process (bench.py reports it as gpu_launches). */
int64_t cnrf_launch_count(void);

/* ------------------------------------------------------------------------------------------------------
 * Ray generation.  RaySampler.__init__ / get_bundle / sample, view_synthesis/nerf/ray_sampler.py:31-51,84-99,66-80.
 * directions[h,w] = ((w-cx)/f, -(h-cy)/f, -1) is evaluated on the fly (same IEEE ops as :44-51);
 * rd = R_b * directions, ro = t_b.  `select` == NULL renders all H*W pixels of every image (get_bundle),
 * otherwise select[b*n + i] is the flat pixel index h*W+w of ray i of image b (the numpy permutation of :71-75
 * stays on the host so the RNG stream is the reference's).  ro may be NULL (get_bundle returns a broadcast).
 * ---------------------------------------------------------------------------------------------------- */
int cnrf_ray_bundle_fwd(const float* pose /*[B,4,4]*/, int64_t B, int32_t H, int32_t W, float focal, float cx,
                        float cy, const int64_t* select /*[B,n] or NULL*/, int64_t n,
                        float* ro /*[B*n,3] or NULL*/, float* rd /*[B*n,3]*/, cnrf_stream_t stream);
/* d pose[b, :3, :3] and d pose[b, :3, 3]; g_pose [B,4,4] is fully written (row 3 = 0).  g_ro may be NULL. */
int cnrf_ray_bundle_bwd(const float* g_ro, const float* g_rd, int64_t B, int32_t H, int32_t W, float focal,
                        float cx, float cy, const int64_t* select, int64_t n, float* g_pose /*[B,4,4]*/,
                        cnrf_stream_t stream);

/* ------------------------------------------------------------------------------------------------------
 * Stratified sampling.  PointSampler.sample_uniform, view_synthesis/nerf/point_sampler.py:60-71.
 * z = lower + (upper-lower)*t_rand (t_rand != NULL, the rand_like draw of :64 supplied by the caller) or the
 * z table (perturb off); pts = ro + rd*z.
 * ---------------------------------------------------------------------------------------------------- */
int cnrf_sample_uniform_fwd(const float* ro, const float* rd, int64_t N, int32_t S, const float* z_table,
                            const float* lower, const float* upper, const float* t_rand /*[N,S] or NULL*/,
                            float* pts /*[N,S,3]*/, float* z /*[N,S]*/, cnrf_stream_t stream);
/* Backward of pts = ro + rd*z w.r.t. ro, rd (z carries no gradient: SURVEY F9).  Used by both samplers. */
int cnrf_points_bwd(const float* g_pts /*[N,S,3]*/, const float* z /*[N,S]*/, int64_t N, int32_t S,
                    float* g_ro /*[N,3]*/, float* g_rd /*[N,3]*/, cnrf_stream_t stream);

/* ------------------------------------------------------------------------------------------------------
 * Hierarchical sampling.  PointSampler.sample_pdf, view_synthesis/nerf/point_sampler.py:84-120.
 * weights: (N, Sc-2) with row stride w_stride (the caller passes the view weights[..., 1:-1], nerf/__init__.py:87).
 * u: (N, Nf) uniform draws of :93, or NULL for linspace(0,1,Nf) (:95-96).
 * Outputs: z [N,Sc+Nf] sorted merge of z_coarse and the new samples (:116), pts = ro + rd*z (:118);
 * optional debug outputs cdf [N,Sc-1] (:86-89) and inds [N,Nf] = searchsorted(cdf,u,right=True) (:101).
 * ---------------------------------------------------------------------------------------------------- */
int cnrf_sample_pdf_fwd(const float* ro, const float* rd, const float* weights, int64_t w_stride,
                        const float* z_coarse /*[N,Sc]*/, int64_t N, int32_t Sc, int32_t Nf, const float* u,
                        float* pts, float* z, float* cdf_out /*or NULL*/, int64_t* inds_out /*or NULL*/,
                        cnrf_stream_t stream);
/* The inverse-CDF step alone on a caller-supplied CDF tensor (point_sampler.py:98-113): lets tests check the
 * bin indices bit-exactly "given an identical CDF tensor".  cdf, bins: [N,C]; u, samples, inds: [N,K]. */
int cnrf_invert_cdf(const float* cdf, const float* bins, const float* u, int64_t N, int32_t C, int32_t K,
                    float* samples, int64_t* inds, cnrf_stream_t stream);

/* ------------------------------------------------------------------------------------------------------
 * Positional encoding.  PositionalEmbedder.embed, view_synthesis/nerf/position_embed.py:44-53.
 * out[m] = [x (if include_input) | sin(f0 x) | cos(f0 x) | sin(f1 x) | ...], every block D wide.
 * ---------------------------------------------------------------------------------------------------- */
int cnrf_pos_embed_fwd(const float* x /*[M,D]*/, int64_t M, int32_t D, const float* freqs /*[L]*/, int32_t L,
                       int32_t include_input, float* out /*[M, D*(include_input+2L)]*/, cnrf_stream_t stream);
int cnrf_pos_embed_bwd(const float* x, const float* g_out, int64_t M, int32_t D, const float* freqs, int32_t L,
                       int32_t include_input, float* g_x /*[M,D]*/, cnrf_stream_t stream);

/* ------------------------------------------------------------------------------------------------------
 * Alpha compositing.  volume_render, view_synthesis/nerf/volumetric_render.py:36-66.
 * raw [N,S,4] = (r,g,b,sigma) pre-activation; any output pointer may be NULL.
 * Backward: any upstream gradient may be NULL (treated as zero); g_z / g_rd may be NULL when not needed.
 * ---------------------------------------------------------------------------------------------------- */
int cnrf_composite_fwd(const float* raw, const float* z, const float* rd, int64_t N, int32_t S, float* rgb /*[N,3]*/,
                       float* disp /*[N]*/, float* acc /*[N]*/, float* weights /*[N,S]*/, float* depth /*[N]*/,
                       cnrf_stream_t stream);
int cnrf_composite_bwd(const float* raw, const float* z, const float* rd, int64_t N, int32_t S,
                       const float* g_rgb, const float* g_disp, const float* g_acc, const float* g_weights,
                       const float* g_depth, float* g_raw /*[N,S,4]*/, float* g_z /*[N,S] or NULL*/,
                       float* g_rd /*[N,3] or NULL*/, cnrf_stream_t stream);

/* ------------------------------------------------------------------------------------------------------
 * CodeNeRF radiance field.  CodeNeRFModel.forward, view_synthesis/models/model.py:160-194, and forward_pass,
 * view_synthesis/nerf/__init__.py:94-134.
 *
 * Layer order of the weight/bias pointer tables (state_dict names, models/model.py:145-156):
 * ---------------------------------------------------------------------------------------------------- */
#define CNRF_NUM_LAYERS 9
#define CNRF_L_XYZ1 0     /* layer_xyz1          [H, dim_xyz]      */
#define CNRF_L_XYZ2 1     /* layer_xyz2          [H, H+C]          */
#define CNRF_L_FC_OUT 2   /* fc_out              [C+1, H+C]        */
#define CNRF_L_SHAPE1 3   /* shape_code_layer1   [C, C]            */
#define CNRF_L_SHAPE2 4   /* shape_code_layer2   [C, C]            */
#define CNRF_L_TEXTURE1 5 /* texture_code_layer1 [C, C]            */
#define CNRF_L_DIR1 6     /* layer_dir1          [H, C+dim_dir]    */
#define CNRF_L_DIR2 7     /* layer_dir2          [H, H]            */
#define CNRF_L_FC_RGB 8   /* fc_rgb              [3, H+C]          */

#define CNRF_PRECISION_FP32 0 /* fp32 FMA on CUDA cores: the 1e-3 parity mode                                     */
#define CNRF_PRECISION_BF16 1 /* tcgen05.mma kind::f16, bf16 operands / fp32 TMEM accumulators: the throughput mode */
#define CNRF_PRECISION_SPLIT 2 /* the fp32 path (fp32 activations, same saved layout) with every contraction on the tensor
                                 * cores as split-bf16: operands carried as hi + lo bf16 images, hi*hi + hi*lo + lo*hi in fp32
                                 * TMEM (~16 mantissa bits).  Used for camera-pose optimisation (eval.py:129-168), where the
                                 * ray / point gradients need better than bf16 and the fp32 FMA path is 7x slower */

#define CNRF_PAIRING_REFERENCE 0 /* flat row k=n*S+s gets the view direction of ray k mod N (nerf/__init__.py:127) */
#define CNRF_PAIRING_CORRECT 1   /* ... of its own ray k / S                                                               */

typedef struct {
  const float* weight[CNRF_NUM_LAYERS]; /* torch.nn.Linear layout: [out][in] row-major */
  const float* bias[CNRF_NUM_LAYERS];
} cnrf_codenerf_params;

typedef struct {
  float* weight[CNRF_NUM_LAYERS]; /* same shapes as the parameters; every buffer is fully overwritten */
  float* bias[CNRF_NUM_LAYERS];
} cnrf_codenerf_grads;

typedef struct {
  int32_t hidden;       /* H                                                                    */
  int32_t code;         /* C = shape_code_size = texture_code_size (models/model.py:145-156)    */
  int32_t dim_xyz;      /* 3*(include_input + 2*L_xyz), 63 in every live config                 */
  int32_t dim_dir;      /* 27                                                                   */
  int64_t num_rays;     /* N: rows of z_s / z_t / view directions                               */
  int32_t num_samples;  /* S: samples per ray; the field has M = N*S rows                       */
  int32_t pairing;      /* CNRF_PAIRING_*                                                       */
  int32_t precision;    /* CNRF_PRECISION_*                                                     */
  int32_t keep;         /* 1: fill `saved` for a later backward, 0: inference                   */
} cnrf_field_shape;

size_t cnrf_codenerf_saved_bytes(const cnrf_field_shape* s);
size_t cnrf_codenerf_workspace_bytes(const cnrf_field_shape* s, int32_t backward);

/* The operator as the reference states it, on already-embedded inputs:
 *   x_xyz [M, dim_xyz] (row stride ld_xyz), x_dir [N, dim_dir] (row stride ld_dir; row r serves the samples the
 *   pairing assigns to ray r), z_s / z_t [N, C] with row strides (0 = one code broadcast to all rays, eval.py:149).
 * CodeNeRFModel.forward(z_s, z_t, x) with per-sample codes is the case N = M, S = 1 with x_xyz = x[:, :dim_xyz],
 * x_dir = x[:, dim_xyz:] (ld = dim_xyz+dim_dir).  raw [M,4] = [rgb_raw | sigma_raw] (models/model.py:194).
 * The latent-code layers run once per ray and enter the per-sample layers as biases (DESIGN.md "code folding");
 * results equal the as-written per-sample evaluation up to fp32 summation order. */
int cnrf_codenerf_fwd(const cnrf_field_shape* s, const cnrf_codenerf_params* p, const float* x_xyz, int64_t ld_xyz,
                      const float* x_dir, int64_t ld_dir, const float* z_s, int64_t ld_zs, const float* z_t,
                      int64_t ld_zt, float* raw /*[M,4]*/, void* saved, size_t saved_bytes, void* workspace,
                      size_t workspace_bytes, cnrf_stream_t stream);
/* g_z_s, g_z_t: dense [N, C].  g_x_xyz [M, dim_xyz] / g_x_dir [N, dim_dir] may be NULL when not needed.
 * g may be NULL in CNRF_PRECISION_FP32 / _SPLIT: a frozen network (eval.py:121-124 sets requires_grad_(False) on both models
 * before the test-time optimisation) -- no weight-gradient GEMM or reduction is launched; CNRF_PRECISION_BF16 returns CNRF_E_NULL. */
int cnrf_codenerf_bwd(const cnrf_field_shape* s, const cnrf_codenerf_params* p, const float* x_xyz, int64_t ld_xyz,
                      const float* x_dir, int64_t ld_dir, const float* z_s, int64_t ld_zs, const float* z_t,
                      int64_t ld_zt, const float* g_raw /*[M,4]*/, const void* saved, size_t saved_bytes,
                      const cnrf_codenerf_grads* g, float* g_z_s, float* g_z_t, float* g_x_xyz, float* g_x_dir,
                      void* workspace, size_t workspace_bytes, cnrf_stream_t stream);

/* forward_pass (nerf/__init__.py:116-134): positional-encode pts [N,S,3] (freqs_xyz, L_xyz) and the L2-normalised
 * view directions rd [N,3] (freqs_dir, L_dir), pair them (s->pairing), run the field.  include_input_* follow
 * dim = 3*(include_input + 2L).  In CNRF_PRECISION_BF16 the encoding is fused into the first tensor-core layer. */
size_t cnrf_field_workspace_bytes(const cnrf_field_shape* s, int32_t backward);
int cnrf_field_fwd(const cnrf_field_shape* s, const cnrf_codenerf_params* p, const float* pts, const float* rd,
                   const float* freqs_xyz, int32_t L_xyz, const float* freqs_dir, int32_t L_dir, const float* z_s,
                   int64_t ld_zs, const float* z_t, int64_t ld_zt, float* raw /*[N,S,4]*/, void* saved,
                   size_t saved_bytes, void* workspace, size_t workspace_bytes, cnrf_stream_t stream);
/* g_pts [N,S,3] / g_rd [N,3] may be NULL (training: rays carry no gradient; test-time pose optimisation needs them).
 * g may be NULL for a frozen network in CNRF_PRECISION_FP32 / _SPLIT (see cnrf_codenerf_bwd). */
int cnrf_field_bwd(const cnrf_field_shape* s, const cnrf_codenerf_params* p, const float* pts, const float* rd,
                   const float* freqs_xyz, int32_t L_xyz, const float* freqs_dir, int32_t L_dir, const float* z_s,
                   int64_t ld_zs, const float* z_t, int64_t ld_zt, const float* g_raw, const void* saved,
                   size_t saved_bytes, const cnrf_codenerf_grads* g, float* g_z_s, float* g_z_t, float* g_pts,
                   float* g_rd, void* workspace, size_t workspace_bytes, cnrf_stream_t stream);

/* forward_pass WITHOUT a caller workspace, for callers that cannot allocate (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid cudaMalloc).  This entry point has no counterpart in the reference.
 * It is cnrf_field_fwd with s->keep == 0 (inference; keep != 0 -> CNRF_E_UNSUPPORTED):
 * same kernels, same results.  Its scratch is a fixed __device__ array of CNRF_NOALLOC_SCRATCH_BYTES inside the library,
 * held by every GPU that loads libcnrf.so until the process exits; the call performs no cudaMalloc / cudaFree and calls no
 * allocating routine.  A shape with cnrf_field_workspace_bytes(s, 0) > CNRF_NOALLOC_SCRATCH_BYTES is rejected with
 * CNRF_E_SCRATCH_LIMIT before anything is launched (64 MiB holds e.g. 2048 rays x 128 samples in CNRF_PRECISION_BF16, the
 * validation chunk of config/srn-cars-code-3080-val.yml; split larger batches into chunks of rays).  All calls on one device
 * share the array and find the previous call's contents in it -- nothing needs clearing, every byte read is written
 * earlier in the same call -- so calls that OVERLAP on the GPU (two streams, two host threads) race: keep them on one
 * stream or order them with events.  The first call on a device looks the array's address up (cudaGetSymbolAddress): make
 * that call once outside any graph capture. */
#define CNRF_NOALLOC_SCRATCH_BYTES (64 * 1024 * 1024)
int cnrf_field_fwd_noalloc(const cnrf_field_shape* s, const cnrf_codenerf_params* p, const float* pts, const float* rd,
                           const float* freqs_xyz, int32_t L_xyz, const float* freqs_dir, int32_t L_dir, const float* z_s,
                           int64_t ld_zs, const float* z_t, int64_t ld_zt, float* raw /*[N,S,4]*/, cnrf_stream_t stream);

/* Shared latent codes: train.py:78-96 gives every ray of an image the image's object id and looks the codes up per ray
 * (ShapeTextureEmbedding.forward, models/model.py:100-103).  These variants take the ids and the two embedding TABLES
 * instead of per-ray code rows: ids [N] int64 (device), tab_s / tab_t [num_codes, C] (row strides ld_s / ld_t).  The code
 * layers and the gradients of their bias tables are evaluated once per run of equal consecutive ids (found on the device,
 * no host synchronisation).  g_tab_s / g_tab_t [num_codes, C] receive the dense table gradients (overwritten), i.e. what
 * torch.nn.Embedding's backward would have accumulated.  CNRF_PRECISION_BF16 only (CNRF_E_UNSUPPORTED otherwise); saved /
 * workspace sizes are those of cnrf_codenerf_saved_bytes / cnrf_field_workspace_bytes. */
int cnrf_field_fwd_ids(const cnrf_field_shape* s, const cnrf_codenerf_params* p, const float* pts, const float* rd,
                       const float* freqs_xyz, int32_t L_xyz, const float* freqs_dir, int32_t L_dir, const int64_t* ids,
                       int64_t num_codes, const float* tab_s, int64_t ld_s, const float* tab_t, int64_t ld_t,
                       float* raw /*[N,S,4]*/, void* saved, size_t saved_bytes, void* workspace, size_t workspace_bytes,
                       cnrf_stream_t stream);
int cnrf_field_bwd_ids(const cnrf_field_shape* s, const cnrf_codenerf_params* p, const float* pts, const float* rd,
                       const float* freqs_xyz, int32_t L_xyz, const float* freqs_dir, int32_t L_dir, const int64_t* ids,
                       int64_t num_codes, const float* tab_s, int64_t ld_s, const float* tab_t, int64_t ld_t,
                       const float* g_raw, const void* saved, size_t saved_bytes, const cnrf_codenerf_grads* g,
                       float* g_tab_s, float* g_tab_t, void* workspace, size_t workspace_bytes, cnrf_stream_t stream);

/* ------------------------------------------------------------------------------------------------------
 * Vanilla NeRF field.  FlexibleNeRFModel.forward (use_viewdirs=True), view_synthesis/models/model.py:51-74 -- SURVEY
 * section 8 row a15 / f1 (config/lego.yml: num_layers 8, hidden 256, skip_connect_ids [4]).
 *   h = relu(layer1 x_xyz); for i < num_layers-1: h = relu(layers_xyz[i] (skip i ? cat(h, x_xyz) : h));
 *   feat = relu(fc_feat h); sigma = fc_alpha feat; v = relu(layers_dir[0] cat(feat, x_dir)); rgb = fc_rgb v;
 *   raw = [rgb | sigma].
 * Inputs as cnrf_codenerf_fwd: x_xyz [M, dim_xyz] (row stride ld_xyz), x_dir [N, dim_dir] per RAY (row stride ld_dir; the
 * pairing assigns rows to samples -- the view-direction half of layers_dir[0] is evaluated once per ray and enters the
 * per-sample layer as a bias row), M = num_rays * num_samples.  FlexibleNeRFModel.forward(x) on an already-embedded
 * [M, dim_xyz+dim_dir] tensor is the case num_rays = M, num_samples = 1.  Every layer is one GEMM of the per-ray toolkit
 * (sgemm.cu in CNRF_PRECISION_FP32; split-bf16 tcgen05 in CNRF_PRECISION_BF16: gemm_stream.cu for the big per-sample
 * layers and their gradients, gemm_tc.cu for the rest); bf16 inference at hidden 256 runs as one fused tcgen05 chain
 * (flex_fwd_tc.cu).
 * dim_dir == 0 selects the use_viewdirs=False network (models/model.py:47-48,75-76): raw = fc_out h, fc_out [4,H] / [4]
 * passed (and its gradient returned) in w_rgb / b_rgb; x_dir may be NULL and fc_feat / layers_dir / fc_alpha are not read.
 * ---------------------------------------------------------------------------------------------------- */
#define CNRF_FLEX_MAX_XYZ_LAYERS 15
typedef struct {
  int32_t num_layers;   /* constructor argument: layer1 + (num_layers - 1) layers_xyz                    */
  int32_t hidden;       /* H, even (layers_dir[0] has H/2 outputs)                                         */
  int32_t dim_xyz;
  int32_t dim_dir;      /* 0: no view directions (fc_out head)                                             */
  uint32_t skip_mask;   /* bit i set: layers_xyz[i] takes cat(h, x_xyz) (models/model.py:33-37,61-64)      */
  int64_t num_rays;
  int32_t num_samples;
  int32_t pairing;      /* CNRF_PAIRING_*   */
  int32_t precision;    /* CNRF_PRECISION_* */
  int32_t keep;         /* 1: fill `saved` for a later backward */
} cnrf_flex_shape;
typedef struct {   /* torch.nn.Linear layout [out][in]; state_dict order of models/model.py:31-46 */
  const float *w_layer1, *b_layer1;
  const float* w_xyz[CNRF_FLEX_MAX_XYZ_LAYERS];
  const float* b_xyz[CNRF_FLEX_MAX_XYZ_LAYERS];
  const float *w_feat, *b_feat, *w_dir, *b_dir, *w_alpha, *b_alpha, *w_rgb, *b_rgb;
} cnrf_flex_params;
typedef struct {   /* same shapes as the parameters; every buffer is fully overwritten */
  float *w_layer1, *b_layer1;
  float* w_xyz[CNRF_FLEX_MAX_XYZ_LAYERS];
  float* b_xyz[CNRF_FLEX_MAX_XYZ_LAYERS];
  float *w_feat, *b_feat, *w_dir, *b_dir, *w_alpha, *b_alpha, *w_rgb, *b_rgb;
} cnrf_flex_grads;
size_t cnrf_flexnerf_saved_bytes(const cnrf_flex_shape* s);
size_t cnrf_flexnerf_workspace_bytes(const cnrf_flex_shape* s, int32_t backward);
int cnrf_flexnerf_fwd(const cnrf_flex_shape* s, const cnrf_flex_params* p, const float* x_xyz, int64_t ld_xyz,
                      const float* x_dir, int64_t ld_dir, float* raw /*[M,4]*/, void* saved, size_t saved_bytes,
                      void* workspace, size_t workspace_bytes, cnrf_stream_t stream);
/* Inference straight from the sample positions: forward_pass (nerf/__init__.py:116-134) embeds pts [M,3] (row stride
 * ld_pts) with PositionalEmbedder (position_embed.py:44-53: freqs_xyz [num_freqs], include_input) and feeds the result to
 * the model; here gamma(x) is produced inside the fused tcgen05 chain and never exists in HBM.  dim_xyz must equal
 * 3 * (include_input + 2 * num_freqs).  Only where cnrf_flexnerf_fused_inference(s) is 1 (CNRF_PRECISION_BF16, keep = 0,
 * hidden 256, dim_xyz <= 64, dim_dir > 0); CNRF_E_UNSUPPORTED otherwise -- embed and call cnrf_flexnerf_fwd then.
 * Workspace: cnrf_flexnerf_workspace_bytes(s, 0). */
int32_t cnrf_flexnerf_fused_inference(const cnrf_flex_shape* s);
int cnrf_flexnerf_fwd_pts(const cnrf_flex_shape* s, const cnrf_flex_params* p, const float* pts, int64_t ld_pts,
                          const float* freqs_xyz, int32_t num_freqs, int32_t include_input, const float* x_dir,
                          int64_t ld_dir, float* raw /*[M,4]*/, void* workspace, size_t workspace_bytes,
                          cnrf_stream_t stream);
/* g_x_xyz [M, dim_xyz] / g_x_dir [N, dim_dir] (dense) may be NULL when the inputs carry no gradient. */
int cnrf_flexnerf_bwd(const cnrf_flex_shape* s, const cnrf_flex_params* p, const float* x_xyz, int64_t ld_xyz,
                      const float* x_dir, int64_t ld_dir, const float* g_raw /*[M,4]*/, const void* saved,
                      size_t saved_bytes, const cnrf_flex_grads* g, float* g_x_xyz, float* g_x_dir, void* workspace,
                      size_t workspace_bytes, cnrf_stream_t stream);

/* ------------------------------------------------------------------------------------------------------
 * Optimiser step (SURVEY section 8 "next" row f3): torch.optim.AdamW as train.py:111-114 runs it, every parameter tensor
 * of the step in one launch.  `tensors` is a HOST array; all pointers inside are device fp32 buffers of `numel` elements.
 * step >= 1 is the 1-based step count of that tensor (bias correction).  Arithmetic per element:
 *   p -= lr*wd*p;  m += (g-m)(1-beta1);  v = beta2 v + (1-beta2) g^2;
 *   p -= lr/(1-beta1^step) * m / (sqrt(v)/sqrt(1-beta2^step) + eps)
 * ---------------------------------------------------------------------------------------------------- */
typedef struct {
  float* param; const float* grad; float* exp_avg; float* exp_avg_sq;
  int64_t numel;
  float lr, weight_decay, step;
} cnrf_adamw_tensor;
int cnrf_adamw_step(const cnrf_adamw_tensor* tensors, int32_t count, float beta1, float beta2, float eps,
                    cnrf_stream_t stream);
/* The same step with the step count on the DEVICE (what torch.optim.AdamW(capturable=True) does): *step_dev is
 * incremented by a one-thread launch and the bias corrections are computed from it inside the kernel, so the launch
 * sequence can be captured in a CUDA graph and replayed (tensors[i].step is ignored; every tensor shares the count). */
int cnrf_adamw_step_dev(const cnrf_adamw_tensor* tensors, int32_t count, float beta1, float beta2, float eps,
                        int64_t* step_dev, cnrf_stream_t stream);

/* ------------------------------------------------------------------------------------------------------
 * Device-resident dataset stage (reference: view_synthesis/datasets/dataset.py:70-74,90 decode a
 * PNG and compute image / 255.0 in float64 -> float32 on the host for every sample; train.py:77-80 gather the selected
 * target pixels).  The views live in HBM as the uint8 crops the PNGs hold:
 *   cnrf_u8_to_unit        out[i] = (float)((double)in[i] / 255.0), i < n      (bit-identical to the loader's arithmetic)
 *   cnrf_gather_pixels_u8  out[(b R + r) C_out + c] = views[view[b]][select[b R + r]][c] / 255.0, c < out_channels
 *                          (views: [n_views][hw][channels] bytes; select values outside [0, hw) set *bad (may be NULL)
 *                          and leave their rows untouched).
 * ---------------------------------------------------------------------------------------------------- */
int cnrf_u8_to_unit(const uint8_t* in, int64_t n, float* out, cnrf_stream_t stream);
int cnrf_gather_pixels_u8(const uint8_t* views, int64_t n_views, int64_t hw, int32_t channels, const int64_t* view,
                          const int64_t* select, int64_t B, int64_t R, int32_t out_channels, float* out, int32_t* bad,
                          cnrf_stream_t stream);

/* ------------------------------------------------------------------------------------------------------
 * Photometric loss of the training loop (train.py:103-104): with a = rgb_coarse, b = rgb_fine and
 * t = target pixels (rows of ld_* floats, first three columns used; RGBA targets: ld_t = 4),
 *   out3[0] = mean((a - t)^2), out3[1] = mean((b - t)^2), out3[2] = out3[0] + out3[1]          (one launch)
 * replaces 2 x torch.nn.functional.mse_loss + add.  Backward: g_loss_* are device scalars (NULL = zero), g_a / g_b dense
 * [N,3] (either may be NULL):  g_a = (g_loss_a + g_loss_sum) * 2 (a - t) / (3N), g_b likewise              (one launch)
 * ---------------------------------------------------------------------------------------------------- */
int cnrf_mse_pair_fwd(const float* rgb_a, int64_t ld_a, const float* rgb_b, int64_t ld_b, const float* target,
                      int64_t ld_t, int64_t N, float* out3, cnrf_stream_t stream);
int cnrf_mse_pair_bwd(const float* rgb_a, int64_t ld_a, const float* rgb_b, int64_t ld_b, const float* target,
                      int64_t ld_t, int64_t N, const float* g_loss_a, const float* g_loss_b, const float* g_loss_sum,
                      float* g_a, float* g_b, cnrf_stream_t stream);

/* ------------------------------------------------------------------------------------------------------
 * The whole render pipeline in one call per direction.  predict_radiance_and_render, view_synthesis/nerf/__init__.py:74-91:
 *   sample_uniform (:82) -> forward_pass(coarse) (:83) -> volume_render (:84) -> sample_pdf on weights[..., 1:-1] (:87) ->
 *   forward_pass(fine) (:88) -> volume_render (:89) -> (rgb_coarse, rgb_fine),
 * and its backward with respect to both models' parameters and the latent codes (the samples are detached, :87 /
 * point_sampler.py:115, so the two passes only meet in the code gradients).  Rays carry no gradient through these entry
 * points; callers that optimise the camera pose (eval.py:129-168) use the per-stage entry points above.
 * The kernels are exactly those of the per-stage entry points; what this removes is the host: two library calls per
 * chunk instead of ~58 launches driven through Python.
 *   t_rand [N,Sc] / u [N,Nf]: the torch.rand draws of point_sampler.py:64,93 (NULL = perturb off);
 *   codes: ids != NULL: z_s / z_t are the embedding TABLES [num_codes,C] and g_z_s / g_z_t receive dense table gradients
 *          (both models summed); ids == NULL: z_s / z_t are per-ray rows [N,C] (row stride 0 = one broadcast row) and
 *          g_z_s / g_z_t are dense [N,C];
 *   extras_fine: NULL, or [3,N] = disp, acc, depth of the fine pass (volumetric_render.py:62-66).
 * ---------------------------------------------------------------------------------------------------- */
typedef struct {
  int32_t hidden, code, dim_xyz, dim_dir;
  int64_t num_rays;
  int32_t num_coarse, num_fine; /* S_c and N_f (nerf/__init__.py:28-31) */
  int32_t pairing, precision;   /* CNRF_PAIRING_*, CNRF_PRECISION_* */
  int32_t keep;                 /* 1: fill `saved` for cnrf_render_bwd */
} cnrf_render_shape;

typedef struct {
  const float* ro;      /* [N,3] */
  const float* rd;      /* [N,3], not normalised */
  const float* z_table; /* [Sc] PointSampler.z_vals */
  const float* lower;   /* [Sc] */
  const float* upper;   /* [Sc] */
  const float* t_rand;  /* [N,Sc] or NULL */
  const float* u;       /* [N,Nf] or NULL */
  const float* freqs_xyz; int32_t L_xyz;
  const float* freqs_dir; int32_t L_dir;
  const int64_t* ids;   /* [N] or NULL */
  int64_t num_codes;
  const float* z_s; int64_t ld_zs;
  const float* z_t; int64_t ld_zt;
} cnrf_render_inputs;

size_t cnrf_render_saved_bytes(const cnrf_render_shape* s);
size_t cnrf_render_workspace_bytes(const cnrf_render_shape* s, int32_t backward);
int cnrf_render_fwd(const cnrf_render_shape* s, const cnrf_codenerf_params* coarse, const cnrf_codenerf_params* fine,
                    const cnrf_render_inputs* in, float* rgb_coarse /*[N,3]*/, float* rgb_fine /*[N,3]*/,
                    float* extras_fine /*[3,N] or NULL*/, void* saved, size_t saved_bytes, void* workspace,
                    size_t workspace_bytes, cnrf_stream_t stream);
/* g_rgb_coarse / g_rgb_fine [N,3]: either may be NULL (zero).  Every buffer of g_coarse / g_fine is fully overwritten. */
int cnrf_render_bwd(const cnrf_render_shape* s, const cnrf_codenerf_params* coarse, const cnrf_codenerf_params* fine,
                    const cnrf_render_inputs* in, const float* g_rgb_coarse, const float* g_rgb_fine, const void* saved,
                    size_t saved_bytes, const cnrf_codenerf_grads* g_coarse, const cnrf_codenerf_grads* g_fine,
                    float* g_z_s, float* g_z_t, void* workspace, size_t workspace_bytes, cnrf_stream_t stream);

/* ------------------------------------------------------------------------------------------------------
 * predict_radiance_and_render + the photometric loss of the training loop in one call per direction (SURVEY section 8
 * row f2; reference: train.py:77-80 gathers the selected target pixels, train.py:103-104 computes
 * mse_loss(rgb_coarse[..., :3], target[..., :3]) + mse_loss(rgb_fine[..., :3], target[..., :3])).  The squared errors are
 * accumulated in the epilogues of the two compositing kernels (warp sums added atomically: the last bits of the loss
 * depend on the summation order) and the last warp of the fine pass writes
 *   loss3 = (coarse_loss, fine_loss, coarse_loss + fine_loss);
 * the backward forms d loss / d rgb = (g_loss_term + g_loss_sum) * 2 (rgb - target) / (3 N) inside the compositing
 * backward from the forward's rgb outputs, so neither the [N,3] gradients nor (uint8 form) the gathered float targets exist.
 * Target of ray n: rows[n * ld_rows + c], c < 3 (rows already gathered; RGBA rows: ld_rows = 4), or, when views != NULL,
 * views[pixel[n] * channels + c] / 255.0 computed as the loader does (float64 division, float32 cast:
 * datasets/dataset.py:72,90) from uint8 views resident in device memory; pixel[n] = view index * H W + selected pixel.
 * Workspace / saved sizes are those of cnrf_render_fwd / _bwd.
 * ---------------------------------------------------------------------------------------------------- */
typedef struct {
  const float* rows; int64_t ld_rows;
  const uint8_t* views; const int64_t* pixel; int32_t channels;
} cnrf_loss_target;
int cnrf_render_loss_fwd(const cnrf_render_shape* s, const cnrf_codenerf_params* coarse, const cnrf_codenerf_params* fine,
                         const cnrf_render_inputs* in, const cnrf_loss_target* target, float* loss3 /*[3]*/,
                         float* rgb_coarse /*[N,3]*/, float* rgb_fine /*[N,3]*/, void* saved, size_t saved_bytes, void* workspace,
                         size_t workspace_bytes, cnrf_stream_t stream);
/* g_loss_*: device scalars, the gradients of the three numbers of loss3 (any may be NULL = 0, not all) */
int cnrf_render_loss_bwd(const cnrf_render_shape* s, const cnrf_codenerf_params* coarse, const cnrf_codenerf_params* fine,
                         const cnrf_render_inputs* in, const cnrf_loss_target* target, const float* rgb_coarse,
                         const float* rgb_fine, const float* g_loss_coarse, const float* g_loss_fine, const float* g_loss_sum,
                         const void* saved, size_t saved_bytes, const cnrf_codenerf_grads* g_coarse,
                         const cnrf_codenerf_grads* g_fine, float* g_z_s, float* g_z_t, void* workspace, size_t workspace_bytes,
                         cnrf_stream_t stream);

/* ------------------------------------------------------------------------------------------------------
 * Diagnostics (used by tests only).  One-CTA tcgen05 GEMM over the 128-byte-swizzled shared-memory image the fused
 * MLP kernels use.  mode 0: D[128,N] = A[128,K] B[N,K]^T (K-major operands); mode 1: D[128,N] = A[K,128]^T B[K,N]
 * (MN-major operands, the dW = dY^T X form).  *err is set to 1 if the kernel's bounded barrier wait timed out.
 * ---------------------------------------------------------------------------------------------------- */
/* Synchronises the device; returns 1 if a tensor-core kernel of this process hit its bounded barrier-wait timeout. */
int cnrf_debug_tc_error(void);
/* Phase timers of the forward-chain kernel (block 0, clock64 sums; 16 slots) and CUDA-event timing of its launches:
 * cnrf_debug_fwd_kernel_ms(min_rows > 0) starts timing every launch over >= min_rows rows (statistics reset),
 * (0) stops; each call returns the mean duration in ms of the launches timed so far. */
int cnrf_debug_fwd_times(long long* host_out16, int reset);
/* Timer builds (-DCNRF_FWD_TIMERS) only: clock64 stamps [tile 2][layer 4][event 8] of one steady-state tile pair of block 0
 * (events: mlp_fwd_tc.cu, g_fwd_trace); all zero in ordinary builds. */
int cnrf_debug_fwd_trace(long long* host_out128);  /* [64] pair trace, then [16 images][4 events] of its layer 1 */
/* The same for the vanilla-NeRF chain (flex_fwd_tc.cu, -DCNRF_FLEX_TIMERS): [18 layers][40 events] of block 0's 4th tile. */
int cnrf_debug_flex_trace(long long* host_out720);
/* ... and of the streaming GEMM (gemm_stream.cu, -DCNRF_STREAM_TIMERS): 64 stamps of block 0's 6th tile. */
int cnrf_debug_stream_trace(long long* host_out64);
double cnrf_debug_fwd_kernel_ms(int64_t min_rows, int64_t* launches, int64_t* rows);
/* The same for any of the three tensor-core kernels: which = 0 forward chain, 1 backward-data chain, 2 weight gradients;
 * min_rows < 0 only queries. */
double cnrf_debug_kernel_ms(int which, int64_t min_rows, int64_t* launches, int64_t* rows);
/* C[M,N] (row stride ldc) = sum_k A[m*a_rs + k*a_cs] * B[k*b_rs + n*b_cs] through the per-ray GEMM toolkit:
 * tc = 0 the fp32 CUDA-core kernel (sgemm.cu), tc = 1 the tcgen05 kernel (gemm_tc.cu, bf16 operands).  accumulate = 1
 * adds into C.  Lets tests pin every operand-majorness combination the per-ray layers use. */
int cnrf_debug_gemm(int32_t tc, const float* A, int64_t a_rs, int64_t a_cs, const float* B, int64_t b_rs, int64_t b_cs,
                    float* C, int64_t ldc, int64_t M, int32_t N, int64_t K, int32_t accumulate, cnrf_stream_t stream);
int cnrf_debug_umma(int32_t mode, const float* A, const float* B, float* D, int32_t N, int32_t K, int32_t* err,
                    cnrf_stream_t stream);
/* TMEM read-throughput probe: `blocks` CTAs of `warps` warps sweep all 512 TMEM columns of their lane quarter `iters`
 * times with tcgen05.ld (mode 0: .32x32b.x32 one load in flight, 1: .x32 two in flight, 2: .x64); clocks[b] = SM clocks of
 * block b, sink = blocks*warps*32 words of scratch.  Bytes read per block = warps * iters * 64 KB. */
/* Bulk-copy ingest probe: `blocks` CTAs each pull iters x bytes through a ring of `slots` cp.async.bulk copies in flight from
 * src (src_bytes long, walked cyclically; per_block_offset != 0 starts every block 1 MiB apart).  clocks[b] = SM clocks. */
int cnrf_debug_bulk_ld(int32_t blocks, int32_t slots, int32_t bytes, int32_t iters, const void* src, size_t src_bytes,
                       int32_t per_block_offset, long long* clocks, cnrf_stream_t stream);
/* mode 1: nthreads issuing threads with a ring of `slots` each; 2: `slots` copies per round on one mbarrier, one wait per
 * round; 3: as 2 with the wait on another warp.  clocks[b] covers iters copies (mode 1: per issuing thread). */
int cnrf_debug_bulk_ld2(int32_t blocks, int32_t slots, int32_t bytes, int32_t iters, int32_t mode, int32_t nthreads,
                        const void* src, size_t src_bytes, long long* clocks, cnrf_stream_t stream);
/* MMA operand-bandwidth probe: `blocks` CTAs each run `chunks` K-64 x N-256 chunks of bf16 tcgen05.mma (M = 128; n_cols = 128:
 * two MMAs per K step, 256: one) over static shared-memory operands while nprod (0-4) threads stream bulk copies of
 * copy_bytes into shared memory.  out[2b] = SM clocks of the MMA loop, out[2b+1] = copies producer 0 completed meanwhile. */
int cnrf_debug_umma_rate(int32_t blocks, int32_t n_cols, int32_t chunks, int32_t nprod, int32_t copy_bytes, const void* src,
                         size_t src_bytes, long long* out, cnrf_stream_t stream);
int cnrf_debug_tmem_ld(int32_t blocks, int32_t warps, int32_t iters, int32_t mode, long long* clocks, uint32_t* sink,
                       cnrf_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* CNRF_H_ */
