// Fused bf16 tensor-core chain of the CodeNeRF field (CNRF_PRECISION_BF16): the per-sample layers
//   h1 = relu(W1 gamma(x) + b1) -> h2 = relu(W2h h1 + beta2[ray]) -> v1 = relu(Wfh h2 + betaf[ray] + [Wdv|bd1][gamma(d)|1])
//   -> v2 = relu(Wd2 v1 + bd2) -> rgb = Wrv v2 + betar[ray], sigma = w_sigma h2 + betas[ray]
// (models/model.py:179-194, code-folded; fc_out folded into layer_dir1: Wfh = Wdf Wout'[:, :H], see codenerf.cu)
// as ONE persistent kernel per direction.  tcgen05.mma (kind::f16, bf16 operands, fp32 accumulators in TMEM), weights
// streamed by 1-D TMA bulk copies of pre-swizzled 32 KB chunk images into a shared-memory ring, activations never leave
// shared memory between layers (except the bf16 tile images saved for the backward by TMA bulk stores).
//
// Geometry: a CTA owns a PAIR of 128-row sample tiles (T0, T1) that march through the layers in lockstep so every weight
// chunk fetched from L2 feeds 256 rows.  TMEM: 2 x 256 fp32 columns (all 512).  Shared memory: 2 x 64 KB activation
// tiles (4 slabs of [128 rows x 64 bf16], 128-byte swizzle) + 3 x 32 KB weight ring.
// Warp roles: warp 0 = TMA producer, warp 1 = MMA issuer (one elected lane), warps 2-5 = epilogue of T0, 6-9 = epilogue
// of T1 (thread = one sample row; TMEM lane quarter = warp_id % 4).
#include <cstdlib>
#include <cstring>
#include "mlp_tc.cuh"

// ---- weight packing ------------------------------------------------------------------------------------------------
// chunk image element (row r in [0,256), col c in [0,64)) of chunk `ch`.
// wf: the folded layer Wfh = Wdf Wout'h as four K-slice partial products [4][256 x 256] fp32, added here in a fixed order
__device__ __forceinline__ float wfh_at(const float* __restrict__ wf, int i) {
  return ((wf[i] + wf[65536 + i]) + wf[2 * 65536 + i]) + wf[3 * 65536 + i];
}
__device__ __forceinline__ void pack_weights_body(const cnrf_codenerf_params& p, const float* __restrict__ wf, int dx, int dd,
                                                  const Packed& pk) {
  const int ch = blockIdx.x;                  // 0 .. FWD_CHUNKS+BWD_CHUNKS-1
  for (int i = blockIdx.y * blockDim.x + threadIdx.x; i < 256 * 64; i += gridDim.y * blockDim.x) {
    int r = i >> 6, c = i & 63;
    float v;
    uint8_t* base;
    if (ch < FWD_CHUNKS) {
      base = pk.fwd + (size_t)ch * CHUNK;
      if (ch == 0) {                          // layer_xyz1 [256, dx] with the bias in column 63 (the input's ones column)
        v = c < dx ? p.weight[CNRF_L_XYZ1][r * dx + c] : (c == 63 ? p.bias[CNRF_L_XYZ1][r] : 0.f);
      } else if (ch == EXT_CHUNK) {           // view-direction half of layer_dir1 + its bias (the extension's ones column)
        v = c < dd ? p.weight[CNRF_L_DIR1][r * (256 + dd) + 256 + c] : (c == dd ? p.bias[CNRF_L_DIR1][r] : 0.f);
      } else {
        int l = (ch - 1) >> 2, k = ((ch - 1) & 3) * 64 + c;   // l: 0 W2h, 1 Wfh, 2 Wd2
        if (l == 0) v = p.weight[CNRF_L_XYZ2][r * 512 + k];
        else if (l == 1) v = wfh_at(wf, r * 256 + k);
        else v = p.weight[CNRF_L_DIR2][r * 256 + k];
      }
    } else {
      int b = ch - FWD_CHUNKS;
      base = pk.bwd + (size_t)b * CHUNK;
      int l = b >> 2, n = (b & 3) * 64 + c;   // image[k = r][n]: l: 0 Wd2^T, 1 Wfh^T, 2 W2h^T
      if (b == BWD_X_CHUNK) {                 // W1^T: rows 64 q + n (input feature n), column c = hidden unit 64 q + c
        const int q = r >> 6, nin = r & 63;
        v = nin < dx ? p.weight[CNRF_L_XYZ1][(q * 64 + c) * dx + nin] : 0.f;
      } else if (l == 0) v = p.weight[CNRF_L_DIR2][n * 256 + r];
      else if (l == 1) v = wfh_at(wf, n * 256 + r);
      else v = p.weight[CNRF_L_XYZ2][n * 512 + r];
    }
    *reinterpret_cast<__nv_bfloat16*>(base + slab_offset(r, c)) = __float2bfloat16(v);
  }
  if (ch == 0 && blockIdx.y == 0) {
    for (int i = threadIdx.x; i < 256; i += blockDim.x) {
      pk.w_sigma[i] = p.weight[CNRF_L_FC_OUT][i];
      pk.b_dir2[i] = p.bias[CNRF_L_DIR2][i];
      pk.zeros[i] = 0.f;
      reinterpret_cast<__nv_bfloat16*>(pk.wsr_bf16)[i] = __float2bfloat16(p.weight[CNRF_L_FC_OUT][i]);
      for (int c = 0; c < 3; ++c) {
        pk.w_rgb[c * 256 + i] = p.weight[CNRF_L_FC_RGB][c * 512 + i];
        reinterpret_cast<__nv_bfloat16*>(pk.wsr_bf16)[(1 + c) * 256 + i] = __float2bfloat16(p.weight[CNRF_L_FC_RGB][c * 512 + i]);
      }
    }
  }
}
__global__ void pack_weights_kernel(cnrf_codenerf_params p, const float* __restrict__ wf, int dx, int dd, Packed pk) {
  pack_weights_body(p, wf, dx, dd, pk);
}
// both models of a render call in one launch (blockIdx.z = model)
struct PackPair { cnrf_codenerf_params p[2]; const float* wf[2]; Packed pk[2]; };
__global__ void pack_weights_pair_kernel(const __grid_constant__ PackPair a, int dx, int dd) {
  pack_weights_body(a.p[blockIdx.z], a.wf[blockIdx.z], dx, dd, a.pk[blockIdx.z]);
}

// ---- host side --------------------------------------------------------------------------------------------------------
__device__ int g_tc_error = 0;   // raised by a bounded barrier wait that timed out (read by cnrf_debug_tc_error)

struct TcChainArgs {   // mirrors the declaration in codenerf.cu
  int64_t N, M; int S, pairing, dx, dd;
  const cnrf_codenerf_params* p;
  const float* x_xyz; int64_t ld_xyz;
  const float* pts; const float* freqs_xyz; int L_xyz; int inc_xyz;
  const float *b2, *bf, *bos, *br;            // per-ray bias tables [N,256], [N,256] (folded layer_dir1), [N] (sigma), [N,3]
  const float* wf;                            // the four K-slices [4][256 x 256] of this call's fold Wfh = Wdf Wout'h
  const float* x_dir; int64_t ld_dir;         // gamma(d^) per ray [N, dd]
  const int32_t* gid;                         // shared-code mode: row of the bias tables that ray r uses (NULL: r)
  uint8_t* packed = nullptr;                  // != NULL: the weight images were packed there already (cnrf_tc_pack_pair)
  bool prezeroed = false;                     // backward: the caller zero-filled every accumulation target
};

static inline size_t up1k(size_t b) { return (b + 1023) & ~size_t(1023); }
static int sm_count() { return cnrf_sm_count(); }
static Packed carve_packed(uint8_t* base) {
  Packed pk;
  pk.fwd = base;
  pk.bwd = base + (size_t)FWD_CHUNKS * CHUNK;
  float* f = reinterpret_cast<float*>(base + (size_t)(FWD_CHUNKS + BWD_CHUNKS) * CHUNK);
  pk.w_sigma = f; pk.w_rgb = f + 256; pk.b_dir2 = f + 1024; pk.zeros = f + 1280;
  pk.wsr_bf16 = reinterpret_cast<uint16_t*>(f + 1536);
  return pk;
}
static int* tc_error_ptr() {
  static int* p[CNRF_MAX_DEVICES] = {};
  return cnrf_symbol_ptr(p, g_tc_error);
}

static inline int64_t tiles_of(int64_t M) { return (M + TILE - 1) / TILE; }
static inline int64_t pairs_of(int64_t M) { return (tiles_of(M) + 1) / 2; }
// per tile: the activation images; then per tile pair: the view-direction extension slab
// ... then the packed weight images of this call (the backward runs on the same weights and reuses them)
size_t cnrf_tc_saved_bytes(int64_t M) { return (size_t)tiles_of(M) * SAVED_PER_TILE + (size_t)pairs_of(M) * SLAB + PACKED_BYTES; }

// where the packed weight images of a call live inside its saved buffer (what cnrf_tc_chain_fwd / _bwd use by default)
uint8_t* cnrf_tc_packed_in_saved(void* saved, int64_t M) {
  uint8_t* saved_al = reinterpret_cast<uint8_t*>(up1k(reinterpret_cast<size_t>(saved)));
  return saved_al + (size_t)tiles_of(M) * SAVED_PER_TILE + (size_t)pairs_of(M) * SLAB;
}
size_t cnrf_tc_packed_bytes() { return PACKED_BYTES + 1024; }
// the weight images of two models in one launch; at0 / at1: 1024-byte aligned, cnrf_tc_packed_bytes() each
int cnrf_tc_pack_pair(const cnrf_codenerf_params* p0, const float* wf0, uint8_t* at0, const cnrf_codenerf_params* p1,
                      const float* wf1, uint8_t* at1, int dx, int dd, cudaStream_t st) {
  if ((reinterpret_cast<uintptr_t>(at0) | reinterpret_cast<uintptr_t>(at1)) & 1023) return CNRF_E_WORKSPACE;
  PackPair a;
  a.p[0] = *p0; a.p[1] = *p1; a.wf[0] = wf0; a.wf[1] = wf1; a.pk[0] = carve_packed(at0); a.pk[1] = carve_packed(at1);
  pack_weights_pair_kernel<<<dim3(FWD_CHUNKS + BWD_CHUNKS, 8, 2), 256, 0, st>>>(a, dx, dd);
  CNRF_LAUNCH_CHECK();
  return CNRF_OK;
}

// Extension slab of every tile pair (the K = 32 extra input columns of layer_dir1): row r, columns 32t .. 32t+31 =
// [gamma(d^ of the ray paired with sample row r of tile t) | 1 | 0].  The pairing reproduces nerf/__init__.py:127
// (flat row m -> ray m mod N) unless CNRF_PAIRING_CORRECT (m / S).
// One thread per 16-byte chunk (8 bf16 columns) of a row: a warp writes four whole 128-byte rows.
__global__ void build_ext_kernel(const float* __restrict__ x_dir, int64_t ld, int dd, int64_t N, int64_t M, int S,
                                 int pairing, uint8_t* __restrict__ ext, int64_t num_pairs) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= num_pairs * 1024) return;
  const int c = (int)(i & 7), r = (int)((i >> 3) & 127);
  const int t = c >> 2, j = c & 3;                    // tile of the pair, 8-column group of its 32 columns
  const int64_t pr = i >> 10;
  const int64_t m = (pr * 2 + t) * TILE + r;
  uint32_t w[4] = {0u, 0u, 0u, 0u};
  if (m < M) {
    int64_t ray;
    if (M < (1ll << 31)) ray = pairing == CNRF_PAIRING_REFERENCE ? (uint32_t)m % (uint32_t)N : (uint32_t)m / (uint32_t)S;
    else ray = pairing == CNRF_PAIRING_REFERENCE ? m % N : m / S;
    const float* src = x_dir + ray * ld;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int c0 = 8 * j + 2 * q, c1 = c0 + 1;
      const float lo = c0 < dd ? __ldg(src + c0) : (c0 == dd ? 1.0f : 0.0f);
      const float hi = c1 < dd ? __ldg(src + c1) : (c1 == dd ? 1.0f : 0.0f);
      w[q] = pack_bf16(lo, hi);
    }
  }
  uint8_t* row = ext + (size_t)pr * SLAB + r * 128;
  *reinterpret_cast<uint4*>(row + ((c ^ (r & 7)) << 4)) = make_uint4(w[0], w[1], w[2], w[3]);
}

int cnrf_tc_chain_fwd(const TcChainArgs& a, float* raw, void* saved, size_t saved_bytes, void* ws, size_t ws_bytes,
                      cudaStream_t st) {
  uint8_t* w = reinterpret_cast<uint8_t*>(up1k(reinterpret_cast<size_t>(ws)));
  if (w + PACKED_BYTES + 1024 + (size_t)pairs_of(a.M) * SLAB > reinterpret_cast<uint8_t*>(ws) + ws_bytes) return CNRF_E_WORKSPACE;
  if (saved && saved_bytes < cnrf_tc_saved_bytes(a.M) + 1024) return CNRF_E_WORKSPACE;
  if (a.dd > 31) return CNRF_E_UNSUPPORTED;
  uint8_t* saved_al = saved ? reinterpret_cast<uint8_t*>(up1k(reinterpret_cast<size_t>(saved))) : nullptr;
  Packed pk = carve_packed(a.packed ? a.packed
                                    : saved_al ? saved_al + (size_t)tiles_of(a.M) * SAVED_PER_TILE + (size_t)pairs_of(a.M) * SLAB : w);
  // the extension slabs live with the saved images when a backward follows (the weight-gradient kernel reads them)
  uint8_t* ext = saved_al ? saved_al + (size_t)tiles_of(a.M) * SAVED_PER_TILE
                          : reinterpret_cast<uint8_t*>(up1k(reinterpret_cast<size_t>(w + PACKED_BYTES)));
  build_ext_kernel<<<grid_for(pairs_of(a.M) * 1024, 256), 256, 0, st>>>(a.x_dir, a.ld_dir, a.dd, a.N, a.M, a.S, a.pairing,
                                                                      ext, pairs_of(a.M));
  CNRF_LAUNCH_CHECK();
  if (!a.packed) {
    pack_weights_kernel<<<dim3(FWD_CHUNKS + BWD_CHUNKS, 8), 256, 0, st>>>(*a.p, a.wf, a.dx, a.dd, pk);
    CNRF_LAUNCH_CHECK();
  }
  FwdArgs f;
  f.N = a.N; f.M = a.M; f.S = a.S; f.pairing = a.pairing; f.dx = a.dx;
  f.pts = a.pts; f.x_xyz = a.x_xyz; f.ld_xyz = a.ld_xyz; f.freqs = a.freqs_xyz; f.L = a.L_xyz; f.inc = a.inc_xyz;
  f.b2 = a.b2; f.bf = a.bf; f.bos = a.bos; f.br = a.br; f.gid = a.gid; f.ext = ext; f.pk = pk; f.raw = raw;
  f.saved = saved_al;
  f.err = tc_error_ptr();
  const int64_t pairs = ((a.M + TILE - 1) / TILE + 1) / 2;
  const int grid = (int)(pairs < sm_count() ? pairs : sm_count());
  return cnrf_tc_launch_fwd(f, grid, st);
}

// Synchronous diagnostic: 1 if any tensor-core kernel of this process hit its barrier-wait timeout.
extern "C" int cnrf_debug_tc_error(void) {
  int v = 0;
  cudaDeviceSynchronize();
  cudaMemcpyFromSymbol(&v, g_tc_error, sizeof(int));
  return v | cnrf_flex_fused_error() | cnrf_gemm_stream_error();
}

// =====================================================================================================================
// Weight gradients: dW[n][k] = sum_m dY[m][n] X[m][k] on tensor cores, operands read as MN-major straight from the saved
// bf16 images (the sample index is the contraction dimension, so the very same image is its own transpose).
// 1-D grid split over six jobs; a CTA accumulates its share of the tiles in TMEM (2 x [128 x N] fp32) and adds it to
// the fp32 gradient with red.global.add.
// =====================================================================================================================
constexpr int DW_THREADS = 192;
constexpr int DW_STAGE = 8 * 8192;          // 64 rows x (256 + 256) bf16
constexpr int DW_RING = 3;
constexpr int DW_SMEM = DW_RING * DW_STAGE + 256 + 2 * 64 * 16;   // ring + barriers + raw-gradient staging

// Reductions the four otherwise idle epilogue warps run on every stage while it sits in shared memory (thread = column
// pair of the 256-wide A image): column sums (a bias gradient), per-ray sums (the gradient of a folded per-ray bias
// table), the same plus the sigma-weighted column sums of the B image (fc_out row 0), or the three rgb-weighted column
// sums of a streamed-only image (fc_rgb).
enum { RED_NONE = 0, RED_COLSUM = 1, RED_RAY = 2, RED_RAY_SIGMA = 3, RED_RGB = 4 };
constexpr int DW_JOBS = 6;
struct DwJob { int dy_img;   /* >= 0: image of the gradient set; < 0: image -1-dy_img of the saved set */
               int x_img;    /* -1: gamma slab, -2: view-direction extension slab of the tile pair */
               int n_cols;   /* 0: no GEMM, reduction only */
               float* out; int ld; int k_valid; float* bias_out; int bias_col;
               int red; float* red_out; int red_ld; float* red2_out; };
struct DwArgs {
  int64_t M;
  int S;
  const uint8_t* saved;
  const uint8_t* dy;
  const uint8_t* ext;
  const float* g_raw;
  const int32_t* gid;      // shared-code mode: row of the per-ray tables that ray r accumulates into (NULL: r)
  float* d_br;             // [N][3] per-ray sums of the rgb gradients   } accumulated by the RED_RGB job
  float* d_bo;             // [N][257], column 0 = per-ray sums of the sigma gradients
  DwJob job[DW_JOBS];
  int first[DW_JOBS + 1];          // first CTA of each job (1-D grid)
  int* err;
};

struct DwBars { uint64_t full[DW_RING], empty[DW_RING], acc_full; uint32_t tmem_slot; int abort_flag; };

__device__ __forceinline__ bool dw_wait(uint64_t* bar, uint32_t parity, DwBars* b, int* err) {
  if (mbar_try_wait(bar, parity)) return true;
  const long long t0 = clock64();
  while (true) {
    if (mbar_try_wait(bar, parity)) return true;
    if (*(volatile int*)&b->abort_flag) return false;
    if (clock64() - t0 > 2000000000LL) { b->abort_flag = 1; atomicExch(err, 1); tc_fatal(); return false; }
  }
}

// ---- reducer-warp helpers: 8 bf16 columns (one 16-byte chunk) of a row --------------------------------------------------
__device__ __forceinline__ void acc8(float* s, const uint4 v) {
  const uint32_t u[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
  for (int j = 0; j < 4; ++j) { s[2 * j] += __uint_as_float(u[j] << 16); s[2 * j + 1] += __uint_as_float(u[j] & 0xFFFF0000u); }
}
__device__ __forceinline__ void fma8(float* s, float g, const uint4 v) {
  const uint32_t u[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    s[2 * j] = fmaf(g, __uint_as_float(u[j] << 16), s[2 * j]);
    s[2 * j + 1] = fmaf(g, __uint_as_float(u[j] & 0xFFFF0000u), s[2 * j + 1]);
  }
}
// Sum the 8 partial columns of the four row phases (lane bits 3, 4) and leave two of them in every lane; returns the
// column (within the 64-column slab) of o0, o1 = that column + 1.
__device__ __forceinline__ int rs_rows(const float* s, int lane, float& o0, float& o1) {
  const bool b1 = (lane & 16) != 0, b0 = (lane & 8) != 0;
  float k[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const float snd = b1 ? s[j] : s[4 + j], kp = b1 ? s[4 + j] : s[j];
    k[j] = kp + __shfl_xor_sync(0xFFFFFFFFu, snd, 16);
  }
  {
    const float snd = b0 ? k[0] : k[2], kp = b0 ? k[2] : k[0];
    o0 = kp + __shfl_xor_sync(0xFFFFFFFFu, snd, 8);
  }
  {
    const float snd = b0 ? k[1] : k[3], kp = b0 ? k[3] : k[1];
    o1 = kp + __shfl_xor_sync(0xFFFFFFFFu, snd, 8);
  }
  return (lane & 7) * 8 + (b1 ? 4 : 0) + (b0 ? 2 : 0);
}

__global__ void __launch_bounds__(DW_THREADS, 1) mlp_dw_tc_kernel(const DwArgs a) {
  extern __shared__ __align__(1024) uint8_t smem[];
  DwBars* bars = reinterpret_cast<DwBars*>(smem + DW_RING * DW_STAGE);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  int ji = 0;
  while (ji < DW_JOBS - 1 && (int)blockIdx.x >= a.first[ji + 1]) ++ji;
  const DwJob job = a.job[ji];
  const int bx = blockIdx.x - a.first[ji], nb = a.first[ji + 1] - a.first[ji];
  const bool is_ext = job.x_img == -2;
  const bool has_mma = job.n_cols > 0;
  const int nslabB = !has_mma ? 0 : job.x_img < 0 ? 1 : 4;
  const int64_t num_tiles = (a.M + TILE - 1) / TILE;
  const int64_t num_halves = num_tiles * 2;
  if (tid == 0) {
    for (int i = 0; i < DW_RING; ++i) {
      mbar_init(&bars->full[i], 1);
      mbar_init(&bars->empty[i], (has_mma ? 1 : 0) + (job.red != RED_NONE ? 4 : 0));   // MMA commit + one per reducer warp
    }
    mbar_init(&bars->acc_full, 1);
    bars->abort_flag = 0;
    if ((smem_u32(smem) & 1023u) != 0) { bars->abort_flag = 1; atomicExch(a.err, 2); tc_fatal(); }
    mbar_fence_init();
  }
  if (warp == 1) tmem_alloc(&bars->tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = bars->tmem_slot;
  const bool has_work = (int64_t)bx < num_halves;

  if (warp == 0) {
    if (lane == 0) {
      uint32_t g = 0;
      for (int64_t h = bx; h < num_halves; h += nb, ++g) {
        const int slot = g % DW_RING;
        if (!dw_wait(&bars->empty[slot], ((g / DW_RING) & 1) ^ 1, bars, a.err)) goto done;
        const int64_t tile = h >> 1;
        const int half = (int)(h & 1);
        const uint8_t* src_a = job.dy_img >= 0
            ? a.dy + (size_t)tile * DY_PER_TILE + (size_t)job.dy_img * TILE_BYTES + half * 8192
            : a.saved + (size_t)tile * SAVED_PER_TILE + SAVED_GAMMA + (size_t)(-1 - job.dy_img) * TILE_BYTES + half * 8192;
        const uint8_t* src_b = is_ext ? a.ext + (size_t)(tile >> 1) * SLAB + half * 8192
                                      : a.saved + (size_t)tile * SAVED_PER_TILE +
                                            (job.x_img < 0 ? 0 : SAVED_GAMMA + (size_t)job.x_img * TILE_BYTES) + half * 8192;
        uint8_t* dst = smem + slot * DW_STAGE;
        mbar_arrive_expect_tx(&bars->full[slot], (4 + nslabB) * 8192);
        for (int s = 0; s < 4; ++s) bulk_g2s(dst + s * 8192, src_a + (size_t)s * SLAB, 8192, &bars->full[slot]);
        for (int s = 0; s < nslabB; ++s) bulk_g2s(dst + (4 + s) * 8192, src_b + (size_t)s * SLAB, 8192, &bars->full[slot]);
      }
    }
  } else if (warp == 1) {
    if (lane == 0 && has_work && has_mma) {
      const uint32_t idesc = idesc_bf16(128, job.n_cols, 1, 1);
      uint32_t g = 0;
      uint32_t seen = 0;               // extension job: tile parities whose accumulator pair has been started
      for (int64_t h = bx; h < num_halves; h += nb, ++g) {
        const int slot = g % DW_RING;
        if (!dw_wait(&bars->full[slot], (g / DW_RING) & 1, bars, a.err)) goto done;
        tc_fence_after();
        const uint32_t sa = smem_u32(smem + slot * DW_STAGE), sb = sa + 4 * 8192;
        // the extension slab serves both tiles of a pair (columns 32 t ..): one accumulator pair per tile parity
        const uint32_t par = is_ext ? (uint32_t)((h >> 1) & 1) : 0u;
        const bool fresh = is_ext ? ((seen >> par) & 1u) == 0u : g == 0;
        seen |= 1u << par;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const uint64_t bd = smem_desc(sb + k * 2048, 8192, 1024, SWIZZLE_128B);
          umma_bf16(tmem + par * 64, smem_desc(sa + k * 2048, 8192, 1024, SWIZZLE_128B), bd, idesc, !(fresh && k == 0));
          umma_bf16(tmem + 256 + par * 64, smem_desc(sa + 2 * 8192 + k * 2048, 8192, 1024, SWIZZLE_128B), bd, idesc,
                    !(fresh && k == 0));
        }
        umma_commit(&bars->empty[slot]);
      }
      umma_commit(&bars->acc_full);
    }
  } else if (has_work) {
    const int q = warp & 3;
    if (job.red != RED_NONE) {
      // warp = 64-column slab of the image, lane = (16-byte chunk c, row phase ph): a thread reads rows ph, ph+4, ... of
      // its 8 columns with one conflict-free LDS.128 per row; the four row phases meet in a shuffle reduce-scatter
      const int rw = warp - 2, w = tid - 64;
      const int c = lane & 7, ph = lane >> 3;
      const bool needs_g = job.red == RED_RAY_SIGMA || job.red == RED_RGB;
      float4* gbuf = reinterpret_cast<float4*>(smem + DW_RING * DW_STAGE + 256);      // [2][64] raw gradients of a stage
      float c0[8], c1[8], c2[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) c0[j] = c1[j] = c2[j] = 0.f;
      // per-ray sums stay in registers while consecutive rays (and stages) add into the same table row -- one flush per
      // ray in the plain mode, one per run of rays that share a code row in the shared-code mode
      float sum[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) sum[j] = 0.f;
      float tsum = 0.f;                    // RED_RGB: this lane's part of the per-ray raw-gradient sum, component rw
      int64_t cur_row = -1;
      auto flush_row = [&]() {
        if (job.red == RED_RGB) {
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) tsum += __shfl_xor_sync(0xFFFFFFFFu, tsum, o);
          if (lane == 0) atomicAdd(rw < 3 ? a.d_br + cur_row * 3 + rw : a.d_bo + cur_row * 257, tsum);
          tsum = 0.f;
        } else {
          float o0, o1;
          const int col = rw * 64 + rs_rows(sum, lane, o0, o1);
          float* o = job.red_out + cur_row * job.red_ld + col;
          atomicAdd(o, o0); atomicAdd(o + 1, o1);
#pragma unroll
          for (int j = 0; j < 8; ++j) sum[j] = 0.f;
        }
      };
      auto enter_ray = [&](int64_t row) {   // uniform over the warp
        if (row != cur_row) {
          if (cur_row >= 0) flush_row();
          cur_row = row;
        }
      };
      float4 gpre = make_float4(0.f, 0.f, 0.f, 0.f);
      if (needs_g && w < 64 && (int64_t)bx * 64 + w < a.M) gpre = __ldg(reinterpret_cast<const float4*>(a.g_raw) + (int64_t)bx * 64 + w);
      uint32_t g = 0;
      for (int64_t h = bx; h < num_halves; h += nb, ++g) {
        const int slot = g % DW_RING;
        const int64_t m0 = h * 64;
        const int rows = (int)(a.M - m0 < 64 ? (a.M - m0 > 0 ? a.M - m0 : 0) : 64);
        const float4* gs = gbuf + (g & 1) * 64;
        if (needs_g) {       // raw gradients of this stage (fetched one stage ahead) -> shared memory
          if (w < 64) gbuf[(g & 1) * 64 + w] = gpre;
          named_bar_sync(2, 128);
          const int64_t mn = (h + nb) * 64 + w;
          gpre = make_float4(0.f, 0.f, 0.f, 0.f);
          if (w < 64 && h + nb < num_halves && mn < a.M) gpre = __ldg(reinterpret_cast<const float4*>(a.g_raw) + mn);
        }
        // first ray of the stage and the row at which it ends (32-bit division when the row count allows)
        const int64_t ray0 = a.M < (1ll << 31) ? (int64_t)((uint32_t)m0 / (uint32_t)a.S) : m0 / a.S;
        const int seg_end0 = (int)((ray0 + 1) * a.S - m0);
        const int64_t row0 = (a.gid && rows > 0) ? (int64_t)__ldg(a.gid + ray0) : ray0;
        if (!dw_wait(&bars->full[slot], (g / DW_RING) & 1, bars, a.err)) goto done;
        const uint8_t* sa = smem + slot * DW_STAGE + rw * 8192;
        // rows ph + 4 i of this thread: chunk (c ^ (row & 7)) alternates between two values
        const uint8_t* sp = sa + ph * 128;
        const uint32_t x0 = (uint32_t)((c ^ ph) << 4), x1 = x0 ^ 64u;
        const bool full_stage = rows == 64;
#define DW_ROW(i) (sp + (i) * 512 + (((i) & 1) ? x1 : x0))
        if (job.red == RED_COLSUM) {
          if (full_stage) {
#pragma unroll
            for (int i = 0; i < 16; ++i) acc8(c0, *reinterpret_cast<const uint4*>(DW_ROW(i)));
          } else {
            for (int i = 0; ph + 4 * i < rows; ++i) acc8(c0, *reinterpret_cast<const uint4*>(DW_ROW(i)));
          }
        } else if (job.red == RED_RGB) {
          // the rgb-weighted column sums do not care about ray boundaries
          if (full_stage) {
#pragma unroll
            for (int i = 0; i < 16; ++i) {
              const uint4 v = *reinterpret_cast<const uint4*>(DW_ROW(i));
              const float4 gr = gs[ph + 4 * i];
              fma8(c0, gr.x, v); fma8(c1, gr.y, v); fma8(c2, gr.z, v);
            }
          } else {
            for (int i = 0; ph + 4 * i < rows; ++i) {
              const uint4 v = *reinterpret_cast<const uint4*>(DW_ROW(i));
              const float4 gr = gs[ph + 4 * i];
              fma8(c0, gr.x, v); fma8(c1, gr.y, v); fma8(c2, gr.z, v);
            }
          }
          // per-ray sums of the raw gradients themselves: warp rw takes component rw, a lane rows lane and lane + 32
          const float* gf = reinterpret_cast<const float*>(gs);
          int64_t ray = ray0;
          int seg_end = seg_end0;
          int r0 = 0;
          if (rows > 0) enter_ray(row0);
          while (r0 < rows) {
            const int r1 = seg_end < rows ? seg_end : rows;
            if (lane >= r0 && lane < r1) tsum += gf[lane * 4 + rw];
            if (lane + 32 >= r0 && lane + 32 < r1) tsum += gf[(lane + 32) * 4 + rw];
            r0 = r1; ++ray; seg_end += a.S;
            if (r0 < rows) enter_ray(a.gid ? (int64_t)__ldg(a.gid + ray) : ray);
          }
        } else {
          if (job.red == RED_RAY_SIGMA) {       // sigma-weighted column sums of the B image: no ray boundaries either
            if (full_stage) {
#pragma unroll
              for (int i = 0; i < 16; ++i) fma8(c0, gs[ph + 4 * i].w, *reinterpret_cast<const uint4*>(DW_ROW(i) + 4 * 8192));
            } else {
              for (int i = 0; ph + 4 * i < rows; ++i)
                fma8(c0, gs[ph + 4 * i].w, *reinterpret_cast<const uint4*>(DW_ROW(i) + 4 * 8192));
            }
          }
          // per-ray sums of the A image, 16 rows at a time; a ray boundary inside a 16-row group takes the slow path
          int64_t ray = ray0;
          int next_b = seg_end0;
          if (rows > 0) enter_ray(row0);
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const int lo = 16 * q, hi = rows < lo + 16 ? rows : lo + 16;
            if (lo < rows) {
              if (next_b >= hi && hi - lo == 16) {
#pragma unroll
                for (int i = 0; i < 4; ++i) acc8(sum, *reinterpret_cast<const uint4*>(DW_ROW(4 * q + i)));
              } else {
                int cur = lo;
                while (cur < hi) {
                  const int r1 = next_b < hi ? next_b : hi;
#pragma unroll
                  for (int i = 0; i < 4; ++i) {
                    const int r = lo + ph + 4 * i;
                    if (r >= cur && r < r1) acc8(sum, *reinterpret_cast<const uint4*>(DW_ROW(4 * q + i)));
                  }
                  cur = r1;
                  if (cur == next_b && cur < hi) {
                    ++ray; next_b += a.S;
                    enter_ray(a.gid ? (int64_t)__ldg(a.gid + ray) : ray);
                  }
                }
              }
              if (next_b == hi) {                // the ray ends with this group
                ++ray; next_b += a.S;
                if (hi < rows) enter_ray(a.gid ? (int64_t)__ldg(a.gid + ray) : ray);
              }
            }
          }
        }
#undef DW_ROW
        __syncwarp();
        if (lane == 0) mbar_arrive(&bars->empty[slot]);
      }
      if (cur_row >= 0) flush_row();
      float o0, o1;
      const int col = rw * 64 + rs_rows(c0, lane, o0, o1);
      float* fo = job.red == RED_RAY_SIGMA ? job.red2_out : job.red_out;
      if (job.red != RED_RAY) { atomicAdd(fo + col, o0); atomicAdd(fo + col + 1, o1); }
      if (job.red == RED_RGB) {
        rs_rows(c1, lane, o0, o1);
        atomicAdd(fo + job.red_ld + col, o0); atomicAdd(fo + job.red_ld + col + 1, o1);
        rs_rows(c2, lane, o0, o1);
        atomicAdd(fo + 2 * job.red_ld + col, o0); atomicAdd(fo + 2 * job.red_ld + col + 1, o1);
      }
    }
    if (has_mma && dw_wait(&bars->acc_full, 0, bars, a.err)) {
      tc_fence_after();
      uint32_t seen = 0;
      if (is_ext) for (int64_t h = bx; h < num_halves; h += nb) seen |= 1u << ((h >> 1) & 1);
      for (int mh = 0; mh < 2; ++mh) {
        const int n = mh * 128 + q * 32 + lane;
        if (is_ext) {                  // tile parity p: columns 32 p .. 32 p + 31 of its own accumulator pair
          for (int p = 0; p < 2; ++p) {
            if (!((seen >> p) & 1u)) continue;
            uint32_t v[32];
            tmem_ld32(tmem + mh * 256 + ((uint32_t)(q * 32) << 16) + p * 64 + p * 32, v);
            tmem_ld_wait();
#pragma unroll
            for (int k = 0; k < 32; ++k) {
              if (k < job.k_valid) atomicAdd(job.out + (size_t)n * job.ld + k, __uint_as_float(v[k]));
              else if (k == job.bias_col) atomicAdd(job.bias_out + n, __uint_as_float(v[k]));
            }
          }
          continue;
        }
        for (int c0 = 0; c0 < job.n_cols; c0 += 32) {
          uint32_t v[32];
          tmem_ld32(tmem + mh * 256 + ((uint32_t)(q * 32) << 16) + c0, v);
          tmem_ld_wait();
          float* orow = job.out + (size_t)n * job.ld + c0;
          // 16-byte vector reductions (red.global.add.v4.f32) wherever the 32 columns are all real and the row is aligned:
          // a quarter of the L2 atomic operations of this epilogue, which is most of the launch's fixed cost (~65 us of a
          // 160 us coarse launch before; every CTA of a job adds a full [256 x N] partial into the same gradient)
          if (c0 + 32 <= job.k_valid && (reinterpret_cast<uintptr_t>(orow) & 15) == 0) {
#pragma unroll
            for (int j = 0; j < 32; j += 4)
              atomicAdd(reinterpret_cast<float4*>(orow + j), make_float4(__uint_as_float(v[j]), __uint_as_float(v[j + 1]),
                                                                         __uint_as_float(v[j + 2]), __uint_as_float(v[j + 3])));
            continue;
          }
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const int k = c0 + j;
            if (k < job.k_valid) atomicAdd(job.out + (size_t)n * job.ld + k, __uint_as_float(v[j]));
            else if (job.bias_out && k == job.bias_col) atomicAdd(job.bias_out + n, __uint_as_float(v[j]));
          }
        }
      }
    }
  }
done:
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 512);
}

// out[g*ld + col] = sum_{q<Q} img[g*g_stride + q*q_stride][col]: per-ray sums of a gradient image whose rows of one ray are
// NOT consecutive (the reference's view-direction pairing, row m <-> ray m mod N) -- only the pose-optimisation backward
// needs it (gradient of the per-ray view-direction encoding); the consecutive-row sums live in the dW reducers.
__global__ void img_group_sum_kernel(const uint8_t* base, size_t per_tile, size_t off, int64_t G, int64_t g_stride,
                                     int64_t Q, int64_t q_stride, float* out, int64_t ld) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= G * 128) return;
  const int64_t g = i >> 7;
  const int w = (int)(i & 127);                       // column pair 2w, 2w+1
  float sx = 0.f, sy = 0.f;
  for (int64_t q = 0; q < Q; ++q) {
    const int64_t m = g * g_stride + q * q_stride;
    const int r = (int)(m & 127);
    const uint8_t* p = base + (size_t)(m >> 7) * per_tile + off + (w >> 5) * SLAB + r * 128 + ((((w >> 2) & 7) ^ (r & 7)) << 4) +
                       ((w & 3) << 2);
    const uint32_t v = __ldg(reinterpret_cast<const uint32_t*>(p));
    sx += __uint_as_float(v << 16); sy += __uint_as_float(v & 0xFFFF0000u);
  }
  out[g * ld + 2 * w] = sx;
  out[g * ld + 2 * w + 1] = sy;
}

size_t cnrf_tc_workspace_bytes(int64_t N, int64_t M, int backward) {
  (void)N;
  size_t b = PACKED_BYTES + 4096 + (size_t)pairs_of(M) * SLAB;
  if (backward) b += (size_t)((M + TILE - 1) / TILE) * DY_PER_TILE;
  return b;
}

// d_bf [N,256]: per-ray sums of d v1p (gradient of the folded per-ray bias); dwfh [256 x 256]: gradient of the folded matrix
// Wfh -- the caller unfolds both into fc_out and layer_dir1 (codenerf.cu).  d_bo [N,257]: only column 0 (sigma) is written.
int cnrf_tc_chain_bwd(const TcChainArgs& a, const float* g_raw, const void* saved, size_t saved_bytes,
                      const cnrf_codenerf_grads* g, float* d_b2, float* d_bo, float* d_br, float* d_bd, float* d_bf,
                      float* dwfh, float* g_x_xyz, void* ws, size_t ws_bytes, cudaStream_t st) {
  // d_bd != NULL: also the per-ray sums of d v1p (the caller turns them into the gradient of the view-direction encoding);
  // g_x_xyz != NULL: the backward chain runs its fifth stage, d gamma(x) = d h1p W1   (pose optimisation, eval.py:129-168)
  if ((d_bd || g_x_xyz) && a.gid) return CNRF_E_UNSUPPORTED;      // rays with shared codes carry no pose gradient
  const int64_t tiles = (a.M + TILE - 1) / TILE;
  uint8_t* w = reinterpret_cast<uint8_t*>(up1k(reinterpret_cast<size_t>(ws)));
  if (w + PACKED_BYTES + 1024 + (size_t)tiles * DY_PER_TILE > reinterpret_cast<uint8_t*>(ws) + ws_bytes) return CNRF_E_WORKSPACE;
  if (saved_bytes < cnrf_tc_saved_bytes(a.M) + 1024) return CNRF_E_WORKSPACE;
  const uint8_t* sv = reinterpret_cast<const uint8_t*>(up1k(reinterpret_cast<size_t>(saved)));
  // the forward packed both image sets (pack_weights_kernel) next to the saved activations
  Packed pk = carve_packed(const_cast<uint8_t*>(sv) + (size_t)tiles * SAVED_PER_TILE + (size_t)pairs_of(a.M) * SLAB);
  uint8_t* dy = reinterpret_cast<uint8_t*>(up1k(reinterpret_cast<size_t>(w + PACKED_BYTES)));
  int rc;
  if (!a.prezeroed) {
    if ((rc = cnrf_fill_zero(g->bias[CNRF_L_XYZ1], 256 * sizeof(float), st))) return rc;
    if ((rc = cnrf_fill_zero(g->bias[CNRF_L_DIR2], 256 * sizeof(float), st))) return rc;
  }
  {
    static bool attr_set[CNRF_MAX_DEVICES] = {};
    const cudaError_t e = cnrf_once_per_device(attr_set, [] {
      return cudaFuncSetAttribute(mlp_dw_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, DW_SMEM); });
    if (e != cudaSuccess) return (int)e;
  }
  BwdArgs b;
  b.M = a.M; b.g_raw = g_raw; b.saved = sv; b.dy = dy; b.g_x = g_x_xyz; b.dx = a.dx; b.pk = pk; b.err = tc_error_ptr();
  const int64_t pairs = (tiles + 1) / 2;
  if ((rc = cnrf_tc_launch_bwd(b, (int)(pairs < sm_count() ? pairs : sm_count()), st))) return rc;

  DwArgs d;
  d.M = a.M; d.S = a.S; d.saved = sv; d.dy = dy; d.g_raw = g_raw; d.gid = a.gid; d.d_br = d_br; d.d_bo = d_bo; d.err = tc_error_ptr();
  d.ext = sv + (size_t)tiles * SAVED_PER_TILE;
  // the per-ray tables are accumulated by the reducer warps
  if (!a.prezeroed) {
    if ((rc = cnrf_fill_zero(d_b2, (size_t)a.N * 256 * sizeof(float), st))) return rc;
    if ((rc = cnrf_fill_zero(d_bo, (size_t)a.N * 257 * sizeof(float), st))) return rc;
    if ((rc = cnrf_fill_zero(d_br, (size_t)a.N * 3 * sizeof(float), st))) return rc;
    if ((rc = cnrf_fill_zero(d_bf, (size_t)a.N * 256 * sizeof(float), st))) return rc;
    if ((rc = cnrf_fill_zero(dwfh, (size_t)256 * 256 * sizeof(float), st))) return rc;
  }
  float* const nil = nullptr;
  // saved images: 0 h1, 1 h2, 2 v1, 3 v2;  gradient images: 0 d v2p, 1 d v1p, 2 d h2p, 3 d h1p
  d.job[0] = {0, 2, 256, g->weight[CNRF_L_DIR2], 256, 256, nil, -1, RED_COLSUM, g->bias[CNRF_L_DIR2], 0, nil};   // d v2p^T v1
  d.job[1] = {1, 1, 256, dwfh, 256, 256, nil, -1, RED_RAY_SIGMA, d_bf, 256, g->weight[CNRF_L_FC_OUT]};  // d v1p^T h2 -> d Wfh; per-ray
                                                                       // sums of d v1p; sigma-weighted column sums of h2 (fc_out row 0)
  d.job[2] = {2, 0, 256, g->weight[CNRF_L_XYZ2], 512, 256, nil, -1, RED_RAY, d_b2, 256, nil};  // d h2p^T h1
  d.job[3] = {3, -1, 64, g->weight[CNRF_L_XYZ1], a.dx, a.dx, g->bias[CNRF_L_XYZ1], 63, RED_NONE, nil, 0, nil};   // d h1p^T [gamma | 1]
  d.job[4] = {1, -2, 64, g->weight[CNRF_L_DIR1] + 256, 256 + a.dd, a.dd, g->bias[CNRF_L_DIR1], a.dd, RED_NONE, nil, 0,
              nil};                                                                            // d v1p^T [gamma(d) | 1]
  d.job[5] = {-1 - 3, 0, 0, nil, 0, 0, nil, -1, RED_RGB, g->weight[CNRF_L_FC_RGB], 512, nil};  // g_rgb^T v2 (no GEMM)
  {   // one CTA per SM, shared in proportion to the bytes a job streams per tile (128 KB for the three 256-wide jobs,
      // 80, 72 and 64 KB for the other three): the kernel is HBM-bound and a single SM only pulls ~45 GB/s
    const int64_t halves = tiles * 2;
    // measured shares (B200, S = 160): reducer-heavy jobs are bound by their four reducer warps, not by their bytes
    static const int share148[DW_JOBS] = {25, 36, 30, 15, 15, 27};
    int shares[DW_JOBS];
    int job_mask = (1 << DW_JOBS) - 1;
    {
      const char* e = nullptr;
#ifdef CNRF_TUNING                                    // tuning builds only: seven comma-separated CTA counts, and a job subset
      e = getenv("CNRF_DW_SHARES");                   // (wrong gradients, right timing)
      if (const char* m = getenv("CNRF_DW_JOBMASK")) job_mask = atoi(m);
#endif
      for (int j = 0; j < DW_JOBS; ++j) {
        int v = share148[j] * sm_count() / 148;
        if (e) { v = atoi(e); const char* n = strchr(e, ','); e = n ? n + 1 : nullptr; }
        shares[j] = v > 0 ? v : 1;
      }
    }
    int at = 0;
    for (int j = 0; j < DW_JOBS; ++j) {
      d.first[j] = at;
      const int want = !((job_mask >> j) & 1) ? 0 : shares[j];
      at += (int)(halves < want ? halves : want);
    }
    d.first[DW_JOBS] = at;
    const bool timed = cnrf_ktimer_begin(CNRF_KT_DW, a.M, st);
    mlp_dw_tc_kernel<<<at, DW_THREADS, DW_SMEM, st>>>(d);
    CNRF_LAUNCH_CHECK();
    if (timed) cnrf_ktimer_end(CNRF_KT_DW, st);
  }
  // layer_dir1's view-direction half and bias get their gradients from the extension job above; the per-ray sums of
  // d v1p are only needed for the gradient of the view directions themselves
  if (d_bd) {
    const unsigned gs_grid = (unsigned)((a.N * 128 + 255) / 256);
    if (a.pairing == CNRF_PAIRING_REFERENCE)
      img_group_sum_kernel<<<gs_grid, 256, 0, st>>>(dy, DY_PER_TILE, 1 * (size_t)TILE_BYTES, a.N, 1, a.S, a.N, d_bd, 256);
    else
      img_group_sum_kernel<<<gs_grid, 256, 0, st>>>(dy, DY_PER_TILE, 1 * (size_t)TILE_BYTES, a.N, a.S, a.S, 1, d_bd, 256);
    CNRF_LAUNCH_CHECK();
  }
  return CNRF_OK;
}
