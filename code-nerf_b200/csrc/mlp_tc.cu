// Fused bf16 tensor-core chain of the CodeNeRF field (CNRF_PRECISION_BF16): the per-sample layers
//   h1 = relu(W1 gamma(x) + b1) -> h2 = relu(W2h h1 + beta2[ray]) -> o = Woh h2 + betao[ray] -> v1 = relu(Wdf o[1:] +
//   betad[pair]) -> v2 = relu(Wd2 v1 + bd2) -> rgb = Wrv v2 + betar[ray]          (models/model.py:179-194, code-folded)
// as ONE persistent kernel per direction.  tcgen05.mma (kind::f16, bf16 operands, fp32 accumulators in TMEM), weights
// streamed by 1-D TMA bulk copies of pre-swizzled 32 KB chunk images into a shared-memory ring, activations never leave
// shared memory between layers (except the bf16 tile images saved for the backward by TMA bulk stores).
//
// Geometry: a CTA owns a PAIR of 128-row sample tiles (T0, T1) that march through the layers in lockstep so every weight
// chunk fetched from L2 feeds 256 rows.  TMEM: 2 x 256 fp32 columns (all 512).  Shared memory: 2 x 64 KB activation
// tiles (4 slabs of [128 rows x 64 bf16], 128-byte swizzle) + 3 x 32 KB weight ring.
// Warp roles: warp 0 = TMA producer, warp 1 = MMA issuer (one elected lane), warps 2-5 = epilogue of T0, 6-9 = epilogue
// of T1 (thread = one sample row; TMEM lane quarter = warp_id % 4).
#include "common.cuh"
#include "internal.h"
#include "tc_common.cuh"

using namespace tc;

namespace tcc {
constexpr int TILE = 128;                 // rows per tile
constexpr int SLAB = TILE * 128;          // bytes of one [128 x 64] bf16 slab
constexpr int TILE_BYTES = 4 * SLAB;      // 64 KB: one [128 x 256] activation tile image
constexpr int CHUNK = 256 * 128;          // 32 KB: one [256 x 64] weight chunk image
constexpr int RING = 3;
constexpr int NTHREADS = 320;
constexpr int FWD_CHUNKS = 17;            // 1 (layer_xyz1, K=64) + 4 x 4
constexpr int BWD_CHUNKS = 16;            // 4 x 4 transposed
constexpr int CONST_BYTES = 4 * 256 * 2;    // bf16 [w_sigma | w_rgb(3)] staged in shared memory
constexpr int SMEM_BYTES = 2 * TILE_BYTES + RING * CHUNK + CONST_BYTES + 256 /*barriers*/;

// Packed weights (bf16 images + fp32 vectors), rebuilt from the fp32 parameters at every call (they change each step).
struct Packed {
  uint8_t* fwd;      // FWD_CHUNKS x 32 KB: [W1|b1], W2h, Woh[1:], Wdf, Wd2   (rows = out feature, cols = in feature)
  uint8_t* bwd;      // BWD_CHUNKS x 32 KB: Wd2^T, Wdf^T, Woh[1:]^T, W2h^T     (rows = in feature,  cols = out feature)
  float* w_sigma;    // [256]  fc_out row 0, hidden half
  float* w_rgb;      // [3][256] fc_rgb, hidden half
  float* b_dir2;     // [256]
  float* zeros;      // [256]
  uint16_t* wsr_bf16; // [4][256] bf16: w_sigma, w_rgb r/g/b (staged into shared memory by the forward kernel)
};
constexpr size_t PACKED_BYTES = (size_t)(FWD_CHUNKS + BWD_CHUNKS) * CHUNK + (256 + 768 + 256 + 256) * sizeof(float) + 2048;
}  // namespace tcc
using namespace tcc;

// ---- weight packing ------------------------------------------------------------------------------------------------
// chunk image element (row r in [0,256), col c in [0,64)) of chunk `ch`.
__global__ void pack_weights_kernel(cnrf_codenerf_params p, int dx, int dd, Packed pk) {
  const int ch = blockIdx.x;                  // 0 .. FWD_CHUNKS+BWD_CHUNKS-1
  for (int i = threadIdx.x; i < 256 * 64; i += blockDim.x) {
    int r = i >> 6, c = i & 63;
    float v;
    uint8_t* base;
    if (ch < FWD_CHUNKS) {
      base = pk.fwd + (size_t)ch * CHUNK;
      if (ch == 0) {                          // layer_xyz1 [256, dx] with the bias in column 63 (the input's ones column)
        v = c < dx ? p.weight[CNRF_L_XYZ1][r * dx + c] : (c == 63 ? p.bias[CNRF_L_XYZ1][r] : 0.f);
      } else {
        int l = (ch - 1) >> 2, k = ((ch - 1) & 3) * 64 + c;   // l: 0 W2h, 1 Woh[1:], 2 Wdf, 3 Wd2
        if (l == 0) v = p.weight[CNRF_L_XYZ2][r * 512 + k];
        else if (l == 1) v = p.weight[CNRF_L_FC_OUT][(r + 1) * 512 + k];
        else if (l == 2) v = p.weight[CNRF_L_DIR1][r * (256 + dd) + k];
        else v = p.weight[CNRF_L_DIR2][r * 256 + k];
      }
    } else {
      int b = ch - FWD_CHUNKS;
      base = pk.bwd + (size_t)b * CHUNK;
      int l = b >> 2, n = (b & 3) * 64 + c;   // image[k = r][n]: l: 0 Wd2^T, 1 Wdf^T, 2 Woh[1:]^T, 3 W2h^T
      if (l == 0) v = p.weight[CNRF_L_DIR2][n * 256 + r];
      else if (l == 1) v = p.weight[CNRF_L_DIR1][n * (256 + dd) + r];
      else if (l == 2) v = p.weight[CNRF_L_FC_OUT][(n + 1) * 512 + r];
      else v = p.weight[CNRF_L_XYZ2][n * 512 + r];
    }
    *reinterpret_cast<__nv_bfloat16*>(base + slab_offset(r, c)) = __float2bfloat16(v);
  }
  if (ch == 0) {
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

// ---- shared pieces of the fused kernels ----------------------------------------------------------------------------------
struct Bars {
  uint64_t w_full[RING], w_empty[RING];
  uint64_t a_ready[2];     // epilogue -> MMA : tile's operand image written (count 1, elected thread after a named barrier)
  uint64_t acc_full[2];    // MMA -> epilogue : tile's accumulator complete
  uint32_t tmem_slot;
  int abort_flag;
};

__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}
__device__ __forceinline__ uint32_t cvt_relu_bf16x2(float lo, float hi) {
  uint32_t d;
  asm("cvt.rn.relu.bf16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(hi), "f"(lo));
  return d;
}
// write 32 consecutive columns [cb*32, cb*32+32) of row r of a tile image (8 bf16 per 16-byte chunk)
__device__ __forceinline__ void store_row_block(uint8_t* tile, int r, int cb, const uint32_t (&pk)[16]) {
  uint8_t* slab = tile + (cb >> 1) * SLAB + r * 128;
  const int c0 = (cb & 1) * 4;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    uint4 v = make_uint4(pk[4 * j], pk[4 * j + 1], pk[4 * j + 2], pk[4 * j + 3]);
    *reinterpret_cast<uint4*>(slab + (((c0 + j) ^ (r & 7)) << 4)) = v;
  }
}

// Bounded mbarrier wait (about a second): a protocol bug must end the kernel, never hang the GPU.
__device__ __forceinline__ bool wait_bar(uint64_t* bar, uint32_t parity, Bars* b, int* err) {
  if (mbar_try_wait(bar, parity)) return true;
  const long long t0 = clock64();
  while (true) {
    if (mbar_try_wait(bar, parity)) return true;
    if (*(volatile int*)&b->abort_flag) return false;
    if (clock64() - t0 > 2000000000LL) {
      b->abort_flag = 1;
      if (err) atomicExch(err, 1);
      return false;
    }
  }
}

constexpr size_t SAVED_GAMMA = SLAB;                          // gamma(x) slab (K = 64)
constexpr size_t SAVED_PER_TILE = SAVED_GAMMA + 5 * (size_t)TILE_BYTES;   // + h1, h2, feat, v1, v2 images

struct FwdArgs {
  int64_t N, M; int S, pairing, dx;
  const float* pts; const float* x_xyz; int64_t ld_xyz;
  const float* freqs; int L; int inc;
  const float *b2, *bof, *bos, *br, *bd;   // per-ray bias tables: [N,256], [N,256] (fc_out feat), [N] (sigma), [N,3], [N,256]
  Packed pk;
  float* raw;
  uint8_t* saved;      // nullptr: inference
  int* err;
};

// gamma(x) row -> 64 bf16 (col 63 = 1 carries layer_xyz1's bias), written to slab 0 of the tile
__device__ __forceinline__ void put_bf16(uint8_t* row, int r, int i, float v) {
  *reinterpret_cast<__nv_bfloat16*>(row + ((((i >> 3) ^ (r & 7)) << 4) | ((i & 7) << 1))) = __float2bfloat16(v);
}
__device__ __forceinline__ void build_input_row(const FwdArgs& a, int64_t m, bool valid, uint8_t* tile, int r) {
  uint8_t* row = tile + r * 128;
#pragma unroll
  for (int j = 0; j < 8; ++j) *reinterpret_cast<uint4*>(row + (j << 4)) = make_uint4(0u, 0u, 0u, 0u);
  if (!valid) return;
  if (a.pts) {
    float x[3] = {a.pts[m * 3], a.pts[m * 3 + 1], a.pts[m * 3 + 2]};
    int o = 0;
    if (a.inc) { put_bf16(row, r, 0, x[0]); put_bf16(row, r, 1, x[1]); put_bf16(row, r, 2, x[2]); o = 3; }
    float s[3], c[3], fprev = 0.f;
    for (int k = 0; k < a.L; ++k) {
      float f = a.freqs[k];
      if (k > 0 && f == 2.0f * fprev) {   // octave step: angle doubling; the error doubles per octave (<= 2^9 * 1e-7),
#pragma unroll                            // far below the 4e-3 rounding of the bf16 operand it feeds
        for (int d = 0; d < 3; ++d) { float s2 = 2.f * s[d] * c[d]; c[d] = 1.f - 2.f * s[d] * s[d]; s[d] = s2; }
      } else {
#pragma unroll
        for (int d = 0; d < 3; ++d) sincosf(x[d] * f, &s[d], &c[d]);
      }
      fprev = f;
#pragma unroll
      for (int d = 0; d < 3; ++d) { put_bf16(row, r, o + 6 * k + d, s[d]); put_bf16(row, r, o + 6 * k + 3 + d, c[d]); }
    }
  } else {
    for (int i = 0; i < a.dx; ++i) put_bf16(row, r, i, a.x_xyz[m * a.ld_xyz + i]);
  }
  put_bf16(row, r, 63, 1.0f);
}

__global__ void __launch_bounds__(NTHREADS, 1) mlp_fwd_tc_kernel(const FwdArgs a) {
  extern __shared__ __align__(1024) uint8_t smem[];   // 128-byte-swizzled images need 1024-byte alignment
  uint8_t* tiles = smem;                              // 2 x 64 KB
  uint8_t* ring = smem + 2 * TILE_BYTES;              // RING x 32 KB
  const __nv_bfloat162* wsr = reinterpret_cast<const __nv_bfloat162*>(ring + RING * CHUNK);   // [4][128] bf16x2
  Bars* bars = reinterpret_cast<Bars*>(ring + RING * CHUNK + CONST_BYTES);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int64_t num_tiles = (a.M + TILE - 1) / TILE;
  const int64_t num_pairs = (num_tiles + 1) / 2;

  if (tid == 0) {
    for (int i = 0; i < RING; ++i) { mbar_init(&bars->w_full[i], 1); mbar_init(&bars->w_empty[i], 1); }
    for (int t = 0; t < 2; ++t) { mbar_init(&bars->a_ready[t], 1); mbar_init(&bars->acc_full[t], 1); }
    bars->abort_flag = 0;
    if ((smem_u32(smem) & 1023u) != 0) { bars->abort_flag = 1; atomicExch(a.err, 2); }
    mbar_fence_init();
  }
  for (int i = tid; i < CONST_BYTES / 4; i += NTHREADS)
    reinterpret_cast<uint32_t*>(ring + RING * CHUNK)[i] = reinterpret_cast<const uint32_t*>(a.pk.wsr_bf16)[i];
  if (warp == 1) tmem_alloc(&bars->tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = bars->tmem_slot;

  if (warp == 0) {
    // ===== TMA producer: stream the 17 weight chunks of every tile pair through the ring =====
    if (lane == 0) {
      uint32_t g = 0;
      for (int64_t pr = blockIdx.x; pr < num_pairs; pr += gridDim.x) {
        for (int ch = 0; ch < FWD_CHUNKS; ++ch, ++g) {
          const int slot = g % RING;
          const uint32_t use = g / RING;
          if (!wait_bar(&bars->w_empty[slot], (use & 1) ^ 1, bars, a.err)) goto done;
          mbar_arrive_expect_tx(&bars->w_full[slot], CHUNK);
          bulk_g2s(ring + slot * CHUNK, a.pk.fwd + (size_t)ch * CHUNK, CHUNK, &bars->w_full[slot]);
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer =====
    if (lane == 0) {
      const uint32_t idesc = idesc_bf16(128, 256, 0, 0);
      uint32_t g = 0, na = 0;
      for (int64_t pr = blockIdx.x; pr < num_pairs; pr += gridDim.x) {
        for (int layer = 0; layer < 5; ++layer, ++na) {
          const int nch = layer == 0 ? 1 : 4;
          for (int c = 0; c < nch; ++c, ++g) {
            const int slot = g % RING;
            if (!wait_bar(&bars->w_full[slot], (g / RING) & 1, bars, a.err)) goto done;
            for (int t = 0; t < 2; ++t) {
              if (c == 0 && !wait_bar(&bars->a_ready[t], na & 1, bars, a.err)) goto done;
              tc_fence_after();
              const uint32_t a_base = smem_u32(tiles + t * TILE_BYTES + c * SLAB);
              const uint32_t b_base = smem_u32(ring + slot * CHUNK);
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                umma_bf16(tmem + t * 256, smem_desc(a_base + k * 32, 16, 1024, SWIZZLE_128B),
                          smem_desc(b_base + k * 32, 16, 1024, SWIZZLE_128B), idesc, (c | k) != 0);
              }
              if (c == nch - 1) umma_commit(&bars->acc_full[t]);
            }
            umma_commit(&bars->w_empty[slot]);
          }
        }
      }
    }
  } else {
    // ===== epilogue warps: thread = one sample row =====
    const int t = (warp - 2) >> 2;                    // tile of the pair
    const int q = warp & 3;                           // TMEM lane quarter this warp may read
    const int r = q * 32 + lane;                      // row in tile
    const bool elected = ((warp - 2) & 3) == 0 && lane == 0;
    const int bar_id = 1 + t;
    uint8_t* tile = tiles + t * TILE_BYTES;
    const uint32_t tacc = tmem + t * 256 + ((uint32_t)(q * 32) << 16);
    uint32_t na = 0;
    for (int64_t pr = blockIdx.x; pr < num_pairs; pr += gridDim.x) {
      const int64_t tile_idx = pr * 2 + t;
      const int64_t m = tile_idx * TILE + r;
      const bool valid = m < a.M;
      const int64_t ray = valid ? m / a.S : 0;
      const int64_t pair = valid ? (a.pairing == CNRF_PAIRING_REFERENCE ? m % a.N : ray) : 0;
      uint8_t* sv = (a.saved && tile_idx < num_tiles) ? a.saved + (size_t)tile_idx * SAVED_PER_TILE : nullptr;
      // ---- input image: gamma(x) ----
      if (elected) bulk_wait_read0();                 // the previous pair's last image store has finished reading smem
      named_bar_sync(bar_id, 128);
      build_input_row(a, m, valid, tile, r);
      fence_proxy_async();
      named_bar_sync(bar_id, 128);
      if (elected) {
        if (sv) { bulk_s2g(sv, tile, SLAB); bulk_commit(); }
        mbar_arrive(&bars->a_ready[t]);
      }
      float sigma = 0.f;
      for (int e = 0; e < 5; ++e, ++na) {
        if (!wait_bar(&bars->acc_full[t], na & 1, bars, a.err)) goto done;
        tc_fence_after();
        if (elected) bulk_wait_read0();               // image store of the previous layer no longer reads this tile
        named_bar_sync(bar_id, 128);
        const float4* bias = reinterpret_cast<const float4*>(e == 1 ? a.b2 + ray * 256 : e == 2 ? a.bof + ray * 256
                                                             : e == 3 ? a.bd + pair * 256 : e == 4 ? a.pk.b_dir2 : a.pk.zeros);
        float rgb[3] = {0.f, 0.f, 0.f};
        float4 bq[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) bq[j] = __ldg(bias + j);
#pragma unroll 1
        for (int cb = 0; cb < 8; ++cb) {
          uint32_t v[32];
          tmem_ld32(tacc + cb * 32, v);
          float4 bn[8];                               // next block's bias rows travel while this block is converted
          const int nb = cb < 7 ? cb + 1 : 7;
#pragma unroll
          for (int j = 0; j < 8; ++j) bn[j] = __ldg(bias + nb * 8 + j);
          tmem_ld_wait();
          float f[32];
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            f[4 * j] = __uint_as_float(v[4 * j]) + bq[j].x;
            f[4 * j + 1] = __uint_as_float(v[4 * j + 1]) + bq[j].y;
            f[4 * j + 2] = __uint_as_float(v[4 * j + 2]) + bq[j].z;
            f[4 * j + 3] = __uint_as_float(v[4 * j + 3]) + bq[j].w;
          }
          uint32_t pk[16];
          if (e == 2) {
#pragma unroll
            for (int j = 0; j < 16; ++j) pk[j] = pack_bf16(f[2 * j], f[2 * j + 1]);
          } else {
#pragma unroll
            for (int j = 0; j < 16; ++j) pk[j] = cvt_relu_bf16x2(f[2 * j], f[2 * j + 1]);
          }
          if (e == 1) {                               // sigma_raw = w_sigma . h2 + betao[ray][0]   (fc_out row 0)
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const float2 w = __bfloat1622float2(wsr[cb * 16 + j]);
              sigma = fmaf(fmaxf(f[2 * j], 0.f), w.x, sigma);
              sigma = fmaf(fmaxf(f[2 * j + 1], 0.f), w.y, sigma);
            }
          }
          if (e == 4) {                               // rgb_raw = Wrv v2 + betar[ray]                (fc_rgb)
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const float x0 = fmaxf(f[2 * j], 0.f), x1 = fmaxf(f[2 * j + 1], 0.f);
#pragma unroll
              for (int c = 0; c < 3; ++c) {
                const float2 w = __bfloat1622float2(wsr[(1 + c) * 128 + cb * 16 + j]);
                rgb[c] = fmaf(x0, w.x, fmaf(x1, w.y, rgb[c]));
              }
            }
          }
          store_row_block(tile, r, cb, pk);
#pragma unroll
          for (int j = 0; j < 8; ++j) bq[j] = bn[j];
        }
        if (e == 1 && valid) sigma += a.bos[ray];
        if (e == 4 && valid) {
          float4 o = make_float4(rgb[0] + a.br[ray * 3], rgb[1] + a.br[ray * 3 + 1], rgb[2] + a.br[ray * 3 + 2], sigma);
          reinterpret_cast<float4*>(a.raw)[m] = o;   // [rgb_raw | sigma_raw]   models/model.py:194
        }
        fence_proxy_async();
        tc_fence_before();
        named_bar_sync(bar_id, 128);
        if (elected) {
          if (sv) { bulk_s2g(sv + SAVED_GAMMA + (size_t)e * TILE_BYTES, tile, TILE_BYTES); bulk_commit(); }
          if (e < 4) mbar_arrive(&bars->a_ready[t]);
        }
      }
    }
    if (elected) bulk_wait0();
  }
done:
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 512);
}

// ---- host side --------------------------------------------------------------------------------------------------------
__device__ int g_tc_error = 0;   // raised by a bounded barrier wait that timed out (read by cnrf_debug_tc_error)

struct TcChainArgs {   // mirrors the declaration in codenerf.cu
  int64_t N, M; int S, pairing, dx, dd;
  const cnrf_codenerf_params* p;
  const float* x_xyz; int64_t ld_xyz;
  const float* pts; const float* freqs_xyz; int L_xyz; int inc_xyz;
  const float *b2, *bo, *br, *bd;
  const float *bof, *bos;                     // fc_out bias table split into an aligned feature part and sigma
};

static inline size_t up1k(size_t b) { return (b + 1023) & ~size_t(1023); }
static int sm_count() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
  }
  return n;
}
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
  static int* p = nullptr;
  if (!p) cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_tc_error);
  return p;
}

size_t cnrf_tc_saved_bytes(int64_t M) { return (size_t)((M + TILE - 1) / TILE) * SAVED_PER_TILE; }

int cnrf_tc_chain_fwd(const TcChainArgs& a, float* raw, void* saved, size_t saved_bytes, void* ws, size_t ws_bytes,
                      cudaStream_t st) {
  uint8_t* w = reinterpret_cast<uint8_t*>(up1k(reinterpret_cast<size_t>(ws)));
  if (w + PACKED_BYTES > reinterpret_cast<uint8_t*>(ws) + ws_bytes) return CNRF_E_WORKSPACE;
  if (saved && saved_bytes < cnrf_tc_saved_bytes(a.M) + 1024) return CNRF_E_WORKSPACE;
  Packed pk = carve_packed(w);
  pack_weights_kernel<<<FWD_CHUNKS + BWD_CHUNKS, 256, 0, st>>>(*a.p, a.dx, a.dd, pk);
  CNRF_LAUNCH_CHECK();
  FwdArgs f;
  f.N = a.N; f.M = a.M; f.S = a.S; f.pairing = a.pairing; f.dx = a.dx;
  f.pts = a.pts; f.x_xyz = a.x_xyz; f.ld_xyz = a.ld_xyz; f.freqs = a.freqs_xyz; f.L = a.L_xyz; f.inc = a.inc_xyz;
  f.b2 = a.b2; f.bof = a.bof; f.bos = a.bos; f.br = a.br; f.bd = a.bd; f.pk = pk; f.raw = raw;
  f.saved = saved ? reinterpret_cast<uint8_t*>(up1k(reinterpret_cast<size_t>(saved))) : nullptr;
  f.err = tc_error_ptr();
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(mlp_fwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
    if (e != cudaSuccess) return (int)e;
    attr_set = true;
  }
  const int64_t pairs = ((a.M + TILE - 1) / TILE + 1) / 2;
  const int grid = (int)(pairs < sm_count() ? pairs : sm_count());
  mlp_fwd_tc_kernel<<<grid, NTHREADS, SMEM_BYTES, st>>>(f);
  CNRF_LAUNCH_CHECK();
  return CNRF_OK;
}

size_t cnrf_tc_workspace_bytes(int64_t N, int64_t M, int backward) {
  (void)N; (void)M; (void)backward;
  return PACKED_BYTES + 2048;
}

int cnrf_tc_chain_bwd(const TcChainArgs&, const float*, const void*, size_t, const cnrf_codenerf_grads*, float*, float*,
                      float*, float*, float*, void*, size_t, cudaStream_t) {
  return CNRF_E_UNSUPPORTED;
}

// Synchronous diagnostic: 1 if any tensor-core kernel of this process hit its barrier-wait timeout.
extern "C" int cnrf_debug_tc_error(void) {
  int v = 0;
  cudaDeviceSynchronize();
  cudaMemcpyFromSymbol(&v, g_tc_error, sizeof(int));
  return v;
}
