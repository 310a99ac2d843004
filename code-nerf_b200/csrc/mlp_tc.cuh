// Shared declarations of the fused tensor-core chain kernels (mlp_fwd_tc.cu, mlp_tc.cu).
#pragma once
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
// Warps of the chain kernels: 0 and 10 = TMA producers, 1 = MMA issuer, 2-9 = epilogue (4 per tile), 11 = image-store warp.
// ONE thread alternating mbarrier.try_wait / expect_tx / cp.async.bulk completes a copy every ~588 clk whatever its size
// (profiles/r02_bulk_copy_probe*.json: 2 issuing threads 294 clk, 4 threads 147 clk per copy, on 1 or 148 SMs alike), while
// the tensor core consumes a 16 KB weight image in ~280 clk: NPROD producers take the images of the stream round-robin.
// The images the backward needs (activations / masked gradients, 64 KB per layer and tile) are NOT stored by TMA: the
// SM's TMA unit works through its queue in order, and a 64 KB store ahead of the weight loads starves the tensor core for
// ~1.7 k clk (forward: 306 us without the stores, 489 us with bulk stores, whether the destination is HBM or an L2-resident
// scratch).  Warp STORE_WARP copies them shared -> global with LDS.128 / STG.128 instead, hand-shaking with the tile's
// epilogue warps through two mbarriers.
constexpr int NPROD = 2;
constexpr int NTHREADS = 384;
constexpr int STORE_WARP = 11;
__device__ __forceinline__ int producer_index(int warp) { return warp == 0 ? 0 : warp == 10 ? 1 : -1; }
// one warp copies `bytes` (a multiple of 4 KB) of a shared-memory image to global memory
// (-DCNRF_EXP_NO_STG / -DCNRF_EXP_NO_COPY: attribution builds of scripts/exp_fwd_store_attribution.sh -- wrong results, right timing)
__device__ __forceinline__ void warp_copy_image(uint8_t* __restrict__ dst, const uint8_t* src_smem, int bytes, int lane) {
#if defined(CNRF_EXP_NO_COPY) || defined(CNRF_EXP_COALESCED_DIRECT)
  return;
#endif
  for (int off = lane * 16; off < bytes; off += 8 * 512) {
    uint4 v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = *reinterpret_cast<const uint4*>(src_smem + off + j * 512);
#ifdef CNRF_EXP_NO_STG
    uint32_t acc = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) acc |= v[j].x ^ v[j].y ^ v[j].z ^ v[j].w;
    if (acc == 0x12345u) __stcs(reinterpret_cast<uint4*>(dst + off), v[0]);
#else
#pragma unroll
    for (int j = 0; j < 8; ++j) __stcs(reinterpret_cast<uint4*>(dst + off + j * 512), v[j]);
#endif
  }
}
// fc_out is FOLDED into layer_dir1 (models/model.py:182-189 apply no activation between them): the per-sample chain has four
// 256-wide tensor-core layers  W1 | W2h | Wfh = Wdf Wout'[:, :H] | Wd2  (the fold is rebuilt from the fp32 parameters at
// every call, codenerf.cu)
constexpr int NUM_TC_LAYERS = 4;
constexpr int FWD_CHUNKS = 14;            // 1 (layer_xyz1, K=64) + 3 x 4 + 1 (view-direction extension of layer_dir1)
constexpr int EXT_CHUNK = 13;             // [Wdv | b_dir1 | 0]: K = dim_dir + 1 <= 32 columns of a [256 x 64] image
constexpr int BWD_CHUNKS = 13;            // 3 x 4 transposed + W1^T (four [64 in x 64 hidden] K-chunks stacked in one image)
constexpr int BWD_X_CHUNK = 12;           // ... its index: B operand of the input-gradient stage d gamma(x) = d h1p W1
constexpr int CONST_BYTES = 4 * 256 * 2;    // bf16 [w_sigma | w_rgb(3)] staged in shared memory
constexpr int SMEM_BYTES = 2 * TILE_BYTES + RING * CHUNK + CONST_BYTES + 256 /*barriers*/;

// Packed weights (bf16 images + fp32 vectors), rebuilt from the fp32 parameters at every call (they change each step).
struct Packed {
  uint8_t* fwd;      // FWD_CHUNKS x 32 KB: [W1|b1], W2h, Wfh, Wd2, [Wdv|bd1]   (rows = out feature, cols = in feature)
  uint8_t* bwd;      // BWD_CHUNKS x 32 KB: Wd2^T, Wfh^T, W2h^T, W1^T          (rows = in feature,  cols = out feature)
  float* w_sigma;    // [256]  fc_out row 0, hidden half
  float* w_rgb;      // [3][256] fc_rgb, hidden half
  float* b_dir2;     // [256]
  float* zeros;      // [256]
  uint16_t* wsr_bf16; // [4][256] bf16: w_sigma, w_rgb r/g/b (staged into shared memory by the forward kernel)
};
constexpr size_t PACKED_BYTES = (size_t)(FWD_CHUNKS + BWD_CHUNKS) * CHUNK + (256 + 768 + 256 + 256) * sizeof(float) + 2048;
}  // namespace tcc
using namespace tcc;

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
__device__ __forceinline__ void prefetch_l1(const void* p) { asm volatile("prefetch.global.L1 [%0];" ::"l"(p)); }
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
      tc_fatal();
      return false;
    }
  }
}

constexpr size_t SAVED_GAMMA = SLAB;                          // gamma(x) slab (K = 64)
constexpr size_t SAVED_MASKS = SAVED_GAMMA + 4 * (size_t)TILE_BYTES;      // after the h1, h2, v1, v2 images
constexpr size_t MASK_BYTES = 128 * 32;                                   // one layer's sign bits: [8 column blocks][128 rows] words
// 4 x 4 KB of sign-bit words (h1, h2, v1, v2 pre-activations; bit 31-j of word cb = column 32 cb + j is <= 0): the
// backward reads 32 B per row and layer instead of re-reading the 64 KB activation image for its ReLU mask
constexpr size_t SAVED_PER_TILE = SAVED_MASKS + 4 * MASK_BYTES;

struct FwdArgs {
  int64_t N, M; int S, pairing, dx;
  const float* pts; const float* x_xyz; int64_t ld_xyz;
  const float* freqs; int L; int inc;
  const float *b2, *bf, *bos, *br;         // per-ray bias tables: [N,256] layer_xyz2, [N,256] folded layer_dir1, [N] (sigma), [N,3]
  const int32_t* gid;                      // row of those tables that ray r uses (NULL: r) -- shared-code mode
  const uint8_t* ext;                      // per tile PAIR: [128 x 64] bf16 slab, columns 32t..32t+31 = [gamma(d_pair) | 1 | 0] of tile t
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
// The same row for the standard encoding [x | sin(2^k x), cos(2^k x), k < L] (position_embed.py:35-53 with include_input
// and octave bands): all 64 values stay in registers (compile-time indices) and leave as eight 16-byte stores.
template <int L>
__device__ __forceinline__ void build_input_row_octaves(const float x[3], bool valid, uint8_t* tile, int r) {
  static_assert(3 + 6 * L <= 63, "row holds 63 inputs + the bias column");
  uint8_t* row = tile + r * 128;
  float v[64];
#pragma unroll
  for (int i = 0; i < 64; ++i) v[i] = 0.f;
  if (valid) {
    float s[3], c[3];
#pragma unroll
    for (int d = 0; d < 3; ++d) { v[d] = x[d]; sincosf(x[d], &s[d], &c[d]); }
#pragma unroll
    for (int k = 0; k < L; ++k) {
      if (k > 0) {                         // octave step by angle doubling (error analysis: build_input_row)
#pragma unroll
        for (int d = 0; d < 3; ++d) { const float s2 = 2.f * s[d] * c[d]; c[d] = 1.f - 2.f * s[d] * s[d]; s[d] = s2; }
      }
#pragma unroll
      for (int d = 0; d < 3; ++d) { v[3 + 6 * k + d] = s[d]; v[3 + 6 * k + 3 + d] = c[d]; }
    }
    v[63] = 1.0f;
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    uint4 q;
    q.x = pack_bf16(v[8 * j], v[8 * j + 1]); q.y = pack_bf16(v[8 * j + 2], v[8 * j + 3]);
    q.z = pack_bf16(v[8 * j + 4], v[8 * j + 5]); q.w = pack_bf16(v[8 * j + 6], v[8 * j + 7]);
    *reinterpret_cast<uint4*>(row + ((j ^ (r & 7)) << 4)) = q;
  }
}

// mlp_fwd_tc.cu
int cnrf_tc_launch_fwd(const FwdArgs& f, int grid, cudaStream_t st);

// Backward-data chain (mlp_bwd_tc.cu)
constexpr size_t DY_PER_TILE = 4 * (size_t)TILE_BYTES;      // d v2p, d v1p, d h2p, d h1p
struct BwdArgs {
  int64_t M;
  const float* g_raw;          // [M,4]
  const uint8_t* saved;        // forward images + sign-bit masks
  uint8_t* dy;                 // gradient images out
  float* g_x;                  // [M, dx] gradient of the embedded points (pose optimisation, eval.py:129-168) or NULL
  int dx;
  Packed pk;
  int* err;
};
int cnrf_tc_launch_bwd(const BwdArgs& b, int grid, cudaStream_t st);
