// Backward-data chain of the fused bf16 tensor-core field (fc_out folded into layer_dir1, mlp_tc.cuh).  Per tile pair:
// stage "-1" on CUDA cores (d v2 = g_rgb Wrv, masked), then three tensor-core stages
//   d v1 = (d v2p Wd2) * [v1>0],  d h2 = (d v1p Wfh + d sigma w_sigma) * [h2>0],  d h1 = (d h2p W2h) * [h1>0]
// with W^T streamed as B operands (16 KB N-half images through a 5-slot TMA ring, as in the forward).  Every masked
// gradient tile is TMA-stored as a bf16 image (same layout as the forward's activation images) for the weight-gradient
// kernel.  The ReLU masks are the 1-bit sign words the forward saved (32 B per row and layer), requested before the
// accumulator wait; the epilogue is specialised per stage at compile time.
// Geometry, warp roles and barrier protocol are those of mlp_fwd_tc.cu.
#include "mlp_tc.cuh"

namespace {
constexpr int HALF = 128 * 128;             // bytes of one [128 x 64] bf16 weight image
constexpr int BRING = 5;
constexpr int HEAD_FLOATS = 256 + 768;      // w_sigma | w_rgb[3]   (contiguous in Packed)
constexpr int OFF_RING = 2 * TILE_BYTES;
constexpr int OFF_HEADS = OFF_RING + BRING * HALF;
constexpr int OFF_BARS = OFF_HEADS + HEAD_FLOATS * 4;
constexpr int BWD_SMEM = OFF_BARS + 256;

struct BBars {
  uint64_t w_full[BRING], w_empty[BRING];
  uint64_t a_ready[2], acc_full[2];
  uint64_t img_ready[2], img_saved[2];   // epilogue <-> image-store warp (mlp_tc.cuh)
  uint32_t tmem_slot;
  int abort_flag;
};

__device__ __forceinline__ bool bwait(uint64_t* bar, uint32_t parity, BBars* b, int* err) {
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

struct BEpi {
  BBars* bars;
  uint8_t* tile;
  const float* heads;      // smem: w_sigma[256] | w_rgb[3][256]
  uint32_t tacc;
  int r, bar_id;
  uint32_t nimg;           // gradient images of this tile handed to the store warp so far
  bool elected, tile_ok;
  const uint8_t* masks;    // this tile's sign words: [4 layers][8 column blocks][128 rows]
  float4 g;                // upstream gradient of this row's (r, g, b, sigma)
  int64_t exp_tile;        // (attribution builds)
};

// warp-wide wait whose result the compiler can prove uniform (keeps the MMA issuer's state in uniform registers)
__device__ __forceinline__ bool uwait(uint64_t* bar, uint32_t parity, BBars* b, int* err) {
  return __all_sync(0xFFFFFFFFu, bwait(bar, parity, b, err));
}

// S = -1: d v2p from g_rgb on CUDA cores; S = 0..2: the tensor-core stages.  Output image index = S + 1.
// The 32-column blocks are processed two per loop iteration with the next block's TMEM load (and mask word) in flight;
// the loop is deliberately not unrolled further (instruction-cache footprint of five specialised stages).
template <int S, bool XG>
__device__ __forceinline__ bool bwd_stage(const BwdArgs& a, BEpi& c, int t, uint32_t parity) {
  constexpr int MK = S == -1 ? 3 : S == 0 ? 2 : S == 1 ? 1 : 0;                 // v2, v1, h2, h1
  // sign words, layout [layer][32-column block][row] (mlp_fwd_tc.cu): a warp reads 128 contiguous bytes per block
  const uint32_t* mp = reinterpret_cast<const uint32_t*>(c.masks + (size_t)(MK >= 0 ? MK : 0) * MASK_BYTES) + c.r;
  // all eight words are requested before the accumulator wait (one L2 round trip per stage, not one per block); the loop
  // below picks its pair with selects -- the loop is not unrolled, and a dynamically indexed array would live in local memory
  uint32_t mk[8] = {0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};
  if (MK >= 0) {
#pragma unroll
    for (int j = 0; j < 8; ++j) mk[j] = __ldg(mp + j * TILE);
  }
  if (S >= 0) {
    if (!bwait(&c.bars->acc_full[t], parity, c.bars, a.err)) return false;
    tc_fence_after();
  }
  // the store warp has read the previous image out of this tile (a failed wait has raised the abort flag)
  if (c.elected && c.nimg) (void)bwait(&c.bars->img_saved[t], (c.nimg - 1) & 1u, c.bars, a.err);
  named_bar_sync(c.bar_id, 128);
  const float4* hw = reinterpret_cast<const float4*>(c.heads);
  auto block = [&](int cb, uint32_t (&v)[32], uint32_t w) {
    float f[32];
    if (S >= 0) {
#pragma unroll
      for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]);
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float4 w0 = hw[64 + cb * 8 + j], w1 = hw[128 + cb * 8 + j], w2 = hw[192 + cb * 8 + j];
        f[4 * j] = fmaf(c.g.x, w0.x, fmaf(c.g.y, w1.x, c.g.z * w2.x));
        f[4 * j + 1] = fmaf(c.g.x, w0.y, fmaf(c.g.y, w1.y, c.g.z * w2.y));
        f[4 * j + 2] = fmaf(c.g.x, w0.z, fmaf(c.g.y, w1.z, c.g.z * w2.z));
        f[4 * j + 3] = fmaf(c.g.x, w0.w, fmaf(c.g.y, w1.w, c.g.z * w2.w));
      }
    }
    if (S == 1) {                                   // + d sigma * w_sigma  (fc_out row 0)
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float4 ws = hw[cb * 8 + j];
        f[4 * j] = fmaf(c.g.w, ws.x, f[4 * j]); f[4 * j + 1] = fmaf(c.g.w, ws.y, f[4 * j + 1]);
        f[4 * j + 2] = fmaf(c.g.w, ws.z, f[4 * j + 2]); f[4 * j + 3] = fmaf(c.g.w, ws.w, f[4 * j + 3]);
      }
    }
    if (MK >= 0) {                                  // bit 31-j set: pre-activation of column 32 cb + j was <= 0
#pragma unroll
      for (int j = 0; j < 32; ++j) f[j] = (w & (0x80000000u >> j)) ? 0.f : f[j];
    }
    uint32_t pk[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) pk[j] = pack_bf16(f[2 * j], f[2 * j + 1]);
    store_row_block(c.tile, c.r, cb, pk);
#ifdef CNRF_EXP_COALESCED_DIRECT
    if (c.tile_ok) {
      uint8_t* img = a.dy + (size_t)c.exp_tile * DY_PER_TILE + (size_t)(S + 1) * TILE_BYTES;
#pragma unroll
      for (int j = 0; j < 4; ++j)
        __stcs(reinterpret_cast<uint4*>(img + (size_t)((cb * 4 + j) * 128 + c.r) * 16), make_uint4(pk[4 * j], pk[4 * j + 1], pk[4 * j + 2], pk[4 * j + 3]));
    }
#endif
  };
  uint32_t va[32], vb[32];
  if (S >= 0) tmem_ld32(c.tacc, va);
#pragma unroll 1
  for (int cb = 0; cb < 8; cb += 2) {
    if (S >= 0) { tmem_ld_wait(); tmem_ld32(c.tacc + (cb + 1) * 32, vb); }
    const uint32_t m_a = (cb & 4) ? ((cb & 2) ? mk[6] : mk[4]) : ((cb & 2) ? mk[2] : mk[0]);
    const uint32_t m_b = (cb & 4) ? ((cb & 2) ? mk[7] : mk[5]) : ((cb & 2) ? mk[3] : mk[1]);
    block(cb, va, m_a);
    if (S >= 0) tmem_ld_wait();
    if (cb + 2 < 8 && S >= 0) tmem_ld32(c.tacc + (cb + 2) * 32, va);
    block(cb + 1, vb, m_b);
  }
  fence_proxy_async();
  tc_fence_before();
  named_bar_sync(c.bar_id, 128);
  if (c.elected) {
    mbar_arrive(&c.bars->img_ready[t]); ++c.nimg;
    if (S < 2 || XG) mbar_arrive(&c.bars->a_ready[t]);     // XG: the input-gradient stage reads d h1p as its A operand
  }
  return true;
}

// Input-gradient stage (pose optimisation, eval.py:129-168): d gamma(x)[m][0..dx) = d h1p[m] W1, 64 accumulator columns,
// written as fp32 rows of g_x [M, dx] for the positional-encoding backward.
__device__ __forceinline__ bool bwd_stage_x(const BwdArgs& a, BEpi& c, int t, uint32_t parity, int64_t m) {
  if (!bwait(&c.bars->acc_full[t], parity, c.bars, a.err)) return false;
  tc_fence_after();
  uint32_t va[32], vb[32];
  tmem_ld32(c.tacc, va);
  tmem_ld32(c.tacc + 32, vb);
  tmem_ld_wait();
  if (m < a.M) {
    float* o = a.g_x + m * a.dx;
#pragma unroll
    for (int j = 0; j < 32; ++j) if (j < a.dx) o[j] = __uint_as_float(va[j]);
#pragma unroll
    for (int j = 0; j < 32; ++j) if (32 + j < a.dx) o[32 + j] = __uint_as_float(vb[j]);
  }
  tc_fence_before();
  named_bar_sync(c.bar_id, 128);          // every row of the tile has drained its accumulator columns
  return true;
}

template <bool XG>
__global__ void __launch_bounds__(NTHREADS, 1) mlp_bwd_tc_kernel(const BwdArgs a) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* tiles = smem;
  uint8_t* ring = smem + OFF_RING;
  float* heads = reinterpret_cast<float*>(smem + OFF_HEADS);
  BBars* bars = reinterpret_cast<BBars*>(smem + OFF_BARS);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int64_t num_tiles = (a.M + TILE - 1) / TILE;
  const int64_t num_pairs = (num_tiles + 1) / 2;

  if (tid == 0) {
    for (int i = 0; i < BRING; ++i) { mbar_init(&bars->w_full[i], 1); mbar_init(&bars->w_empty[i], 1); }
    for (int t = 0; t < 2; ++t) {
      mbar_init(&bars->a_ready[t], 1); mbar_init(&bars->acc_full[t], 1);
      mbar_init(&bars->img_ready[t], 1); mbar_init(&bars->img_saved[t], 1);
    }
    bars->abort_flag = 0;
    if ((smem_u32(smem) & 1023u) != 0) { bars->abort_flag = 1; atomicExch(a.err, 2); tc_fatal(); }
    mbar_fence_init();
  }
  for (int i = tid; i < HEAD_FLOATS; i += NTHREADS) heads[i] = a.pk.w_sigma[i];   // w_sigma | w_rgb are contiguous
  if (warp == 1) tmem_alloc(&bars->tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = bars->tmem_slot;

  const int prod = producer_index(warp);
  if (prod >= 0) {
    // ===== TMA producers: per (stage, tile) the 4 transposed K-chunks x 2 N-halves of that stage =====
    // The two tiles of a pair run half a stage apart (one tile's MMAs under the other's epilogue), so the L2-resident
    // weight images pass through the ring once per tile.  NPROD producers share the image sequence round-robin (one
    // issuing thread manages a bulk copy per ~588 clk, the tensor core eats an image in ~280: mlp_tc.cuh).
    if (lane == 0) {
      uint32_t slot = 0, epar = 1;
      uint32_t turn = 0;
      bool ok = true;
      for (int64_t pr = blockIdx.x; pr < num_pairs && ok; pr += gridDim.x) {
        for (int stage = 0; stage < (XG ? 4 : 3) && ok; ++stage) {
          for (int t = 0; t < 2 && ok; ++t) {
            for (int i = 0; i < (stage == 3 ? 2 : 8); ++i) {      // stage 3: W1^T, one 32 KB image = two ring slots
              if ((int)turn == prod) {
                if (!bwait(&bars->w_empty[slot], epar, bars, a.err)) { ok = false; break; }
                mbar_arrive_expect_tx(&bars->w_full[slot], HALF);
                bulk_g2s(ring + slot * HALF, a.pk.bwd + (size_t)(stage * 8 + i) * HALF, HALF, &bars->w_full[slot]);
              }
              if (++turn == NPROD) turn = 0;
              if (++slot == BRING) { slot = 0; epar ^= 1u; }
            }
          }
        }
        if (*(volatile int*)&bars->abort_flag) break;
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer: warp-uniform control flow, one elected lane issues (see mlp_fwd_tc.cu) =====
    {
      const uint32_t tmem_u = __shfl_sync(0xFFFFFFFFu, tmem, 0);
      // two N-halves in adjacent ring slots = one [256 x 64] image: one MMA of N = 256 per K step (see mlp_fwd_tc.cu)
      const uint32_t idesc = idesc_bf16(128, 128, 0, 0), idesc_wide = idesc_bf16(128, 256, 0, 0);
      const uint64_t a_desc0 = smem_desc(smem_u32(tiles), 16, 1024, SWIZZLE_128B);
      const uint64_t b_desc0 = smem_desc(smem_u32(ring), 16, 1024, SWIZZLE_128B);
      uint32_t slot = 0, wpar = 0, na = 0;
      bool ok = true;
      for (int64_t pr = blockIdx.x; pr < num_pairs && ok; pr += gridDim.x) {
#pragma unroll 1
        for (int stage = 0; stage < 3 && ok; ++stage, ++na) {
#pragma unroll 1
          for (int t = 0; t < 2 && ok; ++t) {
            if (!uwait(&bars->a_ready[t], na & 1, bars, a.err)) { ok = false; break; }
            const uint32_t d0 = tmem_u + t * 256;
#pragma unroll 1
            for (int c = 0; c < 4 && ok; ++c) {
              const uint64_t ad = a_desc0 + (uint64_t)((t * TILE_BYTES + c * SLAB) >> 4);
              const uint32_t s0 = slot;
              if (!uwait(&bars->w_full[slot], wpar, bars, a.err)) { ok = false; break; }
              if (++slot == BRING) { slot = 0; wpar ^= 1u; }
              const uint32_t s1 = slot;
              if (!uwait(&bars->w_full[slot], wpar, bars, a.err)) { ok = false; break; }
              if (++slot == BRING) { slot = 0; wpar ^= 1u; }
              tc_fence_after();
              const uint64_t bd0 = b_desc0 + (uint64_t)((s0 * HALF) >> 4), bd1 = b_desc0 + (uint64_t)((s1 * HALF) >> 4);
              if (elect_one()) {
                if (s1 == s0 + 1) {
#pragma unroll
                  for (int k = 0; k < 4; ++k) umma_bf16(d0, ad + 2 * k, bd0 + 2 * k, idesc_wide, (c | k) != 0);
                  umma_commit(&bars->w_empty[s0]);
                } else {                                // the chunk straddles the ring's wrap
#pragma unroll
                  for (int k = 0; k < 4; ++k) umma_bf16(d0, ad + 2 * k, bd0 + 2 * k, idesc, (c | k) != 0);
                  umma_commit(&bars->w_empty[s0]);
#pragma unroll
                  for (int k = 0; k < 4; ++k) umma_bf16(d0 + 128, ad + 2 * k, bd1 + 2 * k, idesc, (c | k) != 0);
                }
                umma_commit(&bars->w_empty[s1]);
              }
              __syncwarp();
            }
            if (ok && elect_one()) umma_commit(&bars->acc_full[t]);
            __syncwarp();
          }
        }
        if (XG) {                          // d gamma(x) = d h1p W1: N = 64, the four K-chunks sit two per ring image
          const uint32_t idesc_x = idesc_bf16(128, 64, 0, 0);
#pragma unroll 1
          for (int t = 0; t < 2 && ok; ++t) {
            if (!uwait(&bars->a_ready[t], na & 1, bars, a.err)) { ok = false; break; }
            const uint32_t d0 = tmem_u + t * 256;
#pragma unroll
            for (int hi = 0; hi < 2; ++hi) {
              if (!ok) continue;
              if (!uwait(&bars->w_full[slot], wpar, bars, a.err)) { ok = false; continue; }
              tc_fence_after();
              if (elect_one()) {
#pragma unroll
                for (int q = 0; q < 2; ++q) {
                  const uint64_t ad = a_desc0 + (uint64_t)((t * TILE_BYTES + (2 * hi + q) * SLAB) >> 4);
                  const uint64_t bd = b_desc0 + (uint64_t)((slot * HALF + q * 8192) >> 4);
#pragma unroll
                  for (int k = 0; k < 4; ++k) umma_bf16(d0, ad + 2 * k, bd + 2 * k, idesc_x, (hi | q | k) != 0);
                }
                umma_commit(&bars->w_empty[slot]);
              }
              __syncwarp();
              if (++slot == BRING) { slot = 0; wpar ^= 1u; }
            }
            if (ok && elect_one()) umma_commit(&bars->acc_full[t]);
            __syncwarp();
          }
          ++na;
        }
      }
    }
  } else if (warp == STORE_WARP) {
    // ===== image-store warp: the four masked gradient images of both tiles, in the order the epilogues finish them =====
    uint32_t par[2] = {0u, 0u};
    bool ok = true;
    for (int64_t pr = blockIdx.x; pr < num_pairs && ok; pr += gridDim.x) {
#pragma unroll 1
      for (int seq = 0; seq < 8 && ok; ++seq) {
        const int k = seq >> 1, t = seq & 1;
        const int64_t tile_idx = pr * 2 + t;
        if (!bwait(&bars->img_ready[t], par[t], bars, a.err)) { ok = false; break; }
        par[t] ^= 1u;
        if (tile_idx < num_tiles)
          warp_copy_image(a.dy + (size_t)tile_idx * DY_PER_TILE + (size_t)k * TILE_BYTES, tiles + t * TILE_BYTES, TILE_BYTES, lane);
        __syncwarp();
        if (lane == 0) mbar_arrive(&bars->img_saved[t]);
      }
    }
  } else if (warp >= 2 && warp < 10) {
    // ===== epilogue warps: thread = one sample row =====
    const int t = (warp - 2) >> 2;
    const int q = warp & 3;
    BEpi c;
    c.bars = bars;
    c.tile = tiles + t * TILE_BYTES;
    c.heads = heads;
    c.r = q * 32 + lane;
    c.bar_id = 1 + t;
    c.elected = ((warp - 2) & 3) == 0 && lane == 0;
    c.tacc = tmem + t * 256 + ((uint32_t)(q * 32) << 16);
    uint32_t na = 0;
    c.nimg = 0;
    for (int64_t pr = blockIdx.x; pr < num_pairs; pr += gridDim.x) {
      const int64_t tile_idx = pr * 2 + t;
      c.tile_ok = tile_idx < num_tiles;
      c.exp_tile = tile_idx;
      const int64_t m = tile_idx * TILE + c.r;
      c.masks = a.saved + (size_t)(c.tile_ok ? tile_idx : 0) * SAVED_PER_TILE + SAVED_MASKS;
      c.g = m < a.M ? __ldg(reinterpret_cast<const float4*>(a.g_raw) + m) : make_float4(0.f, 0.f, 0.f, 0.f);
      bool ok = bwd_stage<-1, XG>(a, c, t, 0);
      if (ok) { ok = bwd_stage<0, XG>(a, c, t, na & 1); ++na; }
      if (ok) { ok = bwd_stage<1, XG>(a, c, t, na & 1); ++na; }
      if (ok) { ok = bwd_stage<2, XG>(a, c, t, na & 1); ++na; }
      if (XG && ok) { ok = bwd_stage_x(a, c, t, na & 1, m); ++na; }
      if (!ok) break;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 512);
}
}  // namespace

int cnrf_tc_launch_bwd(const BwdArgs& b, int grid, cudaStream_t st) {
  {
    static bool attr_set[CNRF_MAX_DEVICES] = {};
    const cudaError_t e = cnrf_once_per_device(attr_set, [] {
      cudaError_t e1 = cudaFuncSetAttribute(mlp_bwd_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, BWD_SMEM);
      if (e1 == cudaSuccess) e1 = cudaFuncSetAttribute(mlp_bwd_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, BWD_SMEM);
      return e1; });
    if (e != cudaSuccess) return (int)e;
  }
  if (b.g_x && (b.dx < 1 || b.dx > 64)) return CNRF_E_UNSUPPORTED;
  const bool timed = cnrf_ktimer_begin(CNRF_KT_BWD, b.M, st);
  if (b.g_x) mlp_bwd_tc_kernel<true><<<grid, NTHREADS, BWD_SMEM, st>>>(b);
  else mlp_bwd_tc_kernel<false><<<grid, NTHREADS, BWD_SMEM, st>>>(b);
  CNRF_LAUNCH_CHECK();
  if (timed) cnrf_ktimer_end(CNRF_KT_BWD, st);
  return CNRF_OK;
}
