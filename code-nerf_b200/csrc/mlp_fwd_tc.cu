// Forward chain of the fused bf16 tensor-core field (CNRF_PRECISION_BF16), models/model.py:179-194 code-folded, with
// fc_out folded into layer_dir1 (no activation between them, :182-189; Wfh = Wdf Wout'[:, :H] rebuilt every call):
//   h1 = relu(W1 gamma(x) + b1) -> h2 = relu(W2h h1 + beta2[ray]), sigma = w_sigma h2 + betas[ray] ->
//   v1 = relu(Wfh h2 + betaf[ray] + [Wdv | bd1] [gamma(d_pair) | 1]) -> v2 = relu(Wd2 v1 + bd2) -> rgb = Wrv v2 + betar[ray]
// One persistent CTA per SM walks PAIRS of 128-row sample tiles (T0, T1) through the four layers half a layer apart: the
// MMAs of one tile run under the epilogue of the other (weights are L2-resident and pass through the ring once per tile).
//   TMEM   : 2 x 256 fp32 accumulator columns (all 512).
//   smem   : 2 x 64 KB activation tiles (4 slabs of [128 x 64] bf16, 128-byte swizzle, overwritten in place layer by
//            layer) + a ring of five 16 KB weight images ([128 out-features x 64 in-features]: the N-half of a K chunk,
//            streamed by TMA bulk copies) + fp32 head vectors (fc_out row 0, fc_rgb, layer_dir2 bias) + a per-tile
//            staging area for the per-ray bias rows of the current layer (cp.async, prefetched one layer ahead).
//   warps  : 0, 10, 11 = TMA producers (one lane each; the images of the stream round-robin), 1 = MMA issuer (one lane),
//            2-5 = epilogue of T0, 6-9 = epilogue of T1 (thread = one sample row, TMEM lane quarter = warp_id % 4).
// The epilogue is specialised per layer at compile time and fully unrolled: TMEM -> registers -> + bias -> ReLU -> bf16
// -> swizzled smem (the next layer's A operand, also TMA-stored as the image the backward needs).
#include <cstdlib>
#include "mlp_tc.cuh"

__device__ long long g_fwd_dbg[16];
// Timer builds only: absolute clock64 stamps of ONE steady-state tile pair of block 0 (its 4th), [tile][layer][event]:
// 0 epilogue starts waiting for the accumulator, 1 accumulator seen, 2 block loop starts, 3 block loop done, 4 hand-off
// (a_ready arrive), 5 MMA warp sees a_ready, 6 MMA warp has issued the layer's last MMA + commit, 7 gamma(x) built (layer 0)
__device__ long long g_fwd_trace[2 * 4 * 8];
// ... and of the weight images of that pair's (layer 1, both tiles) = 16 images: [image][0 producer sees the slot empty,
// 1 producer has issued the copy, 2 MMA warp sees the image, 3 MMA warp has issued its MMAs + commit]
__device__ long long g_fwd_trace_img[16 * 4];
// Phase timers of block 0 (scripts/time_field.py): compiled in only with -DCNRF_FWD_TIMERS, they cost issue slots
#ifdef CNRF_FWD_TIMERS
#define FDBG(i, dt) do { if (blockIdx.x == 0) atomicAdd((unsigned long long*)&g_fwd_dbg[i], (unsigned long long)(dt)); } while (0)
#define FCLK() clock64()
#define FTRACE(it, t, layer, ev, clk) do { if (blockIdx.x == 0 && (it) == 3) g_fwd_trace[((t) * 4 + (layer)) * 8 + (ev)] = (clk); } while (0)
#define FTRACE_IMG(it, layer, img, ev) do { if (blockIdx.x == 0 && (it) == 3 && (layer) == 1 && (img) < 16) g_fwd_trace_img[(img) * 4 + (ev)] = clock64(); } while (0)
#else
#define FDBG(i, dt) do { } while (0)
#define FCLK() 0ll
#define FTRACE(it, t, layer, ev, clk) do { } while (0)
#define FTRACE_IMG(it, layer, img, ev) do { } while (0)
#endif
namespace {
constexpr int HALF = 128 * 128;             // bytes of one [128 x 64] bf16 weight image
constexpr int FRING = 5;
constexpr int RMAX = 4;                     // rays one tile may span for its per-ray biases to be staged in smem
constexpr int HEAD_FLOATS = 256 + 768 + 256;   // w_sigma | w_rgb[3] | b_dir2   (contiguous in Packed)
constexpr int STAGE_FLOATS = RMAX * 256;       // per tile
constexpr int OFF_RING = 2 * TILE_BYTES;
constexpr int OFF_HEADS = OFF_RING + FRING * HALF;
constexpr int OFF_STAGE = OFF_HEADS + HEAD_FLOATS * 4;
constexpr int OFF_BARS = OFF_STAGE + 2 * STAGE_FLOATS * 4;
constexpr int FWD_SMEM = OFF_BARS + 256;

struct FBars {
  uint64_t w_full[FRING], w_empty[FRING];
  uint64_t a_ready[2];     // epilogue -> MMA : tile's operand image written
  uint64_t acc_full[2];    // MMA -> epilogue : tile's accumulator complete
  uint64_t w_peer[FRING];  // CTA pairs: the other CTA's half of the weight image has landed (forwarded by its warp 1)
  uint64_t img_ready[2];   // epilogue -> store warp : the tile holds a finished image the backward needs
  uint64_t img_saved[2];   // store warp -> epilogue : that image has been read out of shared memory
  uint32_t tmem_slot;
  int abort_flag;
};

// Bounded mbarrier wait (about a second): a protocol bug must end the kernel, never hang the GPU.
__device__ __forceinline__ bool fwait(uint64_t* bar, uint32_t parity, FBars* b, int* err) {
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
// the same against arrivals from the other CTA of a pair (cluster-scope acquire)
__device__ __forceinline__ bool fwait_cluster(uint64_t* bar, uint32_t parity, FBars* b, int* err) {
  if (mbar_try_wait_cluster(bar, parity)) return true;
  const long long t0 = clock64();
  while (true) {
    if (mbar_try_wait_cluster(bar, parity)) return true;
    if (*(volatile int*)&b->abort_flag) return false;
    if (clock64() - t0 > 2000000000LL) {
      b->abort_flag = 1;
      if (err) atomicExch(err, 1);
      tc_fatal();
      return false;
    }
  }
}
__device__ __forceinline__ bool uwait_cluster(uint64_t* bar, uint32_t parity, FBars* b, int* err) {
  return __all_sync(0xFFFFFFFFu, fwait_cluster(bar, parity, b, err));
}
// the same for a whole warp, with a result the compiler can prove warp-uniform (keeps the MMA issuer's state in uniform
// registers)
__device__ __forceinline__ bool uwait(uint64_t* bar, uint32_t parity, FBars* b, int* err) {
  return __all_sync(0xFFFFFFFFu, fwait(bar, parity, b, err));
}

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(smem_u32(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

struct Epi {               // per-thread state of an epilogue thread for the current tile
  FBars* bars;
  uint8_t* tile;
  const float* heads;      // smem: w_sigma[256] | w_rgb[3][256] | b_dir2[256]
  float* stage;            // smem: [RMAX][256] bias rows of this tile's rays
  uint32_t tacc;
  int r, bar_id, slot, nrays, it;
  uint32_t nimg;           // images of this tile handed to the store warp so far (training forward)
  bool elected, valid, staged;
  int64_t m, ray0;
  int row;                 // this thread's row of the per-ray bias tables (its ray, or the ray's code run)
  uint8_t* sv;
  float sigma;
  float px, py, pz;        // this row's sample point of the NEXT pair (prefetched while the current pair runs)
  uint32_t ar_cluster;     // CTA pairs: shared::cluster address of the LEADER's a_ready barrier of this tile (0: single CTA)
};

// copy the `nrays` bias rows [ray0, ray0 + nrays) of a [N,256] fp32 table into the tile's staging area (128 threads)
__device__ __forceinline__ void stage_rows(const Epi& c, const float* table, const int32_t* gid) {
  if (c.staged) {
    for (int piece = c.r; piece < c.nrays * 64; piece += 128) {
      const int64_t ray = c.ray0 + (piece >> 6);
      cp_async16(c.stage + piece * 4, table + (gid ? (int64_t)__ldg(gid + ray) : ray) * 256 + (piece & 63) * 4);
    }
  }
  cp_async_commit();
}

// One layer's epilogue for one tile.  E: 0 layer_xyz1 (its bias arrives through the MMA), 1 layer_xyz2 (+ sigma head),
// 2 folded layer_dir1 (per-ray bias here; its constant bias and view-direction half arrive through the MMA),
// 3 layer_dir2 (+ rgb head).  GB: bias row comes from global memory (register double buffer) instead of shared memory.
// SV: the training forward (saved images + sign words); the inference kernel compiles none of that work in.
template <int E, bool GB, bool SV>
__device__ __forceinline__ bool epi_layer(const FwdArgs& a, Epi& c, int t, uint32_t parity) {
  const float4* gb = nullptr;
  const float4* sb = nullptr;
  if (E == 1) { if (GB) gb = reinterpret_cast<const float4*>(a.b2 + (int64_t)c.row * 256); else sb = reinterpret_cast<const float4*>(c.stage + c.slot * 256); }
  if (E == 2) { if (GB) gb = reinterpret_cast<const float4*>(a.bf + (int64_t)c.row * 256); else sb = reinterpret_cast<const float4*>(c.stage + c.slot * 256); }
  if (E == 3) sb = reinterpret_cast<const float4*>(c.heads + 1024);
  float tail0 = 0.f, tail1 = 0.f, tail2 = 0.f;
  if (E == 1 && c.valid) tail0 = __ldg(a.bos + c.row);
  if (E == 3 && c.valid) { const float* brp = a.br + (int64_t)c.row * 3; tail0 = __ldg(brp); tail1 = __ldg(brp + 1); tail2 = __ldg(brp + 2); }
  const long long tw0 = FCLK();
  if (!fwait(&c.bars->acc_full[t], parity, c.bars, a.err)) return false;
  const long long tw1 = FCLK();
  tc_fence_after();
  // the store warp has read the previous image out of this tile
  // (a failed wait has raised the abort flag: every thread leaves at its next wait)
  if (SV && c.elected && c.nimg) (void)fwait(&c.bars->img_saved[t], (c.nimg - 1) & 1u, c.bars, a.err);
  if ((E == 1 || E == 2) && !GB) cp_async_wait_all();
  named_bar_sync(c.bar_id, 128);
  const long long tw2 = FCLK();
  const float4* hw = reinterpret_cast<const float4*>(c.heads);
  float rgb0 = 0.f, rgb1 = 0.f, rgb2 = 0.f, rgb3 = 0.f, rgb4 = 0.f, rgb5 = 0.f;   // two partial sums per channel
  float sg0 = 0.f, sg1 = 0.f, sg2 = 0.f, sg3 = 0.f;                               // four partial sums of the sigma head
  // sign bits of the pre-activations (the backward's ReLU mask): one 32-bit word per 32-column block
  // layout [layer][32-column block][row]: the 32 lanes of a warp write (and the backward reads) 128 contiguous bytes
  uint32_t* mp = (SV && c.sv) ? reinterpret_cast<uint32_t*>(c.sv + SAVED_MASKS + (size_t)E * MASK_BYTES) + c.r : nullptr;
  // One 32-column block: + bias, heads, sign word, ReLU, bf16, swizzled store.  The loop below keeps the next block's
  // TMEM load in flight while this one is processed, and is NOT unrolled further: five specialised layers unrolled eight
  // times each overflowed the instruction cache (16% of the kernel's stall samples were instruction fetches).
  auto block = [&](int cb, uint32_t (&v)[32]) {
    float4 bcur[8];
    if (E != 0) {
#pragma unroll
      for (int j = 0; j < 8; ++j) bcur[j] = GB ? __ldg(gb + cb * 8 + j) : sb[cb * 8 + j];
    }
    float f[32];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      if (E == 0) {
        f[4 * j] = __uint_as_float(v[4 * j]); f[4 * j + 1] = __uint_as_float(v[4 * j + 1]);
        f[4 * j + 2] = __uint_as_float(v[4 * j + 2]); f[4 * j + 3] = __uint_as_float(v[4 * j + 3]);
      } else {
        f[4 * j] = __uint_as_float(v[4 * j]) + bcur[j].x; f[4 * j + 1] = __uint_as_float(v[4 * j + 1]) + bcur[j].y;
        f[4 * j + 2] = __uint_as_float(v[4 * j + 2]) + bcur[j].z; f[4 * j + 3] = __uint_as_float(v[4 * j + 3]) + bcur[j].w;
      }
    }
    if (E == 1) {                                   // sigma_raw = w_sigma . h2 + betao[ray][0]   (fc_out row 0)
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float4 w = hw[cb * 8 + j];
        sg0 = fmaf(fmaxf(f[4 * j], 0.f), w.x, sg0); sg1 = fmaf(fmaxf(f[4 * j + 1], 0.f), w.y, sg1);
        sg2 = fmaf(fmaxf(f[4 * j + 2], 0.f), w.z, sg2); sg3 = fmaf(fmaxf(f[4 * j + 3], 0.f), w.w, sg3);
      }
    }
    if (E == 3) {                                   // rgb_raw = Wrv v2 + betar[ray]                (fc_rgb)
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float x0 = fmaxf(f[4 * j], 0.f), x1 = fmaxf(f[4 * j + 1], 0.f), x2 = fmaxf(f[4 * j + 2], 0.f),
                    x3 = fmaxf(f[4 * j + 3], 0.f);
        const float4 w0 = hw[64 + cb * 8 + j], w1 = hw[128 + cb * 8 + j], w2 = hw[192 + cb * 8 + j];
        rgb0 = fmaf(x0, w0.x, fmaf(x1, w0.y, rgb0)); rgb3 = fmaf(x2, w0.z, fmaf(x3, w0.w, rgb3));
        rgb1 = fmaf(x0, w1.x, fmaf(x1, w1.y, rgb1)); rgb4 = fmaf(x2, w1.z, fmaf(x3, w1.w, rgb4));
        rgb2 = fmaf(x0, w2.x, fmaf(x1, w2.y, rgb2)); rgb5 = fmaf(x2, w2.z, fmaf(x3, w2.w, rgb5));
      }
    }
    if (SV) {
      uint32_t w = 0;
#pragma unroll
      for (int j = 0; j < 32; ++j) w = __funnelshift_l(__float_as_uint(f[j]), w, 1);   // (w << 1) | sign(f[j])
      if (mp) mp[cb * TILE] = w;
    }
    if (E < 3 || SV) {                              // the last layer's image is only ever read by the backward
      uint32_t pk[16];
#pragma unroll
      for (int j = 0; j < 16; ++j) pk[j] = cvt_relu_bf16x2(f[2 * j], f[2 * j + 1]);
      store_row_block(c.tile, c.r, cb, pk);
#ifdef CNRF_EXP_COALESCED_DIRECT
      if (SV && c.sv) {
        uint8_t* img = c.sv + SAVED_GAMMA + (size_t)E * TILE_BYTES;
#pragma unroll
        for (int j = 0; j < 4; ++j)
          __stcs(reinterpret_cast<uint4*>(img + (size_t)((cb * 4 + j) * 128 + c.r) * 16), make_uint4(pk[4 * j], pk[4 * j + 1], pk[4 * j + 2], pk[4 * j + 3]));
      }
#endif
    }
  };
  uint32_t va[32], vb[32];
  tmem_ld32(c.tacc, va);
#pragma unroll 1
  for (int cb = 0; cb < 8; cb += 2) {
    tmem_ld_wait();
    tmem_ld32(c.tacc + (cb + 1) * 32, vb);
    block(cb, va);
    tmem_ld_wait();
    if (cb + 2 < 8) tmem_ld32(c.tacc + (cb + 2) * 32, va);
    block(cb + 1, vb);
  }
  if (E == 1) c.sigma = (sg0 + sg1) + (sg2 + sg3) + tail0;
  if (E == 3 && c.valid)                            // [rgb_raw | sigma_raw]   models/model.py:194
    reinterpret_cast<float4*>(a.raw)[c.m] = make_float4(rgb0 + rgb3 + tail0, rgb1 + rgb4 + tail1, rgb2 + rgb5 + tail2, c.sigma);
  const long long tw3 = FCLK();
  fence_proxy_async();
  tc_fence_before();
  named_bar_sync(c.bar_id, 128);
  const long long tw4 = FCLK();
  if (c.elected && t == 0) { FDBG(10, tw2 - tw1); FDBG(11, tw3 - tw2); FDBG(12, tw4 - tw3); }
  if (c.elected) { FTRACE(c.it, t, E, 0, tw0); FTRACE(c.it, t, E, 1, tw1); FTRACE(c.it, t, E, 2, tw2); FTRACE(c.it, t, E, 3, tw3); FTRACE(c.it, t, E, 4, tw4); }
  if (c.elected) {
    // the hand-off first: a cluster-scope release issued after the bulk store waits for that store (~3.5 k clk)
    if (E < 3) { if (c.ar_cluster) mbar_arrive_cluster(c.ar_cluster); else mbar_arrive(&c.bars->a_ready[t]); }
    if (SV) { mbar_arrive(&c.bars->img_ready[t]); ++c.nimg; }
    if (t == 0) { FDBG(0, tw1 - tw0); FDBG(1 + E, FCLK() - tw1); FDBG(13, FCLK() - tw4); }
  }
  return true;
}

// CG = 1: one CTA per SM on its own.  CG = 2: the CTAs of a cluster of two (the SMs of one TPC) run ONE tcgen05.mma of
// M = 256 per step (cta_group::2): each CTA keeps its own pair of 128-row tiles and streams only HALF of every weight
// image (its 128 of the 256 output features), so a whole layer (4 x 16 KB) fits the 5-slot ring ahead of time -- the
// single-CTA kernel waits one L2 round trip per (layer, tile) for the images the ring cannot hold -- and the tensor core
// reads half as many operand bytes from shared memory per FLOP.  The leader CTA (cluster rank 0) issues; the other CTA's
// warp 1 forwards "my half has landed" to the leader, and both CTAs' epilogues report to the leader's a_ready barriers.
template <int CG, bool SV>
__global__ void __launch_bounds__(NTHREADS, 1) mlp_fwd_tc_kernel(const FwdArgs a) {
  extern __shared__ __align__(1024) uint8_t smem[];   // 128-byte-swizzled images need 1024-byte alignment
  uint8_t* tiles = smem;
  uint8_t* ring = smem + OFF_RING;
  float* heads = reinterpret_cast<float*>(smem + OFF_HEADS);
  float* stage = reinterpret_cast<float*>(smem + OFF_STAGE);
  FBars* bars = reinterpret_cast<FBars*>(smem + OFF_BARS);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int64_t num_tiles = (a.M + TILE - 1) / TILE;
  const int64_t num_pairs = (num_tiles + 1) / 2;
  const uint32_t rank = CG == 2 ? cluster_ctarank() : 0u;
  // both CTAs of a pair run the same number of iterations: pair index = pair_base + rank (possibly past the end: an
  // all-padding pair that computes on zeros and stores nothing)
  const int64_t pair_first = (int64_t)blockIdx.x - rank;

  if (tid == 0) {
    for (int i = 0; i < FRING; ++i) { mbar_init(&bars->w_full[i], 1); mbar_init(&bars->w_empty[i], 1); mbar_init(&bars->w_peer[i], 1); }
    for (int t = 0; t < 2; ++t) {
      mbar_init(&bars->a_ready[t], CG); mbar_init(&bars->acc_full[t], 1);
      mbar_init(&bars->img_ready[t], 1); mbar_init(&bars->img_saved[t], 1);
    }
    bars->abort_flag = 0;
    if ((smem_u32(smem) & 1023u) != 0) { bars->abort_flag = 1; atomicExch(a.err, 2); tc_fatal(); }
    mbar_fence_init();
  }
  for (int i = tid; i < HEAD_FLOATS; i += NTHREADS) heads[i] = a.pk.w_sigma[i];   // w_sigma | w_rgb | b_dir2 are contiguous
  if (warp == 1) { if (CG == 2) tmem_alloc2(&bars->tmem_slot, 512); else tmem_alloc(&bars->tmem_slot, 512); }
  tc_fence_before();
  __syncthreads();
  if (CG == 2) cluster_sync_all();                      // the peer's barriers exist before anyone arrives on them
  tc_fence_after();
  const uint32_t tmem = bars->tmem_slot;

  const int prod = producer_index(warp);
  if (prod >= 0) {
    // ===== TMA producers =====
    // The two tiles of a pair run half a layer apart (T1's MMAs overlap T0's epilogue), so a layer's weight images pass
    // through the ring once per tile: per (layer, tile) the N-halves of its K-chunks in order (CG = 2: only this CTA's
    // half); layer_dir1 is followed by its extension (this pair's [gamma(d) | 1] slab, then [Wdv | b_dir1]).
    // Every producer walks the whole image sequence (slot / parity bookkeeping) and issues the images img % NPROD == prod.
    if (lane == 0) {
      uint32_t slot = 0, epar = 1;          // ring slot and the parity its w_empty barrier is waited with
      uint32_t turn = 0;                    // img % NPROD
      bool ok = true;
      int it_p = -1;
      for (int64_t pb = pair_first; pb < num_pairs && ok; pb += gridDim.x) {
        ++it_p;
        const int64_t pr = pb + rank < num_pairs ? pb + rank : num_pairs - 1;    // padding pair: any valid slab will do
        for (int layer = 0; layer < NUM_TC_LAYERS && ok; ++layer) {
          const int first = layer == 0 ? 0 : 2 + 8 * (layer - 1);          // half-image index in a.pk.fwd
          const int nmain = (layer == 0 ? 2 : 8) / CG;                       // weight images of the layer proper
          const int count = nmain + (layer == 2 ? 1 + 2 / CG : 0);
          for (int t = 0; t < 2 && ok; ++t) {
            for (int i = 0; i < count; ++i) {
              if ((int)turn == prod) {
                if (!fwait(&bars->w_empty[slot], epar, bars, a.err)) { ok = false; break; }
                FTRACE_IMG(it_p, layer, t * count + i, 0);
                const uint8_t* src;
                if (i < nmain) src = a.pk.fwd + (size_t)(first + (CG == 2 ? 2 * i + (int)rank : i)) * HALF;
                else if (i == nmain) src = a.ext + (size_t)pr * SLAB;
                else src = a.pk.fwd + (size_t)EXT_CHUNK * CHUNK + (size_t)(CG == 2 ? (int)rank : i - nmain - 1) * HALF;
                mbar_arrive_expect_tx(&bars->w_full[slot], HALF);
                bulk_g2s(ring + slot * HALF, src, HALF, &bars->w_full[slot]);
                FTRACE_IMG(it_p, layer, t * count + i, 1);
              }
              if (++turn == NPROD) turn = 0;
              if (++slot == FRING) { slot = 0; epar ^= 1u; }
            }
          }
        }
        if (*(volatile int*)&bars->abort_flag) break;
      }
    }
  } else if (warp == 1 && CG == 2 && rank != 0) {
    // ===== CTA pairs, second CTA: tell the leader when this CTA's half of each weight image has landed =====
    {
      uint32_t slot = 0, wpar = 0;
      bool ok = true;
      for (int64_t pb = pair_first; pb < num_pairs && ok; pb += gridDim.x) {
#pragma unroll 1
        for (int layer = 0; layer < NUM_TC_LAYERS && ok; ++layer) {
          const int count = 2 * ((layer == 0 ? 1 : 4) + (layer == 2 ? 2 : 0));     // both tiles
#pragma unroll 1
          for (int i = 0; i < count && ok; ++i) {
            if (!uwait(&bars->w_full[slot], wpar, bars, a.err)) { ok = false; break; }
            if (elect_one()) mbar_arrive_cluster(mapa_u32(smem_u32(&bars->w_peer[slot]), 0));
            __syncwarp();
            if (++slot == FRING) { slot = 0; wpar ^= 1u; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer =====
    // Every lane runs the (warp-uniform) control flow so that ring counters, descriptors and TMEM addresses live in
    // uniform registers; one elected lane issues tcgen05.mma / tcgen05.commit.  (With the whole loop under `if (lane == 0)`
    // ptxas keeps that state in vector registers and pays an R2UR chain per MMA: ~125 instructions per weight image.)
    {
      const uint32_t tmem_u = __shfl_sync(0xFFFFFFFFu, tmem, 0);       // provably warp-uniform copy
      const uint32_t idesc = CG == 2 ? idesc_bf16(256, 256, 0, 0) : idesc_bf16(128, 128, 0, 0);
      // CG = 1: when the two N-halves of a chunk sit in adjacent ring slots (4 chunks of 5: the ring has 5 slots) they form
      // one [256 x 64] image and ONE MMA of N = 256 per K step covers both.  An M128 N128 K16 step reads 8 KB of operands
      // from shared memory for 64 clk of math, which holds the tensor pipe at 75%; N256 reads 12 KB per 128 clk and reaches
      // 92% (scripts/probe_umma_rate.py, profiles/r02_umma_rate_probe.json).
      const uint32_t idesc_wide = idesc_bf16(128, 256, 0, 0);
      // descriptors differ only in their 14-bit start-address field (bytes >> 4): precompute the bases, add offsets
      const uint64_t a_desc0 = smem_desc(smem_u32(tiles), 16, 1024, SWIZZLE_128B);
      const uint64_t b_desc0 = smem_desc(smem_u32(ring), 16, 1024, SWIZZLE_128B);
      uint32_t slot = 0, wpar = 0, na = 0;        // ring slot, its w_full parity, a_ready parity counter
      bool ok = true;
      int it_m = 0;
#define NEXT_SLOT() do { if (++slot == FRING) { slot = 0; wpar ^= 1u; } } while (0)
#define WAIT_IMAGE() (uwait(&bars->w_full[slot], wpar, bars, a.err) && (CG == 1 || uwait_cluster(&bars->w_peer[slot], wpar, bars, a.err)))
#define MMA(d, ad, bd, acc) do { if (CG == 2) umma_bf16_2cta(d, ad, bd, idesc, acc); else umma_bf16(d, ad, bd, idesc, acc); } while (0)
#define COMMIT(bar) do { if (CG == 2) umma_commit_2cta(bar); else umma_commit(bar); } while (0)
      for (int64_t pb = pair_first; pb < num_pairs && ok; pb += gridDim.x) {
#pragma unroll 1
        for (int layer = 0; layer < NUM_TC_LAYERS && ok; ++layer, ++na) {
          const int nch = layer == 0 ? 1 : 4;
#pragma unroll 1
          for (int t = 0; t < 2 && ok; ++t) {
            const long long m1 = FCLK();
            if (!(CG == 2 ? uwait_cluster(&bars->a_ready[t], na & 1, bars, a.err) : uwait(&bars->a_ready[t], na & 1, bars, a.err))) { ok = false; break; }
            FDBG(9, FCLK() - m1);
            if (lane == 0) FTRACE(it_m, t, layer, 5, FCLK());
            const uint32_t d0 = tmem_u + t * 256;
#pragma unroll 1
            for (int c = 0; c < nch && ok; ++c) {
              const uint64_t ad = a_desc0 + (uint64_t)((t * TILE_BYTES + c * SLAB) >> 4);
              if (CG == 1) {
                const long long m0 = FCLK();
                const uint32_t s0 = slot;
                if (!WAIT_IMAGE()) { ok = false; break; }
                if (lane == 0) FTRACE_IMG(it_m, layer, t * 8 + c * 2, 2);
                NEXT_SLOT();
                const uint32_t s1 = slot;
                if (!WAIT_IMAGE()) { ok = false; break; }
                NEXT_SLOT();
                FDBG(8, FCLK() - m0);
                if (lane == 0) FTRACE_IMG(it_m, layer, t * 8 + c * 2 + 1, 2);
                const long long m2 = FCLK();
                tc_fence_after();
                const uint64_t bd0 = b_desc0 + (uint64_t)((s0 * HALF) >> 4), bd1 = b_desc0 + (uint64_t)((s1 * HALF) >> 4);
                if (elect_one()) {
                  if (s1 == s0 + 1) {
#pragma unroll
                    for (int k = 0; k < 4; ++k) umma_bf16(d0, ad + 2 * k, bd0 + 2 * k, idesc_wide, (c | k) != 0);
                    umma_commit(&bars->w_empty[s0]);
                  } else {                              // the chunk straddles the ring's wrap: two MMAs of N = 128 per K step
#pragma unroll
                    for (int k = 0; k < 4; ++k) umma_bf16(d0, ad + 2 * k, bd0 + 2 * k, idesc, (c | k) != 0);
                    umma_commit(&bars->w_empty[s0]);
#pragma unroll
                    for (int k = 0; k < 4; ++k) umma_bf16(d0 + 128, ad + 2 * k, bd1 + 2 * k, idesc, (c | k) != 0);
                  }
                  const long long m3 = FCLK();
                  umma_commit(&bars->w_empty[s1]);
                  FDBG(14, m3 - m2); FDBG(15, FCLK() - m3);
                }
                __syncwarp();
                if (lane == 0) { FTRACE_IMG(it_m, layer, t * 8 + c * 2, 3); FTRACE_IMG(it_m, layer, t * 8 + c * 2 + 1, 3); }
              } else {                                  // CG = 2: one MMA of M = N = 256 over this CTA's and the peer's half
                if (!WAIT_IMAGE()) { ok = false; break; }
                tc_fence_after();
                const uint64_t bd = b_desc0 + (uint64_t)((slot * HALF) >> 4);
                if (elect_one()) {
#pragma unroll
                  for (int k = 0; k < 4; ++k) MMA(d0, ad + 2 * k, bd + 2 * k, (c | k) != 0);
                  COMMIT(&bars->w_empty[slot]);
                }
                __syncwarp();
                NEXT_SLOT();
              }
            }
            if (layer == 2 && ok) {        // += [gamma(d_pair) | 1] [Wdv | b_dir1]^T : K = 32 columns per tile of the shared slab
              const uint32_t slot_x = slot;
              if (!WAIT_IMAGE()) { ok = false; break; }
              NEXT_SLOT();
              const uint64_t xd = b_desc0 + (uint64_t)((slot_x * HALF) >> 4);
              if (CG == 1) {
                const uint32_t s0 = slot;
                if (!WAIT_IMAGE()) { ok = false; break; }
                NEXT_SLOT();
                const uint32_t s1 = slot;
                if (!WAIT_IMAGE()) { ok = false; break; }
                NEXT_SLOT();
                tc_fence_after();
                const uint64_t bd0 = b_desc0 + (uint64_t)((s0 * HALF) >> 4), bd1 = b_desc0 + (uint64_t)((s1 * HALF) >> 4);
                if (elect_one()) {
                  if (s1 == s0 + 1) {
#pragma unroll
                    for (int k = 0; k < 2; ++k) umma_bf16(d0, xd + 2 * (2 * t + k), bd0 + 2 * k, idesc_wide, 1);
                  } else {
#pragma unroll
                    for (int k = 0; k < 2; ++k) umma_bf16(d0, xd + 2 * (2 * t + k), bd0 + 2 * k, idesc, 1);
#pragma unroll
                    for (int k = 0; k < 2; ++k) umma_bf16(d0 + 128, xd + 2 * (2 * t + k), bd1 + 2 * k, idesc, 1);
                  }
                  umma_commit(&bars->w_empty[s0]);
                  umma_commit(&bars->w_empty[s1]);
                  umma_commit(&bars->w_empty[slot_x]);
                }
                __syncwarp();
              } else {
                if (!WAIT_IMAGE()) { ok = false; break; }
                tc_fence_after();
                const uint64_t bd = b_desc0 + (uint64_t)((slot * HALF) >> 4);
                if (elect_one()) {
#pragma unroll
                  for (int k = 0; k < 2; ++k) MMA(d0, xd + 2 * (2 * t + k), bd + 2 * k, 1);
                  COMMIT(&bars->w_empty[slot]);
                  COMMIT(&bars->w_empty[slot_x]);
                }
                __syncwarp();
                NEXT_SLOT();
              }
            }
            if (ok && elect_one()) COMMIT(&bars->acc_full[t]);
            __syncwarp();
            if (lane == 0) FTRACE(it_m, t, layer, 6, FCLK());
          }
        }
        ++it_m;
      }
#undef NEXT_SLOT
#undef WAIT_IMAGE
#undef MMA
#undef COMMIT
    }
  } else if (warp == STORE_WARP) {
    // ===== image-store warp (training forward): per pair gamma(x) of both tiles, then the four layer images of both =====
    if (SV) {
      uint32_t par[2] = {0u, 0u};
      bool ok = true;
      for (int64_t pb = pair_first; pb < num_pairs && ok; pb += gridDim.x) {
        const int64_t pr = pb + rank;
#pragma unroll 1
        for (int seq = 0; seq < 10 && ok; ++seq) {
          const int k = seq >> 1, t = seq & 1;
          const int64_t tile_idx = pr * 2 + t;
          if (!fwait(&bars->img_ready[t], par[t], bars, a.err)) { ok = false; break; }
          par[t] ^= 1u;
          if (a.saved && tile_idx < num_tiles) {
            uint8_t* sv = a.saved + (size_t)tile_idx * SAVED_PER_TILE;
            if (k == 0) warp_copy_image(sv, tiles + t * TILE_BYTES, SLAB, lane);
            else warp_copy_image(sv + SAVED_GAMMA + (size_t)(k - 1) * TILE_BYTES, tiles + t * TILE_BYTES, TILE_BYTES, lane);
          }
          __syncwarp();
          if (lane == 0) mbar_arrive(&bars->img_saved[t]);
        }
      }
    }
  } else if (warp >= 2 && warp < 10) {
    // ===== epilogue warps: thread = one sample row =====
    const int t = (warp - 2) >> 2;                    // tile of the pair
    const int q = warp & 3;                           // TMEM lane quarter this warp may read
    Epi c;
    c.bars = bars;
    c.tile = tiles + t * TILE_BYTES;
    c.heads = heads;
    c.stage = stage + t * STAGE_FLOATS;
    c.r = q * 32 + lane;
    c.bar_id = 1 + t;
    c.elected = ((warp - 2) & 3) == 0 && lane == 0;
    c.tacc = tmem + t * 256 + ((uint32_t)(q * 32) << 16);
    c.ar_cluster = CG == 2 ? mapa_u32(smem_u32(&bars->a_ready[t]), 0) : 0u;
    uint32_t na = 0;
    // the usual encoding (include_input, L = 10 octaves from 1.0) is built in registers from prefetched points
    bool pe_fast = a.pts != nullptr && a.inc && a.L == 10;
    if (pe_fast) {
#pragma unroll
      for (int k = 0; k < 10; ++k) pe_fast = pe_fast && __ldg(a.freqs + k) == (float)(1 << k);
    }
    c.px = c.py = c.pz = 0.f;
    if (pe_fast) {
      const int64_t m1 = ((int64_t)blockIdx.x * 2 + t) * TILE + c.r;      // blockIdx.x = pair_first + rank
      if (m1 < a.M) { c.px = __ldg(a.pts + m1 * 3); c.py = __ldg(a.pts + m1 * 3 + 1); c.pz = __ldg(a.pts + m1 * 3 + 2); }
    }
    c.it = -1;
    c.nimg = 0;
    for (int64_t pb = pair_first; pb < num_pairs; pb += gridDim.x) {
      ++c.it;
      const int64_t pr = pb + rank;
      const int64_t tile_idx = pr * 2 + t;
      const int64_t first = tile_idx * TILE;
      c.m = first + c.r;
      c.valid = c.m < a.M;
      // rays of this tile (32-bit arithmetic whenever the row count allows: these divisions sit on every tile's start)
      int64_t ray, last_ray;
      const int64_t last = (first + TILE - 1 < a.M ? first + TILE - 1 : a.M - 1);
      if (a.M < (1ll << 31)) {
        const uint32_t S = (uint32_t)a.S;
        ray = c.valid ? (uint32_t)c.m / S : 0;
        c.ray0 = first < a.M ? (uint32_t)first / S : 0;
        last_ray = first < a.M ? (uint32_t)last / S : 0;
      } else {
        ray = c.valid ? c.m / a.S : 0;
        c.ray0 = first < a.M ? first / a.S : 0;
        last_ray = first < a.M ? last / a.S : 0;
      }
      c.row = (int)((a.gid && c.valid) ? (int64_t)__ldg(a.gid + ray) : ray);
      c.nrays = first < a.M ? (int)(last_ray - c.ray0) + 1 : 0;
      c.staged = c.nrays <= RMAX;
      c.slot = c.valid ? (int)(ray - c.ray0) : 0;
      c.sv = (SV && a.saved && tile_idx < num_tiles) ? a.saved + (size_t)tile_idx * SAVED_PER_TILE : nullptr;
      c.sigma = 0.f;
      // ---- input image: gamma(x); the bias rows of layer_xyz2 travel to smem meanwhile ----
      const long long tb0 = FCLK();
      // the previous pair's last image has been read out of shared memory
      if (SV && c.elected && c.nimg) (void)fwait(&bars->img_saved[t], (c.nimg - 1) & 1u, bars, a.err);
      named_bar_sync(c.bar_id, 128);
      stage_rows(c, a.b2, a.gid);
      if (pe_fast) {
        const float x[3] = {c.px, c.py, c.pz};
        build_input_row_octaves<10>(x, c.valid, c.tile, c.r);
        const int64_t mn = c.m + (int64_t)gridDim.x * 2 * TILE;          // same row of this CTA's next pair
        if (mn < a.M) { c.px = __ldg(a.pts + mn * 3); c.py = __ldg(a.pts + mn * 3 + 1); c.pz = __ldg(a.pts + mn * 3 + 2); }
      } else {
        build_input_row(a, c.m, c.valid, c.tile, c.r);
      }
      fence_proxy_async();
      named_bar_sync(c.bar_id, 128);
      if (c.elected) {
        if (c.ar_cluster) mbar_arrive_cluster(c.ar_cluster); else mbar_arrive(&bars->a_ready[t]);
        if (SV) { mbar_arrive(&bars->img_ready[t]); ++c.nimg; }
        if (t == 0) { FDBG(6, FCLK() - tb0); FDBG(7, 1); }
        FTRACE(c.it, t, 0, 7, FCLK());
      }
      bool ok = epi_layer<0, false, SV>(a, c, t, na & 1); ++na;
      if (ok) { ok = c.staged ? epi_layer<1, false, SV>(a, c, t, na & 1) : epi_layer<1, true, SV>(a, c, t, na & 1); ++na; }
      if (ok) {
        stage_rows(c, a.bf, a.gid);                        // every thread of the tile is past its reads of the beta2 rows
        ok = c.staged ? epi_layer<2, false, SV>(a, c, t, na & 1) : epi_layer<2, true, SV>(a, c, t, na & 1); ++na;
      }
      if (ok) { ok = epi_layer<3, false, SV>(a, c, t, na & 1); ++na; }
      if (!ok) break;
    }
    cp_async_wait_all();
  }
  tc_fence_before();
  __syncthreads();
  if (CG == 2) cluster_sync_all();                      // nobody leaves while the peer may still arrive on its barriers
  if (warp == 1) { if (CG == 2) tmem_dealloc2(tmem, 512); else tmem_dealloc(tmem, 512); }
}
}  // namespace

int cnrf_tc_launch_fwd(const FwdArgs& f, int grid, cudaStream_t st) {
  // CNRF_FWD_CTA_PAIRS=1 selects the cta_group::2 kernel.  Measured on B200 (fine pass, 655,360 rows): 585 us against
  // 574 us for the single-CTA kernel -- the pair halves the weight traffic and the accumulator wait (37 k -> 20 k clk per
  // tile pair) but pays ~1 k clk per cluster-scope hand-off (five per tile), so it is not the default yet.
  static int pair_mode = -1;
  if (pair_mode < 0) {
    const char* e = getenv("CNRF_FWD_CTA_PAIRS");
    pair_mode = e ? atoi(e) != 0 : 0;
  }
  {
    static bool attr_set[CNRF_MAX_DEVICES] = {};
    const cudaError_t e = cnrf_once_per_device(attr_set, [] {
      cudaError_t e1 = cudaFuncSetAttribute(mlp_fwd_tc_kernel<1, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, FWD_SMEM);
      if (e1 == cudaSuccess) e1 = cudaFuncSetAttribute(mlp_fwd_tc_kernel<1, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, FWD_SMEM);
      if (e1 == cudaSuccess) e1 = cudaFuncSetAttribute(mlp_fwd_tc_kernel<2, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, FWD_SMEM);
      if (e1 == cudaSuccess) e1 = cudaFuncSetAttribute(mlp_fwd_tc_kernel<2, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, FWD_SMEM);
      return e1; });
    if (e != cudaSuccess) return (int)e;
  }
#ifdef CNRF_TUNING
  { static int g_override = -1; if (g_override < 0) { const char* e = getenv("CNRF_FWD_GRID"); g_override = e ? atoi(e) : 0; }
    if (g_override > 0 && g_override < grid) grid = g_override; }       // tuning aid: fewer CTAs
#endif
  const bool timed = cnrf_ktimer_begin(CNRF_KT_FWD, f.M, st);
  if (pair_mode && grid >= 2) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(grid & ~1));         // whole clusters only
    cfg.blockDim = dim3(NTHREADS);
    cfg.dynamicSmemBytes = FWD_SMEM;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    cudaError_t e = f.saved ? cudaLaunchKernelEx(&cfg, mlp_fwd_tc_kernel<2, true>, f) : cudaLaunchKernelEx(&cfg, mlp_fwd_tc_kernel<2, false>, f);
    if (e != cudaSuccess) return (int)e;
    ++g_cnrf_launches;
  } else {
    if (f.saved) mlp_fwd_tc_kernel<1, true><<<grid, NTHREADS, FWD_SMEM, st>>>(f);
    else mlp_fwd_tc_kernel<1, false><<<grid, NTHREADS, FWD_SMEM, st>>>(f);
    CNRF_LAUNCH_CHECK();
  }
  if (timed) cnrf_ktimer_end(CNRF_KT_FWD, st);
  return CNRF_OK;
}

extern "C" int cnrf_debug_fwd_trace(long long* host_out64) {
  cudaDeviceSynchronize();
  // host_out64 holds 128 entries: the pair trace, then the per-image trace
  cudaError_t e = cudaMemcpyFromSymbol(host_out64, g_fwd_trace, 64 * sizeof(long long));
  if (e == cudaSuccess) e = cudaMemcpyFromSymbol(host_out64 + 64, g_fwd_trace_img, 64 * sizeof(long long));
  return (int)e;
}

extern "C" int cnrf_debug_fwd_times(long long* host_out, int reset) {
  cudaDeviceSynchronize();
  int rc = (int)cudaMemcpyFromSymbol(host_out, g_fwd_dbg, 16 * sizeof(long long));
  if (reset) { long long z[16] = {0}; cudaMemcpyToSymbol(g_fwd_dbg, z, sizeof(z)); }
  return rc;
}
