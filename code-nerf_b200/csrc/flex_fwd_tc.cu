// Fused tcgen05 forward chain of the vanilla NeRF field (FlexibleNeRFModel, view_synthesis/models/model.py:51-74) for
// inference: config/lego.yml's full-image render is forward-only, 640,000 rays x 256 samples per frame.
//
// One persistent CTA per SM takes 128-sample tiles through ALL layers without leaving the SM:
//   smem  : five activation slabs [128 x 64] bf16 (128-byte swizzle): slabs 0-3 = the 256 hidden features, overwritten in
//           place layer by layer; slab 4 = gamma(x) (63 columns + a zero), kept for the skip layer, whose K is simply five
//           slabs instead of four -- cat(h, xyz) (models/model.py:63) never exists.  A ring of three 40 KB slots, each a 32 KB weight slab
//           ([256 out x 64 in] bf16, pre-swizzled in global memory by pack_flex_weights_kernel, so one slab is ONE 1-D TMA
//           bulk copy) runs ahead of the MMAs across layers and tiles.
//   TMEM  : two [128 x 256] fp32 accumulators, used by alternate layers.
//   bias  : comes through the tensor core.  Column 63 of the gamma(x) slab is 1 (dim_xyz <= 63): layer1 and the skip layer
//           carry their bias (bf16) in column 63 of that slab's weights; every other hidden layer issues ONE more K16 MMA of
//           gamma(x)'s last K chunk against an 8 KB image [256 x 16] that is zero but for k = 15 (it travels behind the layer's
//           first weight slab in the same bulk copy and opens the layer's accumulation; handing slab 0 over in two halves
//           of 16 columns per thread was measured too: 171.5 against 168.7 ms per lego frame, not kept).  The epilogue no longer reads a bias vector from shared memory (a third of
//           its shared-memory wavefronts, on the pipe the MMAs' operand reads already load to 75 %) nor adds it.
//   warps : 0-7 epilogue (tcgen05.ld -> ReLU -> bf16 -> swizzled smem = next layer's A operand), 8 = TMA
//           producer, 9 = MMA issuer (elect.sync lane, M128 N256 K16).  -DFT_EPI_WARPS_N=16 gives every sample row four
//           epilogue threads instead of two: measured 2.5 % SLOWER on a lego frame (226 vs 221 ms) -- the epilogue is bound
//           by the shared-memory pipe it shares with the MMAs' operand reads, not by latency.
// Layer-to-layer overlap inside a tile: the eight epilogue warps finish the next layer's A operand one 64-column slab at a
// time (slab s = columns 64 s .. 64 s + 63: warp half h converts its 32 of them) and arrive on that slab's barrier; the MMA
// warp starts the next layer's K slab s as soon as slab s is there, accumulating into the OTHER TMEM buffer, so only the last
// K slab's MMAs remain when the epilogue ends (before: MMA ~2.0 k clk, then epilogue ~3.9 k clk, nothing overlapped).
// Heads: fc_alpha rides in the layers_dir[0] GEMM as output row 128 (N = 144), so sigma costs nothing; the view-direction
// half of layers_dir[0] arrives as a per-ray fp32 bias row (pairing as nerf/__init__.py:127); fc_rgb (128 -> 3) is a
// CUDA-core dot product in that layer's epilogue.  bf16 operands / fp32 accumulation, like the CodeNeRF chain.
#include <algorithm>
#include "common.cuh"
#include "internal.h"
#include "tc_common.cuh"

using namespace tc;

// Timer builds only (-DCNRF_FLEX_TIMERS, scripts/trace_flex.py): clock64 stamps of block 0's 4th tile, [layer][event]:
//   0 epilogue sees the accumulator, 1-4 epilogue warp 0 has arrived on slab 0-3 (last layer: 1 = raw written), 5 gamma(x) built,
//   8 + 3 i + {0 A slab seen, 1 weight slab seen, 2 MMAs issued} for K slab i of the MMA warp, 23 accumulator committed,
//   24 + 2 i + {0 ring slot seen empty, 1 copy issued} for K slab i of the producer warp.
__device__ long long g_flex_trace[18 * 40];
#ifdef CNRF_FLEX_TIMERS
#define XTRACE(it, l, ev) do { if (blockIdx.x == 0 && (it) == 3 && (threadIdx.x & 31) == 0) g_flex_trace[(l) * 40 + (ev)] = clock64(); } while (0)
#else
#define XTRACE(it, l, ev) do { } while (0)
#endif

namespace {
#ifndef FT_EPI_WARPS_N
#define FT_EPI_WARPS_N 8
#endif
constexpr int FT_ROWS = 128, FT_EPI_WARPS = FT_EPI_WARPS_N, FT_THREADS = 32 * (FT_EPI_WARPS + 2);
constexpr int EH = FT_EPI_WARPS / 4;           // epilogue threads per sample row (TMEM lane quarter = warp_id % 4)
constexpr int CW = 64 / EH;                    // columns of every 64-column slab one thread converts
constexpr int DW = 128 / EH;                   // columns of the direction layer one thread reduces
static_assert(EH == 2 || EH == 4, "8 or 16 epilogue warps");
constexpr int SLAB_A = FT_ROWS * 128;          // 16 KB activation slab
constexpr int SLAB_W = 256 * 128;              // 32 KB weight slab (N = 256)
constexpr int SLAB_WD = 144 * 128;             // 18 KB weight slab of the direction layer (N = 128 + sigma row, padded to 144)
constexpr int BIAS_IMG = 256 * 32;             // 8 KB: [256 out x 16 in] bf16, no swizzle (core matrices of 8 rows x 16 B)
constexpr int SLOT = SLAB_W + BIAS_IMG;        // a ring slot: a weight slab, + the layer's bias image behind its first one
constexpr int RING = 3;
constexpr int FT_MAX_LAYERS = 18;
constexpr int FT_MAX_FREQS = 16;
constexpr int NSLAB_A = 6;                     // 0-3 hidden features, 4 gamma(x) | 1, 5 gamma(d) | 1
constexpr int FT_SMEM = NSLAB_A * SLAB_A + RING * SLOT + 3 * 128 * 4 /*w_rgb*/ +
                        (EH - 1) * 128 * 4 * 4 /*rgb partials*/ + FT_MAX_FREQS * 4 + 256 /*barriers*/ + 1024 /*alignment*/;

struct FlexLayer {
  uint32_t wofs;      // byte offset of the layer's first weight slab in the packed image
  int16_t kslabs;     // K slabs of 64
  int16_t a_first;    // first activation slab: 0 (hidden) or 4 (layer1 reads gamma(x))
  int16_t skip;       // 1: the layer's last K slab is activation slab `pad` (4: gamma(x), 5: gamma(d)), issued first
  int16_t n_out;      // 256, or 144 for the direction layer
  int16_t bias_img;   // 1: an 8 KB bias image follows the first K slab (one more K16 MMA against gamma(x)'s ones column)
  int16_t pad;        // the extra slab of a skip layer
};
struct FlexFwdArgs {
  FlexLayer layer[FT_MAX_LAYERS];
  int nlayers;                       // layer1, layers_xyz..., fc_feat, layers_dir[0]
  const uint8_t* packed;
  const float* x_xyz; int64_t ld_xyz; int dx;
  // pts != NULL: gamma(x) is computed here from the sample positions [M,3] (position_embed.py:44-53), x_xyz is not read
  const float* pts; int64_t ld_pts; const float* freqs; int nfreq; int inc;
  const float* x_dir; int64_t ld_dir; int dd;   // [N, dd] embedded view directions, one row per RAY (pairing below)
  const float* w_rgb; const float* b_rgb;
  float* raw;                        // [M, 4]
  int64_t M, N; int S, pairing;
  int* err;
};

// bias != NULL: column `bias_col` of the slab holds the layer's bias (multiplied by the ones column of gamma(x));
// kcols < 0: the job is a bias image instead -- [256 x 16] without swizzle, zero but for k = 15
struct PackJob { const float* w; const float* bias; uint32_t ofs; int ldw; int k0; int16_t rows, kcols, nrows_slab, bias_col; };
static_assert(sizeof(PackJob) == 40, "PackArgs must stay under the 4 KB kernel-parameter limit");
struct PackArgs { PackJob job[FT_MAX_LAYERS * 5 + 4]; int count; uint8_t* out; };

// one block per slab: element (n, kk) of the slab <- W[n][k0 + kk] (0 outside), bf16, 128-byte swizzled
__global__ void pack_flex_weights_kernel(const __grid_constant__ PackArgs a) {
  const PackJob& j = a.job[blockIdx.x];
  uint8_t* dst = a.out + j.ofs;
  if (j.kcols < 0) {      // bias image: element (n, k) at (n / 8) * 256 + (k / 8) * 128 + (n % 8) * 16 + (k % 8) * 2
    for (int i = threadIdx.x; i < 256 * 2; i += blockDim.x) {
      const int n = i >> 1, kc = i & 1;
      uint4 p = make_uint4(0u, 0u, 0u, 0u);
      if (kc == 1) p.w = pack_bf16(0.f, __ldg(j.bias + n));                 // k = 15: the high half of the last word
      *reinterpret_cast<uint4*>(dst + (n >> 3) * 256 + kc * 128 + (n & 7) * 16) = p;
    }
    return;
  }
  for (int i = threadIdx.x; i < j.nrows_slab * 8; i += blockDim.x) {       // 16-byte chunks
    const int n = i >> 3, c = i & 7;
    float v[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int kk = c * 8 + e;
      v[e] = (n < j.rows && kk < j.kcols) ? __ldg(j.w + (int64_t)n * j.ldw + j.k0 + kk) : 0.f;
      if (j.bias && kk == j.bias_col && n < j.rows) v[e] = __ldg(j.bias + n);
    }
    const uint4 p = make_uint4(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]), pack_bf16(v[4], v[5]), pack_bf16(v[6], v[7]));
    *reinterpret_cast<uint4*>(dst + n * 128 + ((c ^ (n & 7)) << 4)) = p;
  }
}

// bf16x2 of (max(lo, 0), max(hi, 0)), `lo` in the low half: one F2FP with the .relu modifier
__device__ __forceinline__ uint32_t pack_bf16_relu(float lo, float hi) {
  uint32_t d;
  asm("cvt.rn.relu.bf16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(hi), "f"(lo));
  return d;
}

// gamma(x) columns [CW H, CW H + CW) of one sample: column 3 (INC + 2 k) + d = sin(f_k x_d), three columns on = cos(f_k x_d)
// (position_embed.py:44-53).  One sincosf per (band, channel) the column range touches; every index is static once unrolled.
// sin / cos of t for the bf16 operand: two-constant Cody-Waite reduction to [-pi, pi] (|t| reaches 2^9 |x|, a few thousand
// radians), then the SFU approximations (2^-21.4 absolute there): ~5e-7 absolute, four orders below a bf16 ulp, for about
// 8 instructions instead of the ~45 of sincosf.
__device__ __forceinline__ void sincos_bf16_operand(float t, float* sn, float* cs) {
  const float n = rintf(t * 0.15915494309189535f);
  float y = fmaf(n, -6.2831854820251465f, t);
  y = fmaf(n, 1.7484555e-7f, y);                     // 2 pi = 6.2831854820251465 - 1.7484555e-7
  *sn = __sinf(y);
  *cs = __cosf(y);
}

template <int H, int INC>
__device__ __forceinline__ void gamma_cols(const float (&x)[3], const float* fr, int nfreq, float (&v)[CW]) {
  constexpr int J0 = CW * H, J1 = J0 + CW;
  constexpr int B0 = J0 / 3 - INC, B1 = (J1 - 1) / 3 - INC;
  constexpr int K0 = B0 < 0 ? 0 : B0 >> 1, K1 = B1 >> 1;
#pragma unroll
  for (int j = 0; j < CW; ++j) v[j] = 0.f;
  if (INC && H == 0) { v[0] = x[0]; v[1] = x[1]; v[2] = x[2]; }
#pragma unroll
  for (int k = K0; k <= K1; ++k) {
    if (k < nfreq) {
      const float f = fr[k];
#pragma unroll
      for (int d = 0; d < 3; ++d) {
        float sn, cs;
        sincos_bf16_operand(__fmul_rn(x[d], f), &sn, &cs);
        const int js = 3 * (INC + 2 * k) + d - J0, jc = js + 3;
        if (js >= 0 && js < CW) v[js] = sn;
        if (jc >= 0 && jc < CW) v[jc] = cs;
      }
    }
  }
}
template <int INC>
__device__ __forceinline__ void gamma_row_part(int h, const float (&x)[3], const float* fr, int nfreq, float (&v)[CW]) {
  if (h == 0) gamma_cols<0, INC>(x, fr, nfreq, v);
  else if (h == 1) gamma_cols<1, INC>(x, fr, nfreq, v);
  else if (EH == 4 && h == 2) gamma_cols<2 % EH, INC>(x, fr, nfreq, v);
  else gamma_cols<3 % EH, INC>(x, fr, nfreq, v);
}
template <int W> struct TmemLd;
template <> struct TmemLd<32> { static __device__ __forceinline__ void ld(uint32_t t, uint32_t (&r)[32]) { tmem_ld32(t, r); } };
template <> struct TmemLd<16> { static __device__ __forceinline__ void ld(uint32_t t, uint32_t (&r)[16]) { tmem_ld16(t, r); } };

__device__ __forceinline__ void epi_arrive(uint64_t* bar, int lane) {
  fence_proxy_async();            // generic-proxy smem writes -> visible to the tensor core's async proxy
  tc_fence_before();
  __syncwarp();
  if (lane == 0) mbar_arrive(bar);
}

__global__ void __launch_bounds__(FT_THREADS, 1) flex_fwd_tc_kernel(const __grid_constant__ FlexFwdArgs a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* act = smem;                                     // 5 slabs
  uint8_t* ring = smem + NSLAB_A * SLAB_A;                       // RING x 40 KB
  float* wrgb_sm = reinterpret_cast<float*>(ring + RING * SLOT);
  float* part_sm = wrgb_sm + 3 * 128;                      // [128][4]
  float* freq_sm = part_sm + (EH - 1) * 128 * 4;                      // [FT_MAX_FREQS]
  uint64_t* bars = reinterpret_cast<uint64_t*>(freq_sm + FT_MAX_FREQS);
  uint64_t *full = bars, *empty = bars + RING, *acc_bar = bars + 2 * RING, *act_bar = bars + 2 * RING + 1;
  uint64_t* slab_bar = bars + 2 * RING + 2;            // [4]: the next layer's A slab s is written (8 arrivals: every epilogue warp)
  uint64_t* dir_bar = bars + 2 * RING + 6;             // gamma(d) of the tile is written (the direction layer's extra K slab)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * RING + 7);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (tid == 0) {
    for (int i = 0; i < RING; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    mbar_init(acc_bar, 1);
    mbar_init(act_bar, FT_EPI_WARPS);                  // gamma(x) of the tile is written (layer1's A operand)
    for (int i = 0; i < 4; ++i) mbar_init(&slab_bar[i], FT_EPI_WARPS);
    mbar_init(dir_bar, FT_EPI_WARPS);
    mbar_fence_init();
  }
  for (int i = tid; i < 3 * 128; i += FT_THREADS) wrgb_sm[i] = __ldg(a.w_rgb + i);
  if (a.pts && tid < a.nfreq) freq_sm[tid] = __ldg(a.freqs + tid);
  for (int i = tid; i < SLAB_A / 16; i += FT_THREADS)          // gamma(d) slab: the columns past dd stay zero for good
    reinterpret_cast<uint4*>(act + 5 * SLAB_A)[i] = make_uint4(0u, 0u, 0u, 0u);
  if (warp == FT_EPI_WARPS + 1) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const int64_t ntiles = (a.M + FT_ROWS - 1) / FT_ROWS;
  bool ok = true;

  if (warp == FT_EPI_WARPS) {
    // ---- TMA producer: the global sequence of weight slabs (tile-major, layer, K slab) through the ring
    uint32_t slot = 0, par = 1;
    int it = 0;
    for (int64_t t = blockIdx.x; t < ntiles && ok; t += gridDim.x, ++it) {
      for (int l = 0; l < a.nlayers && ok; ++l) {
        const FlexLayer L = a.layer[l];
        const uint32_t bytes = L.n_out == 256 ? SLAB_W : SLAB_WD;
        for (int i = 0; i < L.kslabs && ok; ++i) {
          const int j = L.skip ? (i == 0 ? L.kslabs - 1 : i - 1) : i;      // K-slab order of the MMA warp (gamma(x) first)
          ok = mbar_wait(&empty[slot], par, a.err);
          XTRACE(it, l, 24 + 2 * i);
          // the bias image lies behind the layer's FIRST slab in the packed stream and travels with the first copy
          const uint32_t nb = bytes + ((L.bias_img && i == 0) ? (uint32_t)BIAS_IMG : 0u);
          const size_t src = L.wofs + (size_t)j * bytes + ((L.bias_img && i > 0) ? (size_t)BIAS_IMG : 0);
          if (ok && elect_one()) {
            mbar_arrive_expect_tx(&full[slot], nb);
            bulk_g2s(ring + slot * SLOT, a.packed + src, nb, &full[slot]);
          }
          __syncwarp();
          XTRACE(it, l, 25 + 2 * i);
          if (++slot == RING) { slot = 0; par ^= 1u; }
        }
      }
    }
  } else if (warp == FT_EPI_WARPS + 1) {
    // ---- MMA issuer.  The 64-bit operand descriptors differ only in their start-address field (bytes >> 4): the two
    // bases are built once and every MMA adds an offset, which keeps the issue sequence of a K slab short.
    const uint32_t tmem_u = __shfl_sync(0xFFFFFFFFu, tmem, 0);          // provably warp-uniform copy
    const uint64_t a_desc0 = smem_desc(smem_u32(act), 16, 1024, SWIZZLE_128B);
    const uint64_t b_desc0 = smem_desc(smem_u32(ring), 16, 1024, SWIZZLE_128B);
    // bias: K chunk 3 of gamma(x) (columns 48-63, column 63 == 1) against the image [256 x 16], zero but for k = 15
    const uint64_t ones_desc = a_desc0 + (uint64_t)((uint32_t)(4 * SLAB_A + 96) >> 4);
    const uint64_t bias_desc0 = smem_desc(smem_u32(ring) + SLAB_W, 128, 256, 0u /*no swizzle*/);
    uint32_t slot = 0, par = 0, act_ph = 0, slab_ph = 0, dir_ph = 0;
    int it = 0;
    for (int64_t t = blockIdx.x; t < ntiles && ok; t += gridDim.x, ++it) {
#pragma unroll 1
      for (int l = 0; l < a.nlayers && ok; ++l) {
        const FlexLayer L = a.layer[l];
        const uint32_t idesc = idesc_bf16(FT_ROWS, L.n_out, 0, 0);
        const uint32_t acc = tmem_u + (uint32_t)(l & 1) * 256u;   // alternate layers, alternate accumulators
        if (l == 0) {                                       // layer1 reads gamma(x): written once per tile
          ok = __all_sync(0xFFFFFFFFu, mbar_wait(act_bar, act_ph, a.err));
          act_ph ^= 1;
          tc_fence_after();
        }
#pragma unroll 1
        for (int i = 0; i < L.kslabs && ok; ++i) {
          const int j = L.skip ? (i == 0 ? L.kslabs - 1 : i - 1) : i;      // K slab (the skip layer starts with gamma(x), always there)
          const int aslab = (L.skip && j == L.kslabs - 1) ? L.pad : L.a_first + j;
          if (l > 0 && aslab < 4) {                         // a hidden slab: the previous layer's epilogue has finished it
            ok = __all_sync(0xFFFFFFFFu, mbar_wait(&slab_bar[aslab], slab_ph, a.err));
            if (!ok) break;
          } else if (aslab == 5) {                          // gamma(d): written once per tile, early
            ok = __all_sync(0xFFFFFFFFu, mbar_wait(dir_bar, dir_ph, a.err));
            dir_ph ^= 1;
            if (!ok) break;
          }
          XTRACE(it, l, 8 + 3 * i);
          ok = __all_sync(0xFFFFFFFFu, mbar_wait(&full[slot], par, a.err));
          tc_fence_after();
          XTRACE(it, l, 9 + 3 * i);
          const uint64_t ad = a_desc0 + (uint64_t)((uint32_t)(aslab * SLAB_A) >> 4);
          const uint64_t bd = b_desc0 + (uint64_t)((slot * (uint32_t)SLOT) >> 4);
          // the bias MMA opens the layer (gamma(x) is always there): nothing but the last slab's four MMAs is left when
          // the previous epilogue ends
          const bool with_bias = L.bias_img && i == 0;
          if (ok && elect_one()) {
            if (with_bias) umma_bf16(acc, ones_desc, bias_desc0 + (uint64_t)((slot * (uint32_t)SLOT) >> 4), idesc, 0u);
#pragma unroll
            for (int k = 0; k < 4; ++k) umma_bf16(acc, ad + 2 * k, bd + 2 * k, idesc, (i | k) != 0 || with_bias);
            umma_commit(&empty[slot]);                      // the slot is free once these MMAs have read it
          }
          __syncwarp();
          XTRACE(it, l, 10 + 3 * i);
          if (++slot == RING) { slot = 0; par ^= 1u; }
        }
        if (l > 0) slab_ph ^= 1;                            // every slab barrier completes once per hidden-layer epilogue
        if (ok && elect_one()) umma_commit(acc_bar);
        __syncwarp();
        XTRACE(it, l, 23);
      }
    }
  } else {
    // ---- epilogue warps: warp w owns TMEM lanes / rows 32 (w & 3) .. +31 and column part h = w >> 2 of EH
    const int q = warp & 3, h = warp >> 2;
    const int row = q * 32 + lane;
    uint32_t acc_ph = 0;
    int it = 0;
    // gamma(x) | 1 of tile t -> slab 4: this thread converts columns [CW h, CW h + CW) of its row.  Runs for the FIRST tile
    // before the loop and for every later tile while the MMA warp works on the previous tile's direction layer (the
    // epilogue warps would idle there; slab 4's last reader, fc_feat's bias MMA, has completed by then).
    auto build_gamma_x = [&](int64_t t) {
      const int64_t m = t * FT_ROWS + row;
      float v[CW];
      if (a.pts) {
        const float* pr = a.pts + (m < a.M ? m : 0) * a.ld_pts;
        const float x[3] = {__ldg(pr), __ldg(pr + 1), __ldg(pr + 2)};
        if (a.inc) gamma_row_part<1>(h, x, freq_sm, a.nfreq, v); else gamma_row_part<0>(h, x, freq_sm, a.nfreq, v);
      } else {
        const float* xr = a.x_xyz + (m < a.M ? m : 0) * a.ld_xyz;
#pragma unroll
        for (int e = 0; e < CW; ++e) {
          const int col = h * CW + e;
          v[e] = (m < a.M && col < a.dx) ? __ldg(xr + col) : 0.f;
        }
      }
      if (h == EH - 1) v[CW - 1] = 1.f;         // column 63: the ones column every layer's bias is multiplied by (dx <= 63)
#pragma unroll
      for (int c = 0; c < CW / 8; ++c) {
        const uint4 p = make_uint4(pack_bf16(v[c * 8], v[c * 8 + 1]), pack_bf16(v[c * 8 + 2], v[c * 8 + 3]),
                                   pack_bf16(v[c * 8 + 4], v[c * 8 + 5]), pack_bf16(v[c * 8 + 6], v[c * 8 + 7]));
        *reinterpret_cast<uint4*>(act + 4 * SLAB_A + row * 128 + (((h * (CW / 8) + c) ^ (row & 7)) << 4)) = p;
      }
      epi_arrive(act_bar, lane);
    };
    // gamma(d) | 1 of tile t -> slab 5: row r gets the embedded direction of ITS ray (reference pairing: sample m belongs
    // to ray m % N, nerf/__init__.py:127; correct pairing: m / S) and a one in column dd, which the direction layer's bias
    // and fc_alpha's are multiplied by.  Runs after the tile's first epilogue, while the MMA warp is busy with layer 2: only
    // the tile's last layer reads the slab.
    // (warp w takes rows w, w + W, ...: one 108-byte row per load instruction, lane = column; the ray of each row is
    // computed by one lane -- a single division -- and handed round by shuffles)
    const int wd = a.dd + 1;
    auto build_gamma_d = [&](int64_t t) {
      if (warp == 0) XTRACE(it, 1, 34);
      constexpr int RPW = FT_ROWS / FT_EPI_WARPS;          // rows per warp
      const int64_t m0 = t * FT_ROWS;
      int my_ray = 0;
      if (lane < RPW) {
        const int64_t mr = m0 + warp + FT_EPI_WARPS * lane;
        const int64_t mm = mr < a.M ? mr : a.M - 1;
        if (a.M < ((int64_t)1 << 31))
          my_ray = a.pairing == CNRF_PAIRING_REFERENCE ? (int)((uint32_t)mm % (uint32_t)a.N) : (int)((uint32_t)mm / (uint32_t)a.S);
        else
          my_ray = a.pairing == CNRF_PAIRING_REFERENCE ? (int)(mm % a.N) : (int)(mm / a.S);
      }
      for (int c0 = 0; c0 < wd; c0 += 32) {
        const int c = c0 + lane;
        float v[RPW];
#pragma unroll
        for (int i = 0; i < RPW; ++i) {
          const int ray = __shfl_sync(0xFFFFFFFFu, my_ray, i);
          v[i] = c < a.dd ? __ldg(a.x_dir + (int64_t)ray * a.ld_dir + c) : 1.f;
        }
        if (warp == 0) XTRACE(it, 1, 35);
        if (c < wd) {
#pragma unroll
          for (int i = 0; i < RPW; ++i) {
            const int r = warp + FT_EPI_WARPS * i;
            *reinterpret_cast<__nv_bfloat16*>(act + 5 * SLAB_A + r * 128 + (((c >> 3) ^ (r & 7)) << 4) + (c & 7) * 2) =
                __float2bfloat16_rn(v[i]);
          }
        }
      }
      if (warp == 0) XTRACE(it, 1, 36);
      epi_arrive(dir_bar, lane);
      if (warp == 0) XTRACE(it, 1, 37);
    };
    if ((int64_t)blockIdx.x < ntiles) build_gamma_x(blockIdx.x);
    for (int64_t t = blockIdx.x; t < ntiles && ok; t += gridDim.x, ++it) {
      const int64_t m = t * FT_ROWS + row;
      if (warp == 0) XTRACE(it, 0, 5);
      for (int l = 0; l < a.nlayers && ok; ++l) {
        ok = mbar_wait(acc_bar, acc_ph, a.err);
        acc_ph ^= 1;
        tc_fence_after();
        if (!ok) break;
        if (warp == 0) XTRACE(it, l, 0);
        const uint32_t acc = tmem + (uint32_t)(l & 1) * 256u;
        if (l + 1 < a.nlayers) {
          // hidden layer: slab s = columns 64 s .. 64 s + 63 of the next A operand; this thread converts CW of them, then the
          // warp arrives on the slab's barrier -- the MMA warp starts the next layer's K slab s right away
          const uint32_t t0 = acc + ((uint32_t)(q * 32) << 16) + h * CW;
          uint32_t v[2][CW];
          TmemLd<CW>::ld(t0, v[0]);
#pragma unroll
          for (int sl = 0; sl < 4; ++sl) {
            if (warp == 0 && sl == 1) XTRACE(it, l, 30);
            tmem_ld_wait();
            if (warp == 0 && sl == 1) XTRACE(it, l, 31);
            if (sl + 1 < 4) TmemLd<CW>::ld(t0 + (sl + 1) * 64, v[(sl + 1) & 1]);      // next slab's block in flight under this one
#pragma unroll
            for (int c = 0; c < CW / 8; ++c) {
              float x[8];         // the bias is already in the accumulator (it came through the tensor core)
#pragma unroll
              for (int e = 0; e < 8; ++e) x[e] = __uint_as_float(v[sl & 1][c * 8 + e]);
              // ReLU rides in the conversion (cvt.rn.relu.bf16x2.f32): no FMNMX per element
              const uint4 p = make_uint4(pack_bf16_relu(x[0], x[1]), pack_bf16_relu(x[2], x[3]), pack_bf16_relu(x[4], x[5]),
                                         pack_bf16_relu(x[6], x[7]));
              *reinterpret_cast<uint4*>(act + sl * SLAB_A + row * 128 + (((h * (CW / 8) + c) ^ (row & 7)) << 4)) = p;
            }
            if (warp == 0 && sl == 1) XTRACE(it, l, 32);
#ifdef CNRF_FLEX_TIMERS
            fence_proxy_async();
            if (warp == 0 && sl == 1) XTRACE(it, l, 33);
#endif
            epi_arrive(&slab_bar[sl], lane);
            if (warp == 0) XTRACE(it, l, 1 + sl);
          }
          if (l == 0) build_gamma_d(t);
          if (l == a.nlayers - 2 && t + gridDim.x < ntiles) build_gamma_x(t + gridDim.x);
        } else {
          // direction layer: columns [DW h, DW h + DW) -> ReLU -> fc_rgb partial sums; column 128 = sigma (the per-ray
          // direction term and both biases are already in the accumulator: they came in as the gamma(d) | 1 K slab)
          float r = 0.f, g = 0.f, bl = 0.f;
          uint32_t v[DW / 32][32];
#pragma unroll
          for (int cb = 0; cb < DW / 32; ++cb) tmem_ld32(acc + ((uint32_t)(q * 32) << 16) + h * DW + cb * 32, v[cb]);
          tmem_ld_wait();
#pragma unroll
          for (int cb = 0; cb < DW / 32; ++cb) {
#pragma unroll
            for (int e = 0; e < 32; e += 4) {
              const int n = h * DW + cb * 32 + e;
              const float4 wr = *reinterpret_cast<const float4*>(wrgb_sm + n);
              const float4 wg = *reinterpret_cast<const float4*>(wrgb_sm + 128 + n);
              const float4 wb = *reinterpret_cast<const float4*>(wrgb_sm + 256 + n);
              const float x0 = fmaxf(__uint_as_float(v[cb][e]), 0.f), x1 = fmaxf(__uint_as_float(v[cb][e + 1]), 0.f);
              const float x2 = fmaxf(__uint_as_float(v[cb][e + 2]), 0.f), x3 = fmaxf(__uint_as_float(v[cb][e + 3]), 0.f);
              r = fmaf(x3, wr.w, fmaf(x2, wr.z, fmaf(x1, wr.y, fmaf(x0, wr.x, r))));
              g = fmaf(x3, wg.w, fmaf(x2, wg.z, fmaf(x1, wg.y, fmaf(x0, wg.x, g))));
              bl = fmaf(x3, wb.w, fmaf(x2, wb.z, fmaf(x1, wb.y, fmaf(x0, wb.x, bl))));
            }
          }
          float sigma = 0.f;
          if (h == 0) {
            uint32_t v[16];
            tmem_ld16(acc + ((uint32_t)(q * 32) << 16) + 128, v);
            tmem_ld_wait();
            sigma = __uint_as_float(v[0]);
          } else {
            float* ps = part_sm + ((h - 1) * 128 + row) * 4;
            ps[0] = r; ps[1] = g; ps[2] = bl;
          }
          tc_fence_before();
          asm volatile("bar.sync 1, %0;" ::"n"(FT_EPI_WARPS * 32) : "memory");     // the epilogue warps only
          if (h == 0 && m < a.M) {
#pragma unroll
            for (int hh = 1; hh < EH; ++hh) {
              const float4 ps = *reinterpret_cast<const float4*>(part_sm + ((hh - 1) * 128 + row) * 4);
              r += ps.x; g += ps.y; bl += ps.z;
            }
            const float4 o = make_float4(r + __ldg(a.b_rgb), g + __ldg(a.b_rgb + 1), bl + __ldg(a.b_rgb + 2), sigma);
            *reinterpret_cast<float4*>(a.raw + m * 4) = o;     // raw = [rgb | sigma]  (models/model.py:73)
          }
          asm volatile("bar.sync 1, %0;" ::"n"(FT_EPI_WARPS * 32) : "memory");
          if (warp == 0) XTRACE(it, l, 1);
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == FT_EPI_WARPS + 1) tmem_dealloc(tmem, 512);
}

__device__ int g_flex_tc_error = 0;
}  // namespace

// 1 if the fused vanilla-field kernel hit a bounded barrier-wait timeout (folded into cnrf_debug_tc_error)
int cnrf_flex_fused_error() {
  int v = 0;
  cudaMemcpyFromSymbol(&v, g_flex_tc_error, sizeof(int));
  return v;
}

extern "C" int cnrf_debug_flex_trace(long long* host_out720) {
  if (!host_out720) return CNRF_E_NULL;
  const cudaError_t e = cudaMemcpyFromSymbol(host_out720, g_flex_trace, sizeof(long long) * 18 * 40);
  return e == cudaSuccess ? CNRF_OK : (int)e;
}

size_t cnrf_flex_fused_packed_bytes(int num_xyz_layers) {
  return (size_t)SLAB_W * (1 + 5 * (size_t)num_xyz_layers + 4) + (size_t)BIAS_IMG * ((size_t)num_xyz_layers + 1) +
         (size_t)SLAB_WD * 5 + 1024;
}

bool cnrf_flex_fused_eligible(const cnrf_flex_shape* s) {
  return s->precision == CNRF_PRECISION_BF16 && !s->keep && s->dim_dir > 0 && s->dim_dir <= 63 && s->num_rays < ((int64_t)1 << 31) && s->hidden == 256 && s->dim_xyz <= 63 && s->num_layers + 1 <= FT_MAX_LAYERS;
}

// x_dir: [N, dim_dir] embedded view directions; packed: cnrf_flex_fused_packed_bytes() of 1024-byte aligned workspace
int cnrf_flex_fused_fwd(const cnrf_flex_shape* s, const cnrf_flex_params* p, const float* x_xyz, int64_t ld_xyz,
                        const float* x_dir, int64_t ld_dir, float* raw, uint8_t* packed, cudaStream_t st, const FlexPts* gp) {
  if (gp && (gp->nfreq < 1 || gp->nfreq > FT_MAX_FREQS || 3 * (gp->inc + 2 * gp->nfreq) != s->dim_xyz)) return CNRF_E_SHAPE;
  const int H = 256, L = s->num_layers - 1, dx = s->dim_xyz, dd = s->dim_dir;
  FlexFwdArgs a = {};
  PackArgs pk = {};
  pk.out = packed;
  uint32_t ofs = 0;
  auto add_job = [&](const float* w, int ldw, int rows, int k0, int kcols, int nrows_slab, const float* bias = nullptr,
                     int bias_col = 63) {
    PackJob& j = pk.job[pk.count++];
    j.w = w; j.ldw = ldw; j.rows = (int16_t)rows; j.k0 = k0; j.kcols = (int16_t)kcols; j.ofs = ofs; j.nrows_slab = (int16_t)nrows_slab;
    j.bias = bias; j.bias_col = (int16_t)bias_col;    // a gamma(x) slab: its bias rides in column 63 (gamma(d): column dd)
    ofs += (uint32_t)nrows_slab * 128;
  };
  auto add_bias_image = [&](const float* bias) {
    PackJob& j = pk.job[pk.count++];
    j = {};
    j.kcols = -1; j.ofs = ofs; j.bias = bias;
    ofs += (uint32_t)BIAS_IMG;
  };
  int nl = 0;
  a.layer[nl] = {ofs, 1, 4, 0, 256, 0, 0}; ++nl;                                           // layer1: A = gamma(x)
  add_job(p->w_layer1, dx, H, 0, dx, 256, p->b_layer1);
  for (int i = 0; i < L; ++i) {
    const bool skip = (s->skip_mask >> i) & 1u;
    const int K = skip ? H + dx : H;
    a.layer[nl] = {ofs, (int16_t)(skip ? 5 : 4), 0, (int16_t)skip, 256, (int16_t)(skip ? 0 : 1), 4}; ++nl;
    for (int j = 0; j < 4; ++j) {
      add_job(p->w_xyz[i], K, H, 64 * j, 64, 256);
      if (j == 0 && !skip) add_bias_image(p->b_xyz[i]);     // behind the first slab: the bias MMA opens the layer
    }
    if (skip) add_job(p->w_xyz[i], K, H, H, dx, 256, p->b_xyz[i]);
  }
  a.layer[nl] = {ofs, 4, 0, 0, 256, 1, 0}; ++nl;
  for (int j = 0; j < 4; ++j) {
    add_job(p->w_feat, H, H, 64 * j, 64, 256);
    if (j == 0) add_bias_image(p->b_feat);
  }
  // layers_dir[0] (+ fc_alpha as output row 128): K slabs 0-3 = feat, K slab 4 = gamma(d) | 1 against [Wdir[:, H:] | b_dir]
  a.layer[nl] = {ofs, 5, 0, 1, 144, 0, 5}; ++nl;
  const uint32_t dir_ofs = ofs;
  for (int j = 0; j < 4; ++j) add_job(p->w_dir, H + dd, 128, 64 * j, 64, 144);
  add_job(p->w_dir, H + dd, 128, H, dd, 144, p->b_dir, dd);
  a.nlayers = nl;
  pack_flex_weights_kernel<<<pk.count, 256, 0, st>>>(pk);
  CNRF_LAUNCH_CHECK();
  {   // fc_alpha as output row 128 of every direction-layer slab
    PackArgs pa = {};
    pa.out = packed;
    for (int j = 0; j < 4; ++j) {
      PackJob& jb = pa.job[pa.count++];
      jb.w = p->w_alpha; jb.ldw = H; jb.rows = 1; jb.k0 = 64 * j; jb.kcols = 64; jb.ofs = dir_ofs + j * SLAB_WD + 128 * 128;
      jb.nrows_slab = 1;
    }
    {   // ... and its bias in column dd of the gamma(d) slab's row 128
      PackJob& jb = pa.job[pa.count++];
      jb.rows = 1; jb.kcols = 0; jb.ofs = dir_ofs + 4 * SLAB_WD + 128 * 128; jb.nrows_slab = 1;
      jb.bias = p->b_alpha; jb.bias_col = (int16_t)dd;
    }
    pack_flex_weights_kernel<<<pa.count, 32, 0, st>>>(pa);
    CNRF_LAUNCH_CHECK();
  }
  a.packed = packed;
  a.x_xyz = x_xyz; a.ld_xyz = ld_xyz; a.dx = dx;
  if (gp) { a.pts = gp->pts; a.ld_pts = gp->ld; a.freqs = gp->freqs; a.nfreq = gp->nfreq; a.inc = gp->inc; }
  a.x_dir = x_dir; a.ld_dir = ld_dir; a.dd = dd; a.w_rgb = p->w_rgb; a.b_rgb = p->b_rgb;
  a.raw = raw; a.M = s->num_rays * (int64_t)s->num_samples; a.N = s->num_rays; a.S = s->num_samples; a.pairing = s->pairing;
  cudaGetSymbolAddress(reinterpret_cast<void**>(&a.err), g_flex_tc_error);
  {
    static bool attr_set[CNRF_MAX_DEVICES] = {};
    const cudaError_t e = cnrf_once_per_device(attr_set, [] {
      return cudaFuncSetAttribute(flex_fwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, FT_SMEM); });
    if (e != cudaSuccess) return (int)e;
  }
  const int sms = cnrf_sm_count();
  const int64_t ntiles = (a.M + FT_ROWS - 1) / FT_ROWS;
  const int grid = (int)std::min<int64_t>(ntiles, sms);
  flex_fwd_tc_kernel<<<grid, FT_THREADS, FT_SMEM, st>>>(a);
  CNRF_LAUNCH_CHECK();
  return CNRF_OK;
}
