// Streaming tcgen05 GEMM for the big per-SAMPLE layers that run outside a fused chain: vanilla-NeRF training (flexnerf.cu,
// models/model.py:58-76 layer by layer) and the `split` pose chain (codenerf.cu, eval.py:129-168).  Same contract as
// cnrf_sgemm / cnrf_gemm_tc (internal.h) for ONE problem
//     C[m, n] = epi( sum_k A[m, k] B(k, n) ),   M = hundreds of thousands of sample rows,  64 <= N <= 256,  K <= 512,
// A row-major (a_cs == 1), B a weight matrix of either majorness, split-bf16 like gemm_tc.cu (hi*hi + hi*lo + lo*hi in fp32
// TMEM: ~16 mantissa bits).  gemm_tc_batch_kernel walks a CTA's K chunks serially -- load round trip, convert, MMA, wait --
// and is latency-bound on these shapes (tensor pipe 10-17 %, ncu).  Here the roles are split and decoupled by rings:
//   B     : packed ONCE per call by pack_b_kernel into bf16 hi/lo images ([256 x 64] per K chunk, 128-byte swizzle) in a
//           device-resident scratch; a TMA warp streams a chunk's 64 KB through a 2-slot ring per tile.
//   A     : eight converter warps load a [128 x 64] fp32 chunk (the NEXT chunk's loads are in flight while this one is
//           converted), write its hi/lo bf16 images into a 2-slot ring.
//   MMA   : one warp issues 12 MMAs (M128 N256 K16) per chunk into one of TWO TMEM accumulators (alternate tiles).
//   epi   : four warps drain an accumulator under the next tile's MMAs: a warp transposes its [32 x 32] block through shared
//           memory so that bias / mask / C are read and written as whole 128-byte lines -- + bias row, ReLU, ReLU mask, or C +=.
// Persistent: one CTA per SM walks row tiles.  The scratch makes the entry single-stream (as every libcnrf call is).
#include <cstdlib>
#include "common.cuh"
#include "internal.h"
#include "tc_common.cuh"

using namespace tc;

namespace {
constexpr int ST_ROWS = 128, ST_KC = 64, ST_MAXCH = 8;            // K <= 512 (the code-conditioned layers of the fp32-layout CodeNeRF chain)
constexpr int A_IMG = ST_ROWS * 128, B_IMG = 256 * 128;           // one bf16 image of an A / B chunk: 16 KB / 32 KB
constexpr int A_SLOT = 2 * A_IMG, B_SLOT = 2 * B_IMG;             // hi + lo
constexpr int NA = 2, NB = 2;
constexpr int ST_EPI = 8, ST_CONV = 8;
constexpr int ST_THREADS = 32 * (ST_EPI + ST_CONV + 2);           // 0-7 epilogue, 8-15 converters, 16 TMA, 17 MMA
constexpr int ST_STG_LD = 20;                                     // floats per staging row of 16 columns (80 B: 16-byte aligned)
constexpr int ST_SMEM = NA * A_SLOT + NB * B_SLOT + 256 + ST_EPI * 32 * ST_STG_LD * 4 + 1024;

__device__ uint8_t g_stream_b[ST_MAXCH * B_SLOT];                 // packed weights of the running call (512 KB)
__device__ int g_stream_error = 0;
// Timer builds only (-DCNRF_STREAM_TIMERS): clock64 stamps of block 0's 6th tile: [0..] converter per chunk {loads issued, slot
// free, arrived} at 0 + 3 ch; MMA per chunk {A seen, B seen, issued} at 16 + 3 ch, 31 accumulator free, 32 committed; TMA per
// chunk {slot free, issued} at 36 + 2 ch; epilogue warp 0 {48 accumulator seen, 49 done}, epilogue warp 4 {50, 51}.
__device__ long long g_stream_trace[64];
#ifdef CNRF_STREAM_TIMERS
#define STRACE(it, ev) do { if (blockIdx.x == 0 && (it) == 5 && (threadIdx.x & 31) == 0) g_stream_trace[ev] = clock64(); } while (0)
#else
#define STRACE(it, ev) do { } while (0)
#endif

struct StreamArgs {
  GemmArgs g;
  int nch;        // K chunks of 64
  int nmma;       // N rounded up to 16
  int a_vec;      // A rows are 16-byte aligned: float4 loads
  int* err;
};

// grid (chunks, 8): block = 32 rows (n) of one chunk; element (n, kk) <- B(k0 + kk, n), hi / lo bf16, 128-byte swizzled
__global__ void pack_b_kernel(const float* __restrict__ B, int64_t b_rs, int64_t b_cs, int N, int64_t K) {
  const int ch = blockIdx.x, n0 = blockIdx.y * 32;
  uint8_t* hi = g_stream_b + (size_t)ch * B_SLOT;
  uint8_t* lo = hi + B_IMG;
  for (int i = threadIdx.x; i < 32 * ST_KC; i += blockDim.x) {
    int n, kk;
    if (b_rs == 1) { n = n0 + (i >> 6); kk = i & 63; }            // k contiguous in memory
    else           { kk = i >> 5; n = n0 + (i & 31); }            // n contiguous in memory
    const int64_t k = (int64_t)ch * ST_KC + kk;
    const float v = (n < N && k < K) ? __ldg(B + k * b_rs + (int64_t)n * b_cs) : 0.f;
    const __nv_bfloat16 h = __float2bfloat16(v);
    const uint32_t off = slab_offset(n, kk);
    *reinterpret_cast<__nv_bfloat16*>(hi + off) = h;
    *reinterpret_cast<__nv_bfloat16*>(lo + off) = __float2bfloat16(v - __bfloat162float(h));
  }
}

// hi = bf16(v) (round to nearest), lo = bf16(v - hi): packed conversions (one F2FP per pair; the scalar F2F.BF16 is a
// quarter-rate instruction and was the converters' hottest line in ncu), hi read back as a float by shifting
__device__ __forceinline__ void store_hi_lo(uint8_t* hi, uint8_t* lo, uint32_t off, const float4 v) {
  const uint32_t h01 = pack_bf16(v.x, v.y), h23 = pack_bf16(v.z, v.w);
  const float hx = __uint_as_float(h01 << 16), hy = __uint_as_float(h01 & 0xFFFF0000u);
  const float hz = __uint_as_float(h23 << 16), hw = __uint_as_float(h23 & 0xFFFF0000u);
  *reinterpret_cast<uint2*>(hi + off) = make_uint2(h01, h23);
  *reinterpret_cast<uint2*>(lo + off) = make_uint2(pack_bf16(v.x - hx, v.y - hy), pack_bf16(v.z - hz, v.w - hw));
}

__global__ void __launch_bounds__(ST_THREADS, 1) gemm_stream_kernel(const __grid_constant__ StreamArgs a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* a_ring = smem;
  uint8_t* b_ring = smem + NA * A_SLOT;
  uint64_t* bars = reinterpret_cast<uint64_t*>(b_ring + NB * B_SLOT);
  uint64_t *a_full = bars, *a_empty = bars + NA, *b_full = bars + 2 * NA, *b_empty = bars + 2 * NA + NB;
  uint64_t *acc_full = bars + 2 * NA + 2 * NB, *acc_empty = acc_full + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const GemmArgs& g = a.g;
  if (tid == 0) {
    for (int i = 0; i < NA; ++i) { mbar_init(&a_full[i], ST_CONV); mbar_init(&a_empty[i], 1); }
    for (int i = 0; i < NB; ++i) { mbar_init(&b_full[i], 1); mbar_init(&b_empty[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&acc_full[i], 1); mbar_init(&acc_empty[i], ST_EPI); }
    mbar_fence_init();
  }
  if (warp == ST_EPI + ST_CONV + 1) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const int64_t ntiles = (g.M + ST_ROWS - 1) / ST_ROWS;
  bool ok = true;

  if (warp >= ST_EPI && warp < ST_EPI + ST_CONV) {
    // ---- converters: thread = (float4 column group cg of the 64-wide chunk, rows r1 + 16 u)
    const int ct = tid - ST_EPI * 32;
    const int cg = ct & 15, r1 = ct >> 4;
    const uint32_t cpart = (uint32_t)((((cg >> 1)) << 4) | ((cg & 1) << 3));     // chunk bits are XOR-ed with the row below
    auto load = [&](int64_t m0, int ch, float4 (&v)[8]) {
      const int64_t k = (int64_t)ch * ST_KC + cg * 4;
      if (a.a_vec && m0 + ST_ROWS <= g.M && (int64_t)(ch + 1) * ST_KC <= g.K) {      // interior chunk: pointer walk, no checks
        const float* p = g.A + (m0 + r1) * g.a_rs + k;
        const int64_t step = 16 * g.a_rs;
#pragma unroll
        for (int u = 0; u < 8; ++u, p += step) v[u] = __ldg(reinterpret_cast<const float4*>(p));
        return;
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int64_t m = m0 + r1 + 16 * u;
        v[u] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (m < g.M && k < g.K) {
          const float* p = g.A + m * g.a_rs + k;
          if (a.a_vec && k + 3 < g.K) {
            v[u] = __ldg(reinterpret_cast<const float4*>(p));
          } else {
            v[u].x = __ldg(p);
            if (k + 1 < g.K) v[u].y = __ldg(p + 1);
            if (k + 2 < g.K) v[u].z = __ldg(p + 2);
            if (k + 3 < g.K) v[u].w = __ldg(p + 3);
          }
        }
      }
    };
    // the next chunk's loads go out before this one is converted (two chunks ahead was measured: no gain -- the output
    // stores, not these loads, bound the kernel)
    uint32_t s = 0;
    float4 cur[8], nxt[8];
    int64_t t = blockIdx.x;
    int ch = 0;
    if (t < ntiles) load(t * ST_ROWS, 0, cur);
    while (t < ntiles && ok) {
      int64_t tn = t; int chn = ch + 1;
      if (chn == a.nch) { chn = 0; tn = t + gridDim.x; }
      if (tn < ntiles) load(tn * ST_ROWS, chn, nxt);
      const int itc = (int)((t - blockIdx.x) / gridDim.x);
      if (warp == ST_EPI) STRACE(itc, 3 * ch);
      const uint32_t slot = s % NA, par = ((s / NA) & 1) ^ 1;
      ok = mbar_wait(&a_empty[slot], par, a.err);
      if (warp == ST_EPI) STRACE(itc, 3 * ch + 1);
      uint8_t* hi = a_ring + slot * A_SLOT;
      uint8_t* lo = hi + A_IMG;
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int r = r1 + 16 * u;
        const uint32_t off = (uint32_t)(r * 128) + (cpart ^ (uint32_t)((r & 7) << 4));
        store_hi_lo(hi, lo, off, cur[u]);
      }
      fence_proxy_async();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&a_full[slot]);
      if (warp == ST_EPI) STRACE(itc, 3 * ch + 2);
#pragma unroll
      for (int u = 0; u < 8; ++u) cur[u] = nxt[u];
      ++s; t = tn; ch = chn;
    }
  } else if (warp == ST_EPI + ST_CONV) {
    // ---- TMA producer: the packed B chunks, once per tile
    uint32_t s = 0;
    int itt = 0;
    for (int64_t t = blockIdx.x; t < ntiles && ok; t += gridDim.x, ++itt) {
      for (int ch = 0; ch < a.nch && ok; ++ch, ++s) {
        const uint32_t slot = s % NB, par = ((s / NB) & 1) ^ 1;
        ok = mbar_wait(&b_empty[slot], par, a.err);
        STRACE(itt, 36 + 2 * ch);
        if (ok && elect_one()) {
          mbar_arrive_expect_tx(&b_full[slot], B_SLOT);
          bulk_g2s(b_ring + slot * B_SLOT, g_stream_b + (size_t)ch * B_SLOT, B_SLOT, &b_full[slot]);
        }
        __syncwarp();
        STRACE(itt, 37 + 2 * ch);
      }
    }
  } else if (warp == ST_EPI + ST_CONV + 1) {
    // ---- MMA issuer
    const uint32_t tmem_u = __shfl_sync(0xFFFFFFFFu, tmem, 0);
    const uint64_t a_desc0 = smem_desc(smem_u32(a_ring), 16, 1024, SWIZZLE_128B);
    const uint64_t b_desc0 = smem_desc(smem_u32(b_ring), 16, 1024, SWIZZLE_128B);
    const uint32_t idesc = idesc_bf16(ST_ROWS, a.nmma, 0, 0);
    uint32_t sa = 0, sb = 0, it = 0;
    for (int64_t t = blockIdx.x; t < ntiles && ok; t += gridDim.x, ++it) {
      const uint32_t buf = it & 1;
      ok = __all_sync(0xFFFFFFFFu, mbar_wait(&acc_empty[buf], ((it >> 1) & 1) ^ 1, a.err));    // the epilogue has drained it
      tc_fence_after();
      STRACE(it, 31);
      const uint32_t acc = tmem_u + buf * 256u;
#pragma unroll 1
      for (int ch = 0; ch < a.nch && ok; ++ch, ++sa, ++sb) {
        const uint32_t slot_a = sa % NA, slot_b = sb % NB;
        ok = __all_sync(0xFFFFFFFFu, mbar_wait(&a_full[slot_a], (sa / NA) & 1, a.err));
        STRACE(it, 16 + 3 * ch);
        ok = ok && __all_sync(0xFFFFFFFFu, mbar_wait(&b_full[slot_b], (sb / NB) & 1, a.err));
        tc_fence_after();
        STRACE(it, 17 + 3 * ch);
        const uint64_t ah = a_desc0 + (uint64_t)((slot_a * (uint32_t)A_SLOT) >> 4), al = ah + (uint64_t)(A_IMG >> 4);
        const uint64_t bh = b_desc0 + (uint64_t)((slot_b * (uint32_t)B_SLOT) >> 4), bl = bh + (uint64_t)(B_IMG >> 4);
        if (ok && elect_one()) {
#pragma unroll
          for (int k = 0; k < 4; ++k) umma_bf16(acc, ah + 2 * k, bh + 2 * k, idesc, (ch | k) != 0);     // hi * hi
#pragma unroll
          for (int k = 0; k < 4; ++k) umma_bf16(acc, ah + 2 * k, bl + 2 * k, idesc, 1u);                // hi * lo
#pragma unroll
          for (int k = 0; k < 4; ++k) umma_bf16(acc, al + 2 * k, bh + 2 * k, idesc, 1u);                // lo * hi
          umma_commit(&a_empty[slot_a]);
          umma_commit(&b_empty[slot_b]);
        }
        __syncwarp();
        STRACE(it, 18 + 3 * ch);
      }
      if (ok && elect_one()) umma_commit(&acc_full[buf]);
      __syncwarp();
      STRACE(it, 32);
    }
  } else {
    // ---- epilogue warps: TMEM lane quarter = warp.  tcgen05.ld hands a thread 32 columns of ONE row; stored like that, a
    // warp's 16-byte stores hit 32 different rows (32 half-used sectors per instruction: the kernel sat at 2 TB/s of output).
    // Eight warps (two per lane quarter, 128 columns each) turn [32 rows x 16 columns] blocks through private staging areas:
    // four lanes then cover 64 contiguous bytes of a row, and bias / mask / C are read and written as whole sectors.  (Measured
    // and not kept: four warps turning [32 x 32] blocks, i.e. whole 128-byte lines -- 445 against 415 us for a 786k x 256 x 256
    // layer; the bias row preloaded into registers with the block loop unrolled -- 443 us.)
    float* stg = reinterpret_cast<float*>(smem + NA * A_SLOT + NB * B_SLOT + 256) + warp * (32 * ST_STG_LD);
    const int q = warp & 3, hf = warp >> 2;                       // TMEM lane quarter, column half [128 hf, 128 hf + 128)
    const int sub_r = lane >> 2, c4 = (lane & 3) * 4;             // after the turn: 4 lanes cover 64 contiguous bytes of a row
    const bool one_bias_row = g.ld_bias == 0;
    // every row pointer of C / mask / bias is 16-byte aligned at every column this lane touches, and there is at most one bias row
    const bool fast_ptrs = ((reinterpret_cast<uintptr_t>(g.C) & 15) == 0) && ((g.ldc & 3) == 0) &&
                           (!g.mask || (((reinterpret_cast<uintptr_t>(g.mask) & 15) == 0) && ((g.ld_mask & 3) == 0))) &&
                           (!g.bias || (one_bias_row && (reinterpret_cast<uintptr_t>(g.bias) & 15) == 0));
    const bool relu = g.relu != 0;
    uint32_t it = 0;
    for (int64_t t = blockIdx.x; t < ntiles && ok; t += gridDim.x, ++it) {
      const uint32_t buf = it & 1;
      ok = mbar_wait(&acc_full[buf], (it >> 1) & 1, a.err);
      tc_fence_after();
      if (!ok) break;
      if (warp == 0) STRACE(it, 48);
      if (warp == 4) STRACE(it, 50);
      const int64_t mw = t * ST_ROWS + q * 32 + sub_r;            // this lane's first row after the turn (then + 8 j)
      const uint32_t t0 = tmem + buf * 256u + ((uint32_t)(q * 32) << 16) + hf * 128;
      const int nb = min(8, max(0, (a.nmma - hf * 128 + 15) / 16));
      uint32_t va[16], vb[16];
      if (nb > 0) tmem_ld16(t0, va);
      auto block = [&](int cb, const uint32_t (&w)[16], uint32_t (&next)[16]) {
        if (warp == 0 && cb == 2) STRACE(it, 52);
        tmem_ld_wait();
        if (warp == 0 && cb == 2) STRACE(it, 53);
#pragma unroll
        for (int j = 0; j < 4; ++j)
          *reinterpret_cast<uint4*>(stg + lane * ST_STG_LD + 4 * j) = make_uint4(w[4 * j], w[4 * j + 1], w[4 * j + 2], w[4 * j + 3]);
        if (cb + 1 < nb) tmem_ld16(t0 + (cb + 1) * 16, next);                  // next block's load under this one's stores
        __syncwarp();
        if (warp == 0 && cb == 2) STRACE(it, 54);
        const int n = hf * 128 + cb * 16 + c4;
        float4 bv = make_float4(0.f, 0.f, 0.f, 0.f);
        const bool n_ok = n + 3 < g.N;
        if (g.bias && one_bias_row && n_ok && (reinterpret_cast<uintptr_t>(g.bias) & 15) == 0)
          bv = __ldg(reinterpret_cast<const float4*>(g.bias + n));
        float* cp = g.C + mw * g.ldc + n;
        const float* mp = g.mask ? g.mask + mw * g.ld_mask + n : nullptr;
        // interior block of aligned operands with one bias row (every block of a 256-wide layer): no per-element checks
        if (fast_ptrs && n_ok && mw - sub_r + 32 <= g.M && !g.accumulate) {
          const int64_t cstep = 8 * g.ldc, mstep = 8 * g.ld_mask;
#pragma unroll
          for (int j = 0; j < 4; ++j, cp += cstep) {
            const float4 acc = *reinterpret_cast<const float4*>(stg + (sub_r + 8 * j) * ST_STG_LD + c4);
            float4 x = make_float4(acc.x + bv.x, acc.y + bv.y, acc.z + bv.z, acc.w + bv.w);
            if (relu) { x.x = fmaxf(x.x, 0.f); x.y = fmaxf(x.y, 0.f); x.z = fmaxf(x.z, 0.f); x.w = fmaxf(x.w, 0.f); }
            if (mp) {
              const float4 k = __ldg(reinterpret_cast<const float4*>(mp));
              x.x = k.x > 0.f ? x.x : 0.f; x.y = k.y > 0.f ? x.y : 0.f; x.z = k.z > 0.f ? x.z : 0.f; x.w = k.w > 0.f ? x.w : 0.f;
              mp += mstep;
            }
            *reinterpret_cast<float4*>(cp) = x;
          }
          if (warp == 0 && cb == 2) STRACE(it, 55);
          __syncwarp();
          if (warp == 0 && cb == 2) STRACE(it, 56);
          return;
        }
#pragma unroll
        for (int j = 0; j < 4; ++j, cp += 8 * g.ldc, mp = mp ? mp + 8 * g.ld_mask : nullptr) {
          const int64_t m = mw + 8 * j;
          const float4 acc = *reinterpret_cast<const float4*>(stg + (sub_r + 8 * j) * ST_STG_LD + c4);
          if (m >= g.M || n >= g.N) continue;
          float x[4] = {acc.x, acc.y, acc.z, acc.w};
          const float* brow = (g.bias && !one_bias_row) ? g.bias + ((m / g.bias_div) % g.bias_mod) * g.ld_bias + n : nullptr;
          const bool al16 = n_ok && ((reinterpret_cast<uintptr_t>(cp) & 15) == 0) && (!mp || (reinterpret_cast<uintptr_t>(mp) & 15) == 0) &&
                            (!brow || (reinterpret_cast<uintptr_t>(brow) & 15) == 0) &&
                            (!(g.bias && one_bias_row) || (reinterpret_cast<uintptr_t>(g.bias) & 15) == 0);
          if (al16) {
            if (g.accumulate) {
              const float4 c = *reinterpret_cast<const float4*>(cp);
              *reinterpret_cast<float4*>(cp) = make_float4(c.x + x[0], c.y + x[1], c.z + x[2], c.w + x[3]);
              continue;
            }
            if (brow) { const float4 b = __ldg(reinterpret_cast<const float4*>(brow)); x[0] += b.x; x[1] += b.y; x[2] += b.z; x[3] += b.w; }
            else { x[0] += bv.x; x[1] += bv.y; x[2] += bv.z; x[3] += bv.w; }
            if (g.relu) { x[0] = fmaxf(x[0], 0.f); x[1] = fmaxf(x[1], 0.f); x[2] = fmaxf(x[2], 0.f); x[3] = fmaxf(x[3], 0.f); }
            if (mp) {
              const float4 k = __ldg(reinterpret_cast<const float4*>(mp));
              x[0] = k.x > 0.f ? x[0] : 0.f; x[1] = k.y > 0.f ? x[1] : 0.f; x[2] = k.z > 0.f ? x[2] : 0.f; x[3] = k.w > 0.f ? x[3] : 0.f;
            }
            *reinterpret_cast<float4*>(cp) = make_float4(x[0], x[1], x[2], x[3]);
          } else {
            const float* b1 = brow ? brow : (g.bias ? g.bias + n : nullptr);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              if (n + i >= g.N) break;
              float y = x[i];
              if (g.accumulate) { cp[i] += y; continue; }
              if (b1) y += b1[i];
              if (g.relu) y = fmaxf(y, 0.f);
              if (mp) y = mp[i] > 0.f ? y : 0.f;
              cp[i] = y;
            }
          }
        }
        __syncwarp();
      };
#pragma unroll 1
      for (int cb = 0; cb < nb; cb += 2) {
        block(cb, va, vb);
        if (cb + 1 < nb) block(cb + 1, vb, va);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&acc_empty[buf]);
      if (warp == 0) STRACE(it, 49);
      if (warp == 4) STRACE(it, 51);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == ST_EPI + ST_CONV + 1) tmem_dealloc(tmem, 512);
}
}  // namespace

// ---- the same idea for weight gradients: C[m, n] += sum over ROWS r of A(m, r) B(r, n), A(m, r) = dY[r][m], B(r, n) = X[r][n] ------
// Both operands stream (hundreds of thousands of rows, <= 256 columns each); the accumulators never leave TMEM: two
// [128 x N] halves of dW, added to C once per CTA at the end (C is zero-filled by the caller, as for every split-K GEMM here).
// Chunks of 32 rows; an operand chunk [32 x 256] is stored as an MN-major image (rows = the contraction index, four slabs of 64
// columns, 128-byte swizzle -- the layout gemm_tc.cu uses for its MN-major operands), hi and lo.
namespace {
constexpr int WG_KC = 32;
constexpr int WG_SLAB = WG_KC * 128;             // [32 rows x 64 columns] bf16
constexpr int WG_IMG = 4 * WG_SLAB;              // 256 columns: 16 KB
constexpr int WG_SLOT = 4 * WG_IMG;              // A hi | A lo | B hi | B lo
constexpr int WG_NS = 3;
constexpr int WG_CONV = 8;
constexpr int WG_THREADS = 32 * (WG_CONV + 1);   // 0-7 converters (and the final reduction), 8 = MMA
constexpr int WG_SMEM = WG_NS * WG_SLOT + 256 + 1024;

struct WgradArgs {
  GemmArgs g;
  int nmma;              // N rounded up to 16
  int a_vec, b_vec;      // operand rows are 16-byte aligned
  int* err;
};

__global__ void __launch_bounds__(WG_THREADS, 1) gemm_stream_wgrad_kernel(const __grid_constant__ WgradArgs a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + WG_NS * WG_SLOT);
  uint64_t *full = bars, *empty = bars + WG_NS, *done = bars + 2 * WG_NS;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(done + 1);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const GemmArgs& g = a.g;
  if (tid == 0) {
    for (int i = 0; i < WG_NS; ++i) { mbar_init(&full[i], WG_CONV); mbar_init(&empty[i], 1); }
    mbar_init(done, 1);
    mbar_fence_init();
  }
  if (warp == WG_CONV) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const int64_t nchunks = (g.K + WG_KC - 1) / WG_KC;
  const int64_t mine = blockIdx.x < nchunks ? (nchunks - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;   // chunks of this CTA
  const int mhalves = g.M > 128 ? 2 : 1;
  bool ok = true;

  if (warp < WG_CONV) {
    // ---- converters: thread = (row rr + 4 u of the chunk, float4 column group c)
    const int rr = tid >> 6, c = (tid & 63) * 4;
    const uint32_t cofs = (uint32_t)(c >> 6) * WG_SLAB;
    auto load = [&](int64_t q, float4 (&va)[8], float4 (&vb)[8]) {
      const int64_t r0 = (blockIdx.x + q * gridDim.x) * WG_KC + rr;
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int64_t r = r0 + 4 * u;
        va[u] = vb[u] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (r < g.K) {
          if (c < g.M) {
            const float* p = g.A + r * g.a_cs + c;
            if (a.a_vec && c + 3 < g.M) va[u] = __ldg(reinterpret_cast<const float4*>(p));
            else { va[u].x = __ldg(p); if (c + 1 < g.M) va[u].y = __ldg(p + 1); if (c + 2 < g.M) va[u].z = __ldg(p + 2); if (c + 3 < g.M) va[u].w = __ldg(p + 3); }
          }
          if (c < g.N) {
            const float* p = g.B + r * g.b_rs + c;
            if (a.b_vec && c + 3 < g.N) vb[u] = __ldg(reinterpret_cast<const float4*>(p));
            else { vb[u].x = __ldg(p); if (c + 1 < g.N) vb[u].y = __ldg(p + 1); if (c + 2 < g.N) vb[u].z = __ldg(p + 2); if (c + 3 < g.N) vb[u].w = __ldg(p + 3); }
          }
        }
      }
    };
    float4 ca[8], cb[8], na[8], nb[8];
    if (mine > 0) load(0, ca, cb);
    for (int64_t q = 0; q < mine && ok; ++q) {
      if (q + 1 < mine) load(q + 1, na, nb);                     // next chunk's loads under this one's conversion
      const uint32_t slot = (uint32_t)(q % WG_NS), par = (uint32_t)((q / WG_NS) & 1) ^ 1;
      ok = mbar_wait(&empty[slot], par, a.err);
      uint8_t* base = smem + slot * WG_SLOT;
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const uint32_t off = cofs + slab_offset(rr + 4 * u, c & 63);
        store_hi_lo(base, base + WG_IMG, off, ca[u]);
        store_hi_lo(base + 2 * WG_IMG, base + 3 * WG_IMG, off, cb[u]);
      }
      fence_proxy_async();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&full[slot]);
#pragma unroll
      for (int u = 0; u < 8; ++u) { ca[u] = na[u]; cb[u] = nb[u]; }
    }
    // ---- final reduction: warp = (TMEM lane quarter, M half); a thread adds one row of its half into C
    if (mine > 0 && ok) {
      ok = mbar_wait(done, 0, a.err);
      tc_fence_after();
      const int q4 = warp & 3, hf = warp >> 2;
      const int64_t m = hf * 128 + q4 * 32 + lane;
      if (ok && hf < mhalves) {
        float* crow = g.C + m * g.ldc;
        const bool al16 = (reinterpret_cast<uintptr_t>(crow) & 15) == 0;
        const uint32_t t0 = tmem + hf * 256u + ((uint32_t)(q4 * 32) << 16);
#pragma unroll 1
        for (int cbk = 0; cbk < (a.nmma + 31) / 32; ++cbk) {
          uint32_t v[32];
          tmem_ld32(t0 + cbk * 32, v);
          tmem_ld_wait();
          if (m < g.M) {
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
              const int n = cbk * 32 + j;
              if (n >= g.N) break;
              if (al16 && n + 3 < g.N) {
                atomicAdd(reinterpret_cast<float4*>(crow + n), make_float4(__uint_as_float(v[j]), __uint_as_float(v[j + 1]),
                                                                         __uint_as_float(v[j + 2]), __uint_as_float(v[j + 3])));
              } else {
#pragma unroll
                for (int i = 0; i < 4; ++i) if (n + i < g.N) atomicAdd(crow + n + i, __uint_as_float(v[j + i]));
              }
            }
          }
        }
      }
    }
  } else {
    // ---- MMA issuer
    const uint32_t tmem_u = __shfl_sync(0xFFFFFFFFu, tmem, 0);
    const uint64_t desc0 = smem_desc(smem_u32(smem), WG_SLAB, 1024, SWIZZLE_128B);       // MN-major: LBO = slab stride
    const uint32_t idesc = idesc_bf16(128, a.nmma, 1, 1);
#pragma unroll 1
    for (int64_t q = 0; q < mine && ok; ++q) {
      const uint32_t slot = (uint32_t)(q % WG_NS);
      ok = __all_sync(0xFFFFFFFFu, mbar_wait(&full[slot], (uint32_t)((q / WG_NS) & 1), a.err));
      tc_fence_after();
      const uint64_t ah = desc0 + (uint64_t)((slot * (uint32_t)WG_SLOT) >> 4), al = ah + (uint64_t)(WG_IMG >> 4);
      const uint64_t bh = ah + (uint64_t)((2 * WG_IMG) >> 4), bl = ah + (uint64_t)((3 * WG_IMG) >> 4);
      if (ok && elect_one()) {
        for (int mh = 0; mh < mhalves; ++mh) {
          const uint32_t acc = tmem_u + mh * 256u;
          const uint64_t mo = (uint64_t)((mh * 2 * WG_SLAB) >> 4);                       // columns 128 mh ..: two slabs on
          const uint32_t first = q == 0 ? 0u : 1u;
#pragma unroll
          for (int j = 0; j < 2; ++j) umma_bf16(acc, ah + mo + 128 * j, bh + 128 * j, idesc, j == 0 ? first : 1u);   // hi * hi
#pragma unroll
          for (int j = 0; j < 2; ++j) umma_bf16(acc, ah + mo + 128 * j, bl + 128 * j, idesc, 1u);                   // hi * lo
#pragma unroll
          for (int j = 0; j < 2; ++j) umma_bf16(acc, al + mo + 128 * j, bh + 128 * j, idesc, 1u);                   // lo * hi
        }
        umma_commit(&empty[slot]);
      }
      __syncwarp();
    }
    if (ok && mine > 0 && elect_one()) umma_commit(done);
    __syncwarp();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == WG_CONV) tmem_dealloc(tmem, 512);
}
}  // namespace

// dW-shaped problems: A(m, k) = A[m + k * a_cs], B(k, n) = B[k * b_rs + n], partial sums added atomically into C
int cnrf_gemm_stream_wgrad(const GemmArgs& g, cudaStream_t st) {
  static const bool off = getenv("CNRF_GEMM_NO_STREAM") != nullptr;
  if (off || !g.A || !g.B || !g.C) return CNRF_E_UNSUPPORTED;
  if (g.rows_dev || g.k_slice > 0 || g.a_rs != 1 || g.b_cs != 1 || !(g.split_k > 1 || g.accumulate)) return CNRF_E_UNSUPPORTED;
  if (g.bias || g.relu || g.mask) return CNRF_E_UNSUPPORTED;
  if (g.M < 16 || g.M > 256 || g.N < 16 || g.N > 256 || g.K < (int64_t)296 * 128) return CNRF_E_UNSUPPORTED;
  WgradArgs a = {};
  a.g = g;
  a.nmma = (g.N + 15) & ~15;
  a.a_vec = ((reinterpret_cast<uintptr_t>(g.A) & 15) == 0) && ((g.a_cs & 3) == 0);
  a.b_vec = ((reinterpret_cast<uintptr_t>(g.B) & 15) == 0) && ((g.b_rs & 3) == 0);
  cudaGetSymbolAddress(reinterpret_cast<void**>(&a.err), g_stream_error);
  {
    static bool attr_set[CNRF_MAX_DEVICES] = {};
    const cudaError_t e = cnrf_once_per_device(attr_set, [] {
      return cudaFuncSetAttribute(gemm_stream_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, WG_SMEM); });
    if (e != cudaSuccess) return (int)e;
  }
  const int64_t nchunks = (g.K + WG_KC - 1) / WG_KC;
  const int grid = (int)std::min<int64_t>(nchunks, cnrf_sm_count());
  gemm_stream_wgrad_kernel<<<grid, WG_THREADS, WG_SMEM, st>>>(a);
  CNRF_LAUNCH_CHECK();
  return CNRF_OK;
}

extern "C" int cnrf_debug_stream_trace(long long* host_out64) {
  if (!host_out64) return CNRF_E_NULL;
  const cudaError_t e = cudaMemcpyFromSymbol(host_out64, g_stream_trace, sizeof(long long) * 64);
  return e == cudaSuccess ? CNRF_OK : (int)e;
}

int cnrf_gemm_stream_error() {
  int v = 0;
  cudaMemcpyFromSymbol(&v, g_stream_error, sizeof(int));
  return v;
}

// CNRF_E_UNSUPPORTED (nothing launched) unless the problem is one this kernel is for; the caller falls back to gemm_tc.
int cnrf_gemm_stream(const GemmArgs& g, cudaStream_t st) {
  static const bool off = getenv("CNRF_GEMM_NO_STREAM") != nullptr;         // A/B switch
  if (off || !g.A || !g.B || !g.C) return CNRF_E_UNSUPPORTED;
  if (g.a_rs == 1 && g.a_cs != 1) return cnrf_gemm_stream_wgrad(g, st);                 // contraction over rows: weight gradients
  if (g.rows_dev || g.split_k > 1 || g.k_slice > 0 || g.a_cs != 1) return CNRF_E_UNSUPPORTED;
  if (g.N < 64 || g.N > 256 || g.K < 16 || g.K > ST_MAXCH * ST_KC || g.M < 296 * 128) return CNRF_E_UNSUPPORTED;
  if (g.b_rs != 1 && g.b_cs != 1) return CNRF_E_UNSUPPORTED;
  if (g.accumulate && (g.bias || g.relu || g.mask)) return CNRF_E_UNSUPPORTED;
  StreamArgs a = {};
  a.g = g;
  a.nch = (int)((g.K + ST_KC - 1) / ST_KC);
  a.nmma = (g.N + 15) & ~15;
  a.a_vec = ((reinterpret_cast<uintptr_t>(g.A) & 15) == 0) && ((g.a_rs & 3) == 0);
  cudaGetSymbolAddress(reinterpret_cast<void**>(&a.err), g_stream_error);
  {
    static bool attr_set[CNRF_MAX_DEVICES] = {};
    const cudaError_t e = cnrf_once_per_device(attr_set, [] {
      return cudaFuncSetAttribute(gemm_stream_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, ST_SMEM); });
    if (e != cudaSuccess) return (int)e;
  }
  pack_b_kernel<<<dim3(a.nch, 8), 256, 0, st>>>(g.B, g.b_rs, g.b_cs, g.N, g.K);
  CNRF_LAUNCH_CHECK();
  const int64_t ntiles = (g.M + ST_ROWS - 1) / ST_ROWS;
  const int grid = (int)std::min<int64_t>(ntiles, cnrf_sm_count());
  gemm_stream_kernel<<<grid, ST_THREADS, ST_SMEM, st>>>(a);
  CNRF_LAUNCH_CHECK();
  return CNRF_OK;
}
