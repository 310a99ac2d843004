// Batched tcgen05 GEMM for the per-RAY layers of the bf16 mode (latent-code layers, folded-bias tables and their
// backward: models/model.py:174-177 and the code halves of :180-192).  Same contract as the SIMT cnrf_sgemm (internal.h):
//     C(m,n) = epi( sum_k A(m,k) B(k,n) ),   A(m,k) = A[m*a_rs + k*a_cs],   B(k,n) = B[k*b_rs + n*b_cs]
// but the contraction runs on the tensor cores as split-bf16: every fp32 operand value is carried as hi + lo (two bf16
// images) and hi*hi + hi*lo + lo*hi is accumulated in fp32 TMEM (~16 mantissa bits, 3 MMAs per K step -- these GEMMs
// are 5% of the step's FLOPs, and fp32-grade code layers keep their ReLU masks identical to the parity mode's).
//
// These problems are small (M = rays of a chunk, N,K <= 320, or K = rays for dW) and latency-bound, so up to 16 mutually
// independent problems share ONE launch (grid = sum of their tiles).  One CTA = one 128 x 64 output tile; K is walked in
// chunks of 64 (long K of the weight-gradient form: split over CTAs, partials added atomically -- its output is
// zero-filled by the caller).  All threads convert the fp32 operands to bf16 straight into the 128-byte-swizzled
// shared-memory image (rows = the strided index, columns = the contiguous index), ONE thread issues the MMAs, the
// operand majorness is chosen by the descriptor (K-major when k is the contiguous index, MN-major otherwise -- no
// transposed copies), 8 warps run the epilogue with 16-byte stores / vector reductions.
#include <cstdlib>
#include "common.cuh"
#include "internal.h"
#include "tc_common.cuh"

using namespace tc;

namespace {
constexpr int GBM = 128, GBN = 64, GKC = 64, GKSLICE = 256, GTHREADS = 256, GOCC = 4;
constexpr int A_BYTES = GBM * GKC * 2;      // one bf16 image of the A tile for a 64-wide K chunk (16 KB)
constexpr int B_BYTES = GBN * GKC * 2;      // ... of the B tile (8 KB)
constexpr int G_SMEM_WIDE = 2 * (A_BYTES + 256 * GKC * 2) + 1024;          // 128 x 256 tile: 97 KB -> two CTAs per SM
constexpr int G_SMEM = 2 * (A_BYTES + B_BYTES) + 1024 /*alignment slack*/;   // hi and lo images: 49 KB -> four CTAs per SM
                                                                          // hide each other's load round trips

struct GemmBatch {
  GemmArgs g[CNRF_GEMM_BATCH_MAX];
  int64_t kslice[CNRF_GEMM_BATCH_MAX];
  int a_mn[CNRF_GEMM_BATCH_MAX], b_mn[CNRF_GEMM_BATCH_MAX];
  int mt[CNRF_GEMM_BATCH_MAX], nt[CNRF_GEMM_BATCH_MAX];
  int first[CNRF_GEMM_BATCH_MAX + 1];   // first CTA of each problem
  int count;
};

// image(r, c) = src[(r0 + r) * row_stride + c0 + c] (0 outside rows_valid x cols_valid); columns are the contiguous index.
// Slab s holds columns [64 s, 64 s + 64) of every row, rows 128 bytes apart, 128-byte swizzle (tc_common.cuh).
// hi = bf16(x), lo = bf16(x - hi).
template <int U>                                       // independent global loads in flight per thread (see below)
__device__ __noinline__ void fill_image(uint8_t* img_hi, uint8_t* img_lo, int rows_tile, int cols_tile,
                                        uint32_t slab_stride, const float* __restrict__ src, int64_t row_stride,
                                        int64_t r0, int64_t c0, int64_t rows_valid, int64_t cols_valid, int tid) {
  const int groups = cols_tile >> 2;
  const int gshift = 31 - __clz(groups);               // cols_tile is 64, 128, 192 or 256: not always a power of two
  const bool pow2 = (groups & (groups - 1)) == 0;
  const int total = rows_tile * groups;
  const bool vec = ((reinterpret_cast<uintptr_t>(src) & 15) == 0) && ((row_stride & 3) == 0) && ((c0 & 3) == 0);
  // Interior tiles of aligned operands (every chunk of a K = 256 layer): a thread keeps ONE column group and walks rows, so
  // a float4 costs a pointer add, the load, the conversion and a store -- the general loop below spends ~50 integer
  // instructions per float4 on index arithmetic and bounds, which is what bound the big per-sample GEMMs.
  if (vec && pow2 && groups <= GTHREADS && r0 + rows_tile <= rows_valid && c0 + cols_tile <= cols_valid) {
    const int rstep = GTHREADS >> gshift;                // rows covered by one pass of the 256 threads
    const int c = (tid & (groups - 1)) << 2;
    int r = tid >> gshift;
    const float* p = src + (r0 + r) * row_stride + c0 + c;
    const int64_t pstep = (int64_t)rstep * row_stride;
    const uint32_t cpart = (uint32_t)(c >> 6) * slab_stride + (uint32_t)((c & 7) << 1);
    const int chunk = (c & 63) >> 3;
#pragma unroll 1
    for (; r < rows_tile; r += rstep * U, p += (int64_t)U * pstep) {
      float4 v[U];
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (r + u * rstep < rows_tile) v[u] = __ldg(reinterpret_cast<const float4*>(p + (int64_t)u * pstep));
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int rr = r + u * rstep;
        if (rr < rows_tile) {
          const float hx = __bfloat162float(__float2bfloat16(v[u].x)), hy = __bfloat162float(__float2bfloat16(v[u].y));
          const float hz = __bfloat162float(__float2bfloat16(v[u].z)), hw = __bfloat162float(__float2bfloat16(v[u].w));
          const uint32_t off = cpart + (uint32_t)(rr * 128 + ((chunk ^ (rr & 7)) << 4));
          *reinterpret_cast<uint2*>(img_hi + off) = make_uint2(pack_bf16(hx, hy), pack_bf16(hz, hw));
          *reinterpret_cast<uint2*>(img_lo + off) =
              make_uint2(pack_bf16(v[u].x - hx, v[u].y - hy), pack_bf16(v[u].z - hz, v[u].w - hw));
        }
      }
    }
    return;
  }
  // U = 8: a whole 128 x 64 chunk is ONE round trip (cold misses cost ~2.5 us each); U = 16: the 256 x 64 chunk of a wide tile
#pragma unroll 1
  for (int base = tid; base < total; base += GTHREADS * U) {
    float4 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int i = base + u * GTHREADS;
      v[u] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (i < total) {
        const int r = pow2 ? (i >> gshift) : (i / groups), c = (i - r * groups) << 2;
        const int64_t gr = r0 + r, gc = c0 + c;
        if (gr < rows_valid && gc < cols_valid) {
          const float* p = src + gr * row_stride + gc;
          if (vec && gc + 3 < cols_valid) {
            v[u] = __ldg(reinterpret_cast<const float4*>(p));
          } else {
            v[u].x = __ldg(p);
            if (gc + 1 < cols_valid) v[u].y = __ldg(p + 1);
            if (gc + 2 < cols_valid) v[u].z = __ldg(p + 2);
            if (gc + 3 < cols_valid) v[u].w = __ldg(p + 3);
          }
        }
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int i = base + u * GTHREADS;
      if (i < total) {
        const int r = pow2 ? (i >> gshift) : (i / groups), c = (i - r * groups) << 2;
        const float hx = __bfloat162float(__float2bfloat16(v[u].x)), hy = __bfloat162float(__float2bfloat16(v[u].y));
        const float hz = __bfloat162float(__float2bfloat16(v[u].z)), hw = __bfloat162float(__float2bfloat16(v[u].w));
        const uint32_t off = (c >> 6) * slab_stride + slab_offset(r, c & 63);
        *reinterpret_cast<uint2*>(img_hi + off) = make_uint2(pack_bf16(hx, hy), pack_bf16(hz, hw));
        *reinterpret_cast<uint2*>(img_lo + off) =
            make_uint2(pack_bf16(v[u].x - hx, v[u].y - hy), pack_bf16(v[u].z - hz, v[u].w - hw));
      }
    }
  }
}


// ---- few-row problems (shared-code mode: one code row per run of rays, typically ONE per chunk) -------------------------
// With a device-side row count of at most SMALL_ROWS the tile is a handful of dot products: plain fp32 on the CUDA cores,
// no operand images, no TMEM -- a launch then costs a few microseconds instead of four MMA round trips.
constexpr int SMALL_ROWS = 8;

__device__ __forceinline__ void small_epilogue(const GemmArgs& g, bool add, int64_t m, int64_t n, float y) {
  float* c = g.C + m * g.ldc + n;
  if (add) { atomicAdd(c, y); return; }
  if (g.bias) y += g.bias[((m / g.bias_div) % g.bias_mod) * g.ld_bias + n];
  if (g.relu) y = fmaxf(y, 0.f);
  if (g.mask) y = g.mask[m * g.ld_mask + n] > 0.f ? y : 0.f;
  *c = y;
}

// rows_dim == 1: M <= SMALL_ROWS output rows.  The launch was sized for the worst case (one row tile per 128 rays), so
// the row tiles mx = 1..3 of a column tile are idle CTAs: they take a quarter of its 64 columns each (nsub = 4), and a
// CTA's 256 threads split K sixteen ways -- every thread issues its <= 16 weight loads in one go (the whole tile is a
// single global-memory round trip) and the partial sums meet in shared memory.
__device__ void small_m_tile(const GemmArgs& g, bool add, int64_t n0, int ncols, int64_t kbeg, int64_t kend, float* sm, int tid) {
  const int M = (int)g.M, KK = (int)(kend - kbeg);
  float* a_s = sm;                                   // [SMALL_ROWS][KK]
  float* red = sm + SMALL_ROWS * KK;                 // [16][SMALL_ROWS][16]
  for (int i = tid; i < SMALL_ROWS * KK; i += GTHREADS) {
    const int m = i / KK, k = i - m * KK;
    a_s[i] = m < M ? __ldg(g.A + (int64_t)m * g.a_rs + (kbeg + k) * g.a_cs) : 0.f;
  }
  __syncthreads();
  constexpr int U = 16;
  for (int c0 = 0; c0 < ncols; c0 += 16) {            // one pass when the column tile is shared by four CTAs
    const int nl = tid & 15, ks = tid >> 4;
    const int64_t n = n0 + c0 + nl;
    float acc[SMALL_ROWS];
#pragma unroll
    for (int m = 0; m < SMALL_ROWS; ++m) acc[m] = 0.f;
    if (nl + c0 < ncols && n < g.N) {
      const float* bp = g.B + n * g.b_cs + kbeg * g.b_rs;
      for (int kb = ks * U; kb < KK; kb += 16 * U) {
        float bv[U];
#pragma unroll
        for (int u = 0; u < U; ++u) bv[u] = kb + u < KK ? __ldg(bp + (int64_t)(kb + u) * g.b_rs) : 0.f;
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int k = kb + u < KK ? kb + u : 0;      // bv is 0 past the end
#pragma unroll
          for (int m = 0; m < SMALL_ROWS; ++m) acc[m] = fmaf(a_s[m * KK + k], bv[u], acc[m]);
        }
      }
    }
#pragma unroll
    for (int m = 0; m < SMALL_ROWS; ++m) red[(ks * SMALL_ROWS + m) * 16 + nl] = acc[m];
    __syncthreads();
    if (tid < SMALL_ROWS * 16) {
      const int m = tid >> 4, c = tid & 15;
      if (m < M && c0 + c < ncols && n0 + c0 + c < g.N) {
        float y = 0.f;
#pragma unroll
        for (int q = 0; q < 16; ++q) y += red[(q * SMALL_ROWS + m) * 16 + c];
        small_epilogue(g, add, m, n0 + c0 + c, y);
      }
    }
    __syncthreads();
  }
}

// rows_dim == 2: K <= SMALL_ROWS contraction rows, this CTA owns the 128 x 64 output tile at (m0, n0)
__device__ void small_k_tile(const GemmArgs& g, bool add, int64_t m0, int64_t n0, int tid) {
  const int K = (int)g.K;
  const int nl = tid & 63, mq = tid >> 6;
  const int64_t n = n0 + nl;
  if (n >= g.N) return;
  // all loads of a k step are issued before they are used: 32 independent rows per thread, no atomics in between
  float y[GBM / 4];
#pragma unroll
  for (int i = 0; i < GBM / 4; ++i) y[i] = 0.f;
  for (int k = 0; k < K; ++k) {
    float av[GBM / 4];
#pragma unroll
    for (int i = 0; i < GBM / 4; ++i) {
      const int64_t m = m0 + mq + 4 * i;
      av[i] = m < g.M ? __ldg(g.A + m * g.a_rs + (int64_t)k * g.a_cs) : 0.f;
    }
    const float b = __ldg(g.B + (int64_t)k * g.b_rs + n * g.b_cs);
#pragma unroll
    for (int i = 0; i < GBM / 4; ++i) y[i] = fmaf(av[i], b, y[i]);
  }
#pragma unroll
  for (int i = 0; i < GBM / 4; ++i) {
    const int64_t m = m0 + mq + 4 * i;
    if (m < g.M) small_epilogue(g, add, m, n, y[i]);
  }
}

// BN = 64: the batched per-ray kernel described above.  BN = 256 (one problem per launch, no device-side row counts): the
// per-SAMPLE layers of the vanilla field in training and of the `split` pose chain -- hundreds of thousands of rows against
// a 256-wide weight matrix.  There a 128 x 64 tile converts its 128-row operand tile to bf16 once per 64 output columns and
// issues N = 64 MMAs; the wide tile converts it once and issues N = 256 MMAs (2 CTAs per SM instead of 4: TMEM).  Measured:
// lego.yml's training step 298 -> 234 ms, the C3 pose-optimisation iteration 11.9 -> 10.5 ms.  Both kernels walk a CTA's K
// chunks serially (load round trip -> convert -> MMA -> wait); the wide one is bound by its three load round trips per
// chunk (an L2 prefetch of the whole tile changed nothing: the latency is the dependent chain, not HBM).
template <int BN>
__global__ void __launch_bounds__(GTHREADS, BN == 64 ? GOCC : 2) gemm_tc_batch_kernel(const __grid_constant__ GemmBatch b, int* err) {
  constexpr int GBN = BN;
  constexpr int B_BYTES = BN * GKC * 2;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_slot;
  uint8_t* a_hi = smem;
  uint8_t* a_lo = smem + A_BYTES;
  uint8_t* b_hi = smem + 2 * A_BYTES;
  uint8_t* b_lo = smem + 2 * A_BYTES + B_BYTES;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  int p = 0;
  while (p + 1 < b.count && (int)blockIdx.x >= b.first[p + 1]) ++p;
  GemmArgs g = b.g[p];
  const bool multi_k = (g.K + b.kslice[p] - 1) / b.kslice[p] > 1;      // decided by the worst-case K the host saw
  if (g.rows_dev) {                                    // run-time row count (code_runs.cu): idle CTAs leave at once
    const int64_t rows = *g.rows_dev;
    if (g.rows_dim == 1) g.M = min(g.M, rows); else g.K = min(g.K, rows);
  }
  const int a_mn = b.a_mn[p], b_mn = b.b_mn[p];
  const int local = blockIdx.x - b.first[p];
  // column tiles fastest: the (up to 5) CTAs that share a 128-row A tile are neighbours in the grid, so for the per-SAMPLE
  // layers of the vanilla field (M = millions of rows, flexnerf.cu) A comes from HBM once and from L2 afterwards
  const int ny = local % b.nt[p], rest = local / b.nt[p];
  const int mx = rest % b.mt[p], kz = rest / b.mt[p];
  const int64_t m0 = (int64_t)mx * GBM;
  const int64_t n0 = (int64_t)ny * GBN;
  const int64_t kbeg = (int64_t)kz * b.kslice[p];
  const int64_t kend = min(g.K, kbeg + b.kslice[p]);
  if (BN == 64 && g.rows_dev) {
    const bool add = g.split_k > 1 || g.accumulate || multi_k;
    // the staging buffers of the few-row path must fit the operand-image area: 8 x K + 16 x 8 x 16 floats
    if (g.rows_dim == 1 && g.M <= SMALL_ROWS && kbeg < g.K &&
        (kend - kbeg) * SMALL_ROWS + 16 * SMALL_ROWS * 16 <= (2 * (A_BYTES + B_BYTES)) / 4) {
      const int nsub = b.mt[p] >= 4 ? 4 : 1;           // idle row tiles share the column tile
      if (mx >= nsub) return;
      small_m_tile(g, add, n0 + (nsub == 4 ? 16 * mx : 0), nsub == 4 ? 16 : GBN, kbeg, kend, reinterpret_cast<float*>(smem), tid);
      return;
    }
    if (m0 >= g.M || kbeg >= g.K) return;
    if (g.rows_dim == 2 && g.K <= SMALL_ROWS) { small_k_tile(g, add, m0, n0, tid); return; }
  }
  if (m0 >= g.M || kbeg >= g.K) return;

  if (tid == 0) { mbar_init(&bar, 1); mbar_fence_init(); }
  if (warp == 1) tmem_alloc(&tmem_slot, BN);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  const uint32_t idesc = idesc_bf16(GBM, GBN, a_mn, b_mn);
  bool ok = true;
  uint32_t phase = 0;
#pragma unroll 1
  for (int64_t k0 = kbeg; k0 < kend; k0 += GKC) {
    const int kc = (int)min((int64_t)GKC, kend - k0);
    const int kpad = (kc + 15) & ~15;
    if (k0 > kbeg) {                                   // the previous chunk's MMAs still read the images
      ok = ok && mbar_wait(&bar, phase, err);
      phase ^= 1;
      tc_fence_after();
    }
    // K-major image: [tile rows x kpad cols(k)];  MN-major image: [kpad rows(k) x tile cols]
    const uint32_t a_slab = a_mn ? (uint32_t)kpad * 128u : (uint32_t)GBM * 128u;
    const uint32_t b_slab = b_mn ? (uint32_t)kpad * 128u : (uint32_t)GBN * 128u;
    constexpr int UB = 8;      // 16 loads in flight for the wide B tile were measured: no gain (242 vs 234 ms per lego step)
    if (!a_mn) fill_image<8>(a_hi, a_lo, GBM, (kpad + 63) & ~63, a_slab, g.A, g.a_rs, m0, k0, g.M, k0 + kc, tid);
    else       fill_image<8>(a_hi, a_lo, kpad, GBM, a_slab, g.A, g.a_cs, k0, m0, k0 + kc, g.M, tid);
    if (!b_mn) fill_image<UB>(b_hi, b_lo, GBN, (kpad + 63) & ~63, b_slab, g.B, g.b_cs, n0, k0, g.N, k0 + kc, tid);
    else       fill_image<UB>(b_hi, b_lo, kpad, GBN, b_slab, g.B, g.b_rs, k0, n0, k0 + kc, g.N, tid);
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    if (tid == 0 && ok) {
      tc_fence_after();
      const uint32_t sa[2] = {smem_u32(a_hi), smem_u32(a_lo)}, sb[2] = {smem_u32(b_hi), smem_u32(b_lo)};
#pragma unroll 1
      for (int term = 0; term < 3; ++term) {           // hi*hi, hi*lo, lo*hi
        const uint32_t pa = sa[term == 2], pb = sb[term == 1];
#pragma unroll 1
        for (int j = 0; j < kpad / 16; ++j) {
          const uint64_t ad = a_mn ? smem_desc(pa + j * 2048, a_slab, 1024, SWIZZLE_128B)
                                   : smem_desc(pa + (j >> 2) * a_slab + (j & 3) * 32, 16, 1024, SWIZZLE_128B);
          const uint64_t bd = b_mn ? smem_desc(pb + j * 2048, b_slab, 1024, SWIZZLE_128B)
                                   : smem_desc(pb + (j >> 2) * b_slab + (j & 3) * 32, 16, 1024, SWIZZLE_128B);
          umma_bf16(tmem, ad, bd, idesc, (k0 > kbeg) || term > 0 || j > 0);
        }
      }
      umma_commit(&bar);
    }
  }
  ok = ok && mbar_wait(&bar, phase, err);
  tc_fence_after();
  if (ok) {
    const int q = warp & 3, half = warp >> 2;          // TMEM lane quarter, column half of every 64-wide column group
    const int64_t m = m0 + q * 32 + lane;
    const bool add = g.split_k > 1 || g.accumulate || multi_k;
    const float* brow = (g.bias && m < g.M) ? g.bias + ((m / g.bias_div) % g.bias_mod) * g.ld_bias : nullptr;
    float* crow = g.C + m * g.ldc;
    const float* mrow = g.mask ? g.mask + m * g.ld_mask : nullptr;
    const bool al16 = ((reinterpret_cast<uintptr_t>(crow) & 15) == 0) &&
                      (!brow || (reinterpret_cast<uintptr_t>(brow) & 15) == 0) &&
                      (!mrow || (reinterpret_cast<uintptr_t>(mrow) & 15) == 0);
#pragma unroll 1
    for (int cg = 0; cg < BN / 64; ++cg) {
      if (n0 + cg * 64 >= g.N) break;                  // warp-uniform
      uint32_t v[32];
      tmem_ld32(tmem + ((uint32_t)(q * 32) << 16) + cg * 64 + half * 32, v);
      tmem_ld_wait();
      if (m < g.M) {
#pragma unroll
        for (int j = 0; j < 32; j += 4) {
          const int64_t n = n0 + cg * 64 + half * 32 + j;
          if (n >= g.N) break;
          float x[4] = {__uint_as_float(v[j]), __uint_as_float(v[j + 1]), __uint_as_float(v[j + 2]), __uint_as_float(v[j + 3])};
          if (al16 && n + 3 < g.N) {
            if (add) { atomicAdd(reinterpret_cast<float4*>(crow + n), make_float4(x[0], x[1], x[2], x[3])); continue; }
            if (brow) { const float4 t = __ldg(reinterpret_cast<const float4*>(brow + n)); x[0] += t.x; x[1] += t.y; x[2] += t.z; x[3] += t.w; }
            if (g.relu) { x[0] = fmaxf(x[0], 0.f); x[1] = fmaxf(x[1], 0.f); x[2] = fmaxf(x[2], 0.f); x[3] = fmaxf(x[3], 0.f); }
            if (mrow) {
              const float4 t = __ldg(reinterpret_cast<const float4*>(mrow + n));
              x[0] = t.x > 0.f ? x[0] : 0.f; x[1] = t.y > 0.f ? x[1] : 0.f; x[2] = t.z > 0.f ? x[2] : 0.f; x[3] = t.w > 0.f ? x[3] : 0.f;
            }
            *reinterpret_cast<float4*>(crow + n) = make_float4(x[0], x[1], x[2], x[3]);
          } else {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              if (n + i >= g.N) break;
              float y = x[i];
              if (add) { atomicAdd(crow + n + i, y); continue; }
              if (brow) y += brow[n + i];
              if (g.relu) y = fmaxf(y, 0.f);
              if (mrow) y = mrow[n + i] > 0.f ? y : 0.f;
              crow[n + i] = y;
            }
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, BN);
}
}  // namespace

__device__ int g_gemm_tc_error = 0;
static int* gemm_err_ptr() {
  static int* p[CNRF_MAX_DEVICES] = {};
  return cnrf_symbol_ptr(p, g_gemm_tc_error);
}

// Returns CNRF_E_UNSUPPORTED (nothing launched) when some operand has neither stride equal to 1: the caller then runs
// the problems through cnrf_sgemm.  The problems of one call must not depend on each other.
int cnrf_gemm_tc_batch(const GemmArgs* list, int count, cudaStream_t st) {
  if (count <= 0) return CNRF_OK;
  if (!list || count > CNRF_GEMM_BATCH_MAX) return CNRF_E_SHAPE;
  if (count == 1) {                                    // one big per-sample problem: the streaming kernel (gemm_stream.cu)
    const int rc = cnrf_gemm_stream(list[0], st);
    if (rc != CNRF_E_UNSUPPORTED) return rc;
  }
  GemmBatch b;
  b.count = count;
  int total = 0;
  // one big problem with a wide output (>= 296 row tiles or K slices: two CTAs on every SM): the 128 x 256 tile
  int bn = GBN;
  if (count == 1 && !list[0].rows_dev && list[0].N >= 192 && list[0].K > 0 && list[0].split_k >= 1) {
    static const bool no_wide = getenv("CNRF_GEMM_NO_WIDE") != nullptr;      // A/B switch
    const GemmArgs& a = list[0];
    const bool may_split = a.split_k > 1 || a.accumulate;
    const int64_t ks = a.k_slice > 0 ? a.k_slice : (may_split && a.K > 2 * GKSLICE) ? GKSLICE : a.K;
    if (!no_wide && ((a.M + GBM - 1) / GBM) * ((a.K + ks - 1) / ks) >= 296) bn = 256;
  }
  for (int i = 0; i < count; ++i) {
    const GemmArgs& a = list[i];
    if (!a.A || !a.B || !a.C) return CNRF_E_NULL;
    if (a.M <= 0 || a.N <= 0 || a.K <= 0 || a.split_k < 1) return CNRF_E_SHAPE;
    if (a.a_cs == 1) b.a_mn[i] = 0; else if (a.a_rs == 1) b.a_mn[i] = 1; else return CNRF_E_UNSUPPORTED;
    if (a.b_rs == 1) b.b_mn[i] = 0; else if (a.b_cs == 1) b.b_mn[i] = 1; else return CNRF_E_UNSUPPORTED;
    // split K over CTAs only where the caller already adds into C (weight gradients: K = rays of the chunk)
    const bool may_split = a.split_k > 1 || a.accumulate;
    if (a.k_slice > 0 && (!a.accumulate || a.k_slice % GKC)) return CNRF_E_SHAPE;
    b.kslice[i] = a.k_slice > 0 ? a.k_slice : (may_split && a.K > 2 * GKSLICE) ? GKSLICE : a.K;
    b.g[i] = a;
    b.mt[i] = (int)((a.M + GBM - 1) / GBM);
    b.nt[i] = (a.N + bn - 1) / bn;
    b.first[i] = total;
    total += b.mt[i] * b.nt[i] * (int)((a.K + b.kslice[i] - 1) / b.kslice[i]);
  }
  b.first[count] = total;
  {
    static bool attr_set[CNRF_MAX_DEVICES] = {};
    const cudaError_t e = cnrf_once_per_device(attr_set, [] {
      cudaError_t e1 = cudaFuncSetAttribute(gemm_tc_batch_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, G_SMEM);
      if (e1 == cudaSuccess) e1 = cudaFuncSetAttribute(gemm_tc_batch_kernel<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, G_SMEM_WIDE);
      return e1; });
    if (e != cudaSuccess) return (int)e;
  }
  if (bn == 256) gemm_tc_batch_kernel<256><<<total, GTHREADS, G_SMEM_WIDE, st>>>(b, gemm_err_ptr());
  else gemm_tc_batch_kernel<64><<<total, GTHREADS, G_SMEM, st>>>(b, gemm_err_ptr());
  CNRF_LAUNCH_CHECK();
  return CNRF_OK;
}

int cnrf_gemm_tc(const GemmArgs& a, cudaStream_t st) { return cnrf_gemm_tc_batch(&a, 1, st); }

extern "C" int cnrf_debug_gemm(int32_t tc, const float* A, int64_t a_rs, int64_t a_cs, const float* B, int64_t b_rs,
                               int64_t b_cs, float* C, int64_t ldc, int64_t M, int32_t N, int64_t K, int32_t accumulate,
                               cnrf_stream_t stream) {
  GemmArgs g;
  g.A = A; g.a_rs = a_rs; g.a_cs = a_cs; g.B = B; g.b_rs = b_rs; g.b_cs = b_cs; g.C = C; g.ldc = ldc;
  g.M = M; g.N = N; g.K = K; g.accumulate = accumulate;
  return tc ? cnrf_gemm_tc(g, as_stream(stream)) : cnrf_sgemm(g, as_stream(stream));
}
