// Diagnostic: one-CTA tcgen05 GEMM over the shared-memory image used by the fused MLP kernels, in both operand
// majornesses.  tests/test_gpu_tc.py pins the UMMA descriptor encodings with it before the big kernels are trusted.
#include "common.cuh"
#include "tc_common.cuh"

using namespace tc;

// mode 0 (K-major):  A [128,K], B [N,K] fp32 row-major      D = A B^T
// mode 1 (MN-major): A [K,128], B [K,N] fp32 row-major      D = A^T B      (the dW = dY^T X form)
__global__ void __launch_bounds__(128) umma_probe_kernel(int mode, const float* __restrict__ A,
                                                        const float* __restrict__ B, float* __restrict__ D, int N,
                                                        int K, int* __restrict__ err) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_slot;
  const int tid = threadIdx.x, warp = tid >> 5;
  uint8_t* a_img = smem;
  uint8_t* b_img;
  uint32_t a_slab, b_slab;   // bytes between consecutive 64-column slabs
  if (mode == 0) {
    a_slab = 128 * 128; b_slab = N * 128;
    b_img = a_img + (K / 64) * a_slab;
    for (int i = tid; i < 128 * K; i += 128) {
      int r = i / K, c = i % K;
      *reinterpret_cast<__nv_bfloat16*>(a_img + (c / 64) * a_slab + slab_offset(r, c % 64)) = __float2bfloat16(A[i]);
    }
    for (int i = tid; i < N * K; i += 128) {
      int r = i / K, c = i % K;
      *reinterpret_cast<__nv_bfloat16*>(b_img + (c / 64) * b_slab + slab_offset(r, c % 64)) = __float2bfloat16(B[i]);
    }
  } else {
    a_slab = K * 128; b_slab = K * 128;
    b_img = a_img + 2 * a_slab;
    for (int i = tid; i < K * 128; i += 128) {
      int r = i / 128, c = i % 128;   // r = k index, c = m index
      *reinterpret_cast<__nv_bfloat16*>(a_img + (c / 64) * a_slab + slab_offset(r, c % 64)) = __float2bfloat16(A[i]);
    }
    for (int i = tid; i < K * N; i += 128) {
      int r = i / N, c = i % N;
      *reinterpret_cast<__nv_bfloat16*>(b_img + (c / 64) * b_slab + slab_offset(r, c % 64)) = __float2bfloat16(B[i]);
    }
  }
  if (tid == 0) { mbar_init(&bar, 1); mbar_fence_init(); }
  uint32_t ncols = 32;
  while ((int)ncols < N) ncols <<= 1;
  if (warp == 0) tmem_alloc(&tmem_slot, ncols);
  fence_proxy_async();            // generic-proxy smem writes -> visible to the tensor-core (async) proxy
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  if (tid == 0) {
    const uint32_t idesc = idesc_bf16(128, N, mode, mode);
    for (int j = 0; j < K / 16; ++j) {
      uint64_t ad, bd;
      if (mode == 0) {
        ad = smem_desc(smem_u32(a_img) + (j / 4) * a_slab + (j % 4) * 32, 16, 1024, SWIZZLE_128B);
        bd = smem_desc(smem_u32(b_img) + (j / 4) * b_slab + (j % 4) * 32, 16, 1024, SWIZZLE_128B);
      } else {
        ad = smem_desc(smem_u32(a_img) + j * 2048, a_slab, 1024, SWIZZLE_128B);
        bd = smem_desc(smem_u32(b_img) + j * 2048, b_slab, 1024, SWIZZLE_128B);
      }
      umma_bf16(tmem, ad, bd, idesc, j > 0);
    }
    umma_commit(&bar);
  }
  bool ok = mbar_wait(&bar, 0, err, false);
  tc_fence_after();
  if (ok) {
    for (int c0 = 0; c0 < N; c0 += 32) {
      uint32_t r[32];
      tmem_ld32(tmem + ((uint32_t)(warp * 32) << 16) + c0, r);
      tmem_ld_wait();
      int row = warp * 32 + (tid & 31);
      for (int j = 0; j < 32 && c0 + j < N; ++j) D[row * N + c0 + j] = __uint_as_float(r[j]);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, ncols);
}

extern "C" int cnrf_debug_umma(int32_t mode, const float* A, const float* B, float* D, int32_t N, int32_t K, int32_t* err,
                               cnrf_stream_t stream) {
  CNRF_REQUIRE(A && B && D && err, CNRF_E_NULL);
  CNRF_REQUIRE((mode == 0 || mode == 1) && K > 0 && K % 64 == 0 && N >= 16 && N <= 256, CNRF_E_SHAPE);
  CNRF_REQUIRE(mode == 0 ? N % 16 == 0 : N % 64 == 0, CNRF_E_SHAPE);
  size_t bytes = (size_t)(128 + N) * K * 2 + 1024;
  CNRF_REQUIRE(bytes <= 200 * 1024, CNRF_E_SHAPE);
  cudaError_t e = cudaFuncSetAttribute(umma_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
  if (e != cudaSuccess) return (int)e;
  umma_probe_kernel<<<1, 128, bytes, as_stream(stream)>>>(mode, A, B, D, N, K, err);
  CNRF_LAUNCH_CHECK();
  return CNRF_OK;
}

// ---- TMEM read-throughput probe (scripts/probe_tmem.py): how many bytes per clock tcgen05.ld delivers to the epilogue
// warps of ONE SM, by instruction shape, warps per lane quarter and loads in flight.  The fused MLP kernels' epilogues
// read 128 KB of fp32 accumulators per (layer, tile); DESIGN.md section 5 sizes them against this number.
__device__ __forceinline__ void tmem_ld64_probe(uint32_t taddr, uint32_t (&r)[64]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x64.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32, %33, %34, %35, %36, %37, %38, %39, %40, %41, %42, %43, %44, %45, %46, %47, %48, %49, %50, %51, %52, %53, %54, %55, %56, %57, %58, %59, %60, %61, %62, %63}, [%64];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31]), "=r"(r[32]), "=r"(r[33]), "=r"(r[34]), "=r"(r[35]), "=r"(r[36]), "=r"(r[37]), "=r"(r[38]), "=r"(r[39]), "=r"(r[40]), "=r"(r[41]), "=r"(r[42]), "=r"(r[43]), "=r"(r[44]), "=r"(r[45]), "=r"(r[46]), "=r"(r[47]), "=r"(r[48]), "=r"(r[49]), "=r"(r[50]), "=r"(r[51]), "=r"(r[52]), "=r"(r[53]), "=r"(r[54]), "=r"(r[55]), "=r"(r[56]), "=r"(r[57]), "=r"(r[58]), "=r"(r[59]), "=r"(r[60]), "=r"(r[61]), "=r"(r[62]), "=r"(r[63])
      : "r"(taddr)
      : "memory");
}
template <int mode>
__global__ void __launch_bounds__(512) tmem_ld_probe_kernel(int iters, long long* __restrict__ out,
                                                           uint32_t* __restrict__ sink) {
  __shared__ uint32_t tmem_slot;
  const int tid = threadIdx.x, warp = tid >> 5;
  if (warp == 0) tmem_alloc(&tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t base = tmem_slot + ((uint32_t)((warp & 3) * 32) << 16);
  uint32_t acc = 0;
  __syncthreads();
  const long long t0 = clock64();
  if (mode == 0) {                       // x32, one load in flight
    uint32_t v[32];
    for (int it = 0; it < iters; ++it)
#pragma unroll 2
      for (int cb = 0; cb < 16; ++cb) { tmem_ld32(base + cb * 32, v); tmem_ld_wait(); acc ^= v[0] ^ v[13] ^ v[31]; }
  } else if (mode == 1) {                // x32, two loads in flight (the epilogues' scheme)
    uint32_t va[32], vb[32];
    for (int it = 0; it < iters; ++it) {
      tmem_ld32(base, va);
      for (int cb = 0; cb < 16; cb += 2) {
        tmem_ld_wait(); tmem_ld32(base + (cb + 1) * 32, vb); acc ^= va[0] ^ va[13] ^ va[31];
        tmem_ld_wait(); if (cb + 2 < 16) tmem_ld32(base + (cb + 2) * 32, va); acc ^= vb[0] ^ vb[13] ^ vb[31];
      }
    }
  } else {                               // x64, one load in flight
    uint32_t v[64];
    for (int it = 0; it < iters; ++it)
#pragma unroll 1
      for (int cb = 0; cb < 8; ++cb) { tmem_ld64_probe(base + cb * 64, v); tmem_ld_wait(); acc ^= v[0] ^ v[29] ^ v[63]; }
  }
  __syncthreads();
  const long long t1 = clock64();
  if (tid == 0) out[blockIdx.x] = t1 - t0;
  sink[blockIdx.x * blockDim.x + tid] = acc;
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem_slot, 512);
}

// clocks[b] = SM clocks block b needed for `iters` sweeps of all 512 columns by each of its `warps` warps (every warp reads
// its lane quarter: warps x iters x 64 KB).  mode: 0 = .x32 serial, 1 = .x32 double-buffered, 2 = .x64 serial.
extern "C" int cnrf_debug_tmem_ld(int32_t blocks, int32_t warps, int32_t iters, int32_t mode, long long* clocks, uint32_t* sink,
                                  cnrf_stream_t stream) {
  CNRF_REQUIRE(clocks && sink, CNRF_E_NULL);
  CNRF_REQUIRE(blocks > 0 && warps >= 1 && warps <= 16 && iters > 0 && mode >= 0 && mode <= 2, CNRF_E_SHAPE);
  if (mode == 0) tmem_ld_probe_kernel<0><<<blocks, warps * 32, 0, as_stream(stream)>>>(iters, clocks, sink);
  else if (mode == 1) tmem_ld_probe_kernel<1><<<blocks, warps * 32, 0, as_stream(stream)>>>(iters, clocks, sink);
  else tmem_ld_probe_kernel<2><<<blocks, warps * 32, 0, as_stream(stream)>>>(iters, clocks, sink);
  CNRF_LAUNCH_CHECK();
  return CNRF_OK;
}

// ---- bulk-copy ingest probe (scripts/probe_bulk.py): how fast ONE SM can pull bytes through cp.async.bulk (UBLKCP), by
// copy size, copies in flight, number of SMs pulling at once, and whether all SMs read the same (L2-resident) bytes --
// the forward / backward chains stream their weight images this way, the weight-gradient kernel its operand tiles.
__global__ void __launch_bounds__(64) bulk_ld_probe_kernel(const uint8_t* __restrict__ src, size_t src_bytes, int per_block_offset,
                                                          int bytes, int slots, int iters, long long* __restrict__ out) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  __shared__ uint64_t full[16];
  if (threadIdx.x == 0) {
    for (int i = 0; i < slots; ++i) mbar_init(&full[i], 1);
    mbar_fence_init();
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    const size_t span = src_bytes - (size_t)bytes;
    size_t off = per_block_offset ? ((size_t)blockIdx.x * 1048576u) % (span + 1) : 0;
    off &= ~(size_t)127;
    const long long t0 = clock64();
    int issued = 0, done = 0;
    uint32_t par[16];
    for (int i = 0; i < 16; ++i) par[i] = 0;
    while (done < iters) {
      while (issued < iters && issued - done < slots) {
        const int s = issued % slots;
        mbar_arrive_expect_tx(&full[s], (uint32_t)bytes);
        bulk_g2s(smem_raw + (size_t)s * bytes, src + off, (uint32_t)bytes, &full[s]);
        off += (size_t)bytes;
        if (off > span) off = 0;
        ++issued;
      }
      const int s = done % slots;
      while (!mbar_try_wait(&full[s], par[s])) { }
      par[s] ^= 1u;
      ++done;
    }
    out[blockIdx.x] = clock64() - t0;
  }
}

// Variants of the same question.  mode 1: `nthreads` issuing threads (one per warp), each with its own ring of `slots` copies.
// mode 2: one thread issues `slots` copies back to back on ONE mbarrier (one expect_tx for all of them) and waits once per
// round -- no per-copy wait on the issuing thread.  mode 3: as mode 2 but the waiting is done by a different warp.
__global__ void __launch_bounds__(256) bulk_ld_probe2_kernel(const uint8_t* __restrict__ src, size_t src_bytes, int bytes, int slots,
                                                            int iters, int mode, int nthreads, long long* __restrict__ out) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  __shared__ uint64_t full[64];
  __shared__ uint64_t round_done;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int i = 0; i < 64; ++i) mbar_init(&full[i], 1);
    mbar_init(&round_done, 1);
    mbar_fence_init();
  }
  __syncthreads();
  const size_t span = src_bytes - (size_t)bytes;
  const long long t0 = clock64();
  if (mode == 1) {
    if (lane == 0 && warp < nthreads) {
      size_t off = ((size_t)warp * 262144u) % (span + 1) & ~(size_t)127;
      uint64_t* bar = full + warp * slots;
      uint8_t* buf = smem_raw + (size_t)warp * slots * bytes;
      uint32_t par[16];
      for (int i = 0; i < 16; ++i) par[i] = 0;
      int issued = 0, done = 0;
      while (done < iters) {
        while (issued < iters && issued - done < slots) {
          const int s = issued % slots;
          mbar_arrive_expect_tx(&bar[s], (uint32_t)bytes);
          bulk_g2s(buf + (size_t)s * bytes, src + off, (uint32_t)bytes, &bar[s]);
          off += (size_t)bytes; if (off > span) off = 0;
          ++issued;
        }
        const int s = done % slots;
        while (!mbar_try_wait(&bar[s], par[s])) { }
        par[s] ^= 1u;
        ++done;
      }
    }
  } else {
    const int rounds = iters / slots;
    if (threadIdx.x == 0) {
      size_t off = 0;
      uint32_t par = 0;
      for (int r = 0; r < rounds; ++r) {
        if (mode == 3 && r > 0) { while (!mbar_try_wait(&round_done, (uint32_t)((r - 1) & 1))) { } }
        mbar_arrive_expect_tx(&full[0], (uint32_t)(bytes * slots));
        for (int s = 0; s < slots; ++s) {
          bulk_g2s(smem_raw + (size_t)s * bytes, src + off, (uint32_t)bytes, &full[0]);
          off += (size_t)bytes; if (off > span) off = 0;
        }
        if (mode == 2) { while (!mbar_try_wait(&full[0], par)) { } par ^= 1u; }
      }
    } else if (mode == 3 && threadIdx.x == 32) {
      uint32_t par = 0;
      for (int r = 0; r < rounds; ++r) {
        while (!mbar_try_wait(&full[0], par)) { }
        par ^= 1u;
        mbar_arrive(&round_done);
      }
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) out[blockIdx.x] = clock64() - t0;
}

extern "C" int cnrf_debug_bulk_ld2(int32_t blocks, int32_t slots, int32_t bytes, int32_t iters, int32_t mode, int32_t nthreads,
                                   const void* src, size_t src_bytes, long long* clocks, cnrf_stream_t stream) {
  CNRF_REQUIRE(src && clocks, CNRF_E_NULL);
  CNRF_REQUIRE(blocks > 0 && slots >= 1 && slots <= 16 && bytes >= 128 && bytes % 128 == 0 && iters >= slots, CNRF_E_SHAPE);
  CNRF_REQUIRE(mode >= 1 && mode <= 3 && nthreads >= 1 && nthreads <= 4, CNRF_E_SHAPE);
  const size_t smem = (size_t)slots * bytes * (mode == 1 ? nthreads : 1);
  CNRF_REQUIRE(smem <= 200 * 1024 && src_bytes >= (size_t)bytes, CNRF_E_SHAPE);
  cudaError_t e = cudaFuncSetAttribute(bulk_ld_probe2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return (int)e;
  bulk_ld_probe2_kernel<<<blocks, 256, smem, as_stream(stream)>>>(static_cast<const uint8_t*>(src), src_bytes, bytes, slots, iters, mode,
                                                                 nthreads, clocks);
  CNRF_LAUNCH_CHECK();
  return CNRF_OK;
}

// clocks[b] = SM clocks block b needed to pull iters x bytes through a ring of `slots` bulk copies in flight.
extern "C" int cnrf_debug_bulk_ld(int32_t blocks, int32_t slots, int32_t bytes, int32_t iters, const void* src, size_t src_bytes,
                                  int32_t per_block_offset, long long* clocks, cnrf_stream_t stream) {
  CNRF_REQUIRE(src && clocks, CNRF_E_NULL);
  CNRF_REQUIRE(blocks > 0 && slots >= 1 && slots <= 16 && bytes >= 128 && bytes % 128 == 0 && iters > 0, CNRF_E_SHAPE);
  CNRF_REQUIRE((size_t)slots * bytes <= 200 * 1024 && src_bytes >= (size_t)bytes, CNRF_E_SHAPE);
  cudaError_t e = cudaFuncSetAttribute(bulk_ld_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, slots * bytes);
  if (e != cudaSuccess) return (int)e;
  bulk_ld_probe_kernel<<<blocks, 64, (size_t)slots * bytes, as_stream(stream)>>>(static_cast<const uint8_t*>(src), src_bytes,
                                                                               per_block_offset, bytes, slots, iters, clocks);
  CNRF_LAUNCH_CHECK();
  return CNRF_OK;
}

// ---- MMA operand-bandwidth probe (scripts/probe_umma_rate.py): the rate at which ONE SM executes tcgen05.mma (bf16, M = 128,
// N = 128 or 256, K = 16, both operands from shared memory) over static operand images, alone and while `nprod` threads stream
// bulk copies INTO shared memory.  An M128 N128 K16 step reads 4 KB of A and 4 KB of B for 64 clk of math (128 B/clk), an
// N = 256 step 12 KB for 128 clk (96 B/clk): if the rate drops when copies run beside it, the chain kernels are bound by
// shared-memory bandwidth, not by the tensor pipe.
__global__ void __launch_bounds__(160) umma_rate_probe_kernel(int n_cols, int chunks, int nprod, int copy_bytes, const uint8_t* __restrict__ src,
                                                             size_t src_bytes, long long* __restrict__ out) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t done_bar, full[4][4];
  __shared__ uint32_t tmem_slot;
  __shared__ int stop_flag;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  uint8_t* a_img = smem;                         // [128 x 256] bf16 tile image: 4 slabs of 16 KB
  uint8_t* b_img = smem + 65536;                 // 2 x [256 x 64] chunk images of 32 KB
  uint8_t* sink = smem + 65536 + 65536;          // landing area of the copies: nprod x 2 x copy_bytes <= 96 KB
  for (int i = tid; i < 131072 / 16; i += blockDim.x) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0u, 0u, 0u, 0u);
  if (tid == 0) {
    mbar_init(&done_bar, 1);
    for (int p = 0; p < 4; ++p) for (int s = 0; s < 4; ++s) mbar_init(&full[p][s], 1);
    stop_flag = 0;
    mbar_fence_init();
  }
  if (warp == 0) tmem_alloc(&tmem_slot, 512);
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  if (warp == 0) {
    if (lane == 0) {
      const uint32_t idesc = idesc_bf16(128, n_cols, 0, 0);
      const uint64_t a0 = smem_desc(smem_u32(a_img), 16, 1024, SWIZZLE_128B), b0 = smem_desc(smem_u32(b_img), 16, 1024, SWIZZLE_128B);
      uint32_t par = 0;
      const long long t0 = clock64();
      for (int c = 0; c < chunks; ++c) {           // one chunk = K 64 against N 256 (as one or two MMAs per K step)
        const uint64_t ad = a0 + (uint64_t)(((c & 3) * 16384) >> 4), bd = b0 + (uint64_t)(((c & 1) * 32768) >> 4);
        for (int h = 0; h < 256 / n_cols; ++h)
#pragma unroll
          for (int k = 0; k < 4; ++k) umma_bf16(tmem + h * 128, ad + 2 * k, bd + (uint64_t)((h * 16384) >> 4) + 2 * k, idesc, 1);
        if ((c & 7) == 7 || c == chunks - 1) {       // bound the queue: wait for the tensor pipe every 8 chunks
          umma_commit(&done_bar);
          while (!mbar_try_wait(&done_bar, par)) { }
          par ^= 1u;
        }
      }
      out[blockIdx.x * 2] = clock64() - t0;
      *(volatile int*)&stop_flag = 1;
    }
  } else if (warp - 1 < nprod) {
    if (lane == 0) {
      const int p = warp - 1;
      uint8_t* buf = sink + (size_t)p * 2 * copy_bytes;
      const size_t span = src_bytes - (size_t)copy_bytes;
      size_t off = ((size_t)p * 262144u) % (span + 1) & ~(size_t)127;
      uint32_t par[2] = {0u, 0u};
      long long copies = 0;
      int issued = 0, done = 0;
      while (!*(volatile int*)&stop_flag || done < issued) {
        while (!*(volatile int*)&stop_flag && issued - done < 2) {
          const int s = issued & 1;
          mbar_arrive_expect_tx(&full[p][s], (uint32_t)copy_bytes);
          bulk_g2s(buf + (size_t)s * copy_bytes, src + off, (uint32_t)copy_bytes, &full[p][s]);
          off += (size_t)copy_bytes; if (off > span) off = 0;
          ++issued;
        }
        if (done < issued) {
          const int s = done & 1;
          while (!mbar_try_wait(&full[p][s], par[s])) { }
          par[s] ^= 1u; ++done; ++copies;
        }
      }
      if (p == 0) out[blockIdx.x * 2 + 1] = copies;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
}

// out[2 b] = clocks block b needed for `chunks` K-64 x N-256 chunks of MMA; out[2 b + 1] = bulk copies producer 0 completed meanwhile
extern "C" int cnrf_debug_umma_rate(int32_t blocks, int32_t n_cols, int32_t chunks, int32_t nprod, int32_t copy_bytes, const void* src,
                                    size_t src_bytes, long long* out, cnrf_stream_t stream) {
  CNRF_REQUIRE(src && out, CNRF_E_NULL);
  CNRF_REQUIRE(blocks > 0 && (n_cols == 128 || n_cols == 256) && chunks > 0 && nprod >= 0 && nprod <= 4, CNRF_E_SHAPE);
  CNRF_REQUIRE(copy_bytes >= 128 && copy_bytes % 128 == 0 && copy_bytes <= 16384 && src_bytes >= (size_t)copy_bytes, CNRF_E_SHAPE);
  CNRF_REQUIRE((nprod > 0 ? nprod : 1) * 2 * copy_bytes <= 98304, CNRF_E_SHAPE);
  const int smem = 1024 + 131072 + 98304;
  cudaError_t e = cudaFuncSetAttribute(umma_rate_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  if (e != cudaSuccess) return (int)e;
  umma_rate_probe_kernel<<<blocks, 160, smem, as_stream(stream)>>>(n_cols, chunks, nprod, copy_bytes, static_cast<const uint8_t*>(src),
                                                                   src_bytes, out);
  CNRF_LAUNCH_CHECK();
  return CNRF_OK;
}
