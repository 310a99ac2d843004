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
