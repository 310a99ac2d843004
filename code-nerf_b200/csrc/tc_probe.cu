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
  bool ok = mbar_wait(&bar, 0, err);
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
