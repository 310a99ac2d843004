// Shared latent codes.  train.py:78-96 expands one object id per image to every ray of that image, so the [N, C] code
// tensors the field receives are long runs of identical rows.  With the ids at hand the per-ray code layers, their bias
// tables and the gradients of those tables are evaluated once per RUN of equal consecutive ids instead of once per ray:
//   gid[r]  run index of ray r            rep[g]  first ray of run g            *count  number of runs
// Everything stays on the device (the host never learns the run count): dependent launches are sized for the worst case
// (one run per ray) and read *count.
#include "common.cuh"
#include "internal.h"

namespace {
constexpr int RUN_THREADS = 1024;

__global__ void __launch_bounds__(RUN_THREADS) code_runs_kernel(const int64_t* __restrict__ ids, int64_t N, int64_t num_codes,
                                                                int32_t* __restrict__ gid, int32_t* __restrict__ rep,
                                                                int32_t* __restrict__ count) {
  __shared__ int warp_sum[RUN_THREADS / 32];
  __shared__ int carry_s;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) carry_s = 0;
  __syncthreads();
  for (int64_t base = 0; base < N; base += RUN_THREADS) {
    const int64_t r = base + tid;
    int flag = 0;
    if (r < N) {
      const int64_t id = ids[r];
      if (id < 0 || id >= num_codes) asm volatile("trap;");      // torch.nn.Embedding asserts on the device too
      flag = r == 0 || ids[r - 1] != id;
    }
    int x = flag;                                                // inclusive scan of the run starts
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int y = __shfl_up_sync(0xFFFFFFFFu, x, o);
      if (lane >= o) x += y;
    }
    if (lane == 31) warp_sum[warp] = x;
    __syncthreads();
    if (warp == 0) {
      int w = warp_sum[lane];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int y = __shfl_up_sync(0xFFFFFFFFu, w, o);
        if (lane >= o) w += y;
      }
      warp_sum[lane] = w;
    }
    __syncthreads();
    const int carry = carry_s;
    const int incl = carry + x + (warp ? warp_sum[warp - 1] : 0);
    if (r < N) {
      gid[r] = incl - 1;
      if (flag) rep[incl - 1] = (int32_t)r;
    }
    __syncthreads();
    if (tid == RUN_THREADS - 1) carry_s = incl;
    __syncthreads();
  }
  if (tid == 0) *count = carry_s;
}

// out[g, :] = tab[ids[rep[g]], :]   for g < *count          (C % 4 == 0, rows 16-byte aligned)
__global__ void gather_codes_kernel(const int64_t* __restrict__ ids, const int32_t* __restrict__ rep,
                                    const int32_t* __restrict__ count, int64_t N, int C4, const float* __restrict__ tab,
                                    int64_t ld, float* __restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N * C4) return;
  const int64_t g = i / C4;
  if (g >= *count) return;
  const int c = (int)(i - g * C4);
  const float4 v = __ldg(reinterpret_cast<const float4*>(tab + ids[rep[g]] * ld) + c);
  reinterpret_cast<float4*>(out + g * (int64_t)C4 * 4)[c] = v;
}

// g_tab[ids[rep[g]], :] += g_codes[g, :]   for g < *count   (g_tab zero-filled by the caller; equal ids may recur in
// separate runs, hence the atomics)
__global__ void scatter_code_grads_kernel(const int64_t* __restrict__ ids, const int32_t* __restrict__ rep,
                                          const int32_t* __restrict__ count, int64_t N, int C,
                                          const float* __restrict__ g_codes, float* __restrict__ g_tab, int64_t ld) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N * C) return;
  const int64_t g = i / C;
  if (g >= *count) return;
  const int c = (int)(i - g * C);
  atomicAdd(g_tab + ids[rep[g]] * ld + c, g_codes[g * C + c]);
}
}  // namespace

size_t cnrf_code_runs_bytes(int64_t N) { return 2 * ((size_t)N * sizeof(int32_t) + 255) / 256 * 256 + 256; }

CodeRuns cnrf_code_runs_carve(void* base, int64_t N) {
  CodeRuns r;
  const size_t a = ((size_t)N * sizeof(int32_t) + 255) / 256 * 256;
  char* p = reinterpret_cast<char*>(base);
  r.gid = reinterpret_cast<int32_t*>(p);
  r.rep = reinterpret_cast<int32_t*>(p + a);
  r.count = reinterpret_cast<int32_t*>(p + 2 * a);
  return r;
}

int cnrf_code_runs(const int64_t* ids, int64_t N, int64_t num_codes, const CodeRuns& r, cudaStream_t st) {
  if (N >= (1ll << 31)) return CNRF_E_SHAPE;
  code_runs_kernel<<<1, RUN_THREADS, 0, st>>>(ids, N, num_codes, r.gid, r.rep, r.count);
  CNRF_LAUNCH_CHECK();
  return CNRF_OK;
}

// every ray carries the SAME code row (a stride-0 expand: eval.py:149, nerf/__init__.py:176-177): one run
__global__ void broadcast_runs_kernel(int64_t N, int32_t* __restrict__ gid, int32_t* __restrict__ rep, int32_t* __restrict__ count) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < N) gid[i] = 0;
  if (i == 0) { rep[0] = 0; *count = 1; }
}
int cnrf_code_runs_broadcast(int64_t N, const CodeRuns& r, cudaStream_t st) {
  broadcast_runs_kernel<<<grid_for(N, 256), 256, 0, st>>>(N, r.gid, r.rep, r.count);
  CNRF_LAUNCH_CHECK();
  return CNRF_OK;
}

int cnrf_gather_codes(const int64_t* ids, const CodeRuns& r, int64_t N, int C, const float* tab, int64_t ld, float* out,
                      cudaStream_t st) {
  if (C % 4 || ld % 4 || (reinterpret_cast<uintptr_t>(tab) & 15) || (reinterpret_cast<uintptr_t>(out) & 15)) return CNRF_E_SHAPE;
  gather_codes_kernel<<<grid_for(N * (C / 4), 256), 256, 0, st>>>(ids, r.rep, r.count, N, C / 4, tab, ld, out);
  CNRF_LAUNCH_CHECK();
  return CNRF_OK;
}

int cnrf_scatter_code_grads(const int64_t* ids, const CodeRuns& r, int64_t N, int C, const float* g_codes, float* g_tab,
                            int64_t ld, cudaStream_t st) {
  scatter_code_grads_kernel<<<grid_for(N * C, 256), 256, 0, st>>>(ids, r.rep, r.count, N, C, g_codes, g_tab, ld);
  CNRF_LAUNCH_CHECK();
  return CNRF_OK;
}
