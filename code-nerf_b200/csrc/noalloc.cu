// cnrf_field_fwd_noalloc: forward_pass (inference) for callers that cannot allocate -- code under CUDA graph capture,
// tight serving loops, harnesses that forbid cudaMalloc.  Same kernels and arithmetic as cnrf_field_fwd with keep = 0; the
// only difference is where the scratch lives: instead of a caller-owned `workspace`, a file-scope __device__ array of
// CNRF_NOALLOC_SCRATCH_BYTES (64 MiB) that every GPU which loads the library holds until the process exits.
// No counterpart in the reference.
//
//   * The entry point calls no cudaMalloc / cudaFree / allocating library routine (the rest of libcnrf never allocates
//     either: include/cnrf.h "Conventions").
//   * `g_noalloc_scratch` below is, in HOST code, only a stand-in for the symbol: its device address is obtained once per
//     device, with that device current, by cudaGetSymbolAddress (return code checked) and that address is what the kernels
//     receive.
//   * Shapes whose scratch demand (cnrf_field_workspace_bytes(s, 0)) exceeds the array are rejected with
//     CNRF_E_SCRATCH_LIMIT before anything is launched.
//   * Every call on a device shares the one array and finds the previous call's contents in it.  Nothing in the forward
//     pass needs the scratch to start at zero (every table, image and flag it reads is written earlier in the same call,
//     exactly as with an uninitialised caller workspace), so nothing is cleared.  Calls that overlap on the GPU -- two
//     streams, or two host threads -- RACE on the array: issue them on one stream, or order them with events.
#include "common.cuh"
#include "internal.h"

static __device__ __align__(256) unsigned char g_noalloc_scratch[CNRF_NOALLOC_SCRATCH_BYTES];

namespace {
// device address of the scratch array on the current device (nullptr + *err on failure); cached per device ordinal
void* noalloc_scratch_ptr(cudaError_t* err) {
  static void* cache[CNRF_MAX_DEVICES] = {};
  *err = cudaSuccess;
  const int slot = cnrf_device_slot();
  if (slot >= 0 && cache[slot]) return cache[slot];
  void* p = nullptr;
  const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
  if (e != cudaSuccess || !p) { *err = e != cudaSuccess ? e : cudaErrorInvalidSymbol; return nullptr; }
  if (slot >= 0) cache[slot] = p;
  return p;
}
}  // namespace

extern "C" int cnrf_field_fwd_noalloc(const cnrf_field_shape* s, const cnrf_codenerf_params* p, const float* pts,
                                      const float* rd, const float* freqs_xyz, int32_t L_xyz, const float* freqs_dir,
                                      int32_t L_dir, const float* z_s, int64_t ld_zs, const float* z_t, int64_t ld_zt,
                                      float* raw, cnrf_stream_t stream) {
  if (!s) return CNRF_E_NULL;
  if (s->keep) return CNRF_E_UNSUPPORTED;                 // inference only: a backward needs a caller-owned saved buffer
  const size_t need = cnrf_field_workspace_bytes(s, 0);
  if (need == 0) return CNRF_E_SHAPE;
  if (need > (size_t)CNRF_NOALLOC_SCRATCH_BYTES) return CNRF_E_SCRATCH_LIMIT;
  cudaError_t e;
  void* scratch = noalloc_scratch_ptr(&e);
  if (!scratch) return (int)e;
  return cnrf_field_fwd_impl(s, p, pts, rd, freqs_xyz, L_xyz, freqs_dir, L_dir, nullptr, 0, z_s, ld_zs, z_t, ld_zt, raw, nullptr,
                             0, scratch, (size_t)CNRF_NOALLOC_SCRATCH_BYTES, as_stream(stream));
}
