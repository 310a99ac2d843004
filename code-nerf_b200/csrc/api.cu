// Library-level entry points of libcnrf.so.
#include "common.cuh"

long long g_cnrf_launches = 0;
extern "C" int64_t cnrf_launch_count(void) { return (int64_t)g_cnrf_launches; }

extern "C" int cnrf_abi_version(void) { return CNRF_ABI_VERSION; }

extern "C" const char* cnrf_error_string(int code) {
  switch (code) {
    case CNRF_OK: return "ok";
    case CNRF_E_NULL: return "a required pointer is NULL";
    case CNRF_E_SHAPE: return "a size is out of the supported range";
    case CNRF_E_WORKSPACE: return "workspace or saved buffer missing or too small";
    case CNRF_E_UNSUPPORTED: return "configuration not supported by this precision mode";
    case CNRF_E_DEVICE: return "current device is not an sm_100 (B200) GPU";
    case CNRF_E_SCRATCH_LIMIT: return "shape needs more scratch than the fixed 64 MiB array of cnrf_field_fwd_noalloc: split the rays into smaller chunks";
    default: return code > 0 ? cudaGetErrorString((cudaError_t)code) : "unknown libcnrf error";
  }
}

extern "C" int cnrf_device_check(void) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  int major = 0;
  e = cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
  if (e != cudaSuccess) return (int)e;
  return major == 10 ? CNRF_OK : CNRF_E_DEVICE;
}
