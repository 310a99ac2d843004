// Device-resident dataset stage.  The reference decodes a PNG and divides it by 255.0 on the
// host for every sample of every iteration (datasets/dataset.py:70-74: np.asarray(imageio.imread(f)) / 255.0 in float64,
// .astype(np.float32) at :90) and train.py:77-80 then gathers the selected pixels of the float image on the device.  Here the
// views stay in HBM as the uint8 crops the PNGs hold, and
//   cnrf_u8_to_unit    turns whole views into the float images the reference's loader would have produced, and
//   cnrf_gather_pixels_u8 produces train.py:77-80's target rows directly from the uint8 views (no float image at all);
// both compute (float)((double)v / 255.0) -- the same two roundings as numpy's float64 division followed by the float32
// cast, so the results are bit-identical to the reference's loader (tests/test_gpu_data.py).
#include "common.cuh"
#include "internal.h"

namespace {
__device__ __forceinline__ float unit_of(uint8_t v) { return (float)((double)v / 255.0); }

// out[i] = in[i] / 255.0 over n bytes (n % 4 need not be 0); one uint32 = four pixels' channels per thread
__global__ void u8_to_unit_kernel(const uint8_t* __restrict__ in, int64_t n, float* __restrict__ out) {
  const int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t i = q * 4;
  if (i + 3 < n && ((reinterpret_cast<uintptr_t>(in) & 3) == 0) && ((reinterpret_cast<uintptr_t>(out) & 15) == 0)) {
    const uint32_t w = __ldg(reinterpret_cast<const uint32_t*>(in) + q);
    reinterpret_cast<float4*>(out)[q] = make_float4(unit_of(w & 0xFF), unit_of((w >> 8) & 0xFF), unit_of((w >> 16) & 0xFF), unit_of(w >> 24));
  } else {
    for (int64_t k = i; k < n && k < i + 4; ++k) out[k] = unit_of(__ldg(in + k));
  }
}

// out[(b * R + r) * C_out + c] = views[view[b]][select[b * R + r]][c] / 255.0,  c < C_out <= C  (pixels of C bytes)
__global__ void gather_pixels_u8_kernel(const uint8_t* __restrict__ views, int64_t hw, int C, const int64_t* __restrict__ view,
                                        const int64_t* __restrict__ select, int64_t B, int64_t R, int C_out,
                                        float* __restrict__ out, int* __restrict__ bad) {
  const int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= B * R) return;
  const int64_t b = n / R;
  const int64_t px = __ldg(select + n);
  if (px < 0 || px >= hw) { if (bad) atomicExch(bad, 1); return; }
  const uint8_t* p = views + (__ldg(view + b) * hw + px) * C;
#pragma unroll 4
  for (int c = 0; c < C_out; ++c) out[n * C_out + c] = unit_of(__ldg(p + c));
}
}  // namespace

extern "C" int cnrf_u8_to_unit(const uint8_t* in, int64_t n, float* out, cnrf_stream_t stream) {
  CNRF_REQUIRE(in && out, CNRF_E_NULL);
  CNRF_REQUIRE(n >= 0, CNRF_E_SHAPE);
  if (n == 0) return CNRF_OK;
  const int64_t quads = (n + 3) / 4;
  u8_to_unit_kernel<<<(unsigned)((quads + 255) / 256), 256, 0, as_stream(stream)>>>(in, n, out);
  CNRF_LAUNCH_CHECK();
  return CNRF_OK;
}

extern "C" int cnrf_gather_pixels_u8(const uint8_t* views, int64_t n_views, int64_t hw, int32_t channels, const int64_t* view,
                                     const int64_t* select, int64_t B, int64_t R, int32_t out_channels, float* out, int32_t* bad,
                                     cnrf_stream_t stream) {
  CNRF_REQUIRE(views && view && select && out, CNRF_E_NULL);
  CNRF_REQUIRE(n_views > 0 && hw > 0 && channels >= 1 && channels <= 4 && out_channels >= 1 && out_channels <= channels && B >= 0 && R >= 0,
               CNRF_E_SHAPE);
  if (B * R == 0) return CNRF_OK;
  gather_pixels_u8_kernel<<<(unsigned)((B * R + 255) / 256), 256, 0, as_stream(stream)>>>(views, hw, channels, view, select, B, R,
                                                                                        out_channels, out, bad);
  CNRF_LAUNCH_CHECK();
  return CNRF_OK;
}
