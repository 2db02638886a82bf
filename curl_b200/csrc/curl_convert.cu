// Device-side TF.to_tensor for 8-bit images (reference data.py:152-158, infer.py:38-41): uint8 CHW or HWC
// -> planar fp32 x = u8 / 255, the correctly rounded quotient (curl_vec.cuh::u8_to_unit, no division, no
// I2F).  Lets a training step upload 3 B/px over PCIe instead of 12: the 8-bit data is what exists on the
// host, the fp32 tensor is made next to the kernels that read it.  HBM-bound: 3 B/px in, 12 B/px out.
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/curl_b200.h"
#include "curl_vec.cuh"
#include "curl_host.cuh"

namespace curl {
namespace cvt {

constexpr int THREADS = 256;
constexpr float TWO23 = 8388608.0f;

// byte k of w as the float 2^23 + byte (one PRMT; the subtraction of 2^23 is exact)
__device__ __forceinline__ float byte_of(unsigned w, int k) {
  return __uint_as_float(__byte_perm(w, 0x4B000000u, 0x7650u + k));
}

// four consecutive pixels per thread; HW % 4 == 0 and 4-byte aligned planes (VEC4), else one pixel per thread
template <bool IN_HWC, bool VEC4>
__global__ void __launch_bounds__(THREADS) u8_to_unit_kernel(const unsigned char* __restrict__ in, float* __restrict__ out, int HW) {
  const int b = blockIdx.y;
  const unsigned char* src = in + (size_t)b * 3 * HW;
  float* dst = out + (size_t)b * 3 * HW;
  const int step = VEC4 ? 4 : 1;
  for (int i = (blockIdx.x * THREADS + threadIdx.x) * step; i < HW; i += gridDim.x * THREADS * step) {
    if constexpr (VEC4) {
      float ch[3][4];
      if (IN_HWC) {
        const unsigned* q = reinterpret_cast<const unsigned*>(src + (size_t)3 * i);
        const unsigned w0 = __ldcs(q), w1 = __ldcs(q + 1), w2 = __ldcs(q + 2);
        ch[0][0] = byte_of(w0, 0); ch[1][0] = byte_of(w0, 1); ch[2][0] = byte_of(w0, 2);
        ch[0][1] = byte_of(w0, 3); ch[1][1] = byte_of(w1, 0); ch[2][1] = byte_of(w1, 1);
        ch[0][2] = byte_of(w1, 2); ch[1][2] = byte_of(w1, 3); ch[2][2] = byte_of(w2, 0);
        ch[0][3] = byte_of(w2, 1); ch[1][3] = byte_of(w2, 2); ch[2][3] = byte_of(w2, 3);
      } else {
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          const unsigned w = __ldcs(reinterpret_cast<const unsigned*>(src + (size_t)c * HW + i));
#pragma unroll
          for (int k = 0; k < 4; ++k) ch[c][k] = byte_of(w, k);
        }
      }
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        float4 v;
        v.x = u8_to_unit<float>(ch[c][0] - TWO23); v.y = u8_to_unit<float>(ch[c][1] - TWO23);
        v.z = u8_to_unit<float>(ch[c][2] - TWO23); v.w = u8_to_unit<float>(ch[c][3] - TWO23);
        __stcs(reinterpret_cast<float4*>(dst + (size_t)c * HW + i), v);
      }
    } else {
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const unsigned char u = IN_HWC ? src[(size_t)3 * i + c] : src[(size_t)c * HW + i];
        dst[(size_t)c * HW + i] = u8_to_unit<float>((float)u);
      }
    }
  }
}

using host::sm_count;

}  // namespace cvt
}  // namespace curl

using namespace curl::cvt;

extern "C" int curl_u8_to_unit_f32(const uint8_t* img, int in_layout, float* out, int B, int H, int W, void* stream) {
  if (!img || !out || B <= 0 || H <= 0 || W <= 0 || (long long)H * W > 0x3fffffffLL || B > 65535) return CURL_ERR_BAD_ARG;
  if (in_layout != CURL_LAYOUT_CHW && in_layout != CURL_LAYOUT_HWC) return CURL_ERR_BAD_ARG;
  if (reinterpret_cast<uintptr_t>(out) & 3u) return CURL_ERR_MISALIGNED;
  const int HW = H * W;
  const bool v4 = (HW % 4 == 0) && ((reinterpret_cast<uintptr_t>(img) & 3u) == 0) && ((reinterpret_cast<uintptr_t>(out) & 15u) == 0);
  const long long per = (long long)THREADS * (v4 ? 4 : 1);
  long long gx = ((long long)HW + per - 1) / per;
  const long long cap = ((long long)sm_count() * 16 + B - 1) / B;
  if (gx > cap) gx = cap;
  const dim3 grid((unsigned)(gx < 1 ? 1 : gx), B);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (in_layout == CURL_LAYOUT_HWC) {
    if (v4) u8_to_unit_kernel<true, true><<<grid, THREADS, 0, st>>>(img, out, HW);
    else u8_to_unit_kernel<true, false><<<grid, THREADS, 0, st>>>(img, out, HW);
  } else {
    if (v4) u8_to_unit_kernel<false, true><<<grid, THREADS, 0, st>>>(img, out, HW);
    else u8_to_unit_kernel<false, false><<<grid, THREADS, 0, st>>>(img, out, HW);
  }
  return cudaGetLastError() == cudaSuccess ? CURL_OK : CURL_ERR_LAUNCH;
}
