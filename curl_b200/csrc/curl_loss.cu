// sm_100a kernels + C ABI of the colour terms of the reference's CURLLoss (model.py:36-118).
// One pass over (pred, target, mask) produces, per image, the four sums the loss needs and the two
// Lab L planes MS-SSIM consumes; the backward is one elementwise pass.  Per-pixel math: curl_loss.cuh.
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/curl_b200.h"
#include "curl_loss.cuh"
#include "curl_host.cuh"

namespace curl {
namespace loss {

constexpr int THREADS = 256;
constexpr int NSUM = 5;          // s0..s3 and the number of unmasked pixels

struct LParams {
  const float* pred;     // [B][3][HW]
  const float* target;   // [B][3][HW]
  const void* mask;      // [B][HW] u8 / f32 or null
  float* sums;           // fwd: [B][NSUM]
  float* L_pred;         // fwd: [B][HW] or null
  float* L_target;       // fwd: [B][HW] or null
  float* partials;       // fwd: [B][gridDim.x][NSUM]
  unsigned* tickets;     // fwd: [B]
  const float* w;        // bwd: [B][4] d loss / d sum_k
  const float* gL;       // bwd: [B][HW] or null
  float* gpred;          // bwd: [B][3][HW]
  int HW;
};

using host::sm_count;

template <int MK>
__device__ __forceinline__ float mask1(const void* mask, size_t i) {
  if (MK == MASK_NONE) return 1.0f;
  if (MK == MASK_U8) return reinterpret_cast<const unsigned char*>(mask)[i] ? 1.0f : 0.0f;
  return reinterpret_cast<const float*>(mask)[i];
}

// VEC = 2: one packed pixel pair per thread and iteration (HW even, 8-byte aligned bases); VEC = 1: scalar.
template <int MK, int VEC>
__global__ void __launch_bounds__(THREADS) loss_fwd_kernel(const LParams p) {
  using T = typename std::conditional<VEC == 2, F2, float>::type;
  __shared__ float s_part[(THREADS / 32) * NSUM];
  __shared__ unsigned s_ticket;
  const int b = blockIdx.y, HW = p.HW;
  const float* pr = p.pred + (size_t)b * 3 * HW;
  const float* tg = p.target + (size_t)b * 3 * HW;
  T acc[NSUM];
#pragma unroll
  for (int k = 0; k < NSUM; ++k) acc[k] = T(0.0f);
  const int stride = gridDim.x * THREADS * VEC;
  for (int i = (blockIdx.x * THREADS + threadIdx.x) * VEC; i < HW; i += stride) {
    T x[3], y[3], m, s[4], lp, lt;
    if constexpr (VEC == 2) {
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const float2 a = __ldcs(reinterpret_cast<const float2*>(pr + (size_t)c * HW + i));
        const float2 q = __ldcs(reinterpret_cast<const float2*>(tg + (size_t)c * HW + i));
        x[c] = F2(a.x, a.y); y[c] = F2(q.x, q.y);
      }
      m = F2(mask1<MK>(p.mask, (size_t)b * HW + i), mask1<MK>(p.mask, (size_t)b * HW + i + 1));
    } else {
#pragma unroll
      for (int c = 0; c < 3; ++c) { x[c] = __ldcs(pr + (size_t)c * HW + i); y[c] = __ldcs(tg + (size_t)c * HW + i); }
      m = mask1<MK>(p.mask, (size_t)b * HW + i);
    }
    loss_terms_fwd<T>(x, y, m, s, &lp, &lt);
#pragma unroll
    for (int k = 0; k < 4; ++k) acc[k] = vadd(acc[k], s[k]);
    acc[4] = vadd(acc[4], m);
    if constexpr (VEC == 2) {
      if (p.L_pred) __stcs(reinterpret_cast<float2*>(p.L_pred + (size_t)b * HW + i), make_float2(lp.lo(), lp.hi()));
      if (p.L_target) __stcs(reinterpret_cast<float2*>(p.L_target + (size_t)b * HW + i), make_float2(lt.lo(), lt.hi()));
    } else {
      if (p.L_pred) p.L_pred[(size_t)b * HW + i] = lp;
      if (p.L_target) p.L_target[(size_t)b * HW + i] = lt;
    }
  }
  // registers -> warp shuffles -> shared memory -> one partial row per block -> one ticket per block
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < NSUM; ++k) {
    float v = vhsum(acc[k]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) s_part[warp * NSUM + k] = v;
  }
  __syncthreads();
  float* row = p.partials + ((size_t)b * gridDim.x + blockIdx.x) * NSUM;
  if (threadIdx.x < NSUM) {
    float v = 0.0f;
#pragma unroll
    for (int w = 0; w < THREADS / 32; ++w) v += s_part[w * NSUM + threadIdx.x];
    row[threadIdx.x] = v;
    __threadfence();
  }
  __syncthreads();
  if (threadIdx.x == 0) s_ticket = atomicAdd(&p.tickets[b], 1u);
  __syncthreads();
  if (s_ticket != gridDim.x - 1) return;
  __threadfence();
  if (threadIdx.x < NSUM) {        // last block of the image: fixed-order fold in double
    const float* rows = p.partials + (size_t)b * gridDim.x * NSUM;
    double a = 0.0;
    for (unsigned r = 0; r < gridDim.x; ++r) a += (double)__ldcg(rows + (size_t)r * NSUM + threadIdx.x);
    p.sums[(size_t)b * NSUM + threadIdx.x] = (float)a;
  }
}

template <int MK, int VEC>
__global__ void __launch_bounds__(THREADS) loss_bwd_kernel(const LParams p) {
  using T = typename std::conditional<VEC == 2, F2, float>::type;
  const int b = blockIdx.y, HW = p.HW;
  const float* pr = p.pred + (size_t)b * 3 * HW;
  const float* tg = p.target + (size_t)b * 3 * HW;
  float* gp = p.gpred + (size_t)b * 3 * HW;
  const float w[4] = {p.w[b * 4 + 0], p.w[b * 4 + 1], p.w[b * 4 + 2], p.w[b * 4 + 3]};
  const int stride = gridDim.x * THREADS * VEC;
  for (int i = (blockIdx.x * THREADS + threadIdx.x) * VEC; i < HW; i += stride) {
    T x[3], y[3], m, gl = T(0.0f), g[3];
    if constexpr (VEC == 2) {
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const float2 a = __ldcs(reinterpret_cast<const float2*>(pr + (size_t)c * HW + i));
        const float2 q = __ldcs(reinterpret_cast<const float2*>(tg + (size_t)c * HW + i));
        x[c] = F2(a.x, a.y); y[c] = F2(q.x, q.y);
      }
      m = F2(mask1<MK>(p.mask, (size_t)b * HW + i), mask1<MK>(p.mask, (size_t)b * HW + i + 1));
      if (p.gL) { const float2 q = __ldcs(reinterpret_cast<const float2*>(p.gL + (size_t)b * HW + i)); gl = F2(q.x, q.y); }
    } else {
#pragma unroll
      for (int c = 0; c < 3; ++c) { x[c] = __ldcs(pr + (size_t)c * HW + i); y[c] = __ldcs(tg + (size_t)c * HW + i); }
      m = mask1<MK>(p.mask, (size_t)b * HW + i);
      if (p.gL) gl = __ldcs(p.gL + (size_t)b * HW + i);
    }
    loss_terms_bwd<T>(x, y, m, w, gl, g);
    if constexpr (VEC == 2) {
#pragma unroll
      for (int c = 0; c < 3; ++c)
        __stcs(reinterpret_cast<float2*>(gp + (size_t)c * HW + i), make_float2(g[c].lo(), g[c].hi()));
    } else {
#pragma unroll
      for (int c = 0; c < 3; ++c) __stcs(gp + (size_t)c * HW + i, g[c]);
    }
  }
}

// ------------------------------------------------------------------------------------------------
// Host side
// ------------------------------------------------------------------------------------------------
static bool aligned4(const void* q) { return (reinterpret_cast<uintptr_t>(q) & 3u) == 0; }
static bool aligned8(const void* q) { return (reinterpret_cast<uintptr_t>(q) & 7u) == 0; }
static bool aligned16(const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15u) == 0; }

// blocks per image: whole waves of 8 blocks per SM over the batch, never more blocks than pixel pairs
static int grid_x(int B, int HW) {
  const long long want = ((long long)sm_count() * 8 + B - 1) / B;
  const long long cap = ((long long)HW + 2 * THREADS - 1) / (2 * THREADS);
  long long g = want < cap ? want : cap;
  return (int)(g < 1 ? 1 : g);
}

static size_t tick_bytes(int B) { return ((size_t)B * sizeof(unsigned) + 255) / 256 * 256; }

template <int VEC>
static int launch_fwd(const LParams& p, int B, int mask_kind, cudaStream_t st) {
  const dim3 grid(grid_x(B, p.HW), B), block(THREADS);
  switch (mask_kind) {
    case CURL_MASK_NONE: loss_fwd_kernel<MASK_NONE, VEC><<<grid, block, 0, st>>>(p); break;
    case CURL_MASK_U8: loss_fwd_kernel<MASK_U8, VEC><<<grid, block, 0, st>>>(p); break;
    default: loss_fwd_kernel<MASK_F32, VEC><<<grid, block, 0, st>>>(p); break;
  }
  return cudaGetLastError() == cudaSuccess ? CURL_OK : CURL_ERR_LAUNCH;
}

template <int VEC>
static int launch_bwd(const LParams& p, int B, int mask_kind, cudaStream_t st) {
  const dim3 grid(grid_x(B, p.HW), B), block(THREADS);
  switch (mask_kind) {
    case CURL_MASK_NONE: loss_bwd_kernel<MASK_NONE, VEC><<<grid, block, 0, st>>>(p); break;
    case CURL_MASK_U8: loss_bwd_kernel<MASK_U8, VEC><<<grid, block, 0, st>>>(p); break;
    default: loss_bwd_kernel<MASK_F32, VEC><<<grid, block, 0, st>>>(p); break;
  }
  return cudaGetLastError() == cudaSuccess ? CURL_OK : CURL_ERR_LAUNCH;
}

static int check_common(const float* pred, const float* target, const void* mask, int mask_kind, int B, int H, int W) {
  if (!pred || !target || B <= 0 || H <= 0 || W <= 0) return CURL_ERR_BAD_ARG;
  if ((long long)H * W > 0x3fffffffLL) return CURL_ERR_BAD_ARG;
  if (mask_kind != CURL_MASK_NONE && mask_kind != CURL_MASK_U8 && mask_kind != CURL_MASK_F32) return CURL_ERR_BAD_ARG;
  if ((mask_kind == CURL_MASK_NONE) != (mask == nullptr)) return CURL_ERR_BAD_ARG;
  if (!aligned4(pred) || !aligned4(target)) return CURL_ERR_MISALIGNED;
  if (mask_kind == CURL_MASK_F32 && !aligned4(mask)) return CURL_ERR_MISALIGNED;
  return CURL_OK;
}

}  // namespace loss
}  // namespace curl

using namespace curl;
using namespace curl::loss;

extern "C" {

size_t curl_loss_terms_workspace_bytes(int B, int H, int W) {
  if (B <= 0 || H <= 0 || W <= 0) return 0;
  return tick_bytes(B) + (size_t)B * grid_x(B, H * W) * NSUM * sizeof(float);
}

int curl_loss_terms_fwd_f32(const float* pred, const float* target, const void* mask, int mask_kind, float* sums,
                            float* L_pred, float* L_target, void* workspace, size_t workspace_bytes, int B, int H,
                            int W, void* stream) {
  int rc = check_common(pred, target, mask, mask_kind, B, H, W);
  if (rc != CURL_OK) return rc;
  if (!sums) return CURL_ERR_BAD_ARG;
  if (!aligned4(sums) || !aligned4(L_pred) || !aligned4(L_target)) return CURL_ERR_MISALIGNED;
  const size_t need = curl_loss_terms_workspace_bytes(B, H, W);
  if (!workspace || workspace_bytes < need || !aligned16(workspace)) return CURL_ERR_WORKSPACE;
  LParams p{};
  p.pred = pred; p.target = target; p.mask = mask; p.sums = sums; p.L_pred = L_pred; p.L_target = L_target;
  p.tickets = reinterpret_cast<unsigned*>(workspace);
  p.partials = reinterpret_cast<float*>(reinterpret_cast<char*>(workspace) + tick_bytes(B));
  p.HW = H * W;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (cudaMemsetAsync(workspace, 0, tick_bytes(B), st) != cudaSuccess) return CURL_ERR_LAUNCH;
  const bool v2 = (p.HW % 2 == 0) && aligned8(pred) && aligned8(target) && aligned8(L_pred) && aligned8(L_target) &&
                  (mask_kind != CURL_MASK_F32 || aligned8(mask));
  return v2 ? launch_fwd<2>(p, B, mask_kind, st) : launch_fwd<1>(p, B, mask_kind, st);
}

int curl_loss_terms_bwd_f32(const float* pred, const float* target, const void* mask, int mask_kind, const float* w,
                            const float* gL_pred, float* gpred, int B, int H, int W, void* stream) {
  int rc = check_common(pred, target, mask, mask_kind, B, H, W);
  if (rc != CURL_OK) return rc;
  if (!w || !gpred) return CURL_ERR_BAD_ARG;
  if (!aligned4(w) || !aligned4(gL_pred) || !aligned4(gpred)) return CURL_ERR_MISALIGNED;
  LParams p{};
  p.pred = pred; p.target = target; p.mask = mask; p.w = w; p.gL = gL_pred; p.gpred = gpred;
  p.HW = H * W;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const bool v2 = (p.HW % 2 == 0) && aligned8(pred) && aligned8(target) && aligned8(gL_pred) && aligned8(gpred) &&
                  (mask_kind != CURL_MASK_F32 || aligned8(mask));
  return v2 ? launch_bwd<2>(p, B, mask_kind, st) : launch_bwd<1>(p, B, mask_kind, st);
}

}  // extern "C"
