// sm_100a kernels + C ABI of the layer with segment-clamped curves (CURL paper eq. 1) -- the opt-in
// flag of SURVEY.md section 8(f) row f4.  Per-pixel math: curl_seg.cuh.  Scalar lanes: the segment
// lookup is a per-pixel gather from the knot table in shared memory.
//
// Knot gradients are a per-segment (binned) reduction: every warp owns a private histogram
// [2][total knots] in shared memory and adds to it WITHOUT atomics, in an order fixed by the lane
// numbers (WarpBins below); the block folds its warps' histograms into one partial row, takes an
// atomic ticket, and the image's last block folds all rows in a fixed order in double.  Like the
// default path, the result is bit-reproducible from run to run (round 1 used shared-memory float
// atomics here, which are not).
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/curl_b200.h"
#include "curl_seg.cuh"
#include "curl_host.cuh"

namespace curl {
namespace seg {

constexpr int THREADS = 256;
constexpr int NWARPS = THREADS / 32;

struct SParams {
  const float* img;      // [B][3][HW]
  const void* mask;      // [B][HW] u8 / f32 or null
  const float* knots[3]; // L, R, H rows (log domain)
  long long ld[3];
  int K[3];
  float* out;            // fwd
  float* reg;            // fwd: [B]
  const float* gout;     // bwd
  const float* greg;     // bwd: [B] or null
  float* gimg;           // bwd: or null
  float* gk[3];          // bwd: dense [B][3K | 3K | 4K]
  float* partials;       // bwd: [B][gridDim.x][2*NK]
  unsigned* tickets;     // bwd: [B]
  int HW;
};

__host__ __device__ inline int total_knots(const int K[3]) { return 3 * K[0] + 3 * K[1] + 4 * K[2]; }

using host::sm_count;

template <int MK>
__device__ __forceinline__ float mask1(const void* mask, size_t i) {
  if (MK == MASK_NONE) return 1.0f;
  if (MK == MASK_U8) return reinterpret_cast<const unsigned char*>(mask)[i] ? 1.0f : 0.0f;
  return reinterpret_cast<const float*>(mask)[i];
}

// knots of image b -> shared memory as C = exp(param) (curves.py:53-54,105-106,152-153), curve tables
__device__ __forceinline__ void stage(const SParams& p, int b, float* s_C, int* s_off, SegCurve cv[LAYER_CURVES]) {
  int off = 0;
  for (int g = 0; g < 3; ++g) {
    const int n = (g == 2 ? 4 : 3) * p.K[g];
    const float* src = p.knots[g] + (size_t)b * p.ld[g];
    for (int i = threadIdx.x; i < n; i += blockDim.x) s_C[off + i] = (float)exp((double)src[i]);
    off += n;
  }
  if (threadIdx.x == 0) {
    int o = 0;
    for (int k = 0; k < LAYER_CURVES; ++k) { s_off[k] = o; o += p.K[k < 3 ? 0 : (k < 6 ? 1 : 2)]; }
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < LAYER_CURVES; ++k) cv[k] = SegCurve{s_C + s_off[k], p.K[k < 3 ? 0 : (k < 6 ? 1 : 2)]};
}

template <int MK>
__global__ void __launch_bounds__(THREADS) seg_fwd_kernel(const SParams p) {
  extern __shared__ __align__(16) float smem[];
  __shared__ int s_off[LAYER_CURVES];
  float* s_C = smem;
  const int b = blockIdx.y, HW = p.HW;
  SegCurve cv[LAYER_CURVES];
  stage(p, b, s_C, s_off, cv);
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    double r = 0.0;
#pragma unroll
    for (int k = 0; k < LAYER_CURVES; ++k) r += curve_regulariser(cv[k].C, cv[k].K);
    p.reg[b] = (float)r;
  }
  const float* in = p.img + (size_t)b * 3 * HW;
  float* out = p.out + (size_t)b * 3 * HW;
  for (int i = blockIdx.x * THREADS + threadIdx.x; i < HW; i += gridDim.x * THREADS) {
    const float x[3] = {__ldcs(in + i), __ldcs(in + HW + i), __ldcs(in + 2 * (size_t)HW + i)};
    const float m = mask1<MK>(p.mask, (size_t)b * HW + i);
    float y[3];
    seg_layer_fwd<MK>(x, m, cv, y);
#pragma unroll
    for (int c = 0; c < 3; ++c) __stcs(out + (size_t)c * HW + i, y[c]);
  }
}

// Segment bins of one WARP (a private copy in shared memory), updated WITHOUT atomics and in an order that depends only
// on which lane holds which pixel -- the knot gradients are bit-reproducible from run to run (round 1 used
// shared-memory float atomics, whose order is not).  add() is called by all 32 lanes together.  Lanes that hit the
// same bin (match_any) are ranked by lane number; then either
//   * the groups are small (random content: 32 lanes over K - 1 bins): rank r of every group adds in round r --
//     within a round each bin has at most one writer -- max-group-size rounds of one plain read-modify-write, or
//   * a few bins hold most lanes (smooth content): one butterfly reduction per distinct bin, written by its first lane.
// Both orders are fixed by the lane numbers; which one runs (the cheaper by instruction count) depends on the data only.
struct WarpBins {
  float2* AB;        // this warp's histogram: {A_j, B_j} per bin, one 64-bit read-modify-write per update
  const int* off;
  int lane;
#ifndef SEG_ROUND_COST
#define SEG_ROUND_COST 8
#endif
#ifndef SEG_LEADER_COST
#define SEG_LEADER_COST 26
#endif
  static constexpr int ROUND_COST = SEG_ROUND_COST, LEADER_COST = SEG_LEADER_COST;
  __device__ __forceinline__ void add(int curve, int j, float a, float b) {
    constexpr unsigned FULL = 0xffffffffu;
    const bool active = (a != 0.0f) || (b != 0.0f);
    if (!__any_sync(FULL, active)) return;
    const int idx = off[curve] + j;
    const unsigned peers = __match_any_sync(FULL, active ? idx : 0x10000 + lane);   // idle lanes: groups of one
    const int rank = __popc(peers & ((1u << lane) - 1u));
    const int maxg = __reduce_max_sync(FULL, active ? __popc(peers) : 0);
    const unsigned leaders = __ballot_sync(FULL, active && rank == 0);          // the first lane of every group
    // a round costs about 8 instructions, a butterfly reduction about 26: take the cheaper order
    if (maxg * ROUND_COST <= __popc(leaders) * LEADER_COST) {
      for (int r = 0; r < maxg; ++r) {
        if (active && rank == r) { float2 v = AB[idx]; v.x += a; v.y += b; AB[idx] = v; }
        __syncwarp();
      }
    } else {
      unsigned todo = leaders;
      while (todo) {
        const int leader = __ffs(todo) - 1;
        const int key = __shfl_sync(FULL, idx, leader);
        const bool mine = active && idx == key;
        float va = mine ? a : 0.0f, vb = mine ? b : 0.0f;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) { va += __shfl_xor_sync(FULL, va, o); vb += __shfl_xor_sync(FULL, vb, o); }
        if (lane == leader) { float2 v = AB[key]; v.x += va; v.y += vb; AB[key] = v; }
        todo &= todo - 1;
      }
      __syncwarp();
    }
  }
};

template <int MK, bool WANT_GIN>
__global__ void __launch_bounds__(THREADS) seg_bwd_kernel(const SParams p) {
  extern __shared__ __align__(16) float smem[];
  __shared__ int s_off[LAYER_CURVES];
  __shared__ unsigned s_ticket;
  const int NK = total_knots(p.K), NB = 2 * NK;
  double* s_bins = reinterpret_cast<double*>(smem);          // [NB], used by the image's last block
  float* s_hist = smem + 2 * NB;                             // [NWARPS][NK] {A, B} pairs (8-byte aligned for any NK)
  float* s_C = s_hist + (size_t)NWARPS * NB;                 // [NK]
  const int NH = NWARPS;
  const int b = blockIdx.y, HW = p.HW;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  SegCurve cv[LAYER_CURVES];
  for (int i = threadIdx.x; i < NH * NB; i += THREADS) s_hist[i] = 0.0f;
  stage(p, b, s_C, s_off, cv);                               // ends with a barrier
  WarpBins bins{reinterpret_cast<float2*>(s_hist + (size_t)warp * NB), s_off, lane};
  const float* in = p.img + (size_t)b * 3 * HW;
  const float* go = p.gout + (size_t)b * 3 * HW;
  float* gin = WANT_GIN ? p.gimg + (size_t)b * 3 * HW : nullptr;
  // the loop is warp-uniform (bins.add is a warp-wide operation): lanes past the end work on a grey pixel with a
  // zero upstream gradient, which contributes nothing and is not stored
  for (int i0 = blockIdx.x * THREADS + warp * 32; i0 < HW; i0 += gridDim.x * THREADS) {
    const int i = i0 + lane;
    const bool live = i < HW;
    float x[3] = {0.5f, 0.5f, 0.5f}, g[3] = {0.0f, 0.0f, 0.0f}, m = 1.0f;
    if (live) {
#pragma unroll
      for (int c = 0; c < 3; ++c) { x[c] = __ldcs(in + (size_t)c * HW + i); g[c] = __ldcs(go + (size_t)c * HW + i); }
      m = mask1<MK>(p.mask, (size_t)b * HW + i);
    }
    float gx[3];
    seg_layer_bwd<MK, WANT_GIN>(x, m, cv, g, gx, bins);
    if (WANT_GIN && live) {
#pragma unroll
      for (int c = 0; c < 3; ++c) __stcs(gin + (size_t)c * HW + i, gx[c]);
    }
  }
  __syncthreads();
  // one partial row per block (fixed order over the warps), one ticket per block
  float* row = p.partials + ((size_t)b * gridDim.x + blockIdx.x) * NB;
  for (int k = threadIdx.x; k < NB; k += THREADS) {
    float v = 0.0f;
    const int at = k < NK ? 2 * k : 2 * (k - NK) + 1;          // row: A[NK] then B[NK]; histograms: {A, B} pairs
    for (int w = 0; w < NH; ++w) v += s_hist[w * NB + at];
    row[k] = v;
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) s_ticket = atomicAdd(&p.tickets[b], 1u);
  __syncthreads();
  if (s_ticket != gridDim.x - 1) return;
  __threadfence();
  // last block of image b: fold the rows in a fixed order in double, then expand to knot gradients
  const float* rows = p.partials + (size_t)b * gridDim.x * NB;
  for (int k = warp; k < NB; k += NWARPS) {
    double acc = 0.0;
    for (unsigned r = lane; r < gridDim.x; r += 32) acc += (double)__ldcg(rows + (size_t)r * NB + k);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (lane == 0) s_bins[k] = acc;
  }
  __syncthreads();
  const double greg = p.greg ? (double)p.greg[b] : 0.0;
#pragma unroll
  for (int k = 0; k < LAYER_CURVES; ++k) {
    const int g = k < 3 ? 0 : (k < 6 ? 1 : 2), ci = k - (g == 0 ? 0 : (g == 1 ? 3 : 6));
    const int K = p.K[g];
    float* dst = p.gk[g] + (size_t)b * (g == 2 ? 4 : 3) * K + (size_t)ci * K;
    for (int i = threadIdx.x; i < K; i += THREADS)
      dst[i] = seg_knot_grad(cv[k].C, K, i, s_bins + s_off[k], s_bins + NK + s_off[k], greg, true);
  }
}

// ------------------------------------------------------------------------------------------------
// Host side
// ------------------------------------------------------------------------------------------------
static bool aligned4(const void* q) { return (reinterpret_cast<uintptr_t>(q) & 3u) == 0; }
static bool aligned16(const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15u) == 0; }

static int grid_x(int B, int HW, int blocks_per_sm) {
  const long long want = ((long long)sm_count() * blocks_per_sm + B - 1) / B;
  const long long cap = ((long long)HW + THREADS - 1) / THREADS;
  long long g = want < cap ? want : cap;
  return (int)(g < 1 ? 1 : g);
}
static size_t tick_bytes(int B) { return ((size_t)B * sizeof(unsigned) + 255) / 256 * 256; }
static size_t fwd_smem(int NK) { return (size_t)NK * sizeof(float); }
static size_t bwd_smem(int NK) { return ((size_t)4 * NK + NK + (size_t)NWARPS * 2 * NK) * sizeof(float); }
constexpr int BWD_BLOCKS_PER_SM = 4;

static int fill(SParams& p, const float* img, const void* mask, int mask_kind, const float* L, int64_t ldL,
                const float* R, int64_t ldR, const float* Hk, int64_t ldH, int B, int H, int W, int K_lab, int K_rgb,
                int K_hsv) {
  if (B <= 0 || H <= 0 || W <= 0 || (long long)H * W > 0x3fffffffLL || B > 65535) return CURL_ERR_BAD_ARG;
  if (!img || !L || !R || !Hk) return CURL_ERR_BAD_ARG;
  if (K_lab < 3 || K_rgb < 3 || K_hsv < 3) return CURL_ERR_BAD_ARG;
  if (mask_kind < CURL_MASK_NONE || mask_kind > CURL_MASK_F32) return CURL_ERR_BAD_ARG;
  if ((mask_kind != CURL_MASK_NONE) != (mask != nullptr)) return CURL_ERR_BAD_ARG;
  if (ldL < 3 * (int64_t)K_lab || ldR < 3 * (int64_t)K_rgb || ldH < 4 * (int64_t)K_hsv) return CURL_ERR_BAD_ARG;
  if (3 * K_lab + 3 * K_rgb + 4 * K_hsv > CURL_MAX_KNOTS) return CURL_ERR_TOO_MANY_KNOTS;
  if (!aligned4(img) || !aligned4(L) || !aligned4(R) || !aligned4(Hk)) return CURL_ERR_MISALIGNED;
  if (mask_kind == CURL_MASK_F32 && !aligned4(mask)) return CURL_ERR_MISALIGNED;
  p = SParams{};
  p.img = img; p.mask = mask; p.HW = H * W;
  p.knots[0] = L; p.knots[1] = R; p.knots[2] = Hk;
  p.ld[0] = ldL; p.ld[1] = ldR; p.ld[2] = ldH;
  p.K[0] = K_lab; p.K[1] = K_rgb; p.K[2] = K_hsv;
  return CURL_OK;
}

template <int MK>
static int launch_bwd(SParams& p, int B, cudaStream_t st) {
  const int NK = total_knots(p.K);
  const dim3 grid(grid_x(B, p.HW, BWD_BLOCKS_PER_SM), B);
  auto kern = p.gimg ? seg_bwd_kernel<MK, true> : seg_bwd_kernel<MK, false>;
  const bool ok = p.gimg ? host::ensure_dyn_smem<seg_bwd_kernel<MK, true>>(bwd_smem(NK))
                         : host::ensure_dyn_smem<seg_bwd_kernel<MK, false>>(bwd_smem(NK));
  if (!ok) return CURL_ERR_LAUNCH;
  kern<<<grid, THREADS, bwd_smem(NK), st>>>(p);
  return cudaGetLastError() == cudaSuccess ? CURL_OK : CURL_ERR_LAUNCH;
}

}  // namespace seg
}  // namespace curl

using namespace curl;
using namespace curl::seg;

extern "C" {

int curl_layer_seg_fwd_f32(const float* img, const void* mask, int mask_kind, const float* L, int64_t ldL,
                           const float* R, int64_t ldR, const float* Hk, int64_t ldH, float* out, float* reg, int B,
                           int H, int W, int K_lab, int K_rgb, int K_hsv, void* stream) {
  SParams p;
  int rc = fill(p, img, mask, mask_kind, L, ldL, R, ldR, Hk, ldH, B, H, W, K_lab, K_rgb, K_hsv);
  if (rc != CURL_OK) return rc;
  if (!out || !reg) return CURL_ERR_BAD_ARG;
  if (!aligned4(out)) return CURL_ERR_MISALIGNED;
  p.out = out; p.reg = reg;
  const dim3 grid(grid_x(B, p.HW, 8), B);
  const size_t sm = fwd_smem(total_knots(p.K));
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  switch (mask_kind) {
    case CURL_MASK_NONE: seg_fwd_kernel<MASK_NONE><<<grid, THREADS, sm, st>>>(p); break;
    case CURL_MASK_U8: seg_fwd_kernel<MASK_U8><<<grid, THREADS, sm, st>>>(p); break;
    default: seg_fwd_kernel<MASK_F32><<<grid, THREADS, sm, st>>>(p); break;
  }
  return cudaGetLastError() == cudaSuccess ? CURL_OK : CURL_ERR_LAUNCH;
}

size_t curl_layer_seg_bwd_workspace_bytes(int B, int H, int W, int K_lab, int K_rgb, int K_hsv) {
  if (B <= 0 || H <= 0 || W <= 0 || K_lab < 3 || K_rgb < 3 || K_hsv < 3) return 0;
  const int K[3] = {K_lab, K_rgb, K_hsv};
  return tick_bytes(B) + (size_t)B * grid_x(B, H * W, BWD_BLOCKS_PER_SM) * 2 * total_knots(K) * sizeof(float);
}

int curl_layer_seg_bwd_f32(const float* img, const void* mask, int mask_kind, const float* L, int64_t ldL,
                           const float* R, int64_t ldR, const float* Hk, int64_t ldH, const float* gout,
                           const float* greg, float* gimg, float* gL, float* gR, float* gH, void* workspace,
                           size_t workspace_bytes, int B, int H, int W, int K_lab, int K_rgb, int K_hsv, void* stream) {
  SParams p;
  int rc = fill(p, img, mask, mask_kind, L, ldL, R, ldR, Hk, ldH, B, H, W, K_lab, K_rgb, K_hsv);
  if (rc != CURL_OK) return rc;
  if (!gout || !gL || !gR || !gH) return CURL_ERR_BAD_ARG;
  if (!aligned4(gout) || !aligned4(gimg) || !aligned4(greg)) return CURL_ERR_MISALIGNED;
  const size_t need = curl_layer_seg_bwd_workspace_bytes(B, H, W, K_lab, K_rgb, K_hsv);
  if (!workspace || workspace_bytes < need || !aligned16(workspace)) return CURL_ERR_WORKSPACE;
  p.gout = gout; p.greg = greg; p.gimg = gimg; p.gk[0] = gL; p.gk[1] = gR; p.gk[2] = gH;
  p.tickets = reinterpret_cast<unsigned*>(workspace);
  p.partials = reinterpret_cast<float*>(reinterpret_cast<char*>(workspace) + tick_bytes(B));
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (cudaMemsetAsync(workspace, 0, tick_bytes(B), st) != cudaSuccess) return CURL_ERR_LAUNCH;
  switch (mask_kind) {
    case CURL_MASK_NONE: return launch_bwd<MASK_NONE>(p, B, st);
    case CURL_MASK_U8: return launch_bwd<MASK_U8>(p, B, st);
    default: return launch_bwd<MASK_F32>(p, B, st);
  }
}

}  // extern "C"
