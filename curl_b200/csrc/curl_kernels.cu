// sm_100a kernels + C ABI of the CURL curve layer (see include/curl_b200.h for the contract).
//
// One generic forward kernel and one generic recompute-backward kernel, instantiated per "op":
// the fused layer (reference model.py:137-176), the four colour conversions (colors.py) and the
// three curve groups (curves.py).  Data layout: planar NCHW fp32; one block never straddles images
// (blockIdx.y = image), so the per-image curve coefficients live in shared memory.  The path is
// elementwise: each pixel is read once and written once with 128-bit accesses, and nothing else
// touches HBM (no Lab/HSV intermediates).
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/curl_b200.h"
#include "curl_pixel.cuh"
#include "curl_host.cuh"

namespace curl {

#ifndef CURL_THREADS
#define CURL_THREADS 128   // forward: 128 x 64 blocks per SM measured 1.7-2 % faster than 256 x 16 at 4K (3.215 -> 3.160 ms at batch 64)
#endif
constexpr int THREADS = CURL_THREADS;
// Backward (aligned path) launch shape: three independent 128-thread blocks per SM (144 registers
// each) instead of one 384-thread block.  Same 12 warps per SM, but each scheduler holds one warp of
// each block, the per-stage barrier spans 4 warps instead of 12 and the blocks drift apart, so their
// MUFU / gate / FMA phases interleave: 4K x 16 backward 2.357 -> 2.165 ms; with several sub-tiles per
// stage (BWD_PAIRS, one barrier + one refill per 768 px, two stages) 2.13 ms.  Measured alternatives, ms:
// 192x2 2.343, 96x4 2.377, 64x6 2.234, 32x12 2.317, 128x4 at 128 registers (BWD_SCALAR_MOM) 2.121.
#ifndef BWD_THREADS
#define BWD_THREADS 128
#endif
#ifndef BWD_SPILL
#define BWD_SPILL (6 * CURL_PARK_STAGES)
#endif
// Shipped configuration of the aligned backward: Jac records AND the 20 moment accumulators parked in tensor
// memory (TmemSpill below), which brings the kernel to 128 registers = four 128-thread blocks per SM.
// -DCURL_PARK_SMEM selects the round-1 configuration (records in shared memory, moments in registers, 3 blocks).
#if !defined(CURL_PARK_SMEM) && !defined(CURL_PARK_TMEM)
#define CURL_PARK_TMEM
#define CURL_PARK_MOM
#endif
#ifndef BWD_MIN_BLOCKS
#ifdef CURL_PARK_MOM
#define BWD_MIN_BLOCKS 4
#else
#define BWD_MIN_BLOCKS 3
#endif
#endif
constexpr int MAX_SEGS = 3;
constexpr int MAX_CURVES = 10;

struct KnotSeg {
  const float* p;   // [B][ld] log-domain knots
  float* g;         // [B][nc*K] dense gradient (backward only)
  long long ld;
  int nc, K;
};

struct KParams {
  const float* in;     // [B][3][HW]
  const void* mask;    // [B][HW] u8 or f32, or null
  const float* gout;   // backward: upstream gradient [B][3][HW]
  float* out;          // forward: result; backward: gradient w.r.t. `in` (nullable)
  float* reg;          // forward: [B]
  const float* greg;   // backward: [B] or null
  float* partials;     // backward: [B][gridDim.x][2*NC]
  unsigned* tickets;   // backward: [B], zero on entry
  KnotSeg seg[MAX_SEGS];
  int nseg;
  int HW;
  int linear_knots;    // 1: seg[].p holds C itself (apply_curve), 0: log-domain (C = exp(p))
  int aux;             // op-specific: apply_curve packs channel_in | channel_out << 2
  const unsigned char* in8;   // 8-bit forward: image in, CHW or HWC
  unsigned char* out8;        //                image out
};

// ---------------------------------------------------------------------------------------------
// Ops.  T is the lane type: F2 (two pixels per value, packed FFMA2/FMUL2/FADD2) on the aligned
// path, float on the unaligned tail path.
// ---------------------------------------------------------------------------------------------
// FOLD: the op's reverse sweep works in the scaled units of curl_pixel.cuh ("Units of the fused layer's reverse
// sweep"): the moments are rescaled after the reduction (curve_fold_m) and s_ab[2 NC] holds layer_beta1.
struct OpLayer {
  static constexpr int NC = LAYER_CURVES;
  static constexpr bool FOLD = true;
  template <int MK, class T>
  static __device__ __forceinline__ void fwd(const T x[3], T m, const float* ab, int, T y[3]) {
    layer_fwd<T, MK>(x, m, ab, y);
  }
  // The Jac records of the first four stages (24 packed values) are parked in shared memory during
  // the turn of the reverse sweep: 192 -> 148 registers, which lets 12 warps share an SM.
  static constexpr int SPILL = BWD_SPILL;
  template <int MK, bool WANT_GX, class T, class SP>
  static __device__ __forceinline__ void bwd(const T x[3], T m, const float* ab, int, const T gy[3], T gx[3], T* mom,
                                             const SP& sp) {
    layer_bwd<T, MK, WANT_GX, SP>(x, m, ab, gy, gx, mom, sp);
  }
};

template <int CONV>
struct OpColor {
  static constexpr int NC = 0;
  static constexpr bool FOLD = false;
  template <int MK, class T>
  static __device__ __forceinline__ void fwd(const T x[3], T, const float*, int, T y[3]) {
    if (CONV == CURL_RGB2LAB) rgb2lab_fwd<T, false>(x, y, nullptr);
    if (CONV == CURL_LAB2RGB) lab2rgb_fwd<T, false>(x, y, nullptr);
    if (CONV == CURL_RGB2HSV) rgb2hsv_fwd<T, false, false>(x, y, nullptr);
    if (CONV == CURL_HSV2RGB) hsv2rgb_fwd<T, false, false>(x, y, nullptr);
  }
  static constexpr int SPILL = 0;
  template <int MK, bool WANT_GX, class T, class SP>
  static __device__ __forceinline__ void bwd(const T x[3], T, const float*, int, const T gy[3], T gx[3], T*, const SP&) {
    T y[3];
    if (CONV == CURL_RGB2LAB) { LabJac<T> j; rgb2lab_fwd<T, true>(x, y, &j); rgb2lab_bwd(j, gy, gx); }
    if (CONV == CURL_LAB2RGB) { RgbJac<T> j; lab2rgb_fwd<T, true>(x, y, &j); lab2rgb_bwd(j, gy, gx); }
    if (CONV == CURL_RGB2HSV) { HsvJac<T> j; rgb2hsv_fwd<T, true, false>(x, y, &j); rgb2hsv_bwd(j, gy, gx); }
    if (CONV == CURL_HSV2RGB) { Hsv2RgbJac<T> j; hsv2rgb_fwd<T, true, false>(x, y, &j); hsv2rgb_bwd<T, false>(j, gy, gx); }
  }
};

struct OpAdjustSelf3 {  // adjust_lab, adjust_rgb
  static constexpr int NC = 3;
  static constexpr bool FOLD = false;
  template <int MK, class T>
  static __device__ __forceinline__ void fwd(const T x[3], T, const float* ab, int, T y[3]) {
    adjust_self3_fwd<T, false>(x, ab, y, nullptr);
  }
  static constexpr int SPILL = 0;
  template <int MK, bool WANT_GX, class T, class SP>
  static __device__ __forceinline__ void bwd(const T x[3], T, const float* ab, int, const T gy[3], T gx[3], T* mom,
                                             const SP&) {
    T y[3]; Self3Jac<T> j;
    adjust_self3_fwd<T, true>(x, ab, y, &j);
    adjust_self3_bwd(j, gy, gx, mom);
  }
};

struct OpAdjustHsv {
  static constexpr int NC = 4;
  static constexpr bool FOLD = false;
  template <int MK, class T>
  static __device__ __forceinline__ void fwd(const T x[3], T, const float* ab, int, T y[3]) {
    adjust_hsv_fwd<T, false, false>(x, ab, y, nullptr);
  }
  static constexpr int SPILL = 0;
  template <int MK, bool WANT_GX, class T, class SP>
  static __device__ __forceinline__ void bwd(const T x[3], T, const float* ab, int, const T gy[3], T gx[3], T* mom,
                                             const SP&) {
    T y[3]; HsvCurveJac<T> j;
    adjust_hsv_fwd<T, true, false>(x, ab, y, &j);
    adjust_hsv_bwd(j, ab[3], gy, gx, mom);
  }
};

struct OpCurve1 {  // curves.apply_curve: one curve channel_in -> channel_out, knots given as C
  static constexpr int NC = 1;
  static constexpr bool FOLD = false;
  template <int MK, class T>
  static __device__ __forceinline__ void fwd(const T x[3], T, const float* ab, int aux, T y[3]) {
    single_curve_fwd<T>(x, aux & 3, (aux >> 2) & 3, ab[0], ab[1], y);
  }
  static constexpr int SPILL = 0;
  template <int MK, bool WANT_GX, class T, class SP>
  static __device__ __forceinline__ void bwd(const T x[3], T, const float* ab, int aux, const T gy[3], T gx[3], T* mom,
                                             const SP&) {
    single_curve_bwd<T>(x, aux & 3, (aux >> 2) & 3, ab[0], ab[1], gy, gx, mom);
  }
};

// ---------------------------------------------------------------------------------------------
// Pixel groups: what one thread handles per loop iteration (streaming 128-bit accesses: every
// byte is touched exactly once).  Group<4>: four consecutive pixels as two packed F2 lanes.
// ---------------------------------------------------------------------------------------------
template <int VEC> struct Group;
template <> struct Group<4> {
  using T = F2;
  static constexpr int LANES = 2;   // values of type T per channel
  T v[2];
  static __device__ __forceinline__ Group ld(const float* p) {
    const float4 t = __ldcs(reinterpret_cast<const float4*>(p));
    Group r; r.v[0] = F2(t.x, t.y); r.v[1] = F2(t.z, t.w); return r;
  }
  __device__ __forceinline__ void st(float* p) const {
    __stcs(reinterpret_cast<float4*>(p), make_float4(v[0].lo(), v[0].hi(), v[1].lo(), v[1].hi()));
  }
};
template <> struct Group<2> {   // two consecutive pixels: one packed lane, 64-bit accesses
  using T = F2;
  static constexpr int LANES = 1;
  T v[1];
  static __device__ __forceinline__ Group ld(const float* p) {
    const float2 t = __ldcs(reinterpret_cast<const float2*>(p));
    Group r; r.v[0] = F2(t.x, t.y); return r;
  }
  __device__ __forceinline__ void st(float* p) const {
    __stcs(reinterpret_cast<float2*>(p), make_float2(v[0].lo(), v[0].hi()));
  }
};
template <> struct Group<1> {
  using T = float;
  static constexpr int LANES = 1;
  T v[1];
  static __device__ __forceinline__ Group ld(const float* p) { Group r; r.v[0] = __ldcs(p); return r; }
  __device__ __forceinline__ void st(float* p) const { __stcs(p, v[0]); }
};

template <int MK, int VEC>
__device__ __forceinline__ Group<VEC> load_mask(const void* mask, size_t idx) {
  Group<VEC> g;
  if (MK == MASK_NONE) {
#pragma unroll
    for (int i = 0; i < Group<VEC>::LANES; ++i) g.v[i] = typename Group<VEC>::T(1.0f);
  } else if (MK == MASK_U8) {
    const unsigned char* p = reinterpret_cast<const unsigned char*>(mask) + idx;
    if constexpr (VEC == 4) {
      const uchar4 t = __ldcs(reinterpret_cast<const uchar4*>(p));
      g.v[0] = F2(t.x ? 1.0f : 0.0f, t.y ? 1.0f : 0.0f);
      g.v[1] = F2(t.z ? 1.0f : 0.0f, t.w ? 1.0f : 0.0f);
    } else if constexpr (VEC == 2) {
      const uchar2 t = __ldcs(reinterpret_cast<const uchar2*>(p));
      g.v[0] = F2(t.x ? 1.0f : 0.0f, t.y ? 1.0f : 0.0f);
    } else {
      g.v[0] = __ldcs(p) ? 1.0f : 0.0f;
    }
  } else {
    g = Group<VEC>::ld(reinterpret_cast<const float*>(mask) + idx);
  }
  return g;
}

// ---------------------------------------------------------------------------------------------
// Block prologue: exp() the knots of image b into shared memory, derive (alpha, beta) per curve
// (and, for the block that owns it, the regulariser).  curves.py:19-24,53-54,105-106,152-153.
// ---------------------------------------------------------------------------------------------
// BWD (backward kernels of a FOLD op): s_ab[2 NC] = layer_beta1(beta of the hue -> sat curve, curve 7).
template <class Op, bool BWD>
__device__ __forceinline__ void stage_curves(const KParams& p, int b, float* s_C, float* s_ab, double* s_reg,
                                             bool want_reg) {
  constexpr int NC = Op::NC;
  int off = 0;
  for (int s = 0; s < p.nseg; ++s) {
    const int n = p.seg[s].nc * p.seg[s].K;
    const float* src = p.seg[s].p + (size_t)b * p.seg[s].ld;
    for (int i = threadIdx.x; i < n; i += blockDim.x)
      s_C[off + i] = p.linear_knots ? src[i] : (float)exp((double)src[i]);
    off += n;
  }
  __syncthreads();
  if (threadIdx.x < NC) {
    int j = threadIdx.x, o = 0, K = 0;
    for (int s = 0; s < p.nseg; ++s) {
      K = p.seg[s].K;
      if (j < p.seg[s].nc) { o += j * K; break; }
      j -= p.seg[s].nc;
      o += p.seg[s].nc * K;
    }
    curve_alpha_beta(s_C + o, K, &s_ab[2 * threadIdx.x], &s_ab[2 * threadIdx.x + 1]);
    if (BWD && Op::FOLD && threadIdx.x == 7) s_ab[2 * NC] = layer_beta1(s_ab[2 * 7 + 1]);
    if (want_reg) s_reg[threadIdx.x] = curve_regulariser(s_C + o, K);
  }
  __syncthreads();
}

// ---------------------------------------------------------------------------------------------
// Forward
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void l2_prefetch(const void* gptr, unsigned bytes) {   // bytes: multiple of 16, gptr 16-byte aligned
  asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(gptr), "r"(bytes) : "memory");
}
#ifdef FWD_MIN_BLOCKS
#define FWD_BOUNDS __launch_bounds__(THREADS, FWD_MIN_BLOCKS)
#else
#define FWD_BOUNDS __launch_bounds__(THREADS)
#endif
template <class Op, int MK, int VEC>
__global__ void FWD_BOUNDS fwd_kernel(const KParams p) {
  constexpr int NC = Op::NC;
  __shared__ float s_C[NC > 0 ? CURL_MAX_KNOTS : 1];
  __shared__ float s_ab[NC > 0 ? 2 * NC : 1];
  __shared__ double s_reg[NC > 0 ? NC : 1];
  const int b = blockIdx.y;
  if (NC > 0) {
    const bool owner = (blockIdx.x == 0) && (p.reg != nullptr);
    stage_curves<Op, false>(p, b, s_C, s_ab, s_reg, owner);
    if (owner && threadIdx.x == 0) {
      double r = 0.0;
      for (int j = 0; j < NC; ++j) r += s_reg[j];
      p.reg[b] = (float)r;
    }
  }
  const int HW = p.HW;
  const float* in = p.in + (size_t)b * 3 * HW;
  float* out = p.out + (size_t)b * 3 * HW;
  const int stride = gridDim.x * THREADS * VEC;
  using G = Group<VEC>;
  using T = typename G::T;
#ifdef FWD_PREFETCH
  if constexpr (VEC == 4 && Op::NC == LAYER_CURVES) {
    // The next group's loads are issued between the two lanes' arithmetic (pinned there by the fences: left alone
    // the compiler hoists them to the top of the iteration, where they are live across the whole chain and cost
    // occupancy -- the round-1 prefetch).  They are in flight for half an iteration of every resident warp.
    int i = (blockIdx.x * THREADS + threadIdx.x) * VEC;
    if (i >= HW) return;
    G c0 = G::ld(in + i), c1 = G::ld(in + HW + i), c2 = G::ld(in + 2 * (size_t)HW + i);
    G m = load_mask<MK, VEC>(p.mask, (size_t)b * HW + i);
    while (true) {
      G o0, o1, o2, n0, n1, n2, nm;
      {
        const T x[3] = {c0.v[0], c1.v[0], c2.v[0]};
        T y[3];
        Op::template fwd<MK, T>(x, m.v[0], s_ab, p.aux, y);
        o0.v[0] = y[0]; o1.v[0] = y[1]; o2.v[0] = y[2];
      }
      const int inext = i + stride;
      const bool more = inext < HW;
      asm volatile("" ::: "memory");
      if (more) {
        n0 = G::ld(in + inext); n1 = G::ld(in + HW + inext); n2 = G::ld(in + 2 * (size_t)HW + inext);
        nm = load_mask<MK, VEC>(p.mask, (size_t)b * HW + inext);
      }
      asm volatile("" ::: "memory");
      {
        const T x[3] = {c0.v[1], c1.v[1], c2.v[1]};
        T y[3];
        Op::template fwd<MK, T>(x, m.v[1], s_ab, p.aux, y);
        o0.v[1] = y[0]; o1.v[1] = y[1]; o2.v[1] = y[2];
      }
      o0.st(out + i); o1.st(out + HW + i); o2.st(out + 2 * (size_t)HW + i);
      if (!more) break;
      c0 = n0; c1 = n1; c2 = n2; m = nm; i = inext;
    }
    return;
  }
#endif
  for (int i = (blockIdx.x * THREADS + threadIdx.x) * VEC; i < HW; i += stride) {
    const G c0 = G::ld(in + i), c1 = G::ld(in + HW + i), c2 = G::ld(in + 2 * (size_t)HW + i);
    const G m = load_mask<MK, VEC>(p.mask, (size_t)b * HW + i);
#ifdef FWD_L2PF
    if constexpr (VEC == 4) {
      // the lines this warp reads in its NEXT iteration, asked into L2 now by one lane (no registers held, unlike a
      // register prefetch): 32 lanes x 16 B = 512 B per plane
      const int nb = i - (int)(threadIdx.x & 31) * VEC + stride;
      if ((threadIdx.x & 31) == 0 && nb < HW) {
        const int npx = min(32 * VEC, HW - nb) & ~15;
        if (npx > 0) {
          l2_prefetch(in + nb, npx * 4u); l2_prefetch(in + HW + nb, npx * 4u); l2_prefetch(in + 2 * (size_t)HW + nb, npx * 4u);
          if (MK == MASK_U8) l2_prefetch(reinterpret_cast<const unsigned char*>(p.mask) + (size_t)b * HW + nb, (unsigned)npx);
          if (MK == MASK_F32) l2_prefetch(reinterpret_cast<const float*>(p.mask) + (size_t)b * HW + nb, npx * 4u);
        }
      }
    }
#endif
    G o0, o1, o2;
#pragma unroll
    for (int v = 0; v < G::LANES; ++v) {
      const T x[3] = {c0.v[v], c1.v[v], c2.v[v]};
      T y[3];
      Op::template fwd<MK, T>(x, m.v[v], s_ab, p.aux, y);
      o0.v[v] = y[0]; o1.v[v] = y[1]; o2.v[v] = y[2];
    }
    o0.st(out + i); o1.st(out + HW + i); o2.st(out + 2 * (size_t)HW + i);
  }
}

// ---------------------------------------------------------------------------------------------
// Forward on 8-bit images (curl_layer_fwd_u8): u8 -> x/255 -> layer_fwd -> trunc(out*255) -> u8.
// 3 B/px in, 3 B/px out (+ the mask).  VEC == 4: four consecutive pixels per thread (32-bit
// accesses per plane, or three 32-bit words of an interleaved row); VEC == 1: any size/alignment.
// ---------------------------------------------------------------------------------------------
// An 8-bit channel has 256 possible values, so x = u8 / 255 (u8_to_unit) and its sRGB linearisation
// (srgb_linearise_fwd: 2 of the layer's 19 MUFU per channel and pixel on the kernel's busiest pipe) come from two
// 256-entry tables in shared memory, filled in the block prologue by the SAME device functions the fp32 kernel
// evaluates per pixel -- bit-identical by construction (tests: the 8-bit kernel must equal the quantised fp32 kernel
// exactly).  Per channel and pixel: one PRMT (the byte as an index) and two LDS.32 instead of PRMT + the 2^23 trick +
// u8_to_unit (3 packed) + the linearisation (3 packed, 2 MUFU, compare + select): ~19 instructions and 6 MUFU per
// pixel less.  The look-ups are data-dependent (bank conflicts happen); the LSU pipe is otherwise idle here.
__device__ __forceinline__ unsigned byte_at(unsigned w, int k) { return __byte_perm(w, 0u, 0x4440u + k); }

template <int MK, bool IN_HWC, bool OUT_HWC, int VEC>
__global__ void __launch_bounds__(THREADS) fwd_u8_kernel(const KParams p) {
  constexpr int NC = LAYER_CURVES;
  __shared__ float s_C[CURL_MAX_KNOTS];
  __shared__ float s_ab[2 * NC];
  __shared__ double s_reg[NC];
  __shared__ float s_unit[256], s_lin[256];
  const int b = blockIdx.y;
  for (int t = threadIdx.x; t < 256; t += THREADS) {
    const float x = u8_to_unit((float)t);
    s_unit[t] = x;
    s_lin[t] = srgb_linearise_fwd(x);
  }
  {
    const bool owner = (blockIdx.x == 0) && (p.reg != nullptr);
    stage_curves<OpLayer, false>(p, b, s_C, s_ab, s_reg, owner);      // ends with a barrier: the tables are visible too
    if (owner && threadIdx.x == 0) {
      double r = 0.0;
      for (int j = 0; j < NC; ++j) r += s_reg[j];
      p.reg[b] = (float)r;
    }
  }
  const int HW = p.HW;
  const unsigned char* in = p.in8 + (size_t)b * 3 * HW;
  unsigned char* out = p.out8 + (size_t)b * 3 * HW;
  const int stride = gridDim.x * THREADS * VEC;
  for (int i = (blockIdx.x * THREADS + threadIdx.x) * VEC; i < HW; i += stride) {
    if constexpr (VEC == 4) {
      unsigned ch[3][4];
      if (IN_HWC) {
        const unsigned* w = reinterpret_cast<const unsigned*>(in + 3 * (size_t)i);
        const unsigned w0 = __ldcs(w), w1 = __ldcs(w + 1), w2 = __ldcs(w + 2);   // r0 g0 b0 r1 | g1 b1 r2 g2 | b2 r3 g3 b3
        ch[0][0] = byte_at(w0, 0); ch[1][0] = byte_at(w0, 1); ch[2][0] = byte_at(w0, 2);
        ch[0][1] = byte_at(w0, 3); ch[1][1] = byte_at(w1, 0); ch[2][1] = byte_at(w1, 1);
        ch[0][2] = byte_at(w1, 2); ch[1][2] = byte_at(w1, 3); ch[2][2] = byte_at(w2, 0);
        ch[0][3] = byte_at(w2, 1); ch[1][3] = byte_at(w2, 2); ch[2][3] = byte_at(w2, 3);
      } else {
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          const unsigned w = __ldcs(reinterpret_cast<const unsigned*>(in + (size_t)c * HW + i));
#pragma unroll
          for (int k = 0; k < 4; ++k) ch[c][k] = byte_at(w, k);
        }
      }
      const Group<4> m = load_mask<MK, 4>(p.mask, (size_t)b * HW + i);
      unsigned o[3][4];
#pragma unroll
      for (int v = 0; v < 2; ++v) {
        F2 x[3], lin[3], y[3];
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          x[c] = F2(s_unit[ch[c][2 * v]], s_unit[ch[c][2 * v + 1]]);
          lin[c] = F2(s_lin[ch[c][2 * v]], s_lin[ch[c][2 * v + 1]]);
        }
        layer_fwd<F2, MK, true>(x, m.v[v], s_ab, y, lin);
#pragma unroll
        for (int c = 0; c < 3; ++c) { o[c][2 * v] = unit_to_u8(y[c].lo()); o[c][2 * v + 1] = unit_to_u8(y[c].hi()); }
      }
      if (OUT_HWC) {
        unsigned* w = reinterpret_cast<unsigned*>(out + 3 * (size_t)i);
        __stcs(w, o[0][0] | (o[1][0] << 8) | (o[2][0] << 16) | (o[0][1] << 24));
        __stcs(w + 1, o[1][1] | (o[2][1] << 8) | (o[0][2] << 16) | (o[1][2] << 24));
        __stcs(w + 2, o[2][2] | (o[0][3] << 8) | (o[1][3] << 16) | (o[2][3] << 24));
      } else {
#pragma unroll
        for (int c = 0; c < 3; ++c)
          __stcs(reinterpret_cast<unsigned*>(out + (size_t)c * HW + i),
                 o[c][0] | (o[c][1] << 8) | (o[c][2] << 16) | (o[c][3] << 24));
      }
    } else {
      float x[3], lin[3], y[3];
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const unsigned v8 = __ldcs(in + (IN_HWC ? 3 * (size_t)i + c : (size_t)c * HW + i));
        x[c] = s_unit[v8]; lin[c] = s_lin[v8];
      }
      const Group<1> m = load_mask<MK, 1>(p.mask, (size_t)b * HW + i);
      layer_fwd<float, MK, true>(x, m.v[0], s_ab, y, lin);
#pragma unroll
      for (int c = 0; c < 3; ++c)
        __stcs(out + (OUT_HWC ? 3 * (size_t)i + c : (size_t)c * HW + i), (unsigned char)unit_to_u8(y[c]));
    }
  }
}

// ---------------------------------------------------------------------------------------------
// Knot-gradient reduction shared by the backward kernels: per-thread registers -> warp shuffles ->
// shared memory -> one partial row per block -> ONE atomic ticket per block -> the last block of the
// image folds all partial rows in a fixed order (in double, bitwise deterministic) and expands the
// 2 moments per curve into the knot gradients (curve_knot_grad).
// ---------------------------------------------------------------------------------------------
template <class Op, class T, int NT>
__device__ __forceinline__ void finish_knot_grads(const KParams& p, int b, const T* mom, const float* s_C,
                                                  double* s_mom, float* s_part, unsigned* s_ticket_p) {
  constexpr int NC = Op::NC;
  constexpr int NM = 2 * NC;
  constexpr int NWARPS = NT / 32;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < NM; ++k) {
    float v = vhsum(mom[k]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) s_part[warp * NM + k] = v;
  }
  __syncthreads();
  float* row = p.partials + ((size_t)b * gridDim.x + blockIdx.x) * NM;
  if (threadIdx.x < NM) {
    float v = 0.0f;
#pragma unroll
    for (int w = 0; w < NWARPS; ++w) v += s_part[w * NM + threadIdx.x];
    row[threadIdx.x] = v;
    __threadfence();
  }
  __syncthreads();
  if (threadIdx.x == 0) *s_ticket_p = atomicAdd(&p.tickets[b], 1u);
  __syncthreads();
  if (*s_ticket_p != gridDim.x - 1) return;
  __threadfence();
  // last block of image b: fixed-order fold of the partial rows
  const float* rows = p.partials + (size_t)b * gridDim.x * NM;
  for (int k = warp; k < NM; k += NWARPS) {
    double acc = 0.0;
    for (unsigned r = lane; r < gridDim.x; r += 32) acc += (double)__ldcg(rows + (size_t)r * NM + k);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (lane == 0) s_mom[k] = acc * curve_fold_m(Op::FOLD, k >> 1);   // back from the folded units of the reverse sweep
  }
  __syncthreads();
  const double greg = p.greg ? (double)p.greg[b] : 0.0;
  int cbase = 0, off = 0;
  for (int s = 0; s < p.nseg; ++s) {
    const int K = p.seg[s].K, n = p.seg[s].nc * K;
    float* g = p.seg[s].g + (size_t)b * n;
    for (int i = threadIdx.x; i < n; i += NT) {
      const int c = i / K, kk = i - c * K;
      g[i] = curve_knot_grad(s_C + off + c * K, K, kk, s_mom[2 * (cbase + c)], s_mom[2 * (cbase + c) + 1], greg,
                             p.linear_knots == 0);
    }
    cbase += p.seg[s].nc;
    off += n;
  }
}

// ---------------------------------------------------------------------------------------------
// Backward (recompute).  Knot-gradient reduction: per-thread registers -> warp shuffles -> shared
// memory -> one partial row per block -> one atomic ticket per block -> the last block of the
// image folds all partial rows in a fixed order (in double) and writes the knot gradients.
// ---------------------------------------------------------------------------------------------
template <class Op, int MK, int VEC, bool WANT_GIN>
__global__ void __launch_bounds__(THREADS) bwd_kernel(const KParams p) {
  constexpr int NC = Op::NC;
  constexpr int NM = 2 * NC;
  constexpr int NWARPS = THREADS / 32;
  __shared__ float s_C[NC > 0 ? CURL_MAX_KNOTS : 1];
  __shared__ float s_ab[NC > 0 ? 2 * NC + 1 : 1];
  __shared__ double s_mom[NC > 0 ? NM : 1];
  __shared__ float s_part[NC > 0 ? NWARPS * NM : 1];
  __shared__ unsigned s_ticket;
  const int b = blockIdx.y;
  if (NC > 0) stage_curves<Op, true>(p, b, s_C, s_ab, nullptr, false);

  using G = Group<VEC>;
  using T = typename G::T;
  T mom[NM > 0 ? NM : 1];
#pragma unroll
  for (int k = 0; k < NM; ++k) mom[k] = T(0.0f);

  const int HW = p.HW;
  const float* in = p.in + (size_t)b * 3 * HW;
  const float* go = p.gout + (size_t)b * 3 * HW;
  float* gin = WANT_GIN ? p.out + (size_t)b * 3 * HW : nullptr;
  const int stride = gridDim.x * THREADS * VEC;
  for (int i = (blockIdx.x * THREADS + threadIdx.x) * VEC; i < HW; i += stride) {
    const G c0 = G::ld(in + i), c1 = G::ld(in + HW + i), c2 = G::ld(in + 2 * (size_t)HW + i);
    const G g0 = G::ld(go + i), g1 = G::ld(go + HW + i), g2 = G::ld(go + 2 * (size_t)HW + i);
    G m = load_mask<MK, VEC>(p.mask, (size_t)b * HW + i);
    if constexpr (MK == MASK_U8 && VEC == 1) m.v[0] = bwd_mask_lane_u8(m.v[0] != 0.0f);   // gate offset (layer_bwd)
    static_assert(MK != MASK_U8 || VEC == 1, "the vector backward with a byte mask is bwd_tma_kernel");
    G o0, o1, o2;
#pragma unroll
    for (int v = 0; v < G::LANES; ++v) {
      const T x[3] = {c0.v[v], c1.v[v], c2.v[v]};
      const T gy[3] = {g0.v[v], g1.v[v], g2.v[v]};
      T gx[3];
      Op::template bwd<MK, WANT_GIN, T>(x, m.v[v], s_ab, p.aux, gy, gx, mom, KeepInRegisters());
      o0.v[v] = gx[0]; o1.v[v] = gx[1]; o2.v[v] = gx[2];
    }
    if (WANT_GIN) { o0.st(gin + i); o1.st(gin + HW + i); o2.st(gin + 2 * (size_t)HW + i); }
  }

  if (NC > 0) finish_knot_grads<Op, T, THREADS>(p, b, mom, s_C, s_mom, s_part, &s_ticket);
}

// ---------------------------------------------------------------------------------------------
// TMA (bulk async copy) helpers.  The backward kernel needs 144 registers per thread even with its
// Jac records parked in shared memory, so only 12 warps fit on an SM: too few to hide HBM latency
// with ordinary loads.  Instead one elected thread per block streams each tile's six input planes
// into a shared-memory ring with cp.async.bulk
// (SASS UBLKCP), completion tracked by an mbarrier per stage; the loads of the next STAGES-1 tiles
// are always in flight and cost no registers.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long* bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, unsigned parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, unsigned bytes, unsigned long long* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

#ifndef BWD_PAIRS
#define BWD_PAIRS 3
#endif
constexpr int TMA_SUB = BWD_THREADS * 2;        // pixels per sub-tile: one packed pair per thread
constexpr int TMA_TILE = TMA_SUB * BWD_PAIRS;   // pixels per stage: BWD_PAIRS sub-tiles, one barrier + one refill per stage
#ifndef TMA_STAGES_N
#define TMA_STAGES_N 2
#endif
constexpr int TMA_STAGES = TMA_STAGES_N;
constexpr int TMA_PLANES = 6;           // x0 x1 x2 g0 g1 g2
constexpr size_t TMA_RING_BYTES = (size_t)TMA_STAGES * TMA_PLANES * TMA_TILE * sizeof(float);
#ifdef CURL_PARK_TMEM
constexpr bool PARK_IN_TMEM = true;
#else
constexpr bool PARK_IN_TMEM = false;
#endif
template <class Op> constexpr size_t tma_smem_bytes() {
  return TMA_RING_BYTES + (PARK_IN_TMEM ? 0 : (size_t)Op::SPILL * BWD_THREADS * sizeof(F2));
}

// Shared-memory parking space for Jac records (curl_pixel.cuh "Spill policies"): slot-pair-major, one
// 16-byte column per thread and pair of slots, so a warp's STS.128/LDS.128 touches 512 consecutive bytes
// (conflict-free) and one instruction parks or fetches two packed values.
struct SmemSpill {
  static constexpr bool kActive = true;
  static constexpr bool kMomParked = false;
  unsigned base;   // shared-space byte address, already offset by threadIdx.x * 16
  // asm volatile: the compiler must not forward the stored value to the later load (that would keep
  // the value live in a register, which is exactly what parking is meant to avoid)
  // written as four 32-bit operands: with two 64-bit operands ptxas assembled each register quad with MOVs
  // (30 per pixel pair; backward 2.007 -> 1.951 ms at 4K x 16 without them)
  __device__ __forceinline__ void put2(int slot, F2 a, F2 b) const {
    asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(base + (slot >> 1) * (BWD_THREADS * 16)), "f"(a.lo()), "f"(a.hi()),
                 "f"(b.lo()), "f"(b.hi()) : "memory");
  }
  __device__ __forceinline__ void get2(int slot, F2& a, F2& b) const {
    float x, y, z, w;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(x), "=f"(y), "=f"(z), "=f"(w) : "r"(base + (slot >> 1) * (BWD_THREADS * 16)) : "memory");
    a = F2(x, y); b = F2(z, w);
  }
  __device__ __forceinline__ void before_loads() const {}     // a thread's own STS -> LDS are ordered
  template <class... A> __device__ __forceinline__ void put8(int, A...) const {}      // (tensor-memory parking only)
  template <class... A> __device__ __forceinline__ void get8(int, A&...) const {}
  __device__ __forceinline__ void put6(int o, F2 a, F2 b, F2 c, F2 d, F2 e, F2 f) const {
    put2(o, a, b); put2(o + 2, c, d); put2(o + 4, e, f);
  }
  __device__ __forceinline__ void get6(int o, F2& a, F2& b, F2& c, F2& d, F2& e, F2& f) const {
    get2(o, a, b); get2(o + 2, c, d); get2(o + 4, e, f);
  }
  // split fetch: the LDS results are tracked by the hardware scoreboard, nothing to wait for explicitly
  __device__ __forceinline__ void get6_issue(int o, F2& a, F2& b, F2& c, F2& d, F2& e, F2& f) const { get6(o, a, b, c, d, e, f); }
  __device__ __forceinline__ void get6_wait(F2&, F2&, F2&, F2&, F2&, F2&) const {}
};

// The same parking space in TENSOR MEMORY (-DCURL_PARK_TMEM): each thread owns one TMEM lane (shape 32x32b: lane
// i of a warp addresses TMEM lane 32 (warp % 4) + i) and parks a record of six packed values = 12 consecutive
// 32-bit columns with tcgen05.st (x8 + x4), fetching it with tcgen05.ld + tcgen05.wait::ld.  No shared memory, no
// LSU instructions, no bank conflicts; the block allocates TMEM_COLS columns for its lifetime.
#if defined(CURL_PARK_J5) || defined(CURL_PARK_MOM)
constexpr int TMEM_COLS = 128;      // records: columns 0..75 (24 + 14 slots of two columns); moments: 80..127
#else
constexpr int TMEM_COLS = 64;       // >= 2 * Op::SPILL (48), a power of two >= 32
#endif
#ifdef CURL_PARK_MOM
constexpr bool PARK_MOM = true;     // the 20 moment accumulators live in tensor memory between their updates
#else
constexpr bool PARK_MOM = false;
#endif
constexpr int TMEM_MOM_COL = 80;    // Lab 80..91, RGB 92..103, HSV 104..119 (two columns per accumulator)
struct TmemSpill {
  static constexpr bool kActive = true;
  static constexpr bool kMomParked = PARK_MOM;
  __device__ __forceinline__ void mom_get6(int k, F2* m) const { get6((TMEM_MOM_COL >> 1) + k, m[0], m[1], m[2], m[3], m[4], m[5]); }
  __device__ __forceinline__ void mom_put6(int k, const F2* m) const { put6((TMEM_MOM_COL >> 1) + k, m[0], m[1], m[2], m[3], m[4], m[5]); }
  __device__ __forceinline__ void mom_get8(F2* m) const { get8((TMEM_MOM_COL >> 1) + 12, m[0], m[1], m[2], m[3], m[4], m[5], m[6], m[7]); }
  __device__ __forceinline__ void mom_put8(const F2* m) const { put8((TMEM_MOM_COL >> 1) + 12, m[0], m[1], m[2], m[3], m[4], m[5], m[6], m[7]); }
  unsigned base;   // TMEM address: this warp's lane quarter in bits 31..16, the block's first column in bits 15..0
  // tcgen05.st is asynchronous: the stores of this pixel pair must have completed before they are read back
  __device__ __forceinline__ void before_loads() const { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
  __device__ __forceinline__ void put6(int o, F2 a, F2 b, F2 c, F2 d, F2 e, F2 f) const {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};\n"
        "tcgen05.st.sync.aligned.32x32b.x4.b32 [%0 + 8], {%9, %10, %11, %12};" ::"r"(base + 2 * o),
        "f"(a.lo()), "f"(a.hi()), "f"(b.lo()), "f"(b.hi()), "f"(c.lo()), "f"(c.hi()), "f"(d.lo()), "f"(d.hi()),
        "f"(e.lo()), "f"(e.hi()), "f"(f.lo()), "f"(f.hi()) : "memory");
  }
  // split fetch: the values are exchanged as raw 64-bit registers so that no instruction touches them between the
  // tcgen05.ld and the tcgen05.wait::ld (which takes them as in/out operands: nothing may be moved across it)
  __device__ __forceinline__ void get6_issue(int o, F2& a, F2& b, F2& c, F2& d, F2& e, F2& f) const {
    asm volatile(
        "{\n.reg .b32 t<12>;\n"
        "tcgen05.ld.sync.aligned.32x32b.x8.b32 {t0, t1, t2, t3, t4, t5, t6, t7}, [%6];\n"
        "tcgen05.ld.sync.aligned.32x32b.x4.b32 {t8, t9, t10, t11}, [%6 + 8];\n"
        "mov.b64 %0, {t0, t1};\n mov.b64 %1, {t2, t3};\n mov.b64 %2, {t4, t5};\n"
        "mov.b64 %3, {t6, t7};\n mov.b64 %4, {t8, t9};\n mov.b64 %5, {t10, t11};\n}"
        : "=l"(a.u), "=l"(b.u), "=l"(c.u), "=l"(d.u), "=l"(e.u), "=l"(f.u)
        : "r"(base + 2 * o) : "memory");
  }
  __device__ __forceinline__ void get6_wait(F2& a, F2& b, F2& c, F2& d, F2& e, F2& f) const {
    asm volatile("tcgen05.wait::ld.sync.aligned;" : "+l"(a.u), "+l"(b.u), "+l"(c.u), "+l"(d.u), "+l"(e.u), "+l"(f.u)::"memory");
  }
  __device__ __forceinline__ void put8(int o, F2 a, F2 b, F2 c, F2 d, F2 e, F2 f, F2 g, F2 h) const {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
        ::"r"(base + 2 * o), "f"(a.lo()), "f"(a.hi()), "f"(b.lo()), "f"(b.hi()), "f"(c.lo()), "f"(c.hi()), "f"(d.lo()), "f"(d.hi()),
        "f"(e.lo()), "f"(e.hi()), "f"(f.lo()), "f"(f.hi()), "f"(g.lo()), "f"(g.hi()), "f"(h.lo()), "f"(h.hi()) : "memory");
  }
  __device__ __forceinline__ void get8(int o, F2& a, F2& b, F2& c, F2& d, F2& e, F2& f, F2& g, F2& h) const {
    float v[16];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];\n"
        "tcgen05.wait::ld.sync.aligned;"
        : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7]), "=f"(v[8]),
          "=f"(v[9]), "=f"(v[10]), "=f"(v[11]), "=f"(v[12]), "=f"(v[13]), "=f"(v[14]), "=f"(v[15])
        : "r"(base + 2 * o) : "memory");
    a = F2(v[0], v[1]); b = F2(v[2], v[3]); c = F2(v[4], v[5]); d = F2(v[6], v[7]);
    e = F2(v[8], v[9]); f = F2(v[10], v[11]); g = F2(v[12], v[13]); h = F2(v[14], v[15]);
  }
  __device__ __forceinline__ void get6(int o, F2& a, F2& b, F2& c, F2& d, F2& e, F2& f) const {
    float v[12];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%12];\n"
        "tcgen05.ld.sync.aligned.32x32b.x4.b32 {%8, %9, %10, %11}, [%12 + 8];\n"
        "tcgen05.wait::ld.sync.aligned;"
        : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7]), "=f"(v[8]),
          "=f"(v[9]), "=f"(v[10]), "=f"(v[11])
        : "r"(base + 2 * o) : "memory");
    a = F2(v[0], v[1]); b = F2(v[2], v[3]); c = F2(v[4], v[5]); d = F2(v[6], v[7]); e = F2(v[8], v[9]); f = F2(v[10], v[11]);
  }
};
__device__ __forceinline__ void tmem_alloc(unsigned* smem_slot, unsigned cols) {   // one whole warp
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_slot)), "r"(cols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(unsigned taddr, unsigned cols) {      // the same warp
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
}

// Backward, aligned path (HW % 4 == 0, 16-byte aligned bases).  Same math as bwd_kernel<.., 2, ..>.
template <class Op, int MK, bool WANT_GIN>
__global__ void __launch_bounds__(BWD_THREADS, BWD_MIN_BLOCKS) bwd_tma_kernel(const KParams p) {
  constexpr int NC = Op::NC;
  constexpr int NM = 2 * NC;
  constexpr int NWARPS = BWD_THREADS / 32;
  extern __shared__ __align__(128) float s_ring[];   // [STAGES][PLANES][TILE]
  __shared__ __align__(8) unsigned long long s_full[TMA_STAGES];
  __shared__ float s_C[NC > 0 ? CURL_MAX_KNOTS : 1];
  __shared__ float s_ab[NC > 0 ? 2 * NC + 1 : 1];
  __shared__ double s_mom[NC > 0 ? NM : 1];
  __shared__ float s_part[NC > 0 ? NWARPS * NM : 1];
  __shared__ unsigned s_ticket;
  __shared__ unsigned s_tmem;
  constexpr bool USE_TMEM = PARK_IN_TMEM && Op::SPILL > 0;
  static_assert(!USE_TMEM || (2 * Op::SPILL <= TMEM_COLS && BWD_THREADS == 128), "TMEM parking: 4 warps, 2 columns per slot");
  static_assert(!PARK_J5 || PARK_IN_TMEM, "-DCURL_PARK_J5 needs -DCURL_PARK_TMEM");
  static_assert(!PARK_MOM || PARK_IN_TMEM, "-DCURL_PARK_MOM needs -DCURL_PARK_TMEM");
  const int b = blockIdx.y;
  const int HW = p.HW;
  const float* in = p.in + (size_t)b * 3 * HW;
  const float* go = p.gout + (size_t)b * 3 * HW;
  float* gin = WANT_GIN ? p.out + (size_t)b * 3 * HW : nullptr;
  const int n_tiles = (HW + TMA_TILE - 1) / TMA_TILE;
  const int n_my = ((int)blockIdx.x < n_tiles) ? (n_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;

  auto issue = [&](int it) {   // called by thread 0 only
    const int s = it % TMA_STAGES;
    const int px0 = (blockIdx.x + it * gridDim.x) * TMA_TILE;
    const int n = min(TMA_TILE, HW - px0);
    const unsigned bytes = (unsigned)n * sizeof(float);
    float* dst = s_ring + (size_t)s * TMA_PLANES * TMA_TILE;
    mbar_expect_tx(&s_full[s], bytes * TMA_PLANES);
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      bulk_g2s(dst + c * TMA_TILE, in + (size_t)c * HW + px0, bytes, &s_full[s]);
      bulk_g2s(dst + (3 + c) * TMA_TILE, go + (size_t)c * HW + px0, bytes, &s_full[s]);
    }
  };

  if (threadIdx.x == 0) {
#pragma unroll
    for (int s = 0; s < TMA_STAGES; ++s) mbar_init(&s_full[s], 1);
    mbar_fence_init();
  }
  // The last tile of an image may be partial; its tail pairs read whatever the stage holds.  Zero the
  // ring once so that this is always finite pixel data (their gradients are gated off below).
  for (int i = threadIdx.x; i < (int)(TMA_RING_BYTES / sizeof(float4)); i += BWD_THREADS)
    reinterpret_cast<float4*>(s_ring)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  if constexpr (USE_TMEM) {
    if (threadIdx.x < 32) tmem_alloc(&s_tmem, TMEM_COLS);
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  }
  __syncthreads();
  unsigned tmem_base = 0;
  if constexpr (USE_TMEM) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    tmem_base = s_tmem + (((threadIdx.x >> 5) & 3u) << 21);     // lane quarter of this warp: 32 (warp % 4) << 16
  }
  if (threadIdx.x == 0)
    for (int it = 0; it < TMA_STAGES && it < n_my; ++it) issue(it);
  if (NC > 0) stage_curves<Op, true>(p, b, s_C, s_ab, nullptr, false);
  constexpr bool MOM_TMEM = USE_TMEM && PARK_MOM && NC == LAYER_CURVES;
  if constexpr (MOM_TMEM) {
    const TmemSpill sp{tmem_base};
    const F2 z[8] = {F2(0.0f), F2(0.0f), F2(0.0f), F2(0.0f), F2(0.0f), F2(0.0f), F2(0.0f), F2(0.0f)};
    sp.mom_put6(0, z); sp.mom_put6(6, z); sp.mom_put8(z);
  }

  using T = F2;
#ifdef BWD_SCALAR_MOM
  // moments as scalars (lo + hi folded every iteration): 20 registers instead of 40 across the loop
  float mom[NM > 0 ? NM : 1];
#pragma unroll
  for (int k = 0; k < NM; ++k) mom[k] = 0.0f;
#else
  T mom[NM > 0 ? NM : 1];
  if constexpr (!MOM_TMEM) {
#pragma unroll
    for (int k = 0; k < NM; ++k) mom[k] = T(0.0f);
  }
#endif

  // The mask plane is only 4-byte aligned in general, so it cannot ride the bulk copies; it is
  // fetched with an ordinary load one whole iteration ahead (raw bits kept in one register pair).
  auto mask_raw = [&](int px) -> float2 {   // px: this thread's first pixel of a sub-tile (>= HW past the block's work)
    float2 r = make_float2(1.0f, 1.0f);
    if (MK == MASK_NONE || px >= HW) return r;
    if (MK == MASK_U8) {
      const unsigned short u = __ldcs(reinterpret_cast<const unsigned short*>(
          reinterpret_cast<const unsigned char*>(p.mask) + (size_t)b * HW + px));
      r.x = __uint_as_float((unsigned)u);
    } else {
      r = __ldcs(reinterpret_cast<const float2*>(reinterpret_cast<const float*>(p.mask) + (size_t)b * HW + px));
    }
    return r;
  };
  // Indices advance by addition only (tile by tile, sub-tile by sub-tile): the flat "sub-tile counter" form
  // cost ~40 integer instructions per sub-tile in divisions by the sub-tile and stage counts.
  const int tile_stride = (int)gridDim.x * TMA_TILE;
  int px_tile = (int)blockIdx.x * TMA_TILE + (int)threadIdx.x * 2;   // a block's tile indices never reach n_tiles + gridDim.x: no overflow
  float2 m_next = (n_my > 0) ? mask_raw(px_tile) : make_float2(1.0f, 1.0f);
  int s = 0;
  unsigned parity = 0u;

#pragma unroll 1
  for (int it = 0; it < n_my; ++it, px_tile += tile_stride) {
    // Refill of the stage the block finished with its previous tile.  The barrier sits here, after that
    // tile's arithmetic, and not right behind its shared-memory reads: BAR.SYNC does not wait for a
    // warp's outstanding LDS, so a refill issued straight after such a barrier can land before a
    // congested warp's reads of the old tile are served (seen as rare run-to-run differences of whole
    // warps' pixels with three blocks per SM).  Here every read of the old tile has been consumed.
    if (it > 0) {
      __syncthreads();
      if (threadIdx.x == 0 && it - 1 + TMA_STAGES < n_my) issue(it - 1 + TMA_STAGES);
    }
    mbar_wait(&s_full[s], parity);
    const float* stage = s_ring + (size_t)s * TMA_PLANES * TMA_TILE + threadIdx.x * 2;
    if (++s == TMA_STAGES) { s = 0; parity ^= 1u; }
#ifdef BWD_UNROLL_Q
#pragma unroll
#else
#pragma unroll 1
#endif
    for (int q = 0; q < BWD_PAIRS; ++q) {
      const int px = px_tile + q * TMA_SUB;
      const bool valid = px < HW;     // HW is even: a pair is either fully inside or fully outside
      const float2 m_raw = m_next;
      // next sub-tile of this tile, or the first of the block's next tile (past the image once the block is done)
      m_next = mask_raw(q + 1 < BWD_PAIRS ? px + TMA_SUB : px_tile + tile_stride);
      T m = T(1.0f);
      if (MK == MASK_U8) {     // the backward takes a {0,1} mask as a gate offset: 0 keeps, NaN masks (bwd_mask_lane_u8)
        const unsigned u = __float_as_uint(m_raw.x);
        m = F2(bwd_mask_lane_u8((u & 0xffu) != 0), bwd_mask_lane_u8((u & 0xff00u) != 0));
      } else if (MK == MASK_F32) {
        m = F2(m_raw.x, m_raw.y);
      }
      const float* src = stage + q * TMA_SUB;
      T x[3], gy[3];
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const float2 a = *reinterpret_cast<const float2*>(src + c * TMA_TILE);
        const float2 g = *reinterpret_cast<const float2*>(src + (3 + c) * TMA_TILE);
        x[c] = F2(a.x, a.y);
        // a tail pair's gradient is switched off through m (= 0) where a mask multiplies it anyway
        gy[c] = (MK == MASK_NONE && !valid) ? F2(0.0f) : F2(g.x, g.y);
      }
      if (!valid) m = (MK == MASK_U8) ? T(bwd_mask_lane_u8(false)) : T(0.0f);
      T gx[3];
#ifdef BWD_SCALAR_MOM
      T momi[NM > 0 ? NM : 1];
#pragma unroll
      for (int k = 0; k < NM; ++k) momi[k] = T(0.0f);
#else
      T* const momi = mom;
#endif
      if constexpr (USE_TMEM) {
        const TmemSpill sp{tmem_base};
        Op::template bwd<MK, WANT_GIN, T>(x, m, s_ab, p.aux, gy, gx, momi, sp);
      } else if constexpr (Op::SPILL > 0) {
        const SmemSpill sp{smem_u32(reinterpret_cast<unsigned char*>(s_ring) + TMA_RING_BYTES) + threadIdx.x * 16u};
        Op::template bwd<MK, WANT_GIN, T>(x, m, s_ab, p.aux, gy, gx, momi, sp);
      } else {
        Op::template bwd<MK, WANT_GIN, T>(x, m, s_ab, p.aux, gy, gx, momi, KeepInRegisters());
      }
#ifdef BWD_SCALAR_MOM
#pragma unroll
      for (int k = 0; k < NM; ++k) mom[k] += vhsum(momi[k]);
#endif
      if (WANT_GIN && valid) {
#pragma unroll
        for (int c = 0; c < 3; ++c)
          __stcs(reinterpret_cast<float2*>(gin + (size_t)c * HW + px), make_float2(gx[c].lo(), gx[c].hi()));
      }
    }   // sub-tiles
  }
  if constexpr (MOM_TMEM) {      // the accumulators come back to registers for the reduction
    const TmemSpill sp{tmem_base};
    sp.before_loads();
    sp.mom_get6(0, mom); sp.mom_get6(6, mom + 6); sp.mom_get8(mom + 12);
  }
  if constexpr (USE_TMEM) {
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (threadIdx.x < 32) tmem_dealloc(s_tmem, TMEM_COLS);
  }
#ifdef BWD_SCALAR_MOM
  if (NC > 0) finish_knot_grads<Op, float, BWD_THREADS>(p, b, mom, s_C, s_mom, s_part, &s_ticket);
#else
  if (NC > 0) finish_knot_grads<Op, T, BWD_THREADS>(p, b, mom, s_C, s_mom, s_part, &s_ticket);
#endif
}

// ---------------------------------------------------------------------------------------------
// Forward, aligned path, with the input planes streamed through a shared-memory ring by bulk async copies
// (-DFWD_TMA).  The plain forward kernel leaves 22 % of its warp samples waiting for LDG data: a thread can
// only have the loads of its own next pixels in flight if it holds them in registers, and a register prefetch
// costs occupancy (forward experiments).  Here one elected thread keeps the next FWD_TMA_STAGES - 1
// tiles of the block in flight at no register cost; the arithmetic and the 128-bit streaming stores are those
// of fwd_kernel<Op, MK, 4>.
// ---------------------------------------------------------------------------------------------
#ifndef FWD_TMA_SUB
#define FWD_TMA_SUB 2          // sub-tiles (4 pixels per thread each) per ring stage
#endif
#ifndef FWD_TMA_STAGES_N
#define FWD_TMA_STAGES_N 2
#endif
#ifndef FWD_TMA_BLOCKS
#define FWD_TMA_BLOCKS 9       // resident blocks per SM the grid is sized for (48 registers: 10 fit)
#endif
constexpr int FT_SUB_PX = THREADS * 4;
constexpr int FT_TILE = FT_SUB_PX * FWD_TMA_SUB;
constexpr int FT_STAGES = FWD_TMA_STAGES_N;
constexpr size_t FT_RING_BYTES = (size_t)FT_STAGES * 3 * FT_TILE * sizeof(float);

template <class Op, int MK>
__global__ void __launch_bounds__(THREADS) fwd_tma_kernel(const KParams p) {
  constexpr int NC = Op::NC;
  extern __shared__ __align__(128) float s_ring[];   // [STAGES][3][TILE]
  __shared__ __align__(8) unsigned long long s_full[FT_STAGES];
  __shared__ float s_C[NC > 0 ? CURL_MAX_KNOTS : 1];
  __shared__ float s_ab[NC > 0 ? 2 * NC : 1];
  __shared__ double s_reg[NC > 0 ? NC : 1];
  const int b = blockIdx.y;
  const int HW = p.HW;
  const float* in = p.in + (size_t)b * 3 * HW;
  float* out = p.out + (size_t)b * 3 * HW;
  const int n_tiles = (HW + FT_TILE - 1) / FT_TILE;
  const int n_my = ((int)blockIdx.x < n_tiles) ? (n_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;

  auto issue = [&](int it) {   // thread 0 only
    const int s = it % FT_STAGES;
    const int px0 = (blockIdx.x + it * gridDim.x) * FT_TILE;
    const int n = min(FT_TILE, HW - px0);
    const unsigned bytes = (unsigned)n * sizeof(float);
    float* dst = s_ring + (size_t)s * 3 * FT_TILE;
    mbar_expect_tx(&s_full[s], bytes * 3);
#pragma unroll
    for (int c = 0; c < 3; ++c) bulk_g2s(dst + c * FT_TILE, in + (size_t)c * HW + px0, bytes, &s_full[s]);
  };

  if (threadIdx.x == 0) {
#pragma unroll
    for (int s = 0; s < FT_STAGES; ++s) mbar_init(&s_full[s], 1);
    mbar_fence_init();
  }
  // a partial last tile leaves the tail of its stage untouched: make that finite once (results there are not stored)
  for (int i = threadIdx.x; i < (int)(FT_RING_BYTES / sizeof(float4)); i += THREADS)
    reinterpret_cast<float4*>(s_ring)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncthreads();
  if (threadIdx.x == 0)
    for (int it = 0; it < FT_STAGES && it < n_my; ++it) issue(it);
  if (NC > 0) {
    const bool owner = (blockIdx.x == 0) && (p.reg != nullptr);
    stage_curves<Op, false>(p, b, s_C, s_ab, s_reg, owner);
    if (owner && threadIdx.x == 0) {
      double r = 0.0;
      for (int j = 0; j < NC; ++j) r += s_reg[j];
      p.reg[b] = (float)r;
    }
  }

  using G = Group<4>;
  using T = F2;
  const int tile_stride = (int)gridDim.x * FT_TILE;
  int px_tile = (int)blockIdx.x * FT_TILE + (int)threadIdx.x * 4;
  int s = 0;
  unsigned parity = 0u;
#pragma unroll 1
  for (int it = 0; it < n_my; ++it, px_tile += tile_stride) {
    if (it > 0) {      // refill behind a barrier that follows the previous tile's arithmetic (see bwd_tma_kernel)
      __syncthreads();
      if (threadIdx.x == 0 && it - 1 + FT_STAGES < n_my) issue(it - 1 + FT_STAGES);
    }
    mbar_wait(&s_full[s], parity);
    const float* stage = s_ring + (size_t)s * 3 * FT_TILE + threadIdx.x * 4;
    if (++s == FT_STAGES) { s = 0; parity ^= 1u; }
#ifdef FWD_TMA_UNROLL
#pragma unroll
#else
#pragma unroll 1
#endif
    for (int q = 0; q < FWD_TMA_SUB; ++q) {
      const int px = px_tile + q * FT_SUB_PX;
      if (px < HW) {     // HW % 4 == 0: a thread's four pixels are inside or outside together
        const float* src = stage + q * FT_SUB_PX;
        const float4 a0 = *reinterpret_cast<const float4*>(src);
        const float4 a1 = *reinterpret_cast<const float4*>(src + FT_TILE);
        const float4 a2 = *reinterpret_cast<const float4*>(src + 2 * FT_TILE);
        const G m = load_mask<MK, 4>(p.mask, (size_t)b * HW + px);
        G o0, o1, o2;
        {
          const T x[3] = {F2(a0.x, a0.y), F2(a1.x, a1.y), F2(a2.x, a2.y)};
          T y[3];
          Op::template fwd<MK, T>(x, m.v[0], s_ab, p.aux, y);
          o0.v[0] = y[0]; o1.v[0] = y[1]; o2.v[0] = y[2];
        }
        {
          const T x[3] = {F2(a0.z, a0.w), F2(a1.z, a1.w), F2(a2.z, a2.w)};
          T y[3];
          Op::template fwd<MK, T>(x, m.v[1], s_ab, p.aux, y);
          o0.v[1] = y[0]; o1.v[1] = y[1]; o2.v[1] = y[2];
        }
        o0.st(out + px); o1.st(out + HW + px); o2.st(out + 2 * (size_t)HW + px);
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------
// Host side
// ---------------------------------------------------------------------------------------------
using host::sm_count;
using host::ensure_dyn_smem;

using host::apply_grid_cap;   // test hook, see curl_host.cuh

static int blocks_per_image(int B, int HW, int vec, int blocks_per_sm) {
  const long long tiles = ((long long)HW + (long long)THREADS * vec - 1) / ((long long)THREADS * vec);
  long long cap = ((long long)sm_count() * blocks_per_sm + B - 1) / B;
  if (cap < 1) cap = 1;
  long long n = tiles < cap ? tiles : cap;
  return apply_grid_cap(n);
}

#ifndef BWD_GRID_BPS
#define BWD_GRID_BPS 8
#endif
constexpr int BWD_GRID_BLOCKS_PER_SM = BWD_GRID_BPS;
constexpr int BWD_VEC = 2;   // pixels per thread per iteration on the aligned backward path
#ifndef FWD_GRID_BPS
#define FWD_GRID_BPS 64
#endif
constexpr int FWD_GRID_BLOCKS_PER_SM = FWD_GRID_BPS;

static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }
static bool aligned4(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 3u) == 0; }

static int check_launch() { return cudaGetLastError() == cudaSuccess ? CURL_OK : CURL_ERR_LAUNCH; }

// Grid of the ring forward: whole waves of FWD_TMA_BLOCKS resident blocks per SM, like bwd_grid_x.
static int fwd_tma_grid_x(int B, int HW) {
  const long long tiles = ((long long)HW + FT_TILE - 1) / FT_TILE;
  const int sms = sm_count();
  constexpr int R = FWD_TMA_BLOCKS;
  for (int w = 2 * R; w <= 6 * R; w += R) {
    const long long total = (long long)sms * w;
    if (total % B == 0 && total / B <= tiles) return apply_grid_cap(total / B);
  }
  long long cap = ((long long)sms * 2 * R + B - 1) / B;
  long long n = tiles < cap ? tiles : cap;
  return apply_grid_cap(n);
}

template <class Op, int MK>
static int launch_fwd(const KParams& p, int B, bool vec4, cudaStream_t st) {
#ifdef FWD_TMA
  if (vec4 && Op::NC == LAYER_CURVES) {
    if (!ensure_dyn_smem<fwd_tma_kernel<Op, MK>>(FT_RING_BYTES)) return CURL_ERR_LAUNCH;
    dim3 grid(fwd_tma_grid_x(B, p.HW), B);
    fwd_tma_kernel<Op, MK><<<grid, THREADS, FT_RING_BYTES, st>>>(p);
    return check_launch();
  }
#endif
  const int vec = vec4 ? 4 : 1;
  dim3 grid(blocks_per_image(B, p.HW, vec, FWD_GRID_BLOCKS_PER_SM), B);
  if (vec4) fwd_kernel<Op, MK, 4><<<grid, THREADS, 0, st>>>(p);
  else fwd_kernel<Op, MK, 1><<<grid, THREADS, 0, st>>>(p);
  return check_launch();
}

// Backward grid: BWD_MIN_BLOCKS blocks are resident per SM, so the grid is sized to a whole number of waves
// (B * grid.x a multiple of SM count x resident blocks) whenever some wave count in [W, 2W] allows it; a
// partial last wave costs up to one full wave of time.
static int bwd_grid_x(int B, int HW, int vec) {
  const int px_per_tile = (vec == 1 ? THREADS : TMA_TILE);
  const long long tiles = ((long long)HW + px_per_tile - 1) / px_per_tile;
  const int sms = sm_count();
  // R blocks are resident per SM (BWD_MIN_BLOCKS): a wave is sms * R blocks, so w steps in multiples of R
  constexpr int R = BWD_MIN_BLOCKS;
  for (int w = BWD_GRID_BLOCKS_PER_SM * R; w <= 2 * BWD_GRID_BLOCKS_PER_SM * R; w += R) {
    const long long total = (long long)sms * w;
    if (total % B == 0 && total / B <= tiles) return apply_grid_cap(total / B);
  }
  long long cap = ((long long)sms * BWD_GRID_BLOCKS_PER_SM * R + B - 1) / B;
  long long n = tiles < cap ? tiles : cap;
  return apply_grid_cap(n);
}

template <class Op, int MK>
static int launch_bwd(KParams& p, int B, bool vec4, void* ws, size_t ws_bytes, cudaStream_t st) {
  const int vec = vec4 ? BWD_VEC : 1;
  dim3 grid(bwd_grid_x(B, p.HW, vec), B);
  if (Op::NC > 0) {
    const size_t tick_bytes = ((size_t)B * sizeof(unsigned) + 255) / 256 * 256;
    const size_t need = tick_bytes + (size_t)B * grid.x * 2 * Op::NC * sizeof(float);
    if (ws == nullptr || ws_bytes < need || !aligned16(ws)) return CURL_ERR_WORKSPACE;
    p.tickets = reinterpret_cast<unsigned*>(ws);
    p.partials = reinterpret_cast<float*>(reinterpret_cast<char*>(ws) + tick_bytes);
    if (cudaMemsetAsync(ws, 0, tick_bytes, st) != cudaSuccess) return CURL_ERR_LAUNCH;
  }
  const bool want = p.out != nullptr;
  if (vec4) {
    if (want) {
      if (!ensure_dyn_smem<bwd_tma_kernel<Op, MK, true>>(tma_smem_bytes<Op>())) return CURL_ERR_LAUNCH;
      bwd_tma_kernel<Op, MK, true><<<grid, BWD_THREADS, tma_smem_bytes<Op>(), st>>>(p);
    } else {
      if (!ensure_dyn_smem<bwd_tma_kernel<Op, MK, false>>(tma_smem_bytes<Op>())) return CURL_ERR_LAUNCH;
      bwd_tma_kernel<Op, MK, false><<<grid, BWD_THREADS, tma_smem_bytes<Op>(), st>>>(p);
    }
  } else {
    if (want) bwd_kernel<Op, MK, 1, true><<<grid, THREADS, 0, st>>>(p);
    else bwd_kernel<Op, MK, 1, false><<<grid, THREADS, 0, st>>>(p);
  }
  return check_launch();
}

static bool can_vec4(const KParams& p, int mask_kind) {
  if (p.HW % 4 != 0) return false;
  if (!aligned16(p.in) || (p.out && !aligned16(p.out)) || (p.gout && !aligned16(p.gout))) return false;
  if (mask_kind == CURL_MASK_U8 && !aligned4(p.mask)) return false;
  if (mask_kind == CURL_MASK_F32 && !aligned16(p.mask)) return false;
  return true;
}

static int check_shape(int B, int H, int W) {
  if (B <= 0 || H <= 0 || W <= 0) return CURL_ERR_BAD_ARG;
  if ((long long)H * W > 0x3fffffffLL || B > 65535) return CURL_ERR_BAD_ARG;
  return CURL_OK;
}

}  // namespace curl

using namespace curl;

extern "C" {

int curl_abi_version(void) { return CURL_ABI_VERSION; }

int curl_test_set_grid_cap(int max_blocks_per_image) {
  const int old = host::g_test_grid_cap;
  host::g_test_grid_cap = max_blocks_per_image > 0 ? max_blocks_per_image : 0;
  return old;
}

const char* curl_strerror(int code) {
  switch (code) {
    case CURL_OK: return "success";
    case CURL_ERR_BAD_ARG: return "bad argument (null pointer, non-positive size, K < 3 or unknown enum)";
    case CURL_ERR_TOO_MANY_KNOTS: return "too many knots per image (CURL_MAX_KNOTS)";
    case CURL_ERR_WORKSPACE: return "workspace missing, misaligned or too small (see curl_bwd_workspace_bytes)";
    case CURL_ERR_LAUNCH: return "CUDA launch failed";
    case CURL_ERR_MISALIGNED: return "pointer not 4-byte aligned";
    case CURL_ERR_TOO_LARGE: return "shape needs more scratch than the allocation-free entry point's fixed 1 MiB array "
                                    "(CURL_NOALLOC_SCRATCH_BYTES); use curl_layer_bwd_f32 with a caller-supplied workspace";
    default: return "unknown error";
  }
}

size_t curl_bwd_workspace_bytes(int B, int H, int W) {
  if (check_shape(B, H, W) != CURL_OK) return 0;
  // sized for the scalar path (most blocks) and the widest op (the layer: 20 moments)
  const int HW = H * W;
  int gx = bwd_grid_x(B, HW, 1);
  const int gx4 = bwd_grid_x(B, HW, BWD_VEC);
  if (gx4 > gx) gx = gx4;
  const size_t tick_bytes = ((size_t)B * sizeof(unsigned) + 255) / 256 * 256;
  return tick_bytes + (size_t)B * gx * 2 * LAYER_CURVES * sizeof(float);
}

static int layer_params(KParams& p, const float* img, const void* mask, int mask_kind, const float* L, int64_t ldL,
                        const float* R, int64_t ldR, const float* Hk, int64_t ldH, int B, int H, int W, int K_lab,
                        int K_rgb, int K_hsv) {
  int rc = check_shape(B, H, W);
  if (rc != CURL_OK) return rc;
  if (!img || !L || !R || !Hk) return CURL_ERR_BAD_ARG;
  if (K_lab < 3 || K_rgb < 3 || K_hsv < 3) return CURL_ERR_BAD_ARG;
  if (mask_kind < CURL_MASK_NONE || mask_kind > CURL_MASK_F32) return CURL_ERR_BAD_ARG;
  if ((mask_kind != CURL_MASK_NONE) != (mask != nullptr)) return CURL_ERR_BAD_ARG;
  if (ldL < 3 * (int64_t)K_lab || ldR < 3 * (int64_t)K_rgb || ldH < 4 * (int64_t)K_hsv) return CURL_ERR_BAD_ARG;
  if (3 * K_lab + 3 * K_rgb + 4 * K_hsv > CURL_MAX_KNOTS) return CURL_ERR_TOO_MANY_KNOTS;
  if (!aligned4(img) || !aligned4(L) || !aligned4(R) || !aligned4(Hk)) return CURL_ERR_MISALIGNED;
  p = KParams{};
  p.in = img; p.mask = mask; p.HW = H * W; p.nseg = 3;
  p.seg[0] = KnotSeg{L, nullptr, (long long)ldL, 3, K_lab};
  p.seg[1] = KnotSeg{R, nullptr, (long long)ldR, 3, K_rgb};
  p.seg[2] = KnotSeg{Hk, nullptr, (long long)ldH, 4, K_hsv};
  return CURL_OK;
}

int curl_layer_fwd_f32(const float* img, const void* mask, int mask_kind, const float* L, int64_t ldL,
                       const float* R, int64_t ldR, const float* Hk, int64_t ldH, float* out, float* reg, int B,
                       int H, int W, int K_lab, int K_rgb, int K_hsv, void* stream) {
  KParams p;
  int rc = layer_params(p, img, mask, mask_kind, L, ldL, R, ldR, Hk, ldH, B, H, W, K_lab, K_rgb, K_hsv);
  if (rc != CURL_OK) return rc;
  if (!out || !reg) return CURL_ERR_BAD_ARG;
  if (!aligned4(out)) return CURL_ERR_MISALIGNED;
  p.out = out; p.reg = reg;
  const bool v4 = can_vec4(p, mask_kind);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  switch (mask_kind) {
    case CURL_MASK_NONE: return launch_fwd<OpLayer, MASK_NONE>(p, B, v4, st);
    case CURL_MASK_U8: return launch_fwd<OpLayer, MASK_U8>(p, B, v4, st);
    default: return launch_fwd<OpLayer, MASK_F32>(p, B, v4, st);
  }
}

}  // extern "C"

template <int MK>
static int launch_fwd_u8(const KParams& p, int B, bool vec4, int in_layout, int out_layout, cudaStream_t st) {
  const int vec = vec4 ? 4 : 1;
  dim3 grid(blocks_per_image(B, p.HW, vec, FWD_GRID_BLOCKS_PER_SM), B);
  const int sel = (in_layout == CURL_LAYOUT_HWC ? 2 : 0) | (out_layout == CURL_LAYOUT_HWC ? 1 : 0);
#define CURL_U8_LAUNCH(IH, OH)                                                     \
  do {                                                                             \
    if (vec4) fwd_u8_kernel<MK, IH, OH, 4><<<grid, THREADS, 0, st>>>(p);            \
    else fwd_u8_kernel<MK, IH, OH, 1><<<grid, THREADS, 0, st>>>(p);                 \
  } while (0)
  switch (sel) {
    case 0: CURL_U8_LAUNCH(false, false); break;
    case 1: CURL_U8_LAUNCH(false, true); break;
    case 2: CURL_U8_LAUNCH(true, false); break;
    default: CURL_U8_LAUNCH(true, true); break;
  }
#undef CURL_U8_LAUNCH
  return check_launch();
}

extern "C" {

int curl_layer_fwd_u8(const uint8_t* img, int in_layout, const void* mask, int mask_kind, const float* L,
                      int64_t ldL, const float* R, int64_t ldR, const float* Hk, int64_t ldH, uint8_t* out,
                      int out_layout, float* reg, int B, int H, int W, int K_lab, int K_rgb, int K_hsv,
                      void* stream) {
  KParams p;
  // layer_params validates everything but the image pointers (its float* img is only null-checked here)
  int rc = layer_params(p, reinterpret_cast<const float*>(L), mask, mask_kind, L, ldL, R, ldR, Hk, ldH, B, H, W, K_lab,
                        K_rgb, K_hsv);
  if (rc != CURL_OK) return rc;
  if (!img || !out || !reg) return CURL_ERR_BAD_ARG;
  if (in_layout < CURL_LAYOUT_CHW || in_layout > CURL_LAYOUT_HWC) return CURL_ERR_BAD_ARG;
  if (out_layout < CURL_LAYOUT_CHW || out_layout > CURL_LAYOUT_HWC) return CURL_ERR_BAD_ARG;
  p.in = nullptr; p.in8 = img; p.out8 = out; p.reg = reg;
  bool v4 = (p.HW % 4 == 0) && aligned4(img) && aligned4(out);
  if (mask_kind == CURL_MASK_U8 && !aligned4(mask)) v4 = false;
  if (mask_kind == CURL_MASK_F32 && !aligned16(mask)) v4 = false;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  switch (mask_kind) {
    case CURL_MASK_NONE: return launch_fwd_u8<MASK_NONE>(p, B, v4, in_layout, out_layout, st);
    case CURL_MASK_U8: return launch_fwd_u8<MASK_U8>(p, B, v4, in_layout, out_layout, st);
    default: return launch_fwd_u8<MASK_F32>(p, B, v4, in_layout, out_layout, st);
  }
}

int curl_layer_bwd_f32(const float* img, const void* mask, int mask_kind, const float* L, int64_t ldL,
                       const float* R, int64_t ldR, const float* Hk, int64_t ldH, const float* gout,
                       const float* greg, float* gimg, float* gL, float* gR, float* gH, void* workspace,
                       size_t workspace_bytes, int B, int H, int W, int K_lab, int K_rgb, int K_hsv, void* stream) {
  KParams p;
  int rc = layer_params(p, img, mask, mask_kind, L, ldL, R, ldR, Hk, ldH, B, H, W, K_lab, K_rgb, K_hsv);
  if (rc != CURL_OK) return rc;
  if (!gout || !gL || !gR || !gH) return CURL_ERR_BAD_ARG;
  if (!aligned4(gout) || (gimg && !aligned4(gimg))) return CURL_ERR_MISALIGNED;
  p.gout = gout; p.greg = greg; p.out = gimg;
  p.seg[0].g = gL; p.seg[1].g = gR; p.seg[2].g = gH;
  const bool v4 = can_vec4(p, mask_kind);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  switch (mask_kind) {
    case CURL_MASK_NONE: return launch_bwd<OpLayer, MASK_NONE>(p, B, v4, workspace, workspace_bytes, st);
    case CURL_MASK_U8: return launch_bwd<OpLayer, MASK_U8>(p, B, v4, workspace, workspace_bytes, st);
    default: return launch_bwd<OpLayer, MASK_F32>(p, B, v4, workspace, workspace_bytes, st);
  }
}

int curl_color_fwd_f32(int conversion, const float* in, float* out, int B, int H, int W, void* stream) {
#ifdef CURL_DEV_LAYER_ONLY   /* quick experiment builds: only the layer kernels are instantiated */
  return CURL_ERR_BAD_ARG;
#else
  int rc = check_shape(B, H, W);
  if (rc != CURL_OK) return rc;
  if (!in || !out) return CURL_ERR_BAD_ARG;
  if (!aligned4(in) || !aligned4(out)) return CURL_ERR_MISALIGNED;
  KParams p{};
  p.in = in; p.out = out; p.HW = H * W;
  const bool v4 = can_vec4(p, CURL_MASK_NONE);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  switch (conversion) {
    case CURL_RGB2LAB: return launch_fwd<OpColor<CURL_RGB2LAB>, MASK_NONE>(p, B, v4, st);
    case CURL_LAB2RGB: return launch_fwd<OpColor<CURL_LAB2RGB>, MASK_NONE>(p, B, v4, st);
    case CURL_RGB2HSV: return launch_fwd<OpColor<CURL_RGB2HSV>, MASK_NONE>(p, B, v4, st);
    case CURL_HSV2RGB: return launch_fwd<OpColor<CURL_HSV2RGB>, MASK_NONE>(p, B, v4, st);
    default: return CURL_ERR_BAD_ARG;
  }
#endif
}

int curl_color_bwd_f32(int conversion, const float* in, const float* gout, float* gin, int B, int H, int W,
                       void* stream) {
#ifdef CURL_DEV_LAYER_ONLY   /* quick experiment builds: only the layer kernels are instantiated */
  return CURL_ERR_BAD_ARG;
#else
  int rc = check_shape(B, H, W);
  if (rc != CURL_OK) return rc;
  if (!in || !gout || !gin) return CURL_ERR_BAD_ARG;
  if (!aligned4(in) || !aligned4(gout) || !aligned4(gin)) return CURL_ERR_MISALIGNED;
  KParams p{};
  p.in = in; p.gout = gout; p.out = gin; p.HW = H * W;
  const bool v4 = can_vec4(p, CURL_MASK_NONE);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  switch (conversion) {
    case CURL_RGB2LAB: return launch_bwd<OpColor<CURL_RGB2LAB>, MASK_NONE>(p, B, v4, nullptr, 0, st);
    case CURL_LAB2RGB: return launch_bwd<OpColor<CURL_LAB2RGB>, MASK_NONE>(p, B, v4, nullptr, 0, st);
    case CURL_RGB2HSV: return launch_bwd<OpColor<CURL_RGB2HSV>, MASK_NONE>(p, B, v4, nullptr, 0, st);
    case CURL_HSV2RGB: return launch_bwd<OpColor<CURL_HSV2RGB>, MASK_NONE>(p, B, v4, nullptr, 0, st);
    default: return CURL_ERR_BAD_ARG;
  }
#endif
}

static int adjust_params(KParams& p, int group, const float* in, const float* P, int64_t ldP, int B, int H, int W,
                         int K) {
  int rc = check_shape(B, H, W);
  if (rc != CURL_OK) return rc;
  if (!in || !P || K < 3) return CURL_ERR_BAD_ARG;
  if (group < CURL_ADJUST_LAB || group > CURL_ADJUST_HSV) return CURL_ERR_BAD_ARG;
  const int nc = (group == CURL_ADJUST_HSV) ? 4 : 3;
  if (ldP < (int64_t)nc * K) return CURL_ERR_BAD_ARG;
  if (nc * K > CURL_MAX_KNOTS) return CURL_ERR_TOO_MANY_KNOTS;
  if (!aligned4(in) || !aligned4(P)) return CURL_ERR_MISALIGNED;
  p = KParams{};
  p.in = in; p.HW = H * W; p.nseg = 1;
  p.seg[0] = KnotSeg{P, nullptr, (long long)ldP, nc, K};
  return CURL_OK;
}

int curl_adjust_fwd_f32(int group, const float* in, const float* P, int64_t ldP, float* out, float* reg, int B,
                        int H, int W, int K, void* stream) {
#ifdef CURL_DEV_LAYER_ONLY   /* quick experiment builds: only the layer kernels are instantiated */
  return CURL_ERR_BAD_ARG;
#else
  KParams p;
  int rc = adjust_params(p, group, in, P, ldP, B, H, W, K);
  if (rc != CURL_OK) return rc;
  if (!out || !reg) return CURL_ERR_BAD_ARG;
  if (!aligned4(out)) return CURL_ERR_MISALIGNED;
  p.out = out; p.reg = reg;
  const bool v4 = can_vec4(p, CURL_MASK_NONE);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (group == CURL_ADJUST_HSV) return launch_fwd<OpAdjustHsv, MASK_NONE>(p, B, v4, st);
  return launch_fwd<OpAdjustSelf3, MASK_NONE>(p, B, v4, st);
#endif
}

int curl_adjust_bwd_f32(int group, const float* in, const float* P, int64_t ldP, const float* gout,
                        const float* greg, float* gin, float* gP, void* workspace, size_t workspace_bytes, int B,
                        int H, int W, int K, void* stream) {
#ifdef CURL_DEV_LAYER_ONLY   /* quick experiment builds: only the layer kernels are instantiated */
  return CURL_ERR_BAD_ARG;
#else
  KParams p;
  int rc = adjust_params(p, group, in, P, ldP, B, H, W, K);
  if (rc != CURL_OK) return rc;
  if (!gout || !gP) return CURL_ERR_BAD_ARG;
  if (!aligned4(gout) || (gin && !aligned4(gin))) return CURL_ERR_MISALIGNED;
  p.gout = gout; p.greg = greg; p.out = gin; p.seg[0].g = gP;
  const bool v4 = can_vec4(p, CURL_MASK_NONE);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (group == CURL_ADJUST_HSV) return launch_bwd<OpAdjustHsv, MASK_NONE>(p, B, v4, workspace, workspace_bytes, st);
  return launch_bwd<OpAdjustSelf3, MASK_NONE>(p, B, v4, workspace, workspace_bytes, st);
#endif
}

int curl_apply_curve_fwd_f32(const float* in, const float* C, int64_t ldC, int channel_in, int channel_out,
                             float* out, float* reg, int B, int H, int W, int K, void* stream) {
#ifdef CURL_DEV_LAYER_ONLY   /* quick experiment builds: only the layer kernels are instantiated */
  return CURL_ERR_BAD_ARG;
#else
  int rc = check_shape(B, H, W);
  if (rc != CURL_OK) return rc;
  if (!in || !C || !out || !reg || K < 3 || ldC < K) return CURL_ERR_BAD_ARG;
  if (channel_in < 0 || channel_in > 2 || channel_out < 0 || channel_out > 2) return CURL_ERR_BAD_ARG;
  if (K > CURL_MAX_KNOTS) return CURL_ERR_TOO_MANY_KNOTS;
  if (!aligned4(in) || !aligned4(C) || !aligned4(out)) return CURL_ERR_MISALIGNED;
  KParams p{};
  p.in = in; p.out = out; p.reg = reg; p.HW = H * W; p.nseg = 1; p.linear_knots = 1;
  p.aux = channel_in | (channel_out << 2);
  p.seg[0] = KnotSeg{C, nullptr, (long long)ldC, 1, K};
  return launch_fwd<OpCurve1, MASK_NONE>(p, B, can_vec4(p, CURL_MASK_NONE), reinterpret_cast<cudaStream_t>(stream));
#endif
}

int curl_apply_curve_bwd_f32(const float* in, const float* C, int64_t ldC, int channel_in, int channel_out,
                             const float* gout, const float* greg, float* gin, float* gC, void* workspace,
                             size_t workspace_bytes, int B, int H, int W, int K, void* stream) {
#ifdef CURL_DEV_LAYER_ONLY   /* quick experiment builds: only the layer kernels are instantiated */
  return CURL_ERR_BAD_ARG;
#else
  int rc = check_shape(B, H, W);
  if (rc != CURL_OK) return rc;
  if (!in || !C || !gout || !gC || K < 3 || ldC < K) return CURL_ERR_BAD_ARG;
  if (channel_in < 0 || channel_in > 2 || channel_out < 0 || channel_out > 2) return CURL_ERR_BAD_ARG;
  if (K > CURL_MAX_KNOTS) return CURL_ERR_TOO_MANY_KNOTS;
  if (!aligned4(in) || !aligned4(C) || !aligned4(gout) || (gin && !aligned4(gin))) return CURL_ERR_MISALIGNED;
  KParams p{};
  p.in = in; p.gout = gout; p.greg = greg; p.out = gin; p.HW = H * W; p.nseg = 1; p.linear_knots = 1;
  p.aux = channel_in | (channel_out << 2);
  p.seg[0] = KnotSeg{C, gC, (long long)ldC, 1, K};
  return launch_bwd<OpCurve1, MASK_NONE>(p, B, can_vec4(p, CURL_MASK_NONE), workspace, workspace_bytes,
                                         reinterpret_cast<cudaStream_t>(stream));
#endif
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------
// Allocation-free variant of curl_layer_bwd_f32 (no counterpart in the reference; added
// for callers that cannot allocate: code under CUDA-graph capture, tight serving loops, harnesses
// that forbid cudaMalloc).  The backward's scratch -- per-image tickets + per-block moment partials, what
// curl_bwd_workspace_bytes() sizes -- lives in a file-scope device array instead of a caller-supplied workspace.
//   * every GPU that loads the library holds CURL_NOALLOC_SCRATCH_BYTES (1 MiB) until the process exits;
//   * shapes whose workspace exceeds that are rejected with CURL_ERR_TOO_LARGE before anything is launched;
//   * every call shares the ONE array, which keeps the previous call's contents: the tickets (the only part that
//     must start at zero) are cleared by the memset node launch_bwd enqueues; the partial rows are fully
//     overwritten before they are read.  Two calls that overlap on the GPU (different streams, or one stream per
//     host thread) RACE on it -- serialise them, or use curl_layer_bwd_f32 with one workspace per stream;
//   * calls no cudaMalloc / cudaFree / allocating library routine.  In host code the array's name is only a
//     stand-in for the symbol, not a device address: its address is looked up once per device, with that
//     device current, by cudaGetSymbolAddress and handed to the kernel through the same KParams as a workspace.
// ---------------------------------------------------------------------------------------------
static __device__ __align__(256) unsigned char g_noalloc_scratch[CURL_NOALLOC_SCRATCH_BYTES];

static void* noalloc_scratch_address() {
  static void* cached[host::MAX_DEVICES] = {nullptr};
  const int dev = host::current_device();
  if (dev < 0) return nullptr;
  if (cached[dev] == nullptr) {
    void* ptr = nullptr;
    if (cudaGetSymbolAddress(&ptr, g_noalloc_scratch) != cudaSuccess || ptr == nullptr) return nullptr;
    cached[dev] = ptr;
  }
  return cached[dev];
}

extern "C" int curl_layer_bwd_noalloc_f32(const float* img, const void* mask, int mask_kind, const float* L, int64_t ldL,
                                          const float* R, int64_t ldR, const float* Hk, int64_t ldH, const float* gout,
                                          const float* greg, float* gimg, float* gL, float* gR, float* gH, int B, int H,
                                          int W, int K_lab, int K_rgb, int K_hsv, void* stream) {
  KParams p;
  int rc = layer_params(p, img, mask, mask_kind, L, ldL, R, ldR, Hk, ldH, B, H, W, K_lab, K_rgb, K_hsv);
  if (rc != CURL_OK) return rc;
  if (!gout || !gL || !gR || !gH) return CURL_ERR_BAD_ARG;
  if (!aligned4(gout) || (gimg && !aligned4(gimg))) return CURL_ERR_MISALIGNED;
  const size_t need = curl_bwd_workspace_bytes(B, H, W);
  if (need == 0 || need > (size_t)CURL_NOALLOC_SCRATCH_BYTES) return CURL_ERR_TOO_LARGE;
  void* scratch = noalloc_scratch_address();
  if (scratch == nullptr) return CURL_ERR_LAUNCH;
  p.gout = gout; p.greg = greg; p.out = gimg;
  p.seg[0].g = gL; p.seg[1].g = gR; p.seg[2].g = gH;
  const bool v4 = can_vec4(p, mask_kind);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  switch (mask_kind) {
    case CURL_MASK_NONE: return launch_bwd<OpLayer, MASK_NONE>(p, B, v4, scratch, CURL_NOALLOC_SCRATCH_BYTES, st);
    case CURL_MASK_U8: return launch_bwd<OpLayer, MASK_U8>(p, B, v4, scratch, CURL_NOALLOC_SCRATCH_BYTES, st);
    default: return launch_bwd<OpLayer, MASK_F32>(p, B, v4, scratch, CURL_NOALLOC_SCRATCH_BYTES, st);
  }
}
