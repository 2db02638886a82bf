// sm_100a kernels + C ABI of the polynomial colour regression ("TriSpace" residual), the per-pixel
// path the reference's main.py / infer.py run through TriSpaceRegNet (model.py:411-520):
//   curl_trispace_fwd_f32 : generate_residual (+ generate_image), one pass per pixel, no
//                           [B, NC, H, W] monomial tensor (the reference materialises 126 floats per
//                           pixel and space)
//   curl_trispace_bwd_f32 : gradient w.r.t. the 3 x 3 x NC coefficients per image (recompute)
// Per-pixel math: curl_poly.cuh.  Layout: planar NCHW fp32 like the curve layer; blockIdx.y = image,
// whose 9 * NC coefficients live in shared memory.
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/curl_b200.h"
#include "curl_poly.cuh"
#include "curl_host.cuh"

namespace curl {
namespace tri {

// Forward launch shape: 128 threads x 3 blocks per SM (164 registers, NO spills).  Round 1 shipped 256 x 2, whose
// 128-register cap spilled 28-112 bytes per thread; measured on B200 (ms, 512^2 x 32 / 4K x 8):
//   126 coefficients: 256x2 0.433 / 3.174, 128x3 0.430 / 3.171, 128x2 0.429 / 3.171, 256x1 (no cap) 0.496 / 3.341
//    35 coefficients: 256x2 0.187 / 1.254, 128x3 0.175 / 1.253, 128x2 0.193 / 1.382, 256x1 0.235 / 1.534
#ifndef TRI_FWD_THREADS
#define TRI_FWD_THREADS 128
#endif
#ifndef TRI_FWD_MIN_BLOCKS
#define TRI_FWD_MIN_BLOCKS 3
#endif
constexpr int THREADS = TRI_FWD_THREADS;
constexpr int MAX_DEG = 4;                   // degrees 1..4 are instantiated (the reference's default polynomial_order is 4,
                                             // model.py:413; ChannelPolyLayer's own default degree is 3, model.py:208)

struct TParams {
  const float* img;      // [B][3][HW]
  const float* coef[3];  // R, L, H: [B][3][NC]
  const float* gout;     // backward: [B][3][HW]
  float* out;            // forward: [B][3][HW]
  float* gcoef[3];       // backward: [B][3][NC]
  float* sig;            // [B][9][HW] sigmoid outputs: written by the forward (SAVE), read by the backward (SAVED); or null
  float* partials;       // backward: [B][gridDim.x][9*NC]
  unsigned* tickets;     // backward: [B], zero on entry
  int HW, W, H;
  int clamp;             // apply generate_image: clamp(img + residual, 0, 1)
};

using host::sm_count;

// coefficients of image b -> shared memory as [space][monomial] Coef4 records {c0, c1, c2, 0}
// (the space index is a compile-time constant: indexing p.coef[] with a run-time value made the compiler copy
// the whole parameter block to local memory -- a 112-byte stack frame in every kernel of this file)
template <int NC>
__device__ __forceinline__ void stage_coef(const TParams& p, int b, Coef4* s_coef) {
#pragma unroll
  for (int s = 0; s < 3; ++s) {
    const float* src = p.coef[s] + (size_t)b * 3 * NC;
    for (int r = threadIdx.x; r < NC * 4; r += blockDim.x) {
      const int k = r >> 2, c = r & 3;
      s_coef[s * NC + k].v[c] = c < 3 ? src[c * NC + k] : 0.0f;
    }
  }
  __syncthreads();
}

// normalised pixel coordinates of linear index idx (model.py:480-483): column / width, row / height
__device__ __forceinline__ void coords(int idx, int W, float fW, float fH, float* xs, float* ys) {
  const int row = idx / W, col = idx - row * W;
  *xs = __fdiv_rn((float)col, fW);
  *ys = __fdiv_rn((float)row, fH);
}

// ---------------------------------------------------------------------------------------------
// Forward.  VEC = 2: one packed pixel pair (F2) per thread and iteration, 64-bit accesses -- the
// kernel is FMA-bound (about 1500 multiply-adds per pixel), not HBM-bound, and the fully unrolled
// monomial walk wants its registers; VEC = 1: scalar path for odd HW or unaligned bases.
// ---------------------------------------------------------------------------------------------
template <int NV, int DEG, int VEC, bool SAVE>
__global__ void __launch_bounds__(THREADS, TRI_FWD_MIN_BLOCKS) trispace_fwd_kernel(const TParams p) {
  constexpr int NC = poly_num_coeffs(NV, DEG);
  constexpr bool ROLLED = NV == 5 && DEG == 4 && !SAVE;     // see trispace_fwd (curl_poly.cuh)
  __shared__ Coef4 s_coef[3 * NC];
  const int b = blockIdx.y;
  stage_coef<NC>(p, b, s_coef);
  const int HW = p.HW;
  const float* in = p.img + (size_t)b * 3 * HW;
  float* out = p.out + (size_t)b * 3 * HW;
  float* sg = SAVE ? p.sig + (size_t)b * 9 * HW : nullptr;
  const float fW = (float)p.W, fH = (float)p.H;
  const bool clamp = p.clamp != 0;
  const int stride = gridDim.x * THREADS * VEC;
  for (int i = (blockIdx.x * THREADS + threadIdx.x) * VEC; i < HW; i += stride) {
    if constexpr (VEC == 4) {
      // four pixels per thread: every coefficient record read from shared memory feeds twice the arithmetic
      const float4 c0 = __ldcs(reinterpret_cast<const float4*>(in + i));
      const float4 c1 = __ldcs(reinterpret_cast<const float4*>(in + HW + i));
      const float4 c2 = __ldcs(reinterpret_cast<const float4*>(in + 2 * (size_t)HW + i));
      float xs[4] = {0.f, 0.f, 0.f, 0.f}, ys[4] = {0.f, 0.f, 0.f, 0.f};
      if (NV == 5) {
#pragma unroll
        for (int k = 0; k < 4; ++k) coords(i + k, p.W, fW, fH, &xs[k], &ys[k]);
      }
      const F4 x[3] = {F4(F2(c0.x, c0.y), F2(c0.z, c0.w)), F4(F2(c1.x, c1.y), F2(c1.z, c1.w)), F4(F2(c2.x, c2.y), F2(c2.z, c2.w))};
      // the short polynomial stores its sigmoids after the three spaces (35 coefficients, 4K x 8: 1.248 ms against
      // 1.271 storing as they appear); the long one as they appear (3.30 against 3.34 ms, and 8 B of spills for 60)
      constexpr bool DEFER = SAVE && NV == 3;
      F4 y[3], sv[DEFER ? 9 : 1];
      trispace_fwd<F4, NV, DEG, ROLLED>(x, F4(F2(xs[0], xs[1]), F2(xs[2], xs[3])), F4(F2(ys[0], ys[1]), F2(ys[2], ys[3])), s_coef, clamp, y,
                                       [&](int sp, int k, F4 v) {
                                         if constexpr (DEFER) sv[3 * sp + k] = v;
                                         else if constexpr (SAVE) __stcs(reinterpret_cast<float4*>(sg + (size_t)(3 * sp + k) * HW + i),
                                                                         make_float4(v.a.lo(), v.a.hi(), v.b.lo(), v.b.hi()));
                                       });
#pragma unroll
      for (int c = 0; c < 3; ++c)
        __stcs(reinterpret_cast<float4*>(out + (size_t)c * HW + i), make_float4(y[c].a.lo(), y[c].a.hi(), y[c].b.lo(), y[c].b.hi()));
      if constexpr (DEFER) {
#pragma unroll
        for (int c = 0; c < 9; ++c)
          __stcs(reinterpret_cast<float4*>(sg + (size_t)c * HW + i), make_float4(sv[c].a.lo(), sv[c].a.hi(), sv[c].b.lo(), sv[c].b.hi()));
      }
    } else if constexpr (VEC == 2) {
      const float2 c0 = __ldcs(reinterpret_cast<const float2*>(in + i));
      const float2 c1 = __ldcs(reinterpret_cast<const float2*>(in + HW + i));
      const float2 c2 = __ldcs(reinterpret_cast<const float2*>(in + 2 * (size_t)HW + i));
      float xs[2] = {0.f, 0.f}, ys[2] = {0.f, 0.f};
      if (NV == 5) { coords(i, p.W, fW, fH, &xs[0], &ys[0]); coords(i + 1, p.W, fW, fH, &xs[1], &ys[1]); }
      const F2 x[3] = {F2(c0.x, c0.y), F2(c1.x, c1.y), F2(c2.x, c2.y)};
      F2 y[3];
      trispace_fwd<F2, NV, DEG, ROLLED>(x, F2(xs[0], xs[1]), F2(ys[0], ys[1]), s_coef, clamp, y, [&](int sp, int k, F2 v) {
        if constexpr (SAVE) __stcs(reinterpret_cast<float2*>(sg + (size_t)(3 * sp + k) * HW + i), make_float2(v.lo(), v.hi()));
      });
#pragma unroll
      for (int c = 0; c < 3; ++c)
        __stcs(reinterpret_cast<float2*>(out + (size_t)c * HW + i), make_float2(y[c].lo(), y[c].hi()));
    } else {
      const float x[3] = {__ldcs(in + i), __ldcs(in + HW + i), __ldcs(in + 2 * (size_t)HW + i)};
      float xs = 0.f, ys = 0.f, y[3];
      if (NV == 5) coords(i, p.W, fW, fH, &xs, &ys);
      trispace_fwd<float, NV, DEG, ROLLED>(x, xs, ys, s_coef, clamp, y, [&](int sp, int k, float v) {
        if constexpr (SAVE) __stcs(sg + (size_t)(3 * sp + k) * HW + i, v);
      });
#pragma unroll
      for (int c = 0; c < 3; ++c) __stcs(out + (size_t)c * HW + i, y[c]);
    }
  }
}

// ---------------------------------------------------------------------------------------------
// Backward: coefficient gradients  g[s][c][k] = sum_px gp_s_c(px) * monomial_s_k(px), per image a
// [3 x N] . [N x NC] product for each of the three spaces.  Tiles of 128 pixels:
//   phase A (thread = packed pixel pair, first half of the block): recompute the forward, get gp
//            (trispace_gpoly), write the tile's monomials to shared memory T[k][px] (64-bit stores,
//            consecutive threads -> consecutive pairs: conflict-free) and gp to G[c][px];
//   phase B (thread = monomial k): acc[c] += G[c][px] * T[k][px] over the tile's pixels, two pixels
//            per FFMA2 (row pitch 132 words: a quarter-warp's 128-bit loads fall in 32 distinct banks).
// Nine accumulators per thread live across the whole image share; at the end one partial row per
// block goes to global memory, one atomic ticket per block, and the last block of the image folds
// the rows in a fixed order in double (bitwise deterministic, no float atomics).
// ---------------------------------------------------------------------------------------------
#ifndef TRI_BT
#define TRI_BT 128
#endif
constexpr int BT = TRI_BT;           // threads per block = pixels per tile
constexpr int T_PITCH = BT + 4;

template <int NV, int DEG>
constexpr size_t bwd_smem_bytes() {
  return ((size_t)12 * poly_num_coeffs(NV, DEG) + (size_t)poly_num_coeffs(NV, DEG) * T_PITCH + (size_t)BT * 4) * sizeof(float);
}

#ifndef TRI_BWD_MIN_BLOCKS
#define TRI_BWD_MIN_BLOCKS 3   // measured: 2 -> 3.80, 3 -> 3.94, 4 -> 3.83 Gpx/s (4K x 8, 126 coefficients)
#endif
#ifndef TRI_SKEW_NS
#define TRI_SKEW_NS 0
#endif
template <int NV, int DEG, bool SAVED>
__global__ void __launch_bounds__(BT, TRI_BWD_MIN_BLOCKS) trispace_bwd_kernel(const TParams p) {
  constexpr int NC = poly_num_coeffs(NV, DEG);
  constexpr int NG = 9 * NC;
  extern __shared__ __align__(16) float smem[];
  float* s_G = smem;                        // [3][BT] + pad (16-byte aligned)
  float* s_T = smem + BT * 4;               // [NC][T_PITCH]
  Coef4* s_coef = reinterpret_cast<Coef4*>(s_T + NC * T_PITCH);   // [3 * NC] records (16-byte aligned: T_PITCH % 4 == 0)
  __shared__ unsigned s_ticket;
  const int b = blockIdx.y;
  const int t = threadIdx.x;
  stage_coef<NC>(p, b, s_coef);
#if TRI_SKEW_NS > 0
  {
    // The resident blocks of one SM start together and, their phases taking equally long, stay in
    // lock-step: all in the FMA-bound phase A, then all in the shared-memory-bound phase B.  A
    // one-off start delay per residency slot puts them out of phase for the rest of the kernel.
    unsigned nsm;
    asm("mov.u32 %0, %%nsmid;" : "=r"(nsm));
    const unsigned slot = ((blockIdx.y * gridDim.x + blockIdx.x) / nsm) % TRI_BWD_MIN_BLOCKS;
    if (slot) __nanosleep(slot * TRI_SKEW_NS);
  }
#endif
  const int HW = p.HW;
  const float* in = p.img + (size_t)b * 3 * HW;
  const float* go = p.gout + (size_t)b * 3 * HW;
  const float* sg = SAVED ? p.sig + (size_t)b * 9 * HW : nullptr;
  const float fW = (float)p.W, fH = (float)p.H;
  const bool clamp = p.clamp != 0;
  // Phase B mapping: warp w takes the w-th share of the half tile's pixels, lane l the monomials
  // l, l + 32, ..: every 128-bit load of G (a broadcast, two shared-memory wavefronts) then feeds J
  // monomials instead of one.  Accumulators: [space][monomial slot][channel], packed over (even, odd) pixel.
  constexpr int NWARP = BT / 32;
  constexpr int J = (NC + 31) / 32;               // monomials per thread in phase B
  constexpr int QW = BT / 4 / NWARP;              // groups of four pixels per warp
  const int lane = t & 31, warp = t >> 5;
  F2 acc[3][J][3];
#pragma unroll
  for (int sp = 0; sp < 3; ++sp)
#pragma unroll
    for (int j = 0; j < J; ++j) acc[sp][j][0] = acc[sp][j][1] = acc[sp][j][2] = F2(0.0f);
  const int n_tiles = (HW + 2 * BT - 1) / (2 * BT);
  for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    // phase A, part 1: every thread takes one packed pixel pair of the 2*BT-pixel tile (all four
    // scheduler partitions busy), recomputes the forward and keeps vars / gp in registers.
    F2 vars[3][3], gp[3][3], xs2 = F2(0.0f), ys2 = F2(0.0f);
    {
      const int i = tile * 2 * BT + 2 * t;
      const bool v0 = i < HW, v1 = i + 1 < HW;
      F2 x[3], g[3];
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const float a0 = v0 ? __ldcs(in + (size_t)c * HW + i) : 0.5f, a1 = v1 ? __ldcs(in + (size_t)c * HW + i + 1) : 0.5f;
        const float g0 = v0 ? __ldcs(go + (size_t)c * HW + i) : 0.0f, g1 = v1 ? __ldcs(go + (size_t)c * HW + i + 1) : 0.0f;
        x[c] = F2(a0, a1); g[c] = F2(g0, g1);
      }
      if (NV == 5) {
        float xa, ya, xb, yb;
        coords(v0 ? i : 0, p.W, fW, fH, &xa, &ya);
        coords(v1 ? i + 1 : 0, p.W, fW, fH, &xb, &yb);
        xs2 = F2(xa, xb); ys2 = F2(ya, yb);
      }
      F2 sv[9];
      if constexpr (SAVED) {
#pragma unroll
        for (int c = 0; c < 9; ++c) {
          const float s0 = v0 ? __ldcs(sg + (size_t)c * HW + i) : 0.5f, s1 = v1 ? __ldcs(sg + (size_t)c * HW + i + 1) : 0.5f;
          sv[c] = F2(s0, s1);
        }
      }
      trispace_gpoly<F2, NV, DEG>(x, xs2, ys2, s_coef, clamp, g, vars, gp, SAVED ? sv : nullptr);
    }
    // part 2 + phase B, per space and per half tile: the half's BT/2 threads write their monomials
    // and gp to shared memory, then all threads multiply.  The halves alternate between the lower
    // and the upper warps, so the store work is spread over the partitions as well.
#pragma unroll
    for (int sp = 0; sp < 3; ++sp)
#pragma unroll 1
    for (int half = 0; half < 2; ++half) {
      __syncthreads();                               // the previous product is done with T, G
      if ((t >= BT / 2) == (half == 1)) {
        const int tl = t - half * (BT / 2);
        F2 v[5] = {vars[sp][0], vars[sp][1], vars[sp][2], xs2, ys2};
#pragma unroll
        for (int q = 0; q < 5; ++q) asm volatile("" : "+l"(v[q].u));   // keeps the monomial products inside this
        F2* col = reinterpret_cast<F2*>(s_T + 2 * tl);                  // branch (else hoisted out of the half loop into local memory)
        col[0] = F2(1.0f);                           // the constant term
        MonoStore<F2> st{col, T_PITCH / 2};
        mono_visit<F2, NV, DEG>(v, st);
#pragma unroll
        for (int c = 0; c < 3; ++c) reinterpret_cast<F2*>(s_G + c * BT)[tl] = gp[sp][c];
      }
      __syncthreads();
#pragma unroll 2
      for (int qi = 0; qi < QW; ++qi) {
        const int q = warp * QW + qi;
        const float4 g0 = reinterpret_cast<const float4*>(s_G)[q];                  // channel 0, four pixels
        const float4 g1 = reinterpret_cast<const float4*>(s_G + BT)[q];
        const float4 g2 = reinterpret_cast<const float4*>(s_G + 2 * BT)[q];
#pragma unroll
        for (int j = 0; j < J; ++j) {
          // surplus slots redo a row that exists and falls in other banks than the live rows 32 * j + (0..): row
          // `lane` when there are 32 rows (row 0 would share its banks with row 32: +3 % on the 35-coefficient
          // backward), row 0 for the small polynomials; never stored
          const int k = lane + 32 * j < NC ? lane + 32 * j : (NC >= 32 ? lane : 0);
          const float4 m = reinterpret_cast<const float4*>(s_T + k * T_PITCH)[q];
          const F2 mlo(m.x, m.y), mhi(m.z, m.w);
          acc[sp][j][0] = vfma(F2(g0.x, g0.y), mlo, acc[sp][j][0]);
          acc[sp][j][1] = vfma(F2(g1.x, g1.y), mlo, acc[sp][j][1]);
          acc[sp][j][2] = vfma(F2(g2.x, g2.y), mlo, acc[sp][j][2]);
          acc[sp][j][0] = vfma(F2(g0.z, g0.w), mhi, acc[sp][j][0]);
          acc[sp][j][1] = vfma(F2(g1.z, g1.w), mhi, acc[sp][j][1]);
          acc[sp][j][2] = vfma(F2(g2.z, g2.w), mhi, acc[sp][j][2]);
        }
      }
    }
  }
  // fold the warps' shares: through shared memory (T is free now), fixed order
  __syncthreads();
  float* s_red = s_T;                               // [NWARP][NG]
#pragma unroll
  for (int j = 0; j < J; ++j) {
    const int k = lane + 32 * j;
    if (k < NC) {
#pragma unroll
      for (int sp = 0; sp < 3; ++sp)
#pragma unroll
        for (int c = 0; c < 3; ++c) s_red[warp * NG + (sp * 3 + c) * NC + k] = vhsum(acc[sp][j][c]);
    }
  }
  __syncthreads();
  float* row = p.partials + ((size_t)b * gridDim.x + blockIdx.x) * NG;
  for (int k = t; k < NG; k += BT) {
    float a = 0.0f;
#pragma unroll
    for (int w = 0; w < NWARP; ++w) a += s_red[w * NG + k];
    row[k] = a;
  }
  __threadfence();
  __syncthreads();
  if (t == 0) s_ticket = atomicAdd(&p.tickets[b], 1u);
  __syncthreads();
  if (s_ticket != gridDim.x - 1) return;
  __threadfence();
  const float* rows = p.partials + (size_t)b * gridDim.x * NG;
  for (int k = t; k < NG; k += BT) {
    double a = 0.0;
    for (unsigned r = 0; r < gridDim.x; ++r) a += (double)__ldcg(rows + (size_t)r * NG + k);
    const int s = k / (3 * NC), j = k - s * 3 * NC;
    float* dst = s == 0 ? p.gcoef[0] : (s == 1 ? p.gcoef[1] : p.gcoef[2]);   // no run-time index into the parameter block
    dst[(size_t)b * 3 * NC + j] = (float)a;
  }
}

// ---------------------------------------------------------------------------------------------
// Backward w.r.t. the image: element-wise (no reduction), one packed pixel pair per thread and
// iteration like the forward's VEC = 2 path; VEC = 1 for odd HW or unaligned bases.  p.out = gimg.
// The reference's scripts never ask for this gradient (the image is data), so this kernel is about
// the contract (autograd through model.py:486-508), not a headline: no tuning beyond the forward's.
// ---------------------------------------------------------------------------------------------
template <int NV, int DEG, int VEC, bool SAVED>
__global__ void __launch_bounds__(THREADS, 2) trispace_gimg_kernel(const TParams p) {
  constexpr int NC = poly_num_coeffs(NV, DEG);
  __shared__ Coef4 s_coef[3 * NC];
  const int b = blockIdx.y;
  stage_coef<NC>(p, b, s_coef);
  const int HW = p.HW;
  const float* in = p.img + (size_t)b * 3 * HW;
  const float* go = p.gout + (size_t)b * 3 * HW;
  const float* sg = SAVED ? p.sig + (size_t)b * 9 * HW : nullptr;
  float* gi = p.out + (size_t)b * 3 * HW;
  const float fW = (float)p.W, fH = (float)p.H;
  const bool clamp = p.clamp != 0;
  const int stride = gridDim.x * THREADS * VEC;
  for (int i = (blockIdx.x * THREADS + threadIdx.x) * VEC; i < HW; i += stride) {
    if constexpr (VEC == 2) {
      F2 x[3], g[3], sv[SAVED ? 9 : 1], y[3];
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const float2 a = __ldcs(reinterpret_cast<const float2*>(in + (size_t)c * HW + i));
        const float2 q = __ldcs(reinterpret_cast<const float2*>(go + (size_t)c * HW + i));
        x[c] = F2(a.x, a.y); g[c] = F2(q.x, q.y);
      }
      if constexpr (SAVED) {
#pragma unroll
        for (int c = 0; c < 9; ++c) {
          const float2 a = __ldcs(reinterpret_cast<const float2*>(sg + (size_t)c * HW + i));
          sv[c] = F2(a.x, a.y);
        }
      }
      float xs[2] = {0.f, 0.f}, ys[2] = {0.f, 0.f};
      if (NV == 5) { coords(i, p.W, fW, fH, &xs[0], &ys[0]); coords(i + 1, p.W, fW, fH, &xs[1], &ys[1]); }
      trispace_gimg<F2, NV, DEG>(x, F2(xs[0], xs[1]), F2(ys[0], ys[1]), s_coef, clamp, g, y, SAVED ? sv : nullptr);
#pragma unroll
      for (int c = 0; c < 3; ++c)
        __stcs(reinterpret_cast<float2*>(gi + (size_t)c * HW + i), make_float2(y[c].lo(), y[c].hi()));
    } else {
      float x[3], g[3], sv[SAVED ? 9 : 1], y[3], xs = 0.f, ys = 0.f;
#pragma unroll
      for (int c = 0; c < 3; ++c) { x[c] = __ldcs(in + (size_t)c * HW + i); g[c] = __ldcs(go + (size_t)c * HW + i); }
      if constexpr (SAVED) {
#pragma unroll
        for (int c = 0; c < 9; ++c) sv[c] = __ldcs(sg + (size_t)c * HW + i);
      }
      if (NV == 5) coords(i, p.W, fW, fH, &xs, &ys);
      trispace_gimg<float, NV, DEG>(x, xs, ys, s_coef, clamp, g, y, SAVED ? sv : nullptr);
#pragma unroll
      for (int c = 0; c < 3; ++c) __stcs(gi + (size_t)c * HW + i, y[c]);
    }
  }
}

// ---------------------------------------------------------------------------------------------
// Host side
// ---------------------------------------------------------------------------------------------
static bool aligned16(const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15u) == 0; }
static bool aligned4(const void* q) { return (reinterpret_cast<uintptr_t>(q) & 3u) == 0; }

static int grid_x(int B, int HW, int px_per_block, int blocks_per_sm) {
  const long long tiles = ((long long)HW + px_per_block - 1) / px_per_block;
  long long cap = ((long long)sm_count() * blocks_per_sm + B - 1) / B;
  long long n = tiles < cap ? tiles : cap;
  return host::apply_grid_cap(n);          // test hook (curl_test_set_grid_cap), a no-op by default
}
constexpr int FWD_BLOCKS_PER_SM = 16;
constexpr int BWD_BLOCKS_PER_SM = 12;   // 3 resident blocks per SM (73 KB of shared memory each) x 4 waves
static int bwd_grid_x(int B, int HW) { return grid_x(B, HW, 2 * BT, BWD_BLOCKS_PER_SM); }

static int check_args(const float* img, const float* R, const float* L, const float* Hc, int degree, int spatial,
                      int B, int H, int W) {
  if (B <= 0 || H <= 0 || W <= 0 || (long long)H * W > 0x3fffffffLL || B > 65535) return CURL_ERR_BAD_ARG;
  if (!img || !R || !L || !Hc) return CURL_ERR_BAD_ARG;
  if (degree < 1 || degree > MAX_DEG || (spatial != 0 && spatial != 1)) return CURL_ERR_BAD_ARG;
  if (!aligned4(img) || !aligned4(R) || !aligned4(L) || !aligned4(Hc)) return CURL_ERR_MISALIGNED;
  return CURL_OK;
}

template <int NV, int DEG>
static int launch_fwd(const TParams& p, int B, int vec, cudaStream_t st) {
  // four pixels per thread halve the coefficient reads (the forward's shared-memory pipe is 77 % busy at
  // two).  Measured at 4K x 8: 35 coefficients 46.4 -> 50.6 Gpx/s (53.2 without saving); 126 coefficients
  // 20.2 -> 17.4 (the nested evaluation of 126 terms for four pixels no longer fits 128 registers), so the
  // spatial polynomial stays at two.
#ifndef TRI_SPATIAL_VEC4
  if (NV == 5 && DEG == 4 && vec == 4) vec = 2;
#endif
  dim3 grid(grid_x(B, p.HW, THREADS * vec, FWD_BLOCKS_PER_SM), B);
  if (p.sig) {
    if (vec == 4) trispace_fwd_kernel<NV, DEG, 4, true><<<grid, THREADS, 0, st>>>(p);
    else if (vec == 2) trispace_fwd_kernel<NV, DEG, 2, true><<<grid, THREADS, 0, st>>>(p);
    else trispace_fwd_kernel<NV, DEG, 1, true><<<grid, THREADS, 0, st>>>(p);
  } else {
    if (vec == 4) trispace_fwd_kernel<NV, DEG, 4, false><<<grid, THREADS, 0, st>>>(p);
    else if (vec == 2) trispace_fwd_kernel<NV, DEG, 2, false><<<grid, THREADS, 0, st>>>(p);
    else trispace_fwd_kernel<NV, DEG, 1, false><<<grid, THREADS, 0, st>>>(p);
  }
  return cudaGetLastError() == cudaSuccess ? CURL_OK : CURL_ERR_LAUNCH;
}

template <int NV, int DEG>
static int launch_bwd(const TParams& p, int B, cudaStream_t st) {
  auto kern = p.sig ? trispace_bwd_kernel<NV, DEG, true> : trispace_bwd_kernel<NV, DEG, false>;
  const bool ok = p.sig ? host::ensure_dyn_smem<trispace_bwd_kernel<NV, DEG, true>>(bwd_smem_bytes<NV, DEG>())
                        : host::ensure_dyn_smem<trispace_bwd_kernel<NV, DEG, false>>(bwd_smem_bytes<NV, DEG>());
  if (!ok) return CURL_ERR_LAUNCH;
  dim3 grid(bwd_grid_x(B, p.HW), B);
  kern<<<grid, BT, bwd_smem_bytes<NV, DEG>(), st>>>(p);
  return cudaGetLastError() == cudaSuccess ? CURL_OK : CURL_ERR_LAUNCH;
}

template <int NV, int DEG>
static int launch_gimg(const TParams& p, int B, int vec, cudaStream_t st) {
  dim3 grid(grid_x(B, p.HW, THREADS * vec, FWD_BLOCKS_PER_SM), B);
  if (p.sig) {
    if (vec == 2) trispace_gimg_kernel<NV, DEG, 2, true><<<grid, THREADS, 0, st>>>(p);
    else trispace_gimg_kernel<NV, DEG, 1, true><<<grid, THREADS, 0, st>>>(p);
  } else {
    if (vec == 2) trispace_gimg_kernel<NV, DEG, 2, false><<<grid, THREADS, 0, st>>>(p);
    else trispace_gimg_kernel<NV, DEG, 1, false><<<grid, THREADS, 0, st>>>(p);
  }
  return cudaGetLastError() == cudaSuccess ? CURL_OK : CURL_ERR_LAUNCH;
}

}  // namespace tri
}  // namespace curl

using namespace curl;
using namespace curl::tri;

extern "C" {

int curl_trispace_num_coeffs(int degree, int spatial) {
  if (degree < 0 || degree > 8 || (spatial != 0 && spatial != 1)) return 0;
  return poly_num_coeffs(spatial ? 5 : 3, degree);
}

int curl_trispace_fwd_f32(const float* img, const float* R, const float* L, const float* Hc, int degree,
                          int spatial, int final_clamp, float* out, float* saved, int B, int H, int W, void* stream) {
  int rc = check_args(img, R, L, Hc, degree, spatial, B, H, W);
  if (rc != CURL_OK) return rc;
  if (!out) return CURL_ERR_BAD_ARG;
  if (!aligned4(out) || !aligned4(saved)) return CURL_ERR_MISALIGNED;
  TParams p{};
  p.img = img; p.coef[0] = R; p.coef[1] = L; p.coef[2] = Hc; p.out = out; p.sig = saved;
  p.HW = H * W; p.W = W; p.H = H; p.clamp = final_clamp ? 1 : 0;
  const uintptr_t bits = reinterpret_cast<uintptr_t>(img) | reinterpret_cast<uintptr_t>(out) | reinterpret_cast<uintptr_t>(saved);
  int vec = 1;
  if (p.HW % 2 == 0 && (bits & 7u) == 0) vec = 2;
#ifndef TRI_FWD_NO_VEC4
  if (p.HW % 4 == 0 && (bits & 15u) == 0) vec = 4;
#endif
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
#define CURL_TRI_FWD(D) case D: return spatial ? launch_fwd<5, D>(p, B, vec, st) : launch_fwd<3, D>(p, B, vec, st)
  switch (degree) { CURL_TRI_FWD(1); CURL_TRI_FWD(2); CURL_TRI_FWD(3); default: CURL_TRI_FWD(4); }
#undef CURL_TRI_FWD
}

size_t curl_trispace_bwd_workspace_bytes(int B, int H, int W, int degree, int spatial) {
  if (B <= 0 || H <= 0 || W <= 0 || degree < 1 || degree > MAX_DEG) return 0;
  const size_t tick_bytes = ((size_t)B * sizeof(unsigned) + 255) / 256 * 256;
  return tick_bytes + (size_t)B * bwd_grid_x(B, H * W) * 9 * poly_num_coeffs(spatial ? 5 : 3, degree) * sizeof(float);
}

int curl_trispace_bwd_f32(const float* img, const float* R, const float* L, const float* Hc, int degree,
                          int spatial, int final_clamp, const float* gout, const float* saved, float* gR, float* gL,
                          float* gH, void* workspace, size_t workspace_bytes, int B, int H, int W, void* stream) {
  int rc = check_args(img, R, L, Hc, degree, spatial, B, H, W);
  if (rc != CURL_OK) return rc;
  if (!gout || !gR || !gL || !gH) return CURL_ERR_BAD_ARG;
  if (!aligned4(gout) || !aligned4(saved)) return CURL_ERR_MISALIGNED;
  const size_t need = curl_trispace_bwd_workspace_bytes(B, H, W, degree, spatial);
  if (!workspace || workspace_bytes < need || !aligned16(workspace)) return CURL_ERR_WORKSPACE;
  const size_t tick_bytes = ((size_t)B * sizeof(unsigned) + 255) / 256 * 256;
  TParams p{};
  p.img = img; p.coef[0] = R; p.coef[1] = L; p.coef[2] = Hc; p.gout = gout;
  p.gcoef[0] = gR; p.gcoef[1] = gL; p.gcoef[2] = gH;
  p.sig = const_cast<float*>(saved);
  p.tickets = reinterpret_cast<unsigned*>(workspace);
  p.partials = reinterpret_cast<float*>(reinterpret_cast<char*>(workspace) + tick_bytes);
  p.HW = H * W; p.W = W; p.H = H; p.clamp = final_clamp ? 1 : 0;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (cudaMemsetAsync(workspace, 0, tick_bytes, st) != cudaSuccess) return CURL_ERR_LAUNCH;
#define CURL_TRI_BWD(D) case D: return spatial ? launch_bwd<5, D>(p, B, st) : launch_bwd<3, D>(p, B, st)
  switch (degree) { CURL_TRI_BWD(1); CURL_TRI_BWD(2); CURL_TRI_BWD(3); default: CURL_TRI_BWD(4); }
#undef CURL_TRI_BWD
}

int curl_trispace_bwd_img_f32(const float* img, const float* R, const float* L, const float* Hc, int degree,
                              int spatial, int final_clamp, const float* gout, const float* saved, float* gimg,
                              int B, int H, int W, void* stream) {
  int rc = check_args(img, R, L, Hc, degree, spatial, B, H, W);
  if (rc != CURL_OK) return rc;
  if (!gout || !gimg) return CURL_ERR_BAD_ARG;
  if (!aligned4(gout) || !aligned4(saved) || !aligned4(gimg)) return CURL_ERR_MISALIGNED;
  TParams p{};
  p.img = img; p.coef[0] = R; p.coef[1] = L; p.coef[2] = Hc; p.gout = gout; p.out = gimg;
  p.sig = const_cast<float*>(saved);
  p.HW = H * W; p.W = W; p.H = H; p.clamp = final_clamp ? 1 : 0;
  const uintptr_t bits = reinterpret_cast<uintptr_t>(img) | reinterpret_cast<uintptr_t>(gout) |
                         reinterpret_cast<uintptr_t>(saved) | reinterpret_cast<uintptr_t>(gimg);
  const int vec = (p.HW % 2 == 0 && (bits & 7u) == 0) ? 2 : 1;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
#define CURL_TRI_GIMG(D) case D: return spatial ? launch_gimg<5, D>(p, B, vec, st) : launch_gimg<3, D>(p, B, vec, st)
  switch (degree) { CURL_TRI_GIMG(1); CURL_TRI_GIMG(2); CURL_TRI_GIMG(3); default: CURL_TRI_GIMG(4); }
#undef CURL_TRI_GIMG
}

}  // extern "C"
