// Lane types for the per-pixel arithmetic.
//
//   float : one pixel per value.  On the device the transcendental helpers are single MUFU ops
//           (MUFU.LG2 / MUFU.EX2 / MUFU.RCP); on the host (tests/hostsim only -- never the product)
//           they are double-precision libm rounded to fp32.
//   F2    : two pixels per value in one 64-bit register pair (device only).  sm_100a has packed
//           fp32 arithmetic (PTX fma/mul/add.rn.f32x2 -> SASS FFMA2/FMUL2/FADD2): one issue slot does
//           the work of two scalar instructions, and a scalar register or an immediate is broadcast
//           to both halves for free.  The layer kernels are issue-bound, not HBM-bound, so this is
//           the main lever.  Operations with no packed form (MUFU, min/max, compare/select, .SAT)
//           are done per half.
//
// Everything is written once as templates over the lane type T (curl_pixel.cuh) using the v*
// helpers below; "gates" (clamp-pass indicators, branch selectors) are carried as 0/1 values of
// type T so that applying them is a (packable) multiply.
#pragma once

#if defined(__CUDACC__)
#define CURL_HD __host__ __device__ __forceinline__
#define CURL_D __device__ __forceinline__
#else
#define CURL_HD inline
#endif

#include <math.h>

namespace curl {

// ------------------------------------------------------------------------------------------------
// float lanes
// ------------------------------------------------------------------------------------------------
CURL_HD float vadd(float a, float b) { return a + b; }
CURL_HD float vsub(float a, float b) { return a - b; }
CURL_HD float vmul(float a, float b) { return a * b; }
CURL_HD float vfma(float a, float b, float c) { return fmaf(a, b, c); }
CURL_HD float vneg(float a) { return -a; }

CURL_HD float vlg2(float x) {
#if defined(__CUDA_ARCH__)
  float y; asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y;
#else
  return (float)log2((double)x);
#endif
}
CURL_HD float vex2(float x) {
#if defined(__CUDA_ARCH__)
  float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y;
#else
  return (float)exp2((double)x);
#endif
}
CURL_HD float vrcp(float x) {
#if defined(__CUDA_ARCH__)
  float y; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y;
#else
  return (float)(1.0 / (double)x);
#endif
}
// clamp to [0,1]; NaN -> 0 (FADD.SAT semantics), mirrored on the host
CURL_HD float vsat(float x) {
#if defined(__CUDA_ARCH__)
  return __saturatef(x);
#else
  return x < 0.f ? 0.f : (x > 1.f ? 1.f : (x == x ? x : 0.f));
#endif
}
CURL_HD float vmul_sat(float a, float b) { return vsat(a * b); }
CURL_HD float vadd_sat(float a, float b) { return vsat(a + b); }
CURL_HD float vfma_sat(float a, float b, float c) { return vsat(fmaf(a, b, c)); }
// |a| + c and its clamp: one FADD / FADD.SAT with the |.| operand modifier
CURL_HD float vabs_add(float a, float c) { return fabsf(a) + c; }
CURL_HD float vabs_add_sat(float a, float c) { return vsat(fabsf(a) + c); }
CURL_HD float vnabs_add(float a, float c) { return c - fabsf(a); }
CURL_HD float vnabs_add_sat(float a, float c) { return vsat(c - fabsf(a)); }
// magnitude of g with the sign of t / of -t (one LOP3)
CURL_HD float vcopysign(float g, float t) { return copysignf(g, t); }
CURL_HD float vcopysign_neg(float g, float t) { return copysignf(g, -t); }
CURL_HD float vmax(float a, float b) {
#if defined(__CUDA_ARCH__)
  return fmaxf(a, b);
#else
  return a > b ? a : b;
#endif
}
CURL_HD float vmin(float a, float b) {
#if defined(__CUDA_ARCH__)
  return fminf(a, b);
#else
  return a < b ? a : b;
#endif
}
// 0/1 gates and selects
CURL_HD float vgate_le(float a, float b) { return a <= b ? 1.0f : 0.0f; }
CURL_HD float vgate_lt(float a, float b) { return a < b ? 1.0f : 0.0f; }
CURL_HD float vgate_eq(float a, float b) { return a == b ? 1.0f : 0.0f; }
CURL_HD float vgate_ne(float a, float b) { return a != b ? 1.0f : 0.0f; }
CURL_HD float vsel_le(float a, float b, float t, float f) { return a <= b ? t : f; }   // a <= b ? t : f
CURL_HD float vsel_lt(float a, float b, float t, float f) { return a < b ? t : f; }
CURL_HD float vsel_eq(float a, float b, float t, float f) { return a == b ? t : f; }
CURL_HD float vhsum(float a) { return a; }   // horizontal sum over the lanes of one value

// 8-bit image I/O (reference: torchvision TF.to_tensor = u8 / 255 in fp32; (x * 255) truncated to u8
// on the way out, evaluate.py:63).  u8_to_unit returns the correctly rounded quotient without a
// division: one Newton step on the residual, verified against f / 255.0f for all 256 inputs
// (tests/test_hostsim.py).  unit_to_u8 expects 0 <= y <= 1.
template <class T>
CURL_HD T u8_to_unit(T f) {   // f: an exact integer 0..255 held as a float
  const T r = T((float)(1.0 / 255.0));
  const T q = vmul(f, r);
  const T rem = vfma(q, T(-255.0f), f);
  return vfma(rem, r, q);
}
CURL_HD unsigned unit_to_u8(float y) {
  const float v = y * 255.0f;             // rounded to nearest like torch's mul, then truncated
#if defined(__CUDA_ARCH__)
  return __float_as_uint(__fadd_rd(v, 8388608.0f)) & 0xffu;   // floor via the 2^23 trick: no F2I
#else
  return (unsigned)v;
#endif
}

// ------------------------------------------------------------------------------------------------
// F2 lanes (device only)
// ------------------------------------------------------------------------------------------------
#if defined(__CUDACC__)
struct F2 {
  unsigned long long u;
  F2() = default;
  CURL_D F2(float c) { asm("mov.b64 %0, {%1,%1};" : "=l"(u) : "f"(c)); }          // broadcast
  CURL_D F2(float lo, float hi) { asm("mov.b64 %0, {%1,%2};" : "=l"(u) : "f"(lo), "f"(hi)); }
  CURL_D float lo() const { float a, b; asm("mov.b64 {%0,%1}, %2;" : "=f"(a), "=f"(b) : "l"(u)); return a; }
  CURL_D float hi() const { float a, b; asm("mov.b64 {%0,%1}, %2;" : "=f"(a), "=f"(b) : "l"(u)); return b; }
};
#ifdef ABL_SCALAR
CURL_D F2 vadd(F2 a, F2 b) { return F2(a.lo() + b.lo(), a.hi() + b.hi()); }
CURL_D F2 vmul(F2 a, F2 b) { return F2(a.lo() * b.lo(), a.hi() * b.hi()); }
CURL_D F2 vfma(F2 a, F2 b, F2 c) { return F2(fmaf(a.lo(), b.lo(), c.lo()), fmaf(a.hi(), b.hi(), c.hi())); }
#else
CURL_D F2 vadd(F2 a, F2 b) { F2 r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r.u) : "l"(a.u), "l"(b.u)); return r; }
CURL_D F2 vmul(F2 a, F2 b) { F2 r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r.u) : "l"(a.u), "l"(b.u)); return r; }
CURL_D F2 vfma(F2 a, F2 b, F2 c) {
  F2 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r.u) : "l"(a.u), "l"(b.u), "l"(c.u)); return r;
}
#endif
CURL_D F2 vsub(F2 a, F2 b) { return vfma(b, F2(-1.0f), a); }
CURL_D F2 vneg(F2 a) { return vmul(a, F2(-1.0f)); }

#define CURL_F2_MAP1(name)                                            \
  CURL_D F2 name(F2 a) { return F2(name(a.lo()), name(a.hi())); }
#define CURL_F2_MAP2(name)                                            \
  CURL_D F2 name(F2 a, F2 b) { return F2(name(a.lo(), b.lo()), name(a.hi(), b.hi())); }
#define CURL_F2_MAP3(name)                                            \
  CURL_D F2 name(F2 a, F2 b, F2 c) { return F2(name(a.lo(), b.lo(), c.lo()), name(a.hi(), b.hi(), c.hi())); }
#define CURL_F2_MAP4(name)                                            \
  CURL_D F2 name(F2 a, F2 b, F2 c, F2 d) {                            \
    return F2(name(a.lo(), b.lo(), c.lo(), d.lo()), name(a.hi(), b.hi(), c.hi(), d.hi())); }
#ifdef ABL_NOMUFU
CURL_D F2 vlg2(F2 a) { return a; }
CURL_D F2 vex2(F2 a) { return a; }
CURL_D F2 vrcp(F2 a) { return a; }
#else
CURL_F2_MAP1(vlg2) CURL_F2_MAP1(vex2) CURL_F2_MAP1(vrcp)
#endif
#ifdef ABL_NOSAT
CURL_D F2 vsat(F2 a) { return a; }
#else
CURL_F2_MAP1(vsat)
#endif
CURL_F2_MAP2(vmul_sat) CURL_F2_MAP2(vadd_sat) CURL_F2_MAP3(vfma_sat)
CURL_F2_MAP2(vmax) CURL_F2_MAP2(vmin)
CURL_D F2 vabs_add(F2 a, float c) { return F2(vabs_add(a.lo(), c), vabs_add(a.hi(), c)); }
CURL_D F2 vabs_add_sat(F2 a, float c) { return F2(vabs_add_sat(a.lo(), c), vabs_add_sat(a.hi(), c)); }
CURL_D F2 vnabs_add(F2 a, float c) { return F2(vnabs_add(a.lo(), c), vnabs_add(a.hi(), c)); }
CURL_D F2 vnabs_add_sat(F2 a, float c) { return F2(vnabs_add_sat(a.lo(), c), vnabs_add_sat(a.hi(), c)); }
CURL_F2_MAP2(vcopysign) CURL_F2_MAP2(vcopysign_neg)
#ifdef ABL_NOGATE
CURL_D F2 vgate_le(F2 a, F2 b) { return F2(1.0f); }
CURL_D F2 vgate_lt(F2 a, F2 b) { return F2(1.0f); }
CURL_D F2 vgate_eq(F2 a, F2 b) { return F2(1.0f); }
CURL_D F2 vgate_ne(F2 a, F2 b) { return F2(1.0f); }
#else
CURL_F2_MAP2(vgate_le) CURL_F2_MAP2(vgate_lt) CURL_F2_MAP2(vgate_eq) CURL_F2_MAP2(vgate_ne)
#endif
CURL_F2_MAP4(vsel_le) CURL_F2_MAP4(vsel_lt) CURL_F2_MAP4(vsel_eq)
CURL_D float vhsum(F2 a) { return a.lo() + a.hi(); }

// ------------------------------------------------------------------------------------------------
// F4: two F2 side by side (four pixels per value).  For kernels whose inner loop is bound by
// per-pixel-group shared-memory reads of broadcast operands (the TriSpace coefficients): one read then
// feeds four pixels instead of two.  Every operation is the F2 one on both halves.
// ------------------------------------------------------------------------------------------------
struct F4 {
  F2 a, b;
  F4() = default;
  CURL_D F4(float c) : a(c), b(c) {}
  CURL_D F4(F2 x, F2 y) : a(x), b(y) {}
};
#define CURL_F4_MAP1(name) CURL_D F4 name(F4 x) { return F4(name(x.a), name(x.b)); }
#define CURL_F4_MAP2(name) CURL_D F4 name(F4 x, F4 y) { return F4(name(x.a, y.a), name(x.b, y.b)); }
#define CURL_F4_MAP3(name) CURL_D F4 name(F4 x, F4 y, F4 z) { return F4(name(x.a, y.a, z.a), name(x.b, y.b, z.b)); }
#define CURL_F4_MAP4(name) CURL_D F4 name(F4 x, F4 y, F4 z, F4 w) { return F4(name(x.a, y.a, z.a, w.a), name(x.b, y.b, z.b, w.b)); }
CURL_F4_MAP2(vadd) CURL_F4_MAP2(vsub) CURL_F4_MAP2(vmul) CURL_F4_MAP3(vfma) CURL_F4_MAP1(vneg)
CURL_F4_MAP1(vlg2) CURL_F4_MAP1(vex2) CURL_F4_MAP1(vrcp) CURL_F4_MAP1(vsat)
CURL_F4_MAP2(vmul_sat) CURL_F4_MAP2(vadd_sat) CURL_F4_MAP3(vfma_sat) CURL_F4_MAP2(vmax) CURL_F4_MAP2(vmin)
CURL_F4_MAP2(vgate_le) CURL_F4_MAP2(vgate_lt) CURL_F4_MAP2(vgate_eq) CURL_F4_MAP2(vgate_ne)
CURL_D F4 vabs_add(F4 x, float c) { return F4(vabs_add(x.a, c), vabs_add(x.b, c)); }
CURL_D F4 vabs_add_sat(F4 x, float c) { return F4(vabs_add_sat(x.a, c), vabs_add_sat(x.b, c)); }
CURL_D F4 vnabs_add(F4 x, float c) { return F4(vnabs_add(x.a, c), vnabs_add(x.b, c)); }
CURL_D F4 vnabs_add_sat(F4 x, float c) { return F4(vnabs_add_sat(x.a, c), vnabs_add_sat(x.b, c)); }
CURL_F4_MAP2(vcopysign) CURL_F4_MAP2(vcopysign_neg)
CURL_F4_MAP4(vsel_le) CURL_F4_MAP4(vsel_lt) CURL_F4_MAP4(vsel_eq)
CURL_D float vhsum(F4 x) { return vhsum(x.a) + vhsum(x.b); }
#endif  // __CUDACC__

}  // namespace curl
