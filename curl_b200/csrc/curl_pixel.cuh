// Per-pixel arithmetic of the CURL curve layer: colour conversions, affine knot curves, and their
// hand-derived reverse-mode derivatives.  Everything here is a pure function of one pixel (or, with
// the packed lane type F2, of two pixels) plus the per-image curve coefficients.  It is written once
// as templates over the lane type T (curl_vec.cuh) and instantiated
//   * in the sm_100a kernels (curl_kernels.cu) with T = F2 (FFMA2/FMUL2/FADD2 + MUFU) and
//     T = float (unaligned tail path),
//   * in the host-side derivation check (tests/hostsim) with T = float on the CPU, which is test
//     infrastructure only and is never linked into the product library.
//
// Semantics follow the reference (danielbulhosa/CURL) line by line; citations are file:line in
// that tree.  Where the reference evaluates "both branches * 0/1 mask" the kernel uses a select,
// which is value-identical for finite data.
//
// Backward convention: every forward stage can emit a small "Jac" record holding the local
// derivative factors with all clamp gates (inclusive: torch.clamp passes gradient on [lo, hi]) and
// branch selectors already multiplied in, so the reverse sweep is straight-line packed arithmetic.
#pragma once

#include "curl_vec.cuh"

namespace curl {

// ------------------------------------------------------------------------------------------------
// Constants (reference colors.py).  Folded in double, rounded once to fp32.
// ------------------------------------------------------------------------------------------------
namespace k {
constexpr double EPS = 6.0 / 29.0;                       // colors.py:43,106
constexpr float SRGB_T = 0.04045f;                       // colors.py:37
constexpr float SRGB_INV_SLOPE = (float)(1.0 / 12.92);   // colors.py:37
constexpr float SRGB_A = (float)(1.0 / 1.055);           // colors.py:38  (x + 0.055) / 1.055
constexpr float SRGB_B = (float)(0.055 / 1.055);
constexpr float SRGB_DK = (float)(2.4 / 1.055);          // d/dx u^2.4 = 2.4/1.055 u^1.4
constexpr float F_T = (float)(EPS * EPS * EPS);          // colors.py:45
constexpr float F_SLOPE = (float)(1.0 / (3.0 * EPS * EPS));
constexpr float F_OFF = (float)(4.0 / 29.0);
constexpr float FINV_T = (float)EPS;                     // colors.py:110
constexpr float FINV_SLOPE = (float)(3.0 * EPS * EPS);
constexpr float FINV_OFF = (float)(-3.0 * EPS * EPS * 4.0 / 29.0);
constexpr float LIN_T = 0.0031308f;                      // colors.py:118
constexpr float LIN_SLOPE = 12.92f;
constexpr float GAM_E = (float)(1.0 / 2.4 - 1.0);        // v^(1/2.4) = v * v^(1/2.4 - 1)
constexpr float GAM_DK = (float)(1.055 / 2.4);
// RGB -> XYZ (colors.py:8-13), each column pre-divided by the D65 white (colors.py:24,41).
constexpr double WX = 0.950456, WY = 1.0, WZ = 1.088754;
constexpr float M00 = (float)(0.412453 / WX), M10 = (float)(0.357580 / WX), M20 = (float)(0.180423 / WX);
constexpr float M01 = (float)(0.212671 / WY), M11 = (float)(0.715160 / WY), M21 = (float)(0.072169 / WY);
constexpr float M02 = (float)(0.019334 / WZ), M12 = (float)(0.119193 / WZ), M22 = (float)(0.950227 / WZ);
// f -> normalised Lab (colors.py:16-25,50,57-59):
//   Ln = (116 fy - 16)/100 ; an = (500 (fx-fy)/110 + 1)/2 ; bn = (200 (fy-fz)/110 + 1)/2
constexpr float LN_S = 1.16f, LN_O = -0.16f;
constexpr float AN_S = (float)(500.0 / 220.0), BN_S = (float)(200.0 / 220.0);
// normalised Lab -> f (colors.py:77-86,97-106):
//   fy = (100 Ln + 16)/116 ; fx = fy + (2 an - 1) 110/500 ; fz = fy - (2 bn - 1) 110/200
constexpr float FY_S = (float)(100.0 / 116.0), FY_O = (float)(16.0 / 116.0);
constexpr float FX_S = (float)(220.0 / 500.0), FX_O = (float)(-110.0 / 500.0);
constexpr float FZ_S = (float)(-220.0 / 200.0), FZ_O = (float)(110.0 / 200.0);
// XYZ -> linear RGB (colors.py:69-74), each row pre-multiplied by the white (colors.py:85,114).
constexpr float N00 = (float)(3.2404542 * WX), N10 = (float)(-1.5371385 * WY), N20 = (float)(-0.4985314 * WZ);
constexpr float N01 = (float)(-0.9692660 * WX), N11 = (float)(1.8760108 * WY), N21 = (float)(0.0415560 * WZ);
constexpr float N02 = (float)(0.0556434 * WX), N12 = (float)(-0.2040259 * WY), N22 = (float)(1.0572252 * WZ);
constexpr float HSV_LO = 1e-9f;                          // colors.py:205,240
// Reverse-sweep constants with the factors of the slimmed Jac records folded in (in double):
//   LabJac.d_lin holds (d lin/d x)/SRGB_DK, LabJac.d_f holds 3 (d f/d t);
//   RgbJac.d_finv holds (d finv/d f)/3, RgbJac.d_gam holds (d rgb/d v)/GAM_DK.
constexpr double DK_D = 2.4 / 1.055, GDK_D = 1.055 / 2.4;
constexpr float SRGB_INV_SLOPE_BY_DK = (float)((1.0 / 12.92) / DK_D);
constexpr float F_SLOPE_X3 = (float)(1.0 / (EPS * EPS));
constexpr float FINV_SLOPE_BY3 = (float)(EPS * EPS);
constexpr float LIN_SLOPE_BY_DK = (float)(12.92 / GDK_D);
constexpr float LN_S3 = (float)(1.16 / 3.0), AN_S3 = (float)(500.0 / 220.0 / 3.0), BN_S3 = (float)(200.0 / 220.0 / 3.0);
constexpr float MB00 = (float)(0.412453 / WX * DK_D), MB10 = (float)(0.357580 / WX * DK_D), MB20 = (float)(0.180423 / WX * DK_D);
constexpr float MB01 = (float)(0.212671 / WY * DK_D), MB11 = (float)(0.715160 / WY * DK_D), MB21 = (float)(0.072169 / WY * DK_D);
constexpr float MB02 = (float)(0.019334 / WZ * DK_D), MB12 = (float)(0.119193 / WZ * DK_D), MB22 = (float)(0.950227 / WZ * DK_D);
constexpr float NB00 = (float)(3.2404542 * WX * GDK_D), NB10 = (float)(-1.5371385 * WY * GDK_D), NB20 = (float)(-0.4985314 * WZ * GDK_D);
constexpr float NB01 = (float)(-0.9692660 * WX * GDK_D), NB11 = (float)(1.8760108 * WY * GDK_D), NB21 = (float)(0.0415560 * WZ * GDK_D);
constexpr float NB02 = (float)(0.0556434 * WX * GDK_D), NB12 = (float)(-0.2040259 * WY * GDK_D), NB22 = (float)(1.0572252 * WZ * GDK_D);
constexpr float FY_S3 = (float)(300.0 / 116.0), FX_S3 = (float)(660.0 / 500.0), FZ_S3 = (float)(-660.0 / 200.0);
}  // namespace k

enum MaskKind { MASK_NONE = 0, MASK_U8 = 1, MASK_F32 = 2 };

// ------------------------------------------------------------------------------------------------
// Curve coefficients.  curves.py:29-32 evaluates
//     scale = C[0] + sum_{i=0}^{K-3} (C[i+1]-C[i]) * ((K-1) x - i)
// with no per-segment clamp, i.e. scale = alpha + beta x with
//     beta = (K-1) (C[K-2] - C[0]),   alpha = C[0] - sum_i i (C[i+1]-C[i]).
// C is the fp32-rounded exp of the predicted parameter (curves.py:53-54,105-106,152-153).
// ------------------------------------------------------------------------------------------------
CURL_HD void curve_alpha_beta(const float* C, int K, float* alpha, float* beta) {
  double a = (double)C[0];
  for (int i = 0; i <= K - 3; ++i) a -= (double)i * ((double)C[i + 1] - (double)C[i]);
  *alpha = (float)a;
  *beta = (float)((double)(K - 1) * ((double)C[K - 2] - (double)C[0]));
}

// curves.py:19,24 -- sum_j (slope[j+1]-slope[j])^2, j = 0..K-3
CURL_HD double curve_regulariser(const float* C, int K) {
  double r = 0.0;
  for (int j = 0; j <= K - 3; ++j) {
    double d = (double)C[j + 2] - 2.0 * (double)C[j + 1] + (double)C[j];
    r += d * d;
  }
  return r;
}

// d loss / d param[i] for one curve, given the two pixel moments S0 = sum g, S1 = sum g x
// (g = d loss / d scale) and greg = d loss / d regulariser.  SURVEY.md 8a row a10.
CURL_HD float curve_knot_grad(const float* C, int K, int i, double S0, double S1, double greg, bool log_domain) {
  double dC;
  if (i == 0) dC = S0 - (double)(K - 1) * S1;
  else if (i == K - 2) dC = (double)(K - 1) * S1 - (double)(K - 3) * S0;
  else if (i == K - 1) dC = 0.0;
  else dC = S0;
  // regulariser: d_j = C[j+2] - 2 C[j+1] + C[j], j = 0..K-3 ; d reg / d C[i] = 2 (d_{i-2} - 2 d_{i-1} + d_i)
  double r = 0.0;
  for (int t = 0; t < 3; ++t) {
    int j = i - 2 + t;
    if (j < 0 || j > K - 3) continue;
    double d = (double)C[j + 2] - 2.0 * (double)C[j + 1] + (double)C[j];
    r += (t == 1 ? -2.0 : 1.0) * d;
  }
  dC += 2.0 * r * greg;
  return (float)(log_domain ? dC * (double)C[i] : dC);  // chain through C = exp(param)
}

// ------------------------------------------------------------------------------------------------
// Stage: RGB -> normalised Lab            (reference colors.py:27-62)
// Powers share one MUFU.LG2 + one MUFU.EX2 between value and derivative:
//   u^2.4 = u * u^1.4, u^1.4 = u * 2^(0.4 lg u);   t^(1/3) = t * t^(-2/3), t^(-2/3) = 2^(-2/3 lg t).
// ------------------------------------------------------------------------------------------------
template <class T>
struct LabJac {
  T d_lin[3];  // (d lin_c / d x_c) / SRGB_DK   -- the constant factors are folded into rgb2lab_bwd
  T d_f[3];    // 3 (d f_k / d t_k)
};

// DIRECT (forward-only callers; requires !JAC): the powers are taken as 2^(p lg x) in one step instead of the
// x * 2^((p - 1) lg x) form that shares its LG2 + EX2 with the derivative -- one packed multiply fewer per channel,
// and for the cube root (|p| 1/3 instead of 2/3) a smaller amplification of the LG2 error.  The sRGB power keeps
// direct form too unless CURL_FWD_SHARED_SRGB is defined: 2.4 instead of 0.4 amplifies the LG2 error six times, but
// the shared form pays two more roundings, and measured against the fp64 oracle the direct one is no worse.
#ifdef CURL_FWD_SHARED_SRGB
constexpr bool FWD_DIRECT_SRGB = false;
#else
constexpr bool FWD_DIRECT_SRGB = true;     // measured (dev/fwd_accuracy.py): mean |err| vs fp64 1.21e-7 (shared form 1.23e-7), 2.5 % faster
#endif
// SAT12 (the fused forward only; requires !JAC): a and b leave saturated to [0, 1] -- the pre-clamp of the Lab curves
// that follow (curves.py:36 after the first curve) taken in the producing multiply-add itself (FFMA.SAT per half
// instead of FFMA2 + two FADD.SAT: two issue slots less per channel and pixel pair; sat(fl(x)) either way).
// sRGB -> linear of one channel, forward-only form (colors.py:37-38); also what fills the 256-entry table of the
// 8-bit kernel, so that a table entry and the value computed from the same x are the same bits.
template <class T>
CURL_HD T srgb_linearise_fwd(T x) {
  const T u = vfma(x, T(k::SRGB_A), T(k::SRGB_B));
  T hi;
  if (FWD_DIRECT_SRGB) hi = vex2(vmul(vlg2(u), T(2.4f)));
  else hi = vmul(u, vmul(u, vex2(vmul(vlg2(u), T(0.4f)))));
  return vsel_le(x, T(k::SRGB_T), vmul(x, T(k::SRGB_INV_SLOPE)), hi);
}

// LIN_IN (the 8-bit forward; requires DIRECT): lin_in holds the three linearised channels (table look-ups), x is not read.
template <class T, bool JAC, bool DIRECT = false, bool SAT12 = false, bool LIN_IN = false>
CURL_HD void rgb2lab_fwd(const T x[3], T lab[3], LabJac<T>* jac, const T* lin_in = nullptr) {
  static_assert(!(JAC && DIRECT), "the derivative records need the shared form");
  static_assert(!(JAC && SAT12), "the clamp gates need the unclamped values");
  static_assert(!LIN_IN || DIRECT, "pre-linearised inputs exist for the forward-only form");
  T lin[3];
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    if (LIN_IN) { lin[c] = lin_in[c]; continue; }
    if (DIRECT) { lin[c] = srgb_linearise_fwd(x[c]); continue; }
    // colors.py:37-38.  In the pow branch x > 0.04045 so the reference's clamp(min=1e-4) is inert.
    const T u = vfma(x[c], T(k::SRGB_A), T(k::SRGB_B));
    const T u14 = vmul(u, vex2(vmul(vlg2(u), T(0.4f))));
    const T hi = vmul(u, u14);
    lin[c] = vsel_le(x[c], T(k::SRGB_T), vmul(x[c], T(k::SRGB_INV_SLOPE)), hi);
    if (JAC) jac->d_lin[c] = vsel_le(x[c], T(k::SRGB_T), T(k::SRGB_INV_SLOPE_BY_DK), u14);   // (d lin / d x) / SRGB_DK
  }
  T t[3];
  t[0] = vfma(lin[2], T(k::M20), vfma(lin[1], T(k::M10), vmul(lin[0], T(k::M00))));  // colors.py:40-41
  t[1] = vfma(lin[2], T(k::M21), vfma(lin[1], T(k::M11), vmul(lin[0], T(k::M01))));
  t[2] = vfma(lin[2], T(k::M22), vfma(lin[1], T(k::M12), vmul(lin[0], T(k::M02))));
  T f[3], ef[3];
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    // colors.py:45-47, cube-root branch.  There t > eps^3 > 1e-4 so the reference's clamp is inert.
    if (DIRECT) {
      f[c] = vex2(vmul(vlg2(t[c]), T((float)(1.0 / 3.0))));         // t^(1/3)
    } else {
      ef[c] = vex2(vmul(vlg2(t[c]), T((float)(-2.0 / 3.0))));       // t^(-2/3)
      f[c] = vmul(t[c], ef[c]);
    }
    if (JAC) jac->d_f[c] = ef[c];                                    // 3 (d f / d t)
  }
  // ... and the linear branch t <= eps^3.  (Skipping these selects for warps in which no pixel needs them -- a vote
  // on min(X, Y, Z) <= eps^3 -- was measured slower, 4K x 16 backward 1.806 -> 1.845 ms, forward 0.774 -> 0.783: on
  // uniform inputs 2.5 % of the pixels have a Z that small, i.e. 80 % of the warps take the branch anyway.)
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    f[c] = vsel_le(t[c], T(k::F_T), vfma(t[c], T(k::F_SLOPE), T(k::F_OFF)), f[c]);
    if (JAC) jac->d_f[c] = vsel_le(t[c], T(k::F_T), T(k::F_SLOPE_X3), ef[c]);
  }
  // colors.py:50,57: (116 fy - 16) / 100 = 1.16 (fy - 4/29).  Written as a difference first so that a black
  // pixel (fy == 4/29 exactly) gives L == 0 exactly, as the reference's separately rounded 116 fy - 16 does: L = 0
  // sits on the boundary of every clamp(.., 0, 1) downstream, which passes gradient only on the closed
  // interval -- a fused 1.16 fy - 0.16 leaves -3e-9 there and silently blocks the L gradient of black pixels.
  lab[0] = vmul(vsub(f[1], T(k::F_OFF)), T(k::LN_S));
  if (SAT12) {
    lab[1] = vfma_sat(vsub(f[0], f[1]), T(k::AN_S), T(0.5f));
    lab[2] = vfma_sat(vsub(f[1], f[2]), T(k::BN_S), T(0.5f));
  } else {
    lab[1] = vfma(vsub(f[0], f[1]), T(k::AN_S), T(0.5f));      // colors.py:50,58
    lab[2] = vfma(vsub(f[1], f[2]), T(k::BN_S), T(0.5f));      // colors.py:50,59
  }
}

// FOLD (the fused layer only, see "Units of the fused layer's reverse sweep"): glab arrives as
// (g_L LN_S3, g_a AN_S3, -g_b BN_S3); the sign of the third is absorbed by the MB?2 column.
template <class T, bool FOLD = false>
CURL_HD void rgb2lab_bwd(const LabJac<T>& jac, const T glab[3], T gx[3]) {
  // constants carry the 1/3 of d_f and the SRGB_DK of d_lin; the sign of d f_z is in the MB?2 column
  T gt[3];
  if (FOLD) {
    gt[0] = vmul(glab[1], jac.d_f[0]);
    gt[1] = vmul(vsub(vsub(glab[0], glab[2]), glab[1]), jac.d_f[1]);
    gt[2] = vmul(glab[2], jac.d_f[2]);            // = -(g_b BN_S3) d_f: the column below carries the other sign
  } else {
    const T ga = vmul(glab[1], T(k::AN_S3)), gb = vmul(glab[2], T(k::BN_S3));
    gt[0] = vmul(ga, jac.d_f[0]);
    gt[1] = vmul(vsub(vfma(glab[0], T(k::LN_S3), gb), ga), jac.d_f[1]);
    gt[2] = vmul(gb, jac.d_f[2]);
  }
  constexpr float sg = FOLD ? 1.0f : -1.0f;
  gx[0] = vmul(vfma(gt[2], T(sg * k::MB02), vfma(gt[1], T(k::MB01), vmul(gt[0], T(k::MB00)))), jac.d_lin[0]);
  gx[1] = vmul(vfma(gt[2], T(sg * k::MB12), vfma(gt[1], T(k::MB11), vmul(gt[0], T(k::MB10)))), jac.d_lin[1]);
  gx[2] = vmul(vfma(gt[2], T(sg * k::MB22), vfma(gt[1], T(k::MB21), vmul(gt[0], T(k::MB20)))), jac.d_lin[2]);
}

// ------------------------------------------------------------------------------------------------
// Stage: normalised Lab -> RGB (unclamped)      (reference colors.py:88-123)
// ------------------------------------------------------------------------------------------------
template <class T>
struct RgbJac {
  T d_finv[3];  // (d finv_k / d f_k) / 3          -- the constant factors are folded into lab2rgb_bwd
  T d_gam[3];   // (d rgb_c / d v_c) / GAM_DK
};

// SAT12: as in rgb2lab_fwd, for the RGB curves that follow -- g and b leave saturated (both branches of the select).
template <class T, bool JAC, bool DIRECT = false, bool SAT12 = false>
CURL_HD void lab2rgb_fwd(const T lab[3], T rgb[3], RgbJac<T>* jac) {
  static_assert(!(JAC && DIRECT), "the derivative records need the shared form");
  static_assert(!(JAC && SAT12), "the clamp gates need the unclamped values");
  T f[3];
  f[1] = vfma(lab[0], T(k::FY_S), T(k::FY_O));                    // colors.py:97,104
  f[0] = vfma(lab[1], T(k::FX_S), vadd(f[1], T(k::FX_O)));        // colors.py:98,104
  f[2] = vfma(lab[2], T(k::FZ_S), vadd(f[1], T(k::FZ_O)));        // colors.py:99,104
  T xyz[3], sq[3];
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    // colors.py:110-111, cube branch.  There f > eps > 1e-4 so the clamp is inert.
    sq[c] = vmul(f[c], f[c]);
    xyz[c] = vmul(sq[c], f[c]);
    if (JAC) jac->d_finv[c] = sq[c];                                // (d finv / d f) / 3
  }
  // ... and the linear branch f <= eps
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    xyz[c] = vsel_le(f[c], T(k::FINV_T), vfma(f[c], T(k::FINV_SLOPE), T(k::FINV_OFF)), xyz[c]);
    if (JAC) jac->d_finv[c] = vsel_le(f[c], T(k::FINV_T), T(k::FINV_SLOPE_BY3), sq[c]);
  }
  T v[3];
  v[0] = vfma(xyz[2], T(k::N20), vfma(xyz[1], T(k::N10), vmul(xyz[0], T(k::N00))));  // colors.py:114,117
  v[1] = vfma(xyz[2], T(k::N21), vfma(xyz[1], T(k::N11), vmul(xyz[0], T(k::N01))));
  v[2] = vfma(xyz[2], T(k::N22), vfma(xyz[1], T(k::N12), vmul(xyz[0], T(k::N02))));
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    // colors.py:118-119.  In the pow branch v > 0.0031308 > 1e-4 so the clamp is inert.
    T e, pw;
    if (DIRECT) {
      pw = vex2(vmul(vlg2(v[c]), T((float)(1.0 / 2.4))));          // v^(1/2.4)
    } else {
      e = vex2(vmul(vlg2(v[c]), T(k::GAM_E)));                     // v^(1/2.4 - 1)
      pw = vmul(v[c], e);
    }
    const bool sat = SAT12 && c > 0;
    const T hi = sat ? vfma_sat(pw, T(1.055f), T(-0.055f)) : vfma(pw, T(1.055f), T(-0.055f));
    const T lo = sat ? vmul_sat(v[c], T(k::LIN_SLOPE)) : vmul(v[c], T(k::LIN_SLOPE));
    rgb[c] = vsel_le(v[c], T(k::LIN_T), lo, hi);
    if (JAC) jac->d_gam[c] = vsel_le(v[c], T(k::LIN_T), T(k::LIN_SLOPE_BY_DK), e);               // (d rgb / d v) / GAM_DK
  }
}

// FOLD (the fused layer only): glab leaves WITHOUT its factors (FY_S3, FX_S3, FZ_S3); see "Units of the fused
// layer's reverse sweep" below.
template <class T, bool FOLD = false>
CURL_HD void lab2rgb_bwd(const RgbJac<T>& jac, const T grgb[3], T glab[3]) {
  const T gv0 = vmul(grgb[0], jac.d_gam[0]), gv1 = vmul(grgb[1], jac.d_gam[1]), gv2 = vmul(grgb[2], jac.d_gam[2]);
  const T gf0 = vmul(vfma(gv2, T(k::NB02), vfma(gv1, T(k::NB01), vmul(gv0, T(k::NB00)))), jac.d_finv[0]);
  const T gf1 = vmul(vfma(gv2, T(k::NB12), vfma(gv1, T(k::NB11), vmul(gv0, T(k::NB10)))), jac.d_finv[1]);
  const T gf2 = vmul(vfma(gv2, T(k::NB22), vfma(gv1, T(k::NB21), vmul(gv0, T(k::NB20)))), jac.d_finv[2]);
  if (FOLD) {
    glab[0] = vadd(vadd(gf0, gf1), gf2); glab[1] = gf0; glab[2] = gf2;
  } else {
    glab[0] = vmul(vadd(vadd(gf0, gf1), gf2), T(k::FY_S3));
    glab[1] = vmul(gf0, T(k::FX_S3));
    glab[2] = vmul(gf2, T(k::FZ_S3));
  }
}

// ------------------------------------------------------------------------------------------------
// Stage: RGB -> HSV                             (reference colors.py:195-242)
// Hue uses ADDITIVE selectors: on ties the per-channel terms add (colors.py:221-224).
// max/min route their gradient to the FIRST index attaining them (torch CPU semantics).
// Backward, with gh/gs/gv the upstream gradients:
//   -gdf = gh nkdf_h + gs nkdf_s ;  gmx = gv + gs kmx_s + gdf ;
//   gx_c = pin_c ((gh A) sd_c + gmx amax_c - gdf amin_c)
// The input-clamp pass pin_c and the common hue factor A are applied once in the reverse sweep rather than
// being multiplied into nine record entries in the forward one (11 packed multiplies fewer per pixel pair).
// ------------------------------------------------------------------------------------------------
template <class T>
struct HsvJac {
  T sd[3];      // direct hue terms: selector differences (-1, 0, 1)
  T amax[3];    // one-hot first argmax
  T amin[3];    // one-hot first argmin
  T pin[3];     // input clamp passes
  T A;          // pass_h/6 * (1/df)
  T nkdf_h, nkdf_s, kmx_s;
};

// FOLD (the fused layer only): the hue gradient arrives divided by 6 and the saturation gradient negated (see
// "Units of the fused layer's reverse sweep"), so pass_h carries no 1/6 and nkdf_s, kmx_s no minus sign; with
// IN_UNIT the input-clamp passes pin_c are part of the RGB curves' output gates (self_curve_fwd<.., OUT_PIN>).
template <class T, bool JAC, bool IN_UNIT, bool FOLD = false>
CURL_HD void rgb2hsv_fwd(const T x[3], T hsv[3], HsvJac<T>* jac) {
  T xc[3];
#pragma unroll
  for (int c = 0; c < 3; ++c)  // colors.py:205.  IN_UNIT: caller guarantees 0 <= x <= 1
    xc[c] = IN_UNIT ? vmax(x[c], T(k::HSV_LO)) : vmin(vmax(x[c], T(k::HSV_LO)), T(1.0f));
  const T mx = vmax(xc[0], vmax(xc[1], xc[2]));                // colors.py:211
  const T mn = vmin(xc[0], vmin(xc[1], xc[2]));                // colors.py:212
  const T df = vsub(mx, mn);                                   // colors.py:214
  // 1/df and 1/mx from ONE reciprocal, r = 1/(df mx): MUFU is the forward kernel's busiest pipe.
  // dfc keeps the product away from zero on grey pixels (df == 0), where 1/df is defined as 0
  // (colors.py:186-193) and 1/mx must stay exact: dfc r = 1/mx either way.
  const T dfc = vmax(df, T(1e-20f));
  const T rr = vrcp(vmul(dfc, mx));
  // colors.py:186-193,220: 1/df := 0 on grey pixels.  The forward needs no select for it: there all three channel
  // differences are exactly 0, mx rr = 1e20 is finite, so every d * dinv term is an exact 0 either way, and the sum
  // is multiplied by the [df != 0] gate below (colors.py:221, the `where`).  Only the derivative record needs the 0.
  const T dinv = JAC ? vsel_eq(df, T(0.0f), T(0.0f), vmul(mx, rr)) : vmul(mx, rr);
  const T nz = vgate_ne(df, T(0.0f));
  const T s0 = vgate_eq(xc[0], mx), s1 = vgate_eq(xc[1], mx), s2 = vgate_eq(xc[2], mx);
  const T d12 = vsub(xc[1], xc[2]), d20 = vsub(xc[2], xc[0]), d01 = vsub(xc[0], xc[1]);
  // colors.py:221-224 (grey pixels: `where` replaces the sum by 0)
  T hr = vmul(vmul(d12, dinv), s0);
  hr = vfma(vfma(d20, dinv, T(2.0f)), s1, hr);
  hr = vfma(vfma(d01, dinv, T(4.0f)), s2, hr);
  hr = vmul(hr, nz);
  // colors.py:225-231:  h = (60 hr (+360 if negative)) / 360
  const T h = vfma(hr, T((float)(1.0 / 6.0)), vgate_lt(hr, T(0.0f)));
  const T minv = vmul(dfc, rr);                                // colors.py:234 (mx >= 1e-9, never 0)
  const T s = JAC ? vmul(df, minv) : T(0.0f);                  // colors.py:235-237 (the derivative path needs it unclamped)
  // colors.py:240
  // (the upper bound of the hue clamp is inert: hr <= 5 gives h <= 5/6, and a negative hr gives fl(1 + hr/6) <= 1)
  // (the derivative path keeps the inert min so that the backward kernels' code is unchanged)
  hsv[0] = JAC ? vmin(vmax(h, T(k::HSV_LO)), T(1.0f)) : vmax(h, T(k::HSV_LO));
  // values only: the upper bound inside the multiply (FMUL.SAT per half + one max instead of FMUL2 + min + max);
  // max(sat(s), LO) == min(max(s, LO), 1) for every s, NaN included
  hsv[1] = JAC ? vmin(vmax(s, T(k::HSV_LO)), T(1.0f)) : vmax(vmul_sat(df, minv), T(k::HSV_LO));
  hsv[2] = mx;  // already inside [1e-9, 1]
  if (JAC) {
    const T pass_h6 = FOLD ? vgate_eq(hsv[0], h) : vmul(vgate_eq(hsv[0], h), T((float)(1.0 / 6.0)));   // d h / d hr = 60/360
    const T pass_s = vgate_eq(hsv[1], s);
    const T A = vmul(pass_h6, dinv);      // dinv == 0 on grey pixels blocks the hue gradient
    jac->A = A;
    jac->sd[0] = vsub(s2, s1); jac->sd[1] = vsub(s0, s2); jac->sd[2] = vsub(s1, s0);
    if (!(FOLD && IN_UNIT)) {
      jac->pin[0] = vgate_eq(xc[0], x[0]); jac->pin[1] = vgate_eq(xc[1], x[1]); jac->pin[2] = vgate_eq(xc[2], x[2]);
    }
    const T num = vfma(s2, d01, vfma(s1, d20, vmul(s0, d12)));
    jac->nkdf_h = vmul(vmul(A, num), dinv);                    // -(d h / d df): d(1/df) = -1/df^2
    jac->nkdf_s = FOLD ? vmul(pass_s, minv) : vneg(vmul(pass_s, minv));   // -(d s / d df), s = df / mx
    jac->kmx_s = vmul(jac->nkdf_s, s);                         // d s / d mx = -s / mx
    const T e0 = vgate_eq(xc[0], mn), e1 = vgate_eq(xc[1], mn);
    const T ns0 = vsub(T(1.0f), s0), ne0 = vsub(T(1.0f), e0);
    jac->amax[0] = s0;
    jac->amax[1] = vmul(s1, ns0);
    jac->amax[2] = vmul(ns0, vsub(T(1.0f), s1));
    jac->amin[0] = e0;
    jac->amin[1] = vmul(e1, ne0);
    jac->amin[2] = vmul(ne0, vsub(T(1.0f), e1));
  }
}

template <class T, bool NO_PIN = false>
CURL_HD void rgb2hsv_bwd(const HsvJac<T>& j, const T ghsv[3], T gx[3]) {
  const T ngdf = vfma(ghsv[0], j.nkdf_h, vmul(ghsv[1], j.nkdf_s));
  const T gmx = vsub(vfma(ghsv[1], j.kmx_s, ghsv[2]), ngdf);
  const T ghA = vmul(ghsv[0], j.A);
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const T g = vfma(ghA, j.sd[c], vfma(ngdf, j.amin[c], vmul(gmx, j.amax[c])));
    gx[c] = NO_PIN ? g : vmul(j.pin[c], g);
  }
}

// ------------------------------------------------------------------------------------------------
// Stage: HSV -> RGB                             (reference colors.py:131-177)
//   p = v(1-s), w = v - p, k = 6h.  The reference's pairs of 60-degree ramps are one trapezoid each:
//     r = v - w (ramp60 - ramp240)   = p + w sat(|k-3| - 1)
//     g = p + w (ramp0  - ramp180)   = p + w sat(2 - |k-2|)
//     b = p + w (ramp120 - ramp300)  = p + w sat(2 - |k-4|)
//   (one FFMA2 + one FADD.SAT with |.| per half and channel instead of two of each), with
//   d/dk = +-[the trapezoid is on a slope, ends included] exactly where the reference's clamp gates pass.
//   The output clamp of colors.py:175 is inert here and is not evaluated: 0 <= p <= v <= 1, 0 <= w = fl(v - p) <= v
//   and 0 <= ramp <= 1, so every channel is fl(p + w ramp) with p + w ramp in [0, v + 2^-25] -- it
//   rounds into [0, 1] and the clamp's gradient gate is 1.
// ------------------------------------------------------------------------------------------------
template <class T>
struct Hsv2RgbJac {
  T ramp[3];        // r, g, b = p + w ramp_c
  T hk[3];          // 6 w (d ramp_c / d k) pass_in_h
  T oms, vs;        // 1 - s ; v * pass_in_s
  T pin_v;
};

// FOLD (the fused layer only): hk is formed without its factor 6 and hsv2rgb_bwd returns -d/ds (see "Units of
// the fused layer's reverse sweep").
// poff (the fused layer's backward with a {0,1} mask): added to p, 0 for a kept pixel and NaN for a masked one
// -- the NaN reaches all three outputs, the final clamp gate [sat(x + res) == x + res] is then false, and the
// mask has been applied with one packed add instead of three packed multiplies.  Nothing else reads p.
template <class T, bool JAC, bool IN_UNIT, bool FOLD = false, bool POFF = false>
CURL_HD void hsv2rgb_fwd(const T in[3], T rgb[3], Hsv2RgbJac<T>* jac, T poff = T(0.0f)) {
  const T h = IN_UNIT ? in[0] : vsat(in[0]);                    // colors.py:141
  const T s = IN_UNIT ? in[1] : vsat(in[1]);
  const T v = IN_UNIT ? in[2] : vsat(in[2]);
  const T oms = vsub(T(1.0f), s);
  const T p = vmul(v, oms);
  const T w = vsub(v, p);
  // tg, tb carry the sign of the falling-side convention (2 - k, 4 - k): |.| is the same and the slope sign is theirs
  const T tr = vfma(h, T(6.0f), T(-3.0f)), tg = vfma(h, T(-6.0f), T(2.0f)), tb = vfma(h, T(-6.0f), T(4.0f));
  T ramp[3];
  if (!JAC) {
    ramp[0] = vabs_add_sat(tr, -1.0f); ramp[1] = vnabs_add_sat(tg, 2.0f); ramp[2] = vnabs_add_sat(tb, 2.0f);
  } else {
    const T yr = vabs_add(tr, -1.0f), yg = vnabs_add(tg, 2.0f), yb = vnabs_add(tb, 2.0f);
    ramp[0] = vsat(yr); ramp[1] = vsat(yg); ramp[2] = vsat(yb);
    T w6 = FOLD ? w : vmul(w, T(6.0f));
    if (!IN_UNIT) w6 = vmul(w6, vgate_eq(h, in[0]));
    // on a slope (ends included, as torch.clamp passes gradient on the closed interval): +-1, else 0
    jac->hk[0] = vmul(w6, vcopysign(vgate_eq(ramp[0], yr), tr));
    jac->hk[1] = vmul(w6, vcopysign(vgate_eq(ramp[1], yg), tg));
    jac->hk[2] = vmul(w6, vcopysign(vgate_eq(ramp[2], yb), tb));
    jac->ramp[0] = ramp[0]; jac->ramp[1] = ramp[1]; jac->ramp[2] = ramp[2];
    jac->oms = oms;
    jac->vs = IN_UNIT ? v : vmul(v, vgate_eq(s, in[1]));
    jac->pin_v = IN_UNIT ? T(1.0f) : vgate_eq(v, in[2]);
  }
  const T pp = POFF ? vadd(p, poff) : p;
  rgb[0] = vfma(w, ramp[0], pp);                                // colors.py:143-150
  rgb[1] = vfma(w, ramp[1], pp);                                // colors.py:153-159
  rgb[2] = vfma(w, ramp[2], pp);                                // colors.py:162-168 ; colors.py:175 inert (above)
}

template <class T, bool IN_UNIT, bool FOLD = false>
CURL_HD void hsv2rgb_bwd(const Hsv2RgbJac<T>& j, const T grgb[3], T ghsv[3]) {
  const T gw = vfma(grgb[0], j.ramp[0], vfma(grgb[1], j.ramp[1], vmul(grgb[2], j.ramp[2])));
  ghsv[0] = vfma(grgb[0], j.hk[0], vfma(grgb[1], j.hk[1], vmul(grgb[2], j.hk[2])));
  const T gp = vsub(vadd(vadd(grgb[0], grgb[1]), grgb[2]), gw);   // p enters r, g, b directly and w = v - p
  const T gv = vfma(gp, j.oms, gw);                               // v enters through w and p = v (1 - s)
  ghsv[1] = FOLD ? vmul(gp, j.vs) : vneg(vmul(gp, j.vs));
  ghsv[2] = IN_UNIT ? gv : vmul(gv, j.pin_v);
}

// ------------------------------------------------------------------------------------------------
// Stage: knot curves                            (reference curves.py:4-38)
// One "self" curve: out = clamp(x * (alpha + beta x)); every apply_curve clamps all channels, so a
// channel adjusted by the 2nd/3rd curve of a group sees its input clamped first (PRECLAMP).
// Backward per curve: gs = go * a (d loss / d scale), moments S0 += gs, S1 += gs * xin, gx = go * d.
// ------------------------------------------------------------------------------------------------
template <class T>
struct CurveJac {
  T a;     // pass_out * xin   (xin = the clamped input; pass_out is 0/1, so a * a == pass_out * xin^2)
  T d;     // pass_out * pass_in * (alpha + 2 beta xin)
};

// OUT_PIN (the fused layer's RGB curves): the output gate is [clamp(y, 1e-9, 1) == y], i.e. it includes the pass
// of RGB2HSV's input clamp (colors.py:205) that follows directly, so that stage need not multiply it in; the
// max(o, 1e-9) is the value RGB2HSV forms anyway.
template <class T, bool JAC, bool PRECLAMP, bool OUT_PIN = false>
CURL_HD T self_curve_fwd(T x, float alpha, float beta, CurveJac<T>* jac) {
  const T xin = PRECLAMP ? vsat(x) : x;
  const T s = vfma(T(beta), xin, T(alpha));        // curves.py:31-32 (closed form)
  if (!JAC) return vmul_sat(xin, s);               // curves.py:35-36
  const T y = vmul(xin, s);
  const T o = vsat(y);
  const T po = OUT_PIN ? vgate_eq(vmax(o, T(k::HSV_LO)), y) : vgate_eq(o, y);
  jac->a = vmul(po, xin);
  T d = vmul(po, vfma(T(beta), xin, s));
  if (PRECLAMP) d = vmul(d, vgate_eq(xin, x));
  jac->d = d;
  return o;
}

template <class T>
CURL_HD T self_curve_bwd(const CurveJac<T>& j, T go, T mom[2]) {
  const T gs = vmul(go, j.a);
  mom[0] = vadd(mom[0], gs);
  mom[1] = vfma(gs, j.a, mom[1]);     // gs * xin: gs already carries pass_out, so gs * a == gs * xin
  return vmul(go, j.d);
}

// One general curve cin -> cout on raw (unclamped) inputs, then the whole-pixel clamp: this is
// curves.apply_curve itself (curves.py:4-38) for callers that use it directly.
template <class T>
CURL_HD void single_curve_fwd(const T x[3], int cin, int cout, float alpha, float beta, T y[3]) {
  const T xi = cin == 0 ? x[0] : (cin == 1 ? x[1] : x[2]);
  const T s = vfma(T(beta), xi, T(alpha));
#pragma unroll
  for (int c = 0; c < 3; ++c) y[c] = vsat(c == cout ? vmul(x[c], s) : x[c]);
}

template <class T>
CURL_HD void single_curve_bwd(const T x[3], int cin, int cout, float alpha, float beta, const T gy[3], T gx[3],
                              T mom[2]) {
  const T xi = cin == 0 ? x[0] : (cin == 1 ? x[1] : x[2]);
  const T s = vfma(T(beta), xi, T(alpha));
  T gs = T(0.0f);
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const T pre = (c == cout) ? vmul(x[c], s) : x[c];
    const T g = vmul(gy[c], vgate_eq(vsat(pre), pre));
    gx[c] = (c == cout) ? vmul(g, s) : g;
    if (c == cout) gs = vmul(g, x[c]);
  }
  mom[0] = vadd(mom[0], gs);
  mom[1] = vfma(gs, xi, mom[1]);
#pragma unroll
  for (int c = 0; c < 3; ++c)
    if (c == cin) gx[c] = vfma(gs, T(beta), gx[c]);
}

// adjust_lab / adjust_rgb: three self curves c -> c (curves.py:90-133, 136-180)
template <class T>
struct Self3Jac { CurveJac<T> c[3]; };

// IN_SAT12 (the fused forward): channels 1 and 2 arrive saturated (rgb2lab_fwd / lab2rgb_fwd <.., SAT12>)
template <class T, bool JAC, bool OUT_PIN = false, bool IN_SAT12 = false>
CURL_HD void adjust_self3_fwd(const T x[3], const float* ab, T y[3], Self3Jac<T>* jac) {
  static_assert(!(JAC && IN_SAT12), "the clamp gates need the unclamped values");
  y[0] = self_curve_fwd<T, JAC, false, OUT_PIN>(x[0], ab[0], ab[1], JAC ? &jac->c[0] : nullptr);
  y[1] = self_curve_fwd<T, JAC, !IN_SAT12, OUT_PIN>(x[1], ab[2], ab[3], JAC ? &jac->c[1] : nullptr);
  y[2] = self_curve_fwd<T, JAC, !IN_SAT12, OUT_PIN>(x[2], ab[4], ab[5], JAC ? &jac->c[2] : nullptr);
}

template <class T>
CURL_HD void adjust_self3_bwd(const Self3Jac<T>& j, const T gy[3], T gx[3], T* mom /*6*/) {
  gx[0] = self_curve_bwd(j.c[0], gy[0], mom + 0);
  gx[1] = self_curve_bwd(j.c[1], gy[1], mom + 2);
  gx[2] = self_curve_bwd(j.c[2], gy[2], mom + 4);
}

// adjust_hsv: hue<-f(hue); sat<-f(adjusted hue); sat<-f(sat); val<-f(val)   (curves.py:41-87)
template <class T>
struct HsvCurveJac {
  CurveJac<T> ch;   // curve 0 on hue
  CurveJac<T> cs;   // curve 2 on sat (self)
  CurveJac<T> cv;   // curve 3 on val
  T h1;             // adjusted hue = input of curve 1
  T p;              // pass_y1 * (clamped sat)            : d loss/d scale1 = gs1 * p
  T q;              // pass_y1 * scale1 * pass_sat_in     : d loss/d sat_in = gs1 * q
};

template <class T, bool JAC, bool IN_UNIT>
CURL_HD void adjust_hsv_fwd(const T x[3], const float* ab, T y[3], HsvCurveJac<T>* jac) {
  const T h1 = self_curve_fwd<T, JAC, false>(x[0], ab[0], ab[1], JAC ? &jac->ch : nullptr);
  // curve 1: sat * scale(h1); the first apply_curve already clamped sat and val (curves.py:36)
  const T s_c = IN_UNIT ? x[1] : vsat(x[1]);
  const T sc1 = vfma(T(ab[3]), h1, T(ab[2]));
  T s1;
  if (JAC) {
    const T y1 = vmul(s_c, sc1);
    s1 = vsat(y1);
    const T py = vgate_eq(s1, y1);
    jac->h1 = h1;
    jac->p = vmul(py, s_c);
    T q = vmul(py, sc1);
    if (!IN_UNIT) q = vmul(q, vgate_eq(s_c, x[1]));
    jac->q = q;
  } else {
    s1 = vmul_sat(s_c, sc1);
  }
  y[0] = h1;
  y[1] = self_curve_fwd<T, JAC, false>(s1, ab[4], ab[5], JAC ? &jac->cs : nullptr);
  y[2] = IN_UNIT ? self_curve_fwd<T, JAC, false>(x[2], ab[6], ab[7], JAC ? &jac->cv : nullptr)
                 : self_curve_fwd<T, JAC, true>(x[2], ab[6], ab[7], JAC ? &jac->cv : nullptr);
}

// beta1: the hue -> sat curve's beta (ab[3]); the fused layer passes it in the folded units of its reverse sweep
template <class T>
CURL_HD void adjust_hsv_bwd(const HsvCurveJac<T>& j, float beta1, const T gy[3], T gx[3], T* mom /*8*/) {
  gx[2] = self_curve_bwd(j.cv, gy[2], mom + 6);
  const T gs1 = self_curve_bwd(j.cs, gy[1], mom + 4);
  const T gsc = vmul(gs1, j.p);                    // d loss / d scale of curve 1
  mom[2] = vadd(mom[2], gsc);
  mom[3] = vfma(gsc, j.h1, mom[3]);
  gx[1] = vmul(gs1, j.q);
  const T gh1 = vfma(gsc, T(beta1), gy[0]);        // hue feeds curve 1 through beta_1
  gx[0] = self_curve_bwd(j.ch, gh1, mom + 0);
}

// ------------------------------------------------------------------------------------------------
// Units of the fused layer's reverse sweep (FOLD).  Between two curve groups the gradient passes through
// constant factors: (FY_S3, FX_S3, FZ_S3) leaving lab2rgb_bwd, (LN_S3, AN_S3, BN_S3) entering rgb2lab_bwd, the 6
// of d ramp / d h, the 1/6 of d h / d hr, the signs of d p / d s and d s / d df.  The reverse sweep leaves them
// out (11 packed instructions per pixel pair) and works in scaled units instead:
//   gradient arriving at the Lab curves  = true / (FY_S3, FX_S3, FZ_S3);  leaving them towards rgb2lab_bwd the
//     products FY_S3 LN_S3 = (300/116)(1.16/3) = 1, FX_S3 AN_S3 = 1, FZ_S3 BN_S3 = -1 remain (the two
//     conversions are inverse affine maps): only a sign, which rgb2lab_bwd's matrix column absorbs;
//   hue gradient = true / 6 between hsv2rgb_bwd and rgb2hsv_bwd; sat gradient = -true on the same stretch.
// What is left to do per image, not per pixel:
//   curve_fold_m(c): factor of the two moments of curve c after the reduction;
//   layer_beta1(beta_1) = -beta_1 / 6 for gh1 = gsc beta_1 + g_hue in those units.
// c = 0..2 Lab, 3..5 RGB, 6..9 HSV (hue, hue->sat, sat, val).  The stand-alone ops use 1 everywhere.
// ------------------------------------------------------------------------------------------------
CURL_HD double curve_fold_m(bool layer, int c) {
  if (!layer) return 1.0;
  switch (c) {
    case 0: return 300.0 / 116.0;
    case 1: return 660.0 / 500.0;
    case 2: return -660.0 / 200.0;
    case 6: return 6.0;
    case 7: return -1.0;
    case 8: return -1.0;
    default: return 1.0;
  }
}
CURL_HD float layer_beta1(float beta1) { return (float)((double)beta1 * (-1.0 / 6.0)); }

// ------------------------------------------------------------------------------------------------
// The layer                                     (reference model.py:137-176)
//   rgb2lab -> adjust_lab -> *mask -> lab2rgb -> adjust_rgb -> *mask -> rgb2hsv -> adjust_hsv
//   -> *mask -> hsv2rgb -> clamp(img + residual) * mask
// ab: 10 curves x (alpha, beta): lab 0..2, rgb 3..5, hsv 6..9.
// MK == MASK_F32 multiplies by the (possibly fractional) mask at every stage like the reference.
// For a {0,1} mask (MASK_U8) or no mask the three intermediate products cannot change an unmasked
// pixel and a masked pixel is zeroed by the last product, so only the last one is applied.
// ------------------------------------------------------------------------------------------------
constexpr int LAYER_CURVES = 10;

template <class T>
CURL_HD void vscale3(T a[3], T m) { a[0] = vmul(a[0], m); a[1] = vmul(a[1], m); a[2] = vmul(a[2], m); }

#ifdef CURL_FWD_NO_SAT12
constexpr bool FWD_SAT12 = false;
#else
constexpr bool FWD_SAT12 = true;      // the curves' pre-clamps taken inside the producing conversions (bit-identical)
#endif
template <class T, int MK, bool LIN_IN = false>
CURL_HD void layer_fwd(const T x[3], T m, const float* ab, T out[3], const T* lin_in = nullptr) {
  T lab[3], lab2[3], rgb[3], rgb2[3], hsv[3], hsv2[3], res[3];
  rgb2lab_fwd<T, false, true, FWD_SAT12, LIN_IN>(x, lab, nullptr, lin_in);
  adjust_self3_fwd<T, false, false, FWD_SAT12>(lab, ab, lab2, nullptr);
  if (MK == MASK_F32) vscale3(lab2, m);
  lab2rgb_fwd<T, false, true, FWD_SAT12>(lab2, rgb, nullptr);
  adjust_self3_fwd<T, false, false, FWD_SAT12>(rgb, ab + 6, rgb2, nullptr);
  if (MK == MASK_F32) vscale3(rgb2, m);
  rgb2hsv_fwd<T, false, MK != MASK_F32>(rgb2, hsv, nullptr);
  adjust_hsv_fwd<T, false, true>(hsv, ab + 12, hsv2, nullptr);
  if (MK == MASK_F32) {
    vscale3(hsv2, m);
    hsv2rgb_fwd<T, false, false>(hsv2, res, nullptr);
  } else {
    hsv2rgb_fwd<T, false, true>(hsv2, res, nullptr);
  }
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const T o = vadd_sat(x[c], res[c]);              // model.py:170
    out[c] = (MK == MASK_NONE) ? o : vmul(o, m);
  }
}

// Spill policies for the backward pass.  At the turning point of the reverse sweep every stage's
// Jac record is live (66 values of T), which is what limits the kernel to 8 warps per SM.  The
// records of the first four stages are not needed again until the very end, so the sm_100a kernel
// parks them in shared memory (one 16-byte column per thread and pair of values: conflict-free) and gets them back
// when the reverse sweep reaches their stage.  KeepInRegisters is the no-op policy (host check,
// scalar tail path).
struct KeepInRegisters {
  static constexpr bool kActive = false;
  static constexpr bool kMomParked = false;
};
#ifndef CURL_PARK_STAGES
#define CURL_PARK_STAGES 4         // how many of the first four stages' records are parked (6 slots each)
#endif
constexpr int PARK_STAGES = CURL_PARK_STAGES;
constexpr int SPILL_SLOTS = 6 * PARK_STAGES;   // LabJac 6 + Self3Jac 6 + RgbJac 6 + Self3Jac 6

// Six values per record, parked and fetched as a whole (put6 / get6 at slot o, a multiple of 6).
template <class T, class S> CURL_HD void spill_save(const S& sp, int o, const LabJac<T>& j) {
  sp.put6(o, j.d_lin[0], j.d_lin[1], j.d_lin[2], j.d_f[0], j.d_f[1], j.d_f[2]);
}
template <class T, class S> CURL_HD void spill_load(const S& sp, int o, LabJac<T>& j) {
  sp.get6(o, j.d_lin[0], j.d_lin[1], j.d_lin[2], j.d_f[0], j.d_f[1], j.d_f[2]);
}
template <class T, class S> CURL_HD void spill_save(const S& sp, int o, const RgbJac<T>& j) {
  sp.put6(o, j.d_finv[0], j.d_finv[1], j.d_finv[2], j.d_gam[0], j.d_gam[1], j.d_gam[2]);
}
template <class T, class S> CURL_HD void spill_load(const S& sp, int o, RgbJac<T>& j) {
  sp.get6(o, j.d_finv[0], j.d_finv[1], j.d_finv[2], j.d_gam[0], j.d_gam[1], j.d_gam[2]);
}
template <class T, class S> CURL_HD void spill_save(const S& sp, int o, const Self3Jac<T>& j) {
  sp.put6(o, j.c[0].a, j.c[0].d, j.c[1].a, j.c[1].d, j.c[2].a, j.c[2].d);
}
template <class T, class S> CURL_HD void spill_load(const S& sp, int o, Self3Jac<T>& j) {
  sp.get6(o, j.c[0].a, j.c[0].d, j.c[1].a, j.c[1].d, j.c[2].a, j.c[2].d);
}

// Split fetch: *_issue starts the fetch of a record (the destination registers are pending), *_wait makes them
// readable.  Issued one stage ahead of its use, a fetch costs no stall (-DCURL_PARK_PREFETCH).
template <class T, class S> CURL_HD void spill_load_issue(const S& sp, int o, LabJac<T>& j) {
  sp.get6_issue(o, j.d_lin[0], j.d_lin[1], j.d_lin[2], j.d_f[0], j.d_f[1], j.d_f[2]);
}
template <class T, class S> CURL_HD void spill_load_wait(const S& sp, LabJac<T>& j) {
  sp.get6_wait(j.d_lin[0], j.d_lin[1], j.d_lin[2], j.d_f[0], j.d_f[1], j.d_f[2]);
}
template <class T, class S> CURL_HD void spill_load_issue(const S& sp, int o, RgbJac<T>& j) {
  sp.get6_issue(o, j.d_finv[0], j.d_finv[1], j.d_finv[2], j.d_gam[0], j.d_gam[1], j.d_gam[2]);
}
template <class T, class S> CURL_HD void spill_load_wait(const S& sp, RgbJac<T>& j) {
  sp.get6_wait(j.d_finv[0], j.d_finv[1], j.d_finv[2], j.d_gam[0], j.d_gam[1], j.d_gam[2]);
}
template <class T, class S> CURL_HD void spill_load_issue(const S& sp, int o, Self3Jac<T>& j) {
  sp.get6_issue(o, j.c[0].a, j.c[0].d, j.c[1].a, j.c[1].d, j.c[2].a, j.c[2].d);
}
template <class T, class S> CURL_HD void spill_load_wait(const S& sp, Self3Jac<T>& j) {
  sp.get6_wait(j.c[0].a, j.c[0].d, j.c[1].a, j.c[1].d, j.c[2].a, j.c[2].d);
}
#ifdef CURL_PARK_PREFETCH
constexpr bool PARK_PREFETCH = true;
#else
constexpr bool PARK_PREFETCH = false;
#endif

// RGB -> HSV record (tensor-memory parking only, -DCURL_PARK_J5): the 13 values the fused layer uses (no pin).
#ifdef CURL_PARK_J5
constexpr bool PARK_J5 = true;
#else
constexpr bool PARK_J5 = false;
#endif
constexpr int J5_SLOT = 24, J5_SLOTS = 14;
template <class T, class S> CURL_HD void spill_save(const S& sp, int o, const HsvJac<T>& j) {
  sp.put8(o, j.sd[0], j.sd[1], j.sd[2], j.amax[0], j.amax[1], j.amax[2], j.amin[0], j.amin[1]);
  sp.put6(o + 8, j.amin[2], j.A, j.nkdf_h, j.nkdf_s, j.kmx_s, j.kmx_s);
}
template <class T, class S> CURL_HD void spill_load(const S& sp, int o, HsvJac<T>& j) {
  T dummy;
  sp.get8(o, j.sd[0], j.sd[1], j.sd[2], j.amax[0], j.amax[1], j.amax[2], j.amin[0], j.amin[1]);
  sp.get6(o + 8, j.amin[2], j.A, j.nkdf_h, j.nkdf_s, j.kmx_s, dummy);
}

// The lane value layer_bwd takes as `m`: the mask itself for MASK_F32 (multiplied in at every stage like the
// reference), ignored for MASK_NONE, and for a {0,1} mask (MASK_U8) the GATE OFFSET 0 (kept) / NaN (masked).
CURL_HD float bwd_mask_lane_u8(bool keep) {
#if defined(__CUDA_ARCH__)
  return keep ? 0.0f : __int_as_float(0x7fc00000);
#else
  return keep ? 0.0f : NAN;
#endif
}

// Recompute-forward + reverse.  mom: 20 accumulators (S0,S1 per curve), accumulated in place, in the folded units
// of curve_fold_m.  ab: 10 x (alpha, beta) followed by layer_beta1(beta of the hue -> sat curve).
// SPILL: a policy with put6(slot, 6 x T) / get6(slot, 6 x T&) (see above) or KeepInRegisters.
template <class T, int MK, bool WANT_GX, class SPILL>
CURL_HD void layer_bwd(const T x[3], T m, const float* ab, const T gout[3], T gx[3], T* mom, const SPILL& sp) {
  constexpr bool kSpill = SPILL::kActive;
  constexpr bool UNIT = MK != MASK_F32;          // no fractional mask between the stages: values stay in [0, 1]
  T lab[3], lab2[3], rgb[3], rgb2[3], hsv[3], hsv2[3], res[3];
  LabJac<T> j1; Self3Jac<T> j2; RgbJac<T> j3; Self3Jac<T> j4; HsvJac<T> j5; HsvCurveJac<T> j6; Hsv2RgbJac<T> j7;
  rgb2lab_fwd<T, WANT_GX>(x, lab, &j1);
  if constexpr (kSpill && WANT_GX && PARK_STAGES >= 1) spill_save<T>(sp, 0, j1);
  adjust_self3_fwd<T, true>(lab, ab, lab2, &j2);
  if constexpr (kSpill && PARK_STAGES >= 2) spill_save<T>(sp, 6, j2);
  if (MK == MASK_F32) vscale3(lab2, m);
  lab2rgb_fwd<T, true>(lab2, rgb, &j3);
  if constexpr (kSpill && PARK_STAGES >= 3) spill_save<T>(sp, 12, j3);
  adjust_self3_fwd<T, true, UNIT>(rgb, ab + 6, rgb2, &j4);      // UNIT: output gates include RGB2HSV's input clamp
  if constexpr (kSpill && PARK_STAGES >= 4) spill_save<T>(sp, 18, j4);
  if (MK == MASK_F32) vscale3(rgb2, m);
  rgb2hsv_fwd<T, true, UNIT, true>(rgb2, hsv, &j5);
  if constexpr (kSpill && PARK_J5 && UNIT) spill_save<T>(sp, J5_SLOT, j5);
  adjust_hsv_fwd<T, true, true>(hsv, ab + 12, hsv2, &j6);
  if (MK == MASK_F32) {
    vscale3(hsv2, m);
    hsv2rgb_fwd<T, true, false, true>(hsv2, res, &j7);
  } else if (MK == MASK_U8) {
    hsv2rgb_fwd<T, true, true, true, true>(hsv2, res, &j7, m);   // m is the gate offset here: 0 keeps, NaN masks
  } else {
    hsv2rgb_fwd<T, true, true, true>(hsv2, res, &j7);
  }
  T g[3];
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const T sum = vadd(x[c], res[c]);
    const T pass = vgate_eq(vsat(sum), sum);         // clamp passes on [0,1] inclusive (false for a masked pixel's NaN)
    g[c] = vmul(gout[c], MK == MASK_F32 ? vmul(pass, m) : pass);
  }
  T a[3], b[3];
  if (MK == MASK_F32) hsv2rgb_bwd<T, false, true>(j7, g, a); else hsv2rgb_bwd<T, true, true>(j7, g, a);
  if (MK == MASK_F32) vscale3(a, m);
  if constexpr (SPILL::kMomParked) sp.mom_get8(mom + 12);      // the HSV group's 8 accumulators
  adjust_hsv_bwd(j6, ab[2 * LAYER_CURVES], a, b, mom + 12);
  if constexpr (SPILL::kMomParked) sp.mom_put8(mom + 12);
  if constexpr (kSpill) sp.before_loads();       // parked records are about to be read back
  if constexpr (kSpill && PARK_J5 && UNIT) spill_load<T>(sp, J5_SLOT, j5);
  constexpr bool PF = kSpill && PARK_PREFETCH && PARK_STAGES == 4;      // fetch every record one stage ahead of its use
  if constexpr (PF) spill_load_issue<T>(sp, 18, j4);
  rgb2hsv_bwd<T, UNIT>(j5, b, a);
  if (MK == MASK_F32) vscale3(a, m);
  if constexpr (PF) { spill_load_wait<T>(sp, j4); spill_load_issue<T>(sp, 12, j3); }
  else if constexpr (kSpill && PARK_STAGES >= 4) spill_load<T>(sp, 18, j4);
  if constexpr (SPILL::kMomParked) sp.mom_get6(6, mom + 6);
  adjust_self3_bwd(j4, a, b, mom + 6);
  if constexpr (SPILL::kMomParked) sp.mom_put6(6, mom + 6);
  if constexpr (PF) { spill_load_wait<T>(sp, j3); spill_load_issue<T>(sp, 6, j2); }
  else if constexpr (kSpill && PARK_STAGES >= 3) spill_load<T>(sp, 12, j3);
  lab2rgb_bwd<T, true>(j3, b, a);
  if (MK == MASK_F32) vscale3(a, m);
  if constexpr (PF) { spill_load_wait<T>(sp, j2); if (WANT_GX) spill_load_issue<T>(sp, 0, j1); }
  else if constexpr (kSpill && PARK_STAGES >= 2) spill_load<T>(sp, 6, j2);
  if constexpr (SPILL::kMomParked) sp.mom_get6(0, mom + 0);
  adjust_self3_bwd(j2, a, b, mom + 0);
  if constexpr (SPILL::kMomParked) sp.mom_put6(0, mom + 0);
  if (WANT_GX) {
    if constexpr (PF) spill_load_wait<T>(sp, j1);
    else if constexpr (kSpill && PARK_STAGES >= 1) spill_load<T>(sp, 0, j1);
    rgb2lab_bwd<T, true>(j1, b, a);
    gx[0] = vadd(g[0], a[0]); gx[1] = vadd(g[1], a[1]); gx[2] = vadd(g[2], a[2]);
  }
}

}  // namespace curl
