// Per-pixel arithmetic of the CURL curve layer: colour conversions, affine knot curves, and their
// hand-derived reverse-mode derivatives.  Everything here is a pure function of one pixel (plus
// the per-image curve coefficients), written once and instantiated
//   * in the sm_100a kernels (curl_kernels.cu) with DevMath  -> MUFU.LG2 / MUFU.EX2 / MUFU.RCP,
//   * in the host-side derivation check (tests/hostsim) with HostMath -> double-precision libm,
//     which is test infrastructure only and is never linked into the product library.
//
// Semantics follow the reference (danielbulhosa/CURL) line by line; citations are file:line in
// that tree.  Where the reference evaluates "both branches * 0/1 mask" the kernel uses a select,
// which is value-identical for finite data.
#pragma once

#if defined(__CUDACC__)
#define CURL_HD __host__ __device__ __forceinline__
#else
#define CURL_HD inline
#endif

#include <math.h>

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
constexpr float F_T = (float)(EPS * EPS * EPS);          // colors.py:45
constexpr float F_SLOPE = (float)(1.0 / (3.0 * EPS * EPS));
constexpr float F_OFF = (float)(4.0 / 29.0);
constexpr float FINV_T = (float)EPS;                     // colors.py:110
constexpr float FINV_SLOPE = (float)(3.0 * EPS * EPS);
constexpr float LIN_T = 0.0031308f;                      // colors.py:118
constexpr float LIN_SLOPE = 12.92f;
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
}  // namespace k

// ------------------------------------------------------------------------------------------------
// Math policies
// ------------------------------------------------------------------------------------------------
#if defined(__CUDACC__)
struct DevMath {
  static __device__ __forceinline__ float lg2(float x) {
    float y;
    asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
  }
  static __device__ __forceinline__ float ex2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
  }
  static __device__ __forceinline__ float rcp(float x) {
    float y;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
  }
  static __device__ __forceinline__ float sat(float x) { return __saturatef(x); }
  static __device__ __forceinline__ float fmax_(float a, float b) { return fmaxf(a, b); }
  static __device__ __forceinline__ float fmin_(float a, float b) { return fminf(a, b); }
};
#endif

struct HostMath {  // test-only: correctly rounded stand-ins for the MUFU ops
  static inline float lg2(float x) { return (float)log2((double)x); }
  static inline float ex2(float x) { return (float)exp2((double)x); }
  static inline float rcp(float x) { return (float)(1.0 / (double)x); }
  static inline float sat(float x) { return x < 0.f ? 0.f : (x > 1.f ? 1.f : (x == x ? x : 0.f)); }
  static inline float fmax_(float a, float b) { return a > b ? a : b; }
  static inline float fmin_(float a, float b) { return a < b ? a : b; }
};

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
// ------------------------------------------------------------------------------------------------
struct LabJac {   // local derivatives kept by the backward pass
  float d_lin[3];  // d lin_c / d x_c
  float d_f[3];    // d f_k / d t_k
};

template <class M, bool WANT_JAC>
CURL_HD void rgb2lab_fwd(const float x[3], float lab[3], LabJac* jac) {
  float lin[3];
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    // colors.py:37-38.  In the pow branch x > 0.04045 so the reference's clamp(min=1e-4) is inert.
    const float u = fmaf(x[c], k::SRGB_A, k::SRGB_B);
    const float l = M::lg2(u);
    const float hi = M::ex2(2.4f * l);
    const bool lo_branch = x[c] <= k::SRGB_T;
    lin[c] = lo_branch ? x[c] * k::SRGB_INV_SLOPE : hi;
    if (WANT_JAC) {
      // d/dx ((x+.055)/1.055)^2.4 = 2.4/1.055 * u^1.4
      const float dhi = (float)(2.4 / 1.055) * M::ex2(1.4f * l);
      jac->d_lin[c] = lo_branch ? k::SRGB_INV_SLOPE : dhi;
    }
  }
  float t[3];
  t[0] = fmaf(lin[2], k::M20, fmaf(lin[1], k::M10, lin[0] * k::M00));  // colors.py:40-41
  t[1] = fmaf(lin[2], k::M21, fmaf(lin[1], k::M11, lin[0] * k::M01));
  t[2] = fmaf(lin[2], k::M22, fmaf(lin[1], k::M12, lin[0] * k::M02));
  float f[3];
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    // colors.py:45-47.  In the cube-root branch t > eps^3 > 1e-4 so the clamp is inert.
    const float l = M::lg2(t[c]);
    const float cr = M::ex2(l * (float)(1.0 / 3.0));
    const bool lo_branch = t[c] <= k::F_T;
    f[c] = lo_branch ? fmaf(t[c], k::F_SLOPE, k::F_OFF) : cr;
    if (WANT_JAC) {
      // d/dt t^(1/3) = (1/3) t^(-2/3)
      const float dcr = (float)(1.0 / 3.0) * M::ex2(l * (float)(-2.0 / 3.0));
      jac->d_f[c] = lo_branch ? k::F_SLOPE : dcr;
    }
  }
  lab[0] = fmaf(f[1], k::LN_S, k::LN_O);           // colors.py:50,57
  lab[1] = fmaf(f[0] - f[1], k::AN_S, 0.5f);       // colors.py:50,58
  lab[2] = fmaf(f[1] - f[2], k::BN_S, 0.5f);       // colors.py:50,59
}

CURL_HD void rgb2lab_bwd(const LabJac& jac, const float glab[3], float gx[3]) {
  const float ga = glab[1] * k::AN_S, gb = glab[2] * k::BN_S;
  float gt[3];
  gt[0] = ga * jac.d_f[0];
  gt[1] = (fmaf(glab[0], k::LN_S, gb) - ga) * jac.d_f[1];
  gt[2] = -gb * jac.d_f[2];
  gx[0] = fmaf(gt[2], k::M02, fmaf(gt[1], k::M01, gt[0] * k::M00)) * jac.d_lin[0];
  gx[1] = fmaf(gt[2], k::M12, fmaf(gt[1], k::M11, gt[0] * k::M10)) * jac.d_lin[1];
  gx[2] = fmaf(gt[2], k::M22, fmaf(gt[1], k::M21, gt[0] * k::M20)) * jac.d_lin[2];
}

// ------------------------------------------------------------------------------------------------
// Stage: normalised Lab -> RGB (unclamped)      (reference colors.py:88-123)
// ------------------------------------------------------------------------------------------------
struct RgbJac {
  float d_finv[3];  // d finv_k / d f_k
  float d_gam[3];   // d rgb_c / d v_c
};

template <class M, bool WANT_JAC>
CURL_HD void lab2rgb_fwd(const float lab[3], float rgb[3], RgbJac* jac) {
  float f[3];
  f[1] = fmaf(lab[0], k::FY_S, k::FY_O);           // colors.py:97,104
  f[0] = fmaf(lab[1], k::FX_S, f[1] + k::FX_O);    // colors.py:98,104
  f[2] = fmaf(lab[2], k::FZ_S, f[1] + k::FZ_O);    // colors.py:99,104
  float xyz[3];
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    // colors.py:110-111.  In the cube branch f > eps > 1e-4 so the clamp is inert.
    const bool lo_branch = f[c] <= k::FINV_T;
    const float sq = f[c] * f[c];
    xyz[c] = lo_branch ? k::FINV_SLOPE * (f[c] - k::F_OFF) : sq * f[c];
    if (WANT_JAC) jac->d_finv[c] = lo_branch ? k::FINV_SLOPE : 3.0f * sq;
  }
  float v[3];
  v[0] = fmaf(xyz[2], k::N20, fmaf(xyz[1], k::N10, xyz[0] * k::N00));  // colors.py:114,117
  v[1] = fmaf(xyz[2], k::N21, fmaf(xyz[1], k::N11, xyz[0] * k::N01));
  v[2] = fmaf(xyz[2], k::N22, fmaf(xyz[1], k::N12, xyz[0] * k::N02));
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    // colors.py:118-119.  In the pow branch v > 0.0031308 > 1e-4 so the clamp is inert.
    const float l = M::lg2(v[c]);
    const float pw = M::ex2(l * (float)(1.0 / 2.4));
    const bool lo_branch = v[c] <= k::LIN_T;
    rgb[c] = lo_branch ? v[c] * k::LIN_SLOPE : fmaf(pw, 1.055f, -0.055f);
    if (WANT_JAC) {
      // d/dv 1.055 v^(1/2.4) = (1.055/2.4) v^(1/2.4 - 1)
      const float dpw = (float)(1.055 / 2.4) * M::ex2(l * (float)(1.0 / 2.4 - 1.0));
      jac->d_gam[c] = lo_branch ? k::LIN_SLOPE : dpw;
    }
  }
}

CURL_HD void lab2rgb_bwd(const RgbJac& jac, const float grgb[3], float glab[3]) {
  const float gv0 = grgb[0] * jac.d_gam[0], gv1 = grgb[1] * jac.d_gam[1], gv2 = grgb[2] * jac.d_gam[2];
  const float gf0 = fmaf(gv2, k::N02, fmaf(gv1, k::N01, gv0 * k::N00)) * jac.d_finv[0];
  const float gf1 = fmaf(gv2, k::N12, fmaf(gv1, k::N11, gv0 * k::N10)) * jac.d_finv[1];
  const float gf2 = fmaf(gv2, k::N22, fmaf(gv1, k::N21, gv0 * k::N20)) * jac.d_finv[2];
  glab[0] = (gf0 + gf1 + gf2) * k::FY_S;
  glab[1] = gf0 * k::FX_S;
  glab[2] = gf2 * k::FZ_S;
}

// ------------------------------------------------------------------------------------------------
// Stage: RGB -> HSV                             (reference colors.py:195-242)
// Hue uses ADDITIVE selectors: on ties the per-channel terms add (colors.py:221-224).
// max/min route their gradient to the FIRST index attaining them (torch CPU semantics).
// ------------------------------------------------------------------------------------------------
struct HsvJac {
  float xc[3];       // clamped inputs
  float dinv, minv;  // 1/df (0 if df == 0), 1/mx
  float df;
  float sel[3];      // [x_c == mx] as 0/1
  int imax, imin;
  bool pass_in[3];   // 1e-9 <= x_c <= 1
  bool pass_h, pass_s;
};

template <class M, bool WANT_JAC, bool IN_UNIT>
CURL_HD void rgb2hsv_fwd(const float x[3], float hsv[3], HsvJac* jac) {
  float xc[3];
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    // colors.py:205.  IN_UNIT: caller guarantees 0 <= x <= 1, only the lower bound can bite.
    xc[c] = IN_UNIT ? M::fmax_(x[c], k::HSV_LO) : M::fmin_(M::fmax_(x[c], k::HSV_LO), 1.0f);
    if (WANT_JAC) jac->pass_in[c] = (xc[c] == x[c]);
  }
  const float mx = M::fmax_(xc[0], M::fmax_(xc[1], xc[2]));   // colors.py:211
  const float mn = M::fmin_(xc[0], M::fmin_(xc[1], xc[2]));   // colors.py:212
  const float df = mx - mn;                                    // colors.py:214
  const bool grey = (df == 0.0f);
  const float dinv = grey ? 0.0f : M::rcp(df);                 // colors.py:186-193,220
  const float s0 = (xc[0] == mx) ? 1.0f : 0.0f;
  const float s1 = (xc[1] == mx) ? 1.0f : 0.0f;
  const float s2 = (xc[2] == mx) ? 1.0f : 0.0f;
  // colors.py:221-224 (grey pixels: dinv == 0 gives 0 + 2*s1 + 4*s2, which `where` replaces by 0)
  float hr = ((xc[1] - xc[2]) * dinv) * s0 + fmaf(xc[2] - xc[0], dinv, 2.0f) * s1 +
             fmaf(xc[0] - xc[1], dinv, 4.0f) * s2;
  hr = grey ? 0.0f : hr;
  // colors.py:225-231:  h = (60 hr (+360 if negative)) / 360
  float h = hr * (float)(1.0 / 6.0);
  h = (hr < 0.0f) ? h + 1.0f : h;
  const float minv = M::rcp(mx);                               // colors.py:234 (mx >= 1e-9, never 0)
  const float s = df * minv;                                   // colors.py:235-237
  // colors.py:240
  hsv[0] = M::fmin_(M::fmax_(h, k::HSV_LO), 1.0f);
  hsv[1] = M::fmin_(M::fmax_(s, k::HSV_LO), 1.0f);
  hsv[2] = mx;  // already inside [1e-9, 1]
  if (WANT_JAC) {
    jac->xc[0] = xc[0]; jac->xc[1] = xc[1]; jac->xc[2] = xc[2];
    jac->dinv = dinv; jac->minv = minv; jac->df = df;
    jac->sel[0] = s0; jac->sel[1] = s1; jac->sel[2] = s2;
    jac->imax = (xc[0] == mx) ? 0 : ((xc[1] == mx) ? 1 : 2);
    jac->imin = (xc[0] == mn) ? 0 : ((xc[1] == mn) ? 1 : 2);
    jac->pass_h = (hsv[0] == h);
    jac->pass_s = (hsv[1] == s);
  }
}

CURL_HD void rgb2hsv_bwd(const HsvJac& j, const float ghsv[3], float gx[3]) {
  const float gh = j.pass_h ? ghsv[0] * (float)(1.0 / 6.0) : 0.0f;  // d h / d hr = 60/360
  const float gs = j.pass_s ? ghsv[1] : 0.0f;
  const float gv = ghsv[2];
  // hue terms (zero when df == 0 because dinv == 0 there and `where` blocks the gradient)
  const float A = gh * j.dinv;
  float g0 = A * (j.sel[2] - j.sel[1]);
  float g1 = A * (j.sel[0] - j.sel[2]);
  float g2 = A * (j.sel[1] - j.sel[0]);
  const float num = j.sel[0] * (j.xc[1] - j.xc[2]) + j.sel[1] * (j.xc[2] - j.xc[0]) + j.sel[2] * (j.xc[0] - j.xc[1]);
  float gdf = -(gh * num) * (j.dinv * j.dinv);   // d(1/df) = -1/df^2
  gdf = fmaf(gs, j.minv, gdf);                    // s = df / mx
  const float gmx = fmaf(-gs * j.df, j.minv * j.minv, gv) + gdf;
  const float gmn = -gdf;
  g0 += (j.imax == 0 ? gmx : 0.0f) + (j.imin == 0 ? gmn : 0.0f);
  g1 += (j.imax == 1 ? gmx : 0.0f) + (j.imin == 1 ? gmn : 0.0f);
  g2 += (j.imax == 2 ? gmx : 0.0f) + (j.imin == 2 ? gmn : 0.0f);
  gx[0] = j.pass_in[0] ? g0 : 0.0f;
  gx[1] = j.pass_in[1] ? g1 : 0.0f;
  gx[2] = j.pass_in[2] ? g2 : 0.0f;
}

// ------------------------------------------------------------------------------------------------
// Stage: HSV -> RGB                             (reference colors.py:131-177)
//   p = v(1-s), w = v - p ;  ramp_k = clamp(360 h - k, 0, 60)/60 = sat(6h - k/60)
//   r = v + w (ramp240 - ramp60) ; g = p + w (ramp0 - ramp180) ; b = p + w (ramp120 - ramp300)
// ------------------------------------------------------------------------------------------------
struct Hsv2RgbJac {
  float w, v, one_minus_s;
  float dr, dg, db;          // ramp differences
  float gate[6];             // ramp k unsaturated (inclusive ends), k = 0,60,120,180,240,300
  bool pass_in[3];
  bool pass_out[3];
};

template <class M, bool WANT_JAC, bool IN_UNIT>
CURL_HD void hsv2rgb_fwd(const float hsv_in[3], float rgb[3], Hsv2RgbJac* jac) {
  float q[3];
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    q[c] = IN_UNIT ? hsv_in[c] : M::sat(hsv_in[c]);           // colors.py:141
    if (WANT_JAC) jac->pass_in[c] = IN_UNIT ? true : (q[c] == hsv_in[c]);
  }
  const float h = q[0], s = q[1], v = q[2];
  const float oms = 1.0f - s;
  const float p = v * oms;
  const float w = v - p;
  float t[6], u[6];
#pragma unroll
  for (int i = 0; i < 6; ++i) {
    t[i] = fmaf(h, 6.0f, -(float)i);
    u[i] = M::sat(t[i]);
  }
  const float dr = u[4] - u[1], dg = u[0] - u[3], db = u[2] - u[5];
  const float r = fmaf(w, dr, v);                               // colors.py:143-150
  const float g = fmaf(w, dg, p);                               // colors.py:153-159
  const float b = fmaf(w, db, p);                               // colors.py:162-168
  rgb[0] = M::sat(r); rgb[1] = M::sat(g); rgb[2] = M::sat(b);   // colors.py:175
  if (WANT_JAC) {
    jac->w = w; jac->v = v; jac->one_minus_s = oms;
    jac->dr = dr; jac->dg = dg; jac->db = db;
#pragma unroll
    for (int i = 0; i < 6; ++i) jac->gate[i] = (u[i] == t[i]) ? 1.0f : 0.0f;
    jac->pass_out[0] = (rgb[0] == r); jac->pass_out[1] = (rgb[1] == g); jac->pass_out[2] = (rgb[2] == b);
  }
}

CURL_HD void hsv2rgb_bwd(const Hsv2RgbJac& j, const float grgb[3], float ghsv[3]) {
  const float gr = j.pass_out[0] ? grgb[0] : 0.0f;
  const float gg = j.pass_out[1] ? grgb[1] : 0.0f;
  const float gb = j.pass_out[2] ? grgb[2] : 0.0f;
  const float gw = fmaf(gr, j.dr, fmaf(gg, j.dg, gb * j.db));
  // d/dh: each unsaturated ramp contributes 6 * (+-w) * g
  const float hr = j.gate[4] - j.gate[1], hg = j.gate[0] - j.gate[3], hb = j.gate[2] - j.gate[5];
  const float gh = 6.0f * j.w * fmaf(gr, hr, fmaf(gg, hg, gb * hb));
  const float gp = gg + gb - gw;          // p enters g, b directly and w = v - p
  const float gv = gr + gw + gp * j.one_minus_s;
  const float gs = -gp * j.v;
  ghsv[0] = j.pass_in[0] ? gh : 0.0f;
  ghsv[1] = j.pass_in[1] ? gs : 0.0f;
  ghsv[2] = j.pass_in[2] ? gv : 0.0f;
}

// ------------------------------------------------------------------------------------------------
// Stage: knot curves                            (reference curves.py:4-38)
// One "self" curve: out = clamp(x * (alpha + beta x)); every apply_curve clamps all channels, so a
// channel adjusted by the 2nd/3rd curve of a group sees its input clamped first (PRECLAMP).
// ------------------------------------------------------------------------------------------------
struct CurveJac {
  float xin;       // (clamped) input
  float dydx;      // alpha + 2 beta x
  bool pass_in, pass_out;
};

template <class M, bool WANT_JAC, bool PRECLAMP>
CURL_HD float self_curve_fwd(float x, float alpha, float beta, CurveJac* jac) {
  const float xin = PRECLAMP ? M::sat(x) : x;
  const float s = fmaf(beta, xin, alpha);          // curves.py:31-32 (closed form)
  const float y = xin * s;                         // curves.py:35
  const float o = M::sat(y);                       // curves.py:36
  if (WANT_JAC) {
    jac->xin = xin;
    jac->dydx = fmaf(beta, xin, s);
    jac->pass_in = PRECLAMP ? (xin == x) : true;
    jac->pass_out = (o == y);
  }
  return o;
}

// mom[0] += g, mom[1] += g * xin  with g = d loss / d scale
CURL_HD float self_curve_bwd(const CurveJac& j, float go, float mom[2]) {
  const float gy = j.pass_out ? go : 0.0f;
  const float gs = gy * j.xin;
  mom[0] += gs;
  mom[1] = fmaf(gs, j.xin, mom[1]);
  const float gx = gy * j.dydx;
  return j.pass_in ? gx : 0.0f;
}

// One general curve cin -> cout on raw (unclamped) inputs, then the whole-pixel clamp: this is
// curves.apply_curve itself (curves.py:4-38) for callers that use it directly.
template <class M>
CURL_HD void single_curve_fwd(const float x[3], int cin, int cout, float alpha, float beta, float y[3]) {
  const float xi = cin == 0 ? x[0] : (cin == 1 ? x[1] : x[2]);
  const float s = fmaf(beta, xi, alpha);
#pragma unroll
  for (int c = 0; c < 3; ++c) y[c] = M::sat(c == cout ? x[c] * s : x[c]);
}

template <class M>
CURL_HD void single_curve_bwd(const float x[3], int cin, int cout, float alpha, float beta, const float gy[3],
                              float gx[3], float mom[2]) {
  const float xi = cin == 0 ? x[0] : (cin == 1 ? x[1] : x[2]);
  const float s = fmaf(beta, xi, alpha);
  float gs = 0.0f;
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const float pre = (c == cout) ? x[c] * s : x[c];
    const float g = (M::sat(pre) == pre) ? gy[c] : 0.0f;
    gx[c] = (c == cout) ? g * s : g;
    if (c == cout) gs = g * x[c];
  }
  mom[0] += gs;
  mom[1] = fmaf(gs, xi, mom[1]);
#pragma unroll
  for (int c = 0; c < 3; ++c)
    if (c == cin) gx[c] = fmaf(gs, beta, gx[c]);
}

// adjust_lab / adjust_rgb: three self curves c -> c (curves.py:90-133, 136-180)
struct Self3Jac { CurveJac c[3]; };

template <class M, bool WANT_JAC>
CURL_HD void adjust_self3_fwd(const float x[3], const float* ab, float y[3], Self3Jac* jac) {
  y[0] = self_curve_fwd<M, WANT_JAC, false>(x[0], ab[0], ab[1], WANT_JAC ? &jac->c[0] : nullptr);
  y[1] = self_curve_fwd<M, WANT_JAC, true>(x[1], ab[2], ab[3], WANT_JAC ? &jac->c[1] : nullptr);
  y[2] = self_curve_fwd<M, WANT_JAC, true>(x[2], ab[4], ab[5], WANT_JAC ? &jac->c[2] : nullptr);
}

CURL_HD void adjust_self3_bwd(const Self3Jac& j, const float gy[3], float gx[3], float* mom /*6*/) {
  gx[0] = self_curve_bwd(j.c[0], gy[0], mom + 0);
  gx[1] = self_curve_bwd(j.c[1], gy[1], mom + 2);
  gx[2] = self_curve_bwd(j.c[2], gy[2], mom + 4);
}

// adjust_hsv: hue<-f(hue); sat<-f(adjusted hue); sat<-f(sat); val<-f(val)   (curves.py:41-87)
struct HsvCurveJac {
  CurveJac ch;        // curve 0 on hue
  CurveJac cs;        // curve 2 on sat (self)
  CurveJac cv;        // curve 3 on val
  float s_c, h1, sc1; // clamped sat input, adjusted hue, scale of curve 1
  bool pass_s_in, pass_y1;
};

template <class M, bool WANT_JAC, bool IN_UNIT>
CURL_HD void adjust_hsv_fwd(const float x[3], const float* ab, float y[3], HsvCurveJac* jac) {
  const float h1 = self_curve_fwd<M, WANT_JAC, false>(x[0], ab[0], ab[1], WANT_JAC ? &jac->ch : nullptr);
  // curve 1: sat * scale(h1); the first apply_curve already clamped sat and val (curves.py:36)
  const float s_c = IN_UNIT ? x[1] : M::sat(x[1]);
  const float sc1 = fmaf(ab[3], h1, ab[2]);
  const float y1 = s_c * sc1;
  const float s1 = M::sat(y1);
  const float s2 = self_curve_fwd<M, WANT_JAC, false>(s1, ab[4], ab[5], WANT_JAC ? &jac->cs : nullptr);
  const float v1 = IN_UNIT ? self_curve_fwd<M, WANT_JAC, false>(x[2], ab[6], ab[7], WANT_JAC ? &jac->cv : nullptr)
                           : self_curve_fwd<M, WANT_JAC, true>(x[2], ab[6], ab[7], WANT_JAC ? &jac->cv : nullptr);
  y[0] = h1; y[1] = s2; y[2] = v1;
  if (WANT_JAC) {
    jac->s_c = s_c; jac->h1 = h1; jac->sc1 = sc1;
    jac->pass_s_in = IN_UNIT ? true : (s_c == x[1]);
    jac->pass_y1 = (s1 == y1);
  }
}

CURL_HD void adjust_hsv_bwd(const HsvCurveJac& j, const float* ab, const float gy[3], float gx[3], float* mom /*8*/) {
  gx[2] = self_curve_bwd(j.cv, gy[2], mom + 6);
  const float gs1 = self_curve_bwd(j.cs, gy[1], mom + 4);
  const float gy1 = j.pass_y1 ? gs1 : 0.0f;
  const float gsc = gy1 * j.s_c;                 // d loss / d scale of curve 1
  mom[2] += gsc;
  mom[3] = fmaf(gsc, j.h1, mom[3]);
  const float gs_c = gy1 * j.sc1;
  gx[1] = j.pass_s_in ? gs_c : 0.0f;
  const float gh1 = fmaf(gsc, ab[3], gy[0]);     // hue feeds curve 1 through beta_1
  gx[0] = self_curve_bwd(j.ch, gh1, mom + 0);
}

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

template <class M, int MK>
CURL_HD void layer_fwd(const float x[3], float m, const float* ab, float out[3]) {
  float lab[3], lab2[3], rgb[3], rgb2[3], hsv[3], hsv2[3], res[3];
  rgb2lab_fwd<M, false>(x, lab, nullptr);
  adjust_self3_fwd<M, false>(lab, ab, lab2, nullptr);
  if (MK == MASK_F32) { lab2[0] *= m; lab2[1] *= m; lab2[2] *= m; }
  lab2rgb_fwd<M, false>(lab2, rgb, nullptr);
  adjust_self3_fwd<M, false>(rgb, ab + 6, rgb2, nullptr);
  if (MK == MASK_F32) { rgb2[0] *= m; rgb2[1] *= m; rgb2[2] *= m; }
  rgb2hsv_fwd<M, false, MK != MASK_F32>(rgb2, hsv, nullptr);
  adjust_hsv_fwd<M, false, true>(hsv, ab + 12, hsv2, nullptr);
  if (MK == MASK_F32) {
    hsv2[0] *= m; hsv2[1] *= m; hsv2[2] *= m;
    hsv2rgb_fwd<M, false, false>(hsv2, res, nullptr);
  } else {
    hsv2rgb_fwd<M, false, true>(hsv2, res, nullptr);
  }
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const float o = M::sat(x[c] + res[c]);           // model.py:170
    out[c] = (MK == MASK_NONE) ? o : o * m;
  }
}

// Recompute-forward + reverse.  mom: 20 accumulators (S0,S1 per curve), accumulated in place.
template <class M, int MK>
CURL_HD void layer_bwd(const float x[3], float m, const float* ab, const float gout[3], float gx[3], float* mom) {
  float lab[3], lab2[3], rgb[3], rgb2[3], hsv[3], hsv2[3], res[3];
  LabJac j1; Self3Jac j2; RgbJac j3; Self3Jac j4; HsvJac j5; HsvCurveJac j6; Hsv2RgbJac j7;
  rgb2lab_fwd<M, true>(x, lab, &j1);
  adjust_self3_fwd<M, true>(lab, ab, lab2, &j2);
  if (MK == MASK_F32) { lab2[0] *= m; lab2[1] *= m; lab2[2] *= m; }
  lab2rgb_fwd<M, true>(lab2, rgb, &j3);
  adjust_self3_fwd<M, true>(rgb, ab + 6, rgb2, &j4);
  if (MK == MASK_F32) { rgb2[0] *= m; rgb2[1] *= m; rgb2[2] *= m; }
  rgb2hsv_fwd<M, true, MK != MASK_F32>(rgb2, hsv, &j5);
  adjust_hsv_fwd<M, true, true>(hsv, ab + 12, hsv2, &j6);
  if (MK == MASK_F32) {
    hsv2[0] *= m; hsv2[1] *= m; hsv2[2] *= m;
    hsv2rgb_fwd<M, true, false>(hsv2, res, &j7);
  } else {
    hsv2rgb_fwd<M, true, true>(hsv2, res, &j7);
  }
  float g[3];
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const float sum = x[c] + res[c];
    const float go = (MK == MASK_NONE) ? gout[c] : gout[c] * m;
    g[c] = (M::sat(sum) == sum) ? go : 0.0f;          // clamp passes on [0,1] inclusive
  }
  float a[3], b[3];
  hsv2rgb_bwd(j7, g, a);
  if (MK == MASK_F32) { a[0] *= m; a[1] *= m; a[2] *= m; }
  adjust_hsv_bwd(j6, ab + 12, a, b, mom + 12);
  rgb2hsv_bwd(j5, b, a);
  if (MK == MASK_F32) { a[0] *= m; a[1] *= m; a[2] *= m; }
  adjust_self3_bwd(j4, a, b, mom + 6);
  lab2rgb_bwd(j3, b, a);
  if (MK == MASK_F32) { a[0] *= m; a[1] *= m; a[2] *= m; }
  adjust_self3_bwd(j2, a, b, mom + 0);
  rgb2lab_bwd(j1, b, a);
  gx[0] = g[0] + a[0]; gx[1] = g[1] + a[1]; gx[2] = g[2] + a[2];
}

}  // namespace curl
