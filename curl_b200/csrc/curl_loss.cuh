// Per-pixel arithmetic of the colour terms of the reference's CURLLoss (model.py:36-118): masked L1
// in RGB, cosine similarity in RGB, L1 in clamped Lab, L1 in the HSV cone -- everything in that loss
// except MS-SSIM, which consumes the Lab L channel this code also hands out.  Templates over the lane
// type T like curl_pixel.cuh (F2 on the device, float in the host-side derivation check).
//
//   p = pred * mask, t = target * mask                                   (model.py:89)
//   s0 = sum_c |p_c - t_c|                                               (model.py:91)
//   s1 = cosine_similarity(p, t, dim=1)                                  (model.py:95)
//   s2 = sum_c |clamp(rgb2lab(p))_c - clamp(rgb2lab(t))_c|               (model.py:57-60, 98-100)
//   s3 = sum_c |cone(p)_c - cone(t)_c|,  cone = (v s cos 2 pi h, v s sin 2 pi h, v) of
//        clamp(rgb2hsv(.), 0, 1)                                          (model.py:66-79, 106-108)
#pragma once

#include "curl_pixel.cuh"

namespace curl {

namespace lk {
constexpr float COS_EPS = 1e-8f;                 // torch.nn.functional.cosine_similarity default eps
constexpr float TWO_PI = 6.283185307179586f;     // 2 * math.pi (model.py:71), rounded to fp32 on use
}

// sin / cos of an angle in radians in [0, 2 pi]
CURL_HD float vsin(float x) {
#if defined(__CUDA_ARCH__)
  return __sinf(x);
#else
  return (float)sin((double)x);
#endif
}
CURL_HD float vcos(float x) {
#if defined(__CUDA_ARCH__)
  return __cosf(x);
#else
  return (float)cos((double)x);
#endif
}
CURL_HD float vsqrt(float x) {
#if defined(__CUDA_ARCH__)
  return __fsqrt_rn(x);
#else
  return (float)sqrt((double)x);
#endif
}
CURL_HD float vrsq(float x) {
#if defined(__CUDA_ARCH__)
  return __frsqrt_rn(x);
#else
  return (float)(1.0 / sqrt((double)x));
#endif
}
CURL_HD float vabs(float x) { return fabsf(x); }
CURL_HD float vsign(float x) { return x > 0.f ? 1.f : (x < 0.f ? -1.f : 0.f); }   // torch.sign: 0 at 0
#if defined(__CUDACC__)
CURL_F2_MAP1(vsin) CURL_F2_MAP1(vcos) CURL_F2_MAP1(vsqrt) CURL_F2_MAP1(vrsq) CURL_F2_MAP1(vabs) CURL_F2_MAP1(vsign)
#endif

template <class T>
struct ConeOf {
  T a, b, v;        // v s cos(theta), v s sin(theta), v
  T cs, sn, s;      // kept for the derivative
};

template <class T>
CURL_HD ConeOf<T> hsv_cone(const T hsv[3]) {
  ConeOf<T> c;
  const T th = vmul(hsv[0], T(lk::TWO_PI));                 // model.py:71
  c.cs = vcos(th); c.sn = vsin(th);
  c.s = hsv[1]; c.v = hsv[2];
  const T vs = vmul(hsv[2], hsv[1]);                        // model.py:75-76: val * sat * cos(hue)
  c.a = vmul(vs, c.cs); c.b = vmul(vs, c.sn);
  return c;
}

// cosine_similarity of two 3-vectors the way ATen computes it: each vector divided by
// max(norm, eps) first, then the dot product.  1 / max(sqrt(n2), eps) == rsqrt(max(n2, eps^2)): one
// MUFU.RSQ per vector instead of a square root and a reciprocal.  Returns the squared norm of p and
// 1 / max(|p|, eps) for the backward.
template <class T>
CURL_HD T cos_sim3(const T p[3], const T t[3], T pn[3], T tn[3], T* np2_out, T* ip_out) {
  const T np2 = vfma(p[2], p[2], vfma(p[1], p[1], vmul(p[0], p[0])));
  const T nt2 = vfma(t[2], t[2], vfma(t[1], t[1], vmul(t[0], t[0])));
  const T ip = vrsq(vmax(np2, T(lk::COS_EPS * lk::COS_EPS))), it = vrsq(vmax(nt2, T(lk::COS_EPS * lk::COS_EPS)));
#pragma unroll
  for (int c = 0; c < 3; ++c) { pn[c] = vmul(p[c], ip); tn[c] = vmul(t[c], it); }
  *np2_out = np2; *ip_out = ip;
  return vfma(pn[2], tn[2], vfma(pn[1], tn[1], vmul(pn[0], tn[0])));
}

#ifdef CURL_LOSS_FWD_SHARED
constexpr bool LOSS_FWD_DIRECT = false;
#else
constexpr bool LOSS_FWD_DIRECT = true;
#endif
// Forward: the four per-pixel terms and the two Lab L values (for MS-SSIM, model.py:101-103).
template <class T>
CURL_HD void loss_terms_fwd(const T pred[3], const T target[3], T m, T s[4], T* L_pred, T* L_target) {
  T p[3], t[3];
#pragma unroll
  for (int c = 0; c < 3; ++c) { p[c] = vmul(pred[c], m); t[c] = vmul(target[c], m); }
  s[0] = vadd(vadd(vabs(vsub(p[0], t[0])), vabs(vsub(p[1], t[1]))), vabs(vsub(p[2], t[2])));
  T pn[3], tn[3], np2, ip;
  s[1] = cos_sim3(p, t, pn, tn, &np2, &ip);
  T lp[3], lt[3];
  // forward only: the powers in their direct form 2^(p lg x) (see rgb2lab_fwd), the same for both images so that
  // identical pixels give identical Lab values; the backward keeps the form that shares its powers with the derivative
  rgb2lab_fwd<T, false, LOSS_FWD_DIRECT>(p, lp, nullptr);
  rgb2lab_fwd<T, false, LOSS_FWD_DIRECT>(t, lt, nullptr);
#pragma unroll
  for (int c = 0; c < 3; ++c) { lp[c] = vsat(lp[c]); lt[c] = vsat(lt[c]); }                 // model.py:58
  s[2] = vadd(vadd(vabs(vsub(lp[0], lt[0])), vabs(vsub(lp[1], lt[1]))), vabs(vsub(lp[2], lt[2])));
  *L_pred = lp[0]; *L_target = lt[0];
  T hp[3], ht[3];
  rgb2hsv_fwd<T, false, false>(p, hp, nullptr);             // output already inside [1e-9, 1]: the clamp of
  rgb2hsv_fwd<T, false, false>(t, ht, nullptr);             // model.py:67 changes nothing
  const ConeOf<T> cp = hsv_cone(hp), ct = hsv_cone(ht);
  s[3] = vadd(vadd(vabs(vsub(cp.a, ct.a)), vabs(vsub(cp.b, ct.b))), vabs(vsub(cp.v, ct.v)));
}

// Backward w.r.t. pred.  w[k] = d loss / d (sum of s_k), gL = d loss / d L_pred (from MS-SSIM).
template <class T>
CURL_HD void loss_terms_bwd(const T pred[3], const T target[3], T m, const float w[4], T gL, T gpred[3]) {
  T p[3], t[3];
#pragma unroll
  for (int c = 0; c < 3; ++c) { p[c] = vmul(pred[c], m); t[c] = vmul(target[c], m); }
  T g[3];
  // L1 in RGB
#pragma unroll
  for (int c = 0; c < 3; ++c) g[c] = vmul(vsign(vsub(p[c], t[c])), T(w[0]));
  // cosine: cos = sum pn_c tn_c, pn = p / max(|p|, eps); clamp_min passes gradient where |p| >= eps
  {
    T pn[3], tn[3], np2, ip;
    const T cs = cos_sim3(p, t, pn, tn, &np2, &ip);
    const T big = vgate_le(T(lk::COS_EPS * lk::COS_EPS), np2);   // 1 where the norm itself is used
    // d cos / d p_c = (tn_c - big * cos * p_c / |p|) / max(|p|, eps);  p_c/|p| == pn_c where big
    const T k = vmul(vmul(cs, big), T(w[1]));
#pragma unroll
    for (int c = 0; c < 3; ++c) g[c] = vfma(vmul(vsub(vmul(tn[c], T(w[1])), vmul(k, pn[c])), ip), T(1.0f), g[c]);
  }
  // Lab
  {
    T lp[3], lt[3], gl[3], gx[3];
    LabJac<T> j;
    rgb2lab_fwd<T, true>(p, lp, &j);
    rgb2lab_fwd<T, false>(t, lt, nullptr);
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const T cl = vsat(lp[c]);
      const T pass = vgate_eq(cl, lp[c]);
      T gc = vmul(vsign(vsub(cl, vsat(lt[c]))), T(w[2]));
      if (c == 0) gc = vadd(gc, gL);
      gl[c] = vmul(gc, pass);
    }
    rgb2lab_bwd(j, gl, gx);
#pragma unroll
    for (int c = 0; c < 3; ++c) g[c] = vadd(g[c], gx[c]);
  }
  // HSV cone
  {
    T hp[3], ht[3], gh[3], gx[3];
    HsvJac<T> j;
    rgb2hsv_fwd<T, true, false>(p, hp, &j);
    rgb2hsv_fwd<T, false, false>(t, ht, nullptr);
    const ConeOf<T> cp = hsv_cone(hp), ct = hsv_cone(ht);
    const T da = vmul(vsign(vsub(cp.a, ct.a)), T(w[3])), db = vmul(vsign(vsub(cp.b, ct.b)), T(w[3]));
    const T dv = vmul(vsign(vsub(cp.v, ct.v)), T(w[3]));
    const T rad = vfma(cp.sn, db, vmul(cp.cs, da));          // d / d (v s)
    const T tan_ = vsub(vmul(cp.cs, db), vmul(cp.sn, da));   // d / d theta, per unit (v s)
    gh[0] = vmul(vmul(vmul(cp.v, cp.s), tan_), T(lk::TWO_PI));
    gh[1] = vmul(cp.v, rad);
    gh[2] = vfma(cp.s, rad, dv);
    rgb2hsv_bwd(j, gh, gx);
#pragma unroll
    for (int c = 0; c < 3; ++c) g[c] = vadd(g[c], gx[c]);
  }
#pragma unroll
  for (int c = 0; c < 3; ++c) gpred[c] = vmul(g[c], m);     // p = pred * mask
}

}  // namespace curl
