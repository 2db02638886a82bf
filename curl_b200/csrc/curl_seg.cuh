// Opt-in "segment-clamped" curves: the curve of eq. (1) of the CURL paper (arXiv 1911.13175),
//     scale(x) = k_0 + sum_{i=0}^{K-2} (k_{i+1} - k_i) * clamp((K-1) x - i, 0, 1),
// i.e. the piecewise-linear interpolant through all K knots, constant outside [0, 1].  The reference
// code evaluates something else (curves.py:29-32: no clamp, last slope left out -- affine in x,
// SURVEY.md F5), which is what every other kernel in this library reproduces; this file is the
// SURVEY.md section 8(f) row f4 flag for compatibility with the published method.  Everything
// around the scale is the reference's: out[c_out] = in[c_out] * scale(in[c_in]), whole-pixel clamp
// after every curve (curves.py:34-36), the curve wiring of adjust_lab / adjust_rgb / adjust_hsv, the
// regulariser (curves.py:24), the layer chain (model.py:137-176).
//
// Scalar lanes only (the segment lookup is a per-pixel gather).  Knot gradients need a per-SEGMENT
// reduction: pixel with segment j and fraction f sends  gs * (1 - f) to knot j and gs * f to knot
// j + 1 (gs = d loss / d scale), accumulated as two bins per segment, A_j = sum gs, B_j = sum gs f.
#pragma once

#include "curl_pixel.cuh"

namespace curl {

struct SegCurve {
  const float* C;   // K knots (already exp'ed)
  int K;
};

struct SegRec {     // what the reverse sweep needs from one apply_curve
  float scale, dscale, xout, f;
  int j;
  float pass[3];
};

// derivative semantics are torch.autograd's on the formula above: clamp passes gradient on its
// closed interval, so at an interior knot (t == j exactly) both neighbouring segments contribute.
CURL_HD void seg_eval(const SegCurve& cv, float x, float* scale, float* dscale, int* j_out, float* f_out) {
  const float km1 = (float)(cv.K - 1);
  const float t = x * km1;                                   // curve_steps * x, one rounding (curves.py:31)
  const float fl = floorf(t);
  const int j = (int)fminf(fmaxf(fl, 0.0f), (float)(cv.K - 2));
  const float u = t - (float)j;
  const float f = fminf(fmaxf(u, 0.0f), 1.0f);
  const float c0 = cv.C[j], sl = cv.C[j + 1] - c0;
  *scale = fmaf(sl, f, c0);
  float d = (u >= 0.0f && u <= 1.0f) ? sl : 0.0f;
  if (u == 0.0f && j >= 1) d += c0 - cv.C[j - 1];
  *dscale = d * km1;
  *j_out = j; *f_out = f;
}

// curves.apply_curve with the clamped scale
template <bool REC>
CURL_HD void seg_apply_fwd(const SegCurve& cv, const float x[3], int cin, int cout, float y[3], SegRec* r) {
  float s, ds, f; int j;
  seg_eval(cv, x[cin], &s, &ds, &j, &f);
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const float pre = (c == cout) ? x[c] * s : x[c];
    y[c] = vsat(pre);
    if (REC) r->pass[c] = (y[c] == pre) ? 1.0f : 0.0f;
  }
  if (REC) { r->scale = s; r->dscale = ds; r->xout = x[cout]; r->f = f; r->j = j; }
}

// BINS: void add(int curve, int j, float a, float b)
template <class BINS>
CURL_HD void seg_apply_bwd(const SegRec& r, int curve, int cin, int cout, const float gy[3], float gx[3], BINS& bins) {
  float g[3];
#pragma unroll
  for (int c = 0; c < 3; ++c) g[c] = gy[c] * r.pass[c];
  const float gs = g[cout] * r.xout;                         // d loss / d scale
#pragma unroll
  for (int c = 0; c < 3; ++c) gx[c] = (c == cout) ? g[c] * r.scale : g[c];
  gx[cin] = fmaf(gs, r.dscale, gx[cin]);
  bins.add(curve, r.j, gs, gs * r.f);
}

// the curve wiring: (channel_in, channel_out) of the layer's ten curves (curves.py:41-180)
CURL_HD int seg_cin(int k) { return k < 6 ? k % 3 : (k == 6 ? 0 : (k == 7 ? 0 : k - 7)); }
CURL_HD int seg_cout(int k) { return k < 6 ? k % 3 : (k == 6 ? 0 : (k == 7 ? 1 : k - 7)); }

template <int MK>
CURL_HD void seg_layer_fwd(const float x[3], float m, const SegCurve cv[LAYER_CURVES], float out[3]) {
  float a[3], b[3];
  rgb2lab_fwd<float, false, true>(x, a, nullptr);
#pragma unroll
  for (int k = 0; k < 3; ++k) { seg_apply_fwd<false>(cv[k], a, seg_cin(k), seg_cout(k), b, nullptr); a[0] = b[0]; a[1] = b[1]; a[2] = b[2]; }
  if (MK == MASK_F32) vscale3(a, m);
  lab2rgb_fwd<float, false, true>(a, b, nullptr);
#pragma unroll
  for (int k = 3; k < 6; ++k) { seg_apply_fwd<false>(cv[k], b, seg_cin(k), seg_cout(k), a, nullptr); b[0] = a[0]; b[1] = a[1]; b[2] = a[2]; }
  if (MK == MASK_F32) vscale3(b, m);
  rgb2hsv_fwd<float, false, false>(b, a, nullptr);
#pragma unroll
  for (int k = 6; k < 10; ++k) { seg_apply_fwd<false>(cv[k], a, seg_cin(k), seg_cout(k), b, nullptr); a[0] = b[0]; a[1] = b[1]; a[2] = b[2]; }
  if (MK == MASK_F32) vscale3(a, m);
  hsv2rgb_fwd<float, false, false>(a, b, nullptr);
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const float o = vadd_sat(x[c], b[c]);                    // model.py:170
    out[c] = (MK == MASK_NONE) ? o : o * m;
  }
}

template <int MK, bool WANT_GX, class BINS>
CURL_HD void seg_layer_bwd(const float x[3], float m, const SegCurve cv[LAYER_CURVES], const float gout[3],
                           float gx[3], BINS& bins) {
  float a[3], b[3];
  LabJac<float> j1; RgbJac<float> j3; HsvJac<float> j5; Hsv2RgbJac<float> j7;
  SegRec rec[LAYER_CURVES];
  rgb2lab_fwd<float, WANT_GX>(x, a, &j1);
#pragma unroll
  for (int k = 0; k < 3; ++k) { seg_apply_fwd<true>(cv[k], a, seg_cin(k), seg_cout(k), b, &rec[k]); a[0] = b[0]; a[1] = b[1]; a[2] = b[2]; }
  if (MK == MASK_F32) vscale3(a, m);
  lab2rgb_fwd<float, true>(a, b, &j3);
#pragma unroll
  for (int k = 3; k < 6; ++k) { seg_apply_fwd<true>(cv[k], b, seg_cin(k), seg_cout(k), a, &rec[k]); b[0] = a[0]; b[1] = a[1]; b[2] = a[2]; }
  if (MK == MASK_F32) vscale3(b, m);
  rgb2hsv_fwd<float, true, false>(b, a, &j5);
#pragma unroll
  for (int k = 6; k < 10; ++k) { seg_apply_fwd<true>(cv[k], a, seg_cin(k), seg_cout(k), b, &rec[k]); a[0] = b[0]; a[1] = b[1]; a[2] = b[2]; }
  if (MK == MASK_F32) vscale3(a, m);
  hsv2rgb_fwd<float, true, false>(a, b, &j7);
  float g[3], p[3], q[3];
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const float sum = x[c] + b[c];
    const float pass = (vsat(sum) == sum) ? 1.0f : 0.0f;     // clamp passes on [0,1] inclusive
    g[c] = gout[c] * ((MK == MASK_NONE) ? pass : pass * m);
  }
  hsv2rgb_bwd<float, false>(j7, g, p);
  if (MK == MASK_F32) vscale3(p, m);
#pragma unroll
  for (int k = 9; k >= 6; --k) { seg_apply_bwd(rec[k], k, seg_cin(k), seg_cout(k), p, q, bins); p[0] = q[0]; p[1] = q[1]; p[2] = q[2]; }
  rgb2hsv_bwd(j5, p, q);
  if (MK == MASK_F32) vscale3(q, m);
#pragma unroll
  for (int k = 5; k >= 3; --k) { seg_apply_bwd(rec[k], k, seg_cin(k), seg_cout(k), q, p, bins); q[0] = p[0]; q[1] = p[1]; q[2] = p[2]; }
  lab2rgb_bwd(j3, q, p);
  if (MK == MASK_F32) vscale3(p, m);
#pragma unroll
  for (int k = 2; k >= 0; --k) { seg_apply_bwd(rec[k], k, seg_cin(k), seg_cout(k), p, q, bins); p[0] = q[0]; p[1] = q[1]; p[2] = q[2]; }
  if (WANT_GX) {
    rgb2lab_bwd(j1, p, q);
    gx[0] = g[0] + q[0]; gx[1] = g[1] + q[1]; gx[2] = g[2] + q[2];
  }
}

// d loss / d param[i] of one curve from its segment bins (A_j = sum gs, B_j = sum gs f, j = 0..K-2)
// and greg = d loss / d regulariser; the regulariser part is curve_knot_grad's (curves.py:24).
CURL_HD float seg_knot_grad(const float* C, int K, int i, const double* A, const double* B, double greg, bool log_domain) {
  double dC = 0.0;
  if (i <= K - 2) dC += A[i] - B[i];
  if (i >= 1) dC += B[i - 1];
  double r = 0.0;
  for (int t = 0; t < 3; ++t) {
    int j = i - 2 + t;
    if (j < 0 || j > K - 3) continue;
    double d = (double)C[j + 2] - 2.0 * (double)C[j + 1] + (double)C[j];
    r += (t == 1 ? -2.0 : 1.0) * d;
  }
  dC += 2.0 * r * greg;
  return (float)(log_domain ? dC * (double)C[i] : dC);
}

}  // namespace curl
