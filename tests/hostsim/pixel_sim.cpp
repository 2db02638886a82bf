// TEST INFRASTRUCTURE ONLY -- host instantiation of curl_b200/csrc/curl_pixel.cuh.
//
// The per-pixel forward/backward functions the sm_100a kernels are built from are plain templates
// over a math policy.  This file instantiates them with T = float on the host (double-precision libm
// rounded to fp32 in place of MUFU.LG2/EX2/RCP) behind a small C interface so that the CPU-only test suite
// can check the hand-derived reverse-mode code, the tie/clamp/branch semantics and the knot-gradient
// moment expansion against the oracle without a GPU.  It is compiled by tests/ on demand into
// tests/hostsim/_build/ and is never part of, linked into, or reachable from the product library.
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include <vector>

#include "../../curl_b200/csrc/curl_pixel.cuh"
#include "../../curl_b200/csrc/curl_poly.cuh"
#include "../../curl_b200/csrc/curl_loss.cuh"
#include "../../curl_b200/csrc/curl_seg.cuh"

using namespace curl;

namespace {

struct Seg { const float* p; int64_t ld; int nc, K; float* g; };

struct Curves {
  std::vector<float> C, ab;
  double reg = 0.0;
  bool layer = false;        // the fused layer's folded units (curve_fold_m, layer_beta1)
  // ab: (alpha, beta) per curve, then layer_beta1 -- the layout the kernels stage in shared memory
  void stage(const Seg* segs, int nseg, int b) {
    C.clear(); ab.clear(); reg = 0.0;
    for (int s = 0; s < nseg; ++s) {
      const int n = segs[s].nc * segs[s].K;
      for (int i = 0; i < n; ++i) C.push_back((float)exp((double)segs[s].p[(size_t)b * segs[s].ld + i]));
    }
    int off = 0;
    for (int s = 0; s < nseg; ++s)
      for (int c = 0; c < segs[s].nc; ++c) {
        float a, bt;
        curve_alpha_beta(C.data() + off, segs[s].K, &a, &bt);
        ab.push_back(a); ab.push_back(bt);
        reg += curve_regulariser(C.data() + off, segs[s].K);
        off += segs[s].K;
      }
    if (layer) ab.push_back(layer_beta1(ab[2 * 7 + 1]));
  }
  void grads(const Seg* segs, int nseg, int b, const std::vector<double>& mom, double greg) const {
    int off = 0, cb = 0;
    for (int s = 0; s < nseg; ++s) {
      const int K = segs[s].K;
      for (int c = 0; c < segs[s].nc; ++c) {
        for (int i = 0; i < K; ++i)
          segs[s].g[(size_t)b * segs[s].nc * K + c * K + i] =
              curve_knot_grad(C.data() + off, K, i, mom[2 * (cb + c)] * curve_fold_m(layer, cb + c),
                              mom[2 * (cb + c) + 1] * curve_fold_m(layer, cb + c), greg, true);
        off += K;
      }
      cb += segs[s].nc;
    }
  }
};

float mask_at(const void* mask, int kind, size_t i) {
  if (kind == MASK_NONE) return 1.0f;
  if (kind == MASK_U8) return ((const unsigned char*)mask)[i] ? 1.0f : 0.0f;
  return ((const float*)mask)[i];
}

template <int MK>
void layer_fwd_mk(const float x[3], float m, const float* ab, float y[3]) { layer_fwd<float, MK>(x, m, ab, y); }
template <int MK>
void layer_bwd_mk(const float x[3], float m, const float* ab, const float gy[3], float gx[3], float* mom) {
  layer_bwd<float, MK, true>(x, m, ab, gy, gx, mom, KeepInRegisters());
}

}  // namespace

extern "C" {

int sim_layer(int do_bwd, const float* img, const void* mask, int mask_kind, const float* L, int64_t ldL,
              const float* R, int64_t ldR, const float* Hk, int64_t ldH, float* out, float* reg,
              const float* gout, const float* greg, float* gimg, float* gL, float* gR, float* gH, int B, int HW,
              int Kl, int Kr, int Kh) {
  Seg segs[3] = {{L, ldL, 3, Kl, gL}, {R, ldR, 3, Kr, gR}, {Hk, ldH, 4, Kh, gH}};
  Curves cv;
  cv.layer = true;
  for (int b = 0; b < B; ++b) {
    cv.stage(segs, 3, b);
    if (reg) reg[b] = (float)cv.reg;
    std::vector<float> mom(20, 0.0f);
    std::vector<double> momd(20, 0.0);
    for (int i = 0; i < HW; ++i) {
      const size_t o = (size_t)b * 3 * HW + i;
      const float x[3] = {img[o], img[o + HW], img[o + 2 * (size_t)HW]};
      const float m = mask_at(mask, mask_kind, (size_t)b * HW + i);
      if (!do_bwd) {
        float y[3];
        if (mask_kind == MASK_NONE) layer_fwd_mk<MASK_NONE>(x, m, cv.ab.data(), y);
        else if (mask_kind == MASK_U8) layer_fwd_mk<MASK_U8>(x, m, cv.ab.data(), y);
        else layer_fwd_mk<MASK_F32>(x, m, cv.ab.data(), y);
        out[o] = y[0]; out[o + HW] = y[1]; out[o + 2 * (size_t)HW] = y[2];
      } else {
        const float gy[3] = {gout[o], gout[o + HW], gout[o + 2 * (size_t)HW]};
        float gx[3];
        for (auto& v : mom) v = 0.0f;
        if (mask_kind == MASK_NONE) layer_bwd_mk<MASK_NONE>(x, m, cv.ab.data(), gy, gx, mom.data());
        else if (mask_kind == MASK_U8) layer_bwd_mk<MASK_U8>(x, bwd_mask_lane_u8(m != 0.0f), cv.ab.data(), gy, gx, mom.data());
        else layer_bwd_mk<MASK_F32>(x, m, cv.ab.data(), gy, gx, mom.data());
        for (int k = 0; k < 20; ++k) momd[k] += (double)mom[k];
        if (gimg) { gimg[o] = gx[0]; gimg[o + HW] = gx[1]; gimg[o + 2 * (size_t)HW] = gx[2]; }
      }
    }
    if (do_bwd) cv.grads(segs, 3, b, momd, greg ? (double)greg[b] : 0.0);
  }
  return 0;
}

int sim_color(int do_bwd, int conv, const float* in, const float* gout, float* out, int B, int HW) {
  for (int b = 0; b < B; ++b)
    for (int i = 0; i < HW; ++i) {
      const size_t o = (size_t)b * 3 * HW + i;
      const float x[3] = {in[o], in[o + HW], in[o + 2 * (size_t)HW]};
      float y[3], r[3];
      float gy[3] = {0, 0, 0};
      if (do_bwd) { gy[0] = gout[o]; gy[1] = gout[o + HW]; gy[2] = gout[o + 2 * (size_t)HW]; }
      switch (conv) {
        case 0: { LabJac<float> j; rgb2lab_fwd<float, true>(x, y, &j); rgb2lab_bwd(j, gy, r); break; }
        case 1: { RgbJac<float> j; lab2rgb_fwd<float, true>(x, y, &j); lab2rgb_bwd(j, gy, r); break; }
        case 2: { HsvJac<float> j; rgb2hsv_fwd<float, true, false>(x, y, &j); rgb2hsv_bwd(j, gy, r); break; }
        case 3: { Hsv2RgbJac<float> j; hsv2rgb_fwd<float, true, false>(x, y, &j); hsv2rgb_bwd<float, false>(j, gy, r); break; }
        default: return -1;
      }
      const float* w = do_bwd ? r : y;
      out[o] = w[0]; out[o + HW] = w[1]; out[o + 2 * (size_t)HW] = w[2];
    }
  return 0;
}

int sim_adjust(int do_bwd, int group, const float* in, const float* P, int64_t ldP, float* out, float* reg,
               const float* gout, const float* greg, float* gin, float* gP, int B, int HW, int K) {
  const int nc = group == 2 ? 4 : 3;
  Seg segs[1] = {{P, ldP, nc, K, gP}};
  Curves cv;
  for (int b = 0; b < B; ++b) {
    cv.stage(segs, 1, b);
    if (reg) reg[b] = (float)cv.reg;
    std::vector<double> momd(2 * nc, 0.0);
    for (int i = 0; i < HW; ++i) {
      const size_t o = (size_t)b * 3 * HW + i;
      const float x[3] = {in[o], in[o + HW], in[o + 2 * (size_t)HW]};
      float y[3], gx[3];
      float gy[3] = {0, 0, 0};
      float mom[8] = {0, 0, 0, 0, 0, 0, 0, 0};
      if (do_bwd) { gy[0] = gout[o]; gy[1] = gout[o + HW]; gy[2] = gout[o + 2 * (size_t)HW]; }
      if (group == 2) {
        HsvCurveJac<float> j;
        adjust_hsv_fwd<float, true, false>(x, cv.ab.data(), y, &j);
        adjust_hsv_bwd(j, cv.ab[3], gy, gx, mom);
      } else {
        Self3Jac<float> j;
        adjust_self3_fwd<float, true>(x, cv.ab.data(), y, &j);
        adjust_self3_bwd(j, gy, gx, mom);
      }
      if (!do_bwd) { out[o] = y[0]; out[o + HW] = y[1]; out[o + 2 * (size_t)HW] = y[2]; }
      else {
        for (int k = 0; k < 2 * nc; ++k) momd[k] += (double)mom[k];
        if (gin) { gin[o] = gx[0]; gin[o + HW] = gx[1]; gin[o + 2 * (size_t)HW] = gx[2]; }
      }
    }
    if (do_bwd) cv.grads(segs, 1, b, momd, greg ? (double)greg[b] : 0.0);
  }
  return 0;
}

// TriSpace polynomial residual (curl_poly.cuh), host instantiation.  do_bwd == 0: out[B][3][HW];
// do_bwd == 2: out = image gradient; do_bwd == 1: gR/gL/gH[B][3][NC] accumulated in double over the image (the device kernel's tiling
// and reduction are not modelled -- only the per-pixel math and the monomial order).
}  // extern "C"
template <int NV, int DEG>
static void sim_trispace_nv(int do_bwd, const float* img, const float* R, const float* L, const float* Hc, int clamp,
                            const float* gout, float* out, float* gR, float* gL, float* gH, int B, int H, int W) {
  constexpr int NC = poly_num_coeffs(NV, DEG);
  const int HW = H * W;
  std::vector<Coef4> coef(3 * NC);
  std::vector<float> terms(NC);
  std::vector<double> acc(9 * NC);
  for (int b = 0; b < B; ++b) {
    const float* src[3] = {R, L, Hc};
    for (int s = 0; s < 3; ++s)
      for (int k = 0; k < NC; ++k) {
        for (int c = 0; c < 3; ++c) coef[s * NC + k].v[c] = src[s][((size_t)b * 3 + c) * NC + k];
        coef[s * NC + k].v[3] = 0.f;
      }
    std::fill(acc.begin(), acc.end(), 0.0);
    for (int i = 0; i < HW; ++i) {
      const float x[3] = {img[((size_t)b * 3 + 0) * HW + i], img[((size_t)b * 3 + 1) * HW + i], img[((size_t)b * 3 + 2) * HW + i]};
      const float xs = NV == 5 ? (float)(i % W) / (float)W : 0.f, ys = NV == 5 ? (float)(i / W) / (float)H : 0.f;
      if (!do_bwd) {
        float y[3];
        trispace_fwd<float, NV, DEG>(x, xs, ys, coef.data(), clamp != 0, y);
        for (int c = 0; c < 3; ++c) out[((size_t)b * 3 + c) * HW + i] = y[c];
      } else if (do_bwd == 2) {                        // image gradient -> out
        const float g[3] = {gout[((size_t)b * 3 + 0) * HW + i], gout[((size_t)b * 3 + 1) * HW + i], gout[((size_t)b * 3 + 2) * HW + i]};
        float y[3];
        trispace_gimg<float, NV, DEG>(x, xs, ys, coef.data(), clamp != 0, g, y);
        for (int c = 0; c < 3; ++c) out[((size_t)b * 3 + c) * HW + i] = y[c];
      } else {
        const float g[3] = {gout[((size_t)b * 3 + 0) * HW + i], gout[((size_t)b * 3 + 1) * HW + i], gout[((size_t)b * 3 + 2) * HW + i]};
        float vars[3][3], gp[3][3];
        trispace_gpoly<float, NV, DEG>(x, xs, ys, coef.data(), clamp != 0, g, vars, gp);
        for (int s = 0; s < 3; ++s) {
          const float v[5] = {vars[s][0], vars[s][1], vars[s][2], xs, ys};
          terms[0] = 1.0f;
          MonoStore<float> st{terms.data(), 1};
          mono_visit<float, NV, DEG>(v, st);
          for (int c = 0; c < 3; ++c)
            for (int k = 0; k < NC; ++k) acc[(s * 3 + c) * NC + k] += (double)gp[s][c] * (double)terms[k];
        }
      }
    }
    if (do_bwd == 1) {
      float* dst[3] = {gR, gL, gH};
      for (int s = 0; s < 3; ++s)
        for (int j = 0; j < 3 * NC; ++j) dst[s][(size_t)b * 3 * NC + j] = (float)acc[s * 3 * NC + j];
    }
  }
}
extern "C" {
int sim_trispace_deg(int do_bwd, const float* img, const float* R, const float* L, const float* Hc, int spatial, int degree,
                     int clamp, const float* gout, float* out, float* gR, float* gL, float* gH, int B, int H, int W) {
#define SIM_TRI(D)                                                                                  \
  case D:                                                                                           \
    if (spatial) sim_trispace_nv<5, D>(do_bwd, img, R, L, Hc, clamp, gout, out, gR, gL, gH, B, H, W); \
    else sim_trispace_nv<3, D>(do_bwd, img, R, L, Hc, clamp, gout, out, gR, gL, gH, B, H, W);         \
    return 0;
  switch (degree) {
    SIM_TRI(1) SIM_TRI(2) SIM_TRI(3) SIM_TRI(4)
    default: return -1;
  }
#undef SIM_TRI
}
int sim_trispace(int do_bwd, const float* img, const float* R, const float* L, const float* Hc, int spatial, int clamp,
                 const float* gout, float* out, float* gR, float* gL, float* gH, int B, int H, int W) {
  return sim_trispace_deg(do_bwd, img, R, L, Hc, spatial, 4, clamp, gout, out, gR, gL, gH, B, H, W);
}

// 8-bit I/O conversions of curl_layer_fwd_u8 (curl_vec.cuh), host instantiation: out[i] = u8_to_unit(i)
// for i = 0..255, and unit_to_u8 over an arbitrary array.
void sim_u8_to_unit_table(float* out) {
  for (int i = 0; i < 256; ++i) out[i] = curl::u8_to_unit<float>((float)i);
}
void sim_unit_to_u8(const float* y, unsigned char* out, long long n) {
  for (long long i = 0; i < n; ++i) out[i] = (unsigned char)curl::unit_to_u8(y[i]);
}

// Colour terms of CURLLoss (curl_loss.cuh), host instantiation.  do_bwd = 0: sums[B][5] (double
// accumulation) and the two L planes; do_bwd = 1: gpred given w[B][4] and gL (nullable).
int sim_loss_terms(int do_bwd, const float* pred, const float* target, const float* mask, float* sums, float* L_pred,
                   float* L_target, const float* w, const float* gL, float* gpred, int B, int HW) {
  for (int b = 0; b < B; ++b) {
    double acc[5] = {0, 0, 0, 0, 0};
    for (int i = 0; i < HW; ++i) {
      const float x[3] = {pred[((size_t)b * 3 + 0) * HW + i], pred[((size_t)b * 3 + 1) * HW + i], pred[((size_t)b * 3 + 2) * HW + i]};
      const float y[3] = {target[((size_t)b * 3 + 0) * HW + i], target[((size_t)b * 3 + 1) * HW + i], target[((size_t)b * 3 + 2) * HW + i]};
      const float m = mask ? mask[(size_t)b * HW + i] : 1.0f;
      if (!do_bwd) {
        float s[4], lp, lt;
        loss_terms_fwd<float>(x, y, m, s, &lp, &lt);
        for (int k = 0; k < 4; ++k) acc[k] += (double)s[k];
        acc[4] += (double)m;
        L_pred[(size_t)b * HW + i] = lp; L_target[(size_t)b * HW + i] = lt;
      } else {
        float g[3];
        loss_terms_bwd<float>(x, y, m, w + b * 4, gL ? gL[(size_t)b * HW + i] : 0.0f, g);
        for (int c = 0; c < 3; ++c) gpred[((size_t)b * 3 + c) * HW + i] = g[c];
      }
    }
    if (!do_bwd) for (int k = 0; k < 5; ++k) sums[b * 5 + k] = (float)acc[k];
  }
  return 0;
}

// Segment-clamped layer (curl_seg.cuh), host instantiation with double bins.  knots: per image the
// 10 curves' log-domain parameters, concatenated [3*Kl | 3*Kr | 4*Kh]; mask: float or null.
struct HostBins {
  double* A; double* B; const int* off;
  void add(int curve, int j, float a, float b) { A[off[curve] + j] += (double)a; B[off[curve] + j] += (double)b; }
};
int sim_layer_seg(int do_bwd, const float* img, const float* mask, const float* knots, int Kl, int Kr, int Kh,
                  float* out, float* reg, const float* gout, const float* greg, float* gimg, float* gknots, int B, int HW) {
  const int Ks[10] = {Kl, Kl, Kl, Kr, Kr, Kr, Kh, Kh, Kh, Kh};
  int off[11]; off[0] = 0;
  for (int k = 0; k < 10; ++k) off[k + 1] = off[k] + Ks[k];
  const int NK = off[10];
  std::vector<float> C(NK);
  std::vector<double> A(NK), Bv(NK);
  for (int b = 0; b < B; ++b) {
    for (int i = 0; i < NK; ++i) C[i] = (float)exp((double)knots[(size_t)b * NK + i]);
    SegCurve cv[10];
    for (int k = 0; k < 10; ++k) cv[k] = SegCurve{C.data() + off[k], Ks[k]};
    std::fill(A.begin(), A.end(), 0.0); std::fill(Bv.begin(), Bv.end(), 0.0);
    HostBins bins{A.data(), Bv.data(), off};
    for (int i = 0; i < HW; ++i) {
      const float x[3] = {img[((size_t)b * 3 + 0) * HW + i], img[((size_t)b * 3 + 1) * HW + i], img[((size_t)b * 3 + 2) * HW + i]};
      const float m = mask ? mask[(size_t)b * HW + i] : 1.0f;
      if (!do_bwd) {
        float y[3];
        if (mask) seg_layer_fwd<MASK_F32>(x, m, cv, y); else seg_layer_fwd<MASK_NONE>(x, m, cv, y);
        for (int c = 0; c < 3; ++c) out[((size_t)b * 3 + c) * HW + i] = y[c];
      } else {
        const float g[3] = {gout[((size_t)b * 3 + 0) * HW + i], gout[((size_t)b * 3 + 1) * HW + i], gout[((size_t)b * 3 + 2) * HW + i]};
        float gx[3];
        if (mask) seg_layer_bwd<MASK_F32, true>(x, m, cv, g, gx, bins); else seg_layer_bwd<MASK_NONE, true>(x, m, cv, g, gx, bins);
        for (int c = 0; c < 3; ++c) gimg[((size_t)b * 3 + c) * HW + i] = gx[c];
      }
    }
    if (!do_bwd) {
      double r = 0.0;
      for (int k = 0; k < 10; ++k) r += curve_regulariser(C.data() + off[k], Ks[k]);
      reg[b] = (float)r;
    } else {
      for (int k = 0; k < 10; ++k)
        for (int i = 0; i < Ks[k]; ++i)
          gknots[(size_t)b * NK + off[k] + i] = seg_knot_grad(C.data() + off[k], Ks[k], i, A.data() + off[k], Bv.data() + off[k],
                                                              greg ? (double)greg[b] : 0.0, true);
    }
  }
  return 0;
}

}  // extern "C"
