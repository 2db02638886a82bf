// Per-pixel arithmetic of the reference's polynomial colour regression ("TriSpace" residual):
//   ChannelPolyLayer.forward              (reference model.py:319-333)
//   TriSpaceRegNet.generate_residual      (reference model.py:486-503)
//   TriSpaceRegNet.generate_image         (reference model.py:505-508)
// written, like curl_pixel.cuh, as templates over the lane type T (F2 on the device, float in the
// host-side derivation check).
//
// A degree-D polynomial in NV variables has NC = C(NV + D, D) coefficients per output channel.  The
// reference orders them the way ChannelPolyLayer.generate_powers (model.py:221-243) enumerates the
// exponent tuples: by total degree, and inside one degree lexicographically DESCENDING in
// (e0, e1, ...): 1, x0, x1, .., x0^2, x0 x1, .., x1^2, ...  mono_rank() is that position, computed at
// compile time; mono_visit() walks all monomials depth-first (each is its parent times one
// variable, so one multiply per monomial and at most D partial products alive) and hands every
// (rank, value) to a visitor.  Nothing is tabulated at run time and no [.., NC, H, W] tensor exists.
#pragma once

#include "curl_pixel.cuh"

namespace curl {

// ------------------------------------------------------------------------------------------------
// Compile-time combinatorics
// ------------------------------------------------------------------------------------------------
#if defined(__CUDACC__)
#define CURL_CX __host__ __device__ constexpr
#else
#define CURL_CX constexpr
#endif
CURL_CX int binom(int n, int r) {
  if (r < 0 || r > n) return 0;
  long long v = 1;
  for (int i = 1; i <= r; ++i) v = v * (n - r + i) / i;
  return (int)v;
}
// number of m-tuples of non-negative integers with sum s
CURL_CX int ncomp(int s, int m) { return m == 0 ? (s == 0 ? 1 : 0) : binom(s + m - 1, m - 1); }
CURL_CX int poly_num_coeffs(int nv, int deg) { return binom(nv + deg, deg); }

// position of the exponent tuple e[0..nv) in the reference's coefficient order
CURL_CX int mono_rank(int nv, int e0, int e1, int e2, int e3, int e4) {
  const int e[5] = {e0, e1, e2, e3, e4};
  int d = 0;
  for (int i = 0; i < nv; ++i) d += e[i];
  int r = 0;
  for (int j = 0; j < d; ++j) r += ncomp(j, nv);             // all monomials of lower degree
  int rem = d;
  for (int i = 0; i + 1 < nv; ++i) {                         // tuples of the same degree that come first
    for (int f = e[i] + 1; f <= rem; ++f) r += ncomp(rem - f, nv - 1 - i);
    rem -= e[i];
  }
  return r;
}

// ------------------------------------------------------------------------------------------------
// Depth-first walk over the monomials of degree 1..DEG (the constant term, rank 0, is the caller's).
// Visitor: template <int RANK> void operator()(T value).
// ------------------------------------------------------------------------------------------------
template <class T, int NV, int LEFT, int V, int E0, int E1, int E2, int E3, int E4>
struct MonoWalk {
  template <int v, class Vis>
  static CURL_HD void from(const T* x, T prod, Vis& vis) {
    if constexpr (v < NV) {
      constexpr int F0 = E0 + (v == 0), F1 = E1 + (v == 1), F2_ = E2 + (v == 2), F3 = E3 + (v == 3), F4 = E4 + (v == 4);
      const T p = vmul(prod, x[v]);
      vis.template operator()<mono_rank(NV, F0, F1, F2_, F3, F4)>(p);
      if constexpr (LEFT > 1) MonoWalk<T, NV, LEFT - 1, v, F0, F1, F2_, F3, F4>::template from<v>(x, p, vis);
      from<v + 1>(x, prod, vis);
    }
  }
};

template <class T, int NV, int DEG, class Vis>
CURL_HD void mono_visit(const T* x, Vis& vis) {
  // degree-1 monomials are the variables themselves: start the walk from prod = 1 without the multiply
  MonoWalk<T, NV, DEG, 0, 0, 0, 0, 0, 0>::template from<0>(x, T(1.0f), vis);
}

// out_c = sum_k coef[c][k] * monomial_k   (model.py:331-333).  The coefficients are staged as
// Coef4 records, one per monomial: {c0, c1, c2, -} for the three output channels, so one 128-bit
// shared-memory load (every lane the same address: a single broadcast wavefront) feeds three
// multiply-adds.  With one 32-bit load per multiply-add the forward kernel was bound by
// shared-memory wavefronts (1134 per 64 pixels = exactly its measured 16.4 Gpx/s).
struct alignas(16) Coef4 { float v[4]; };

// Evaluation is nested (multivariate Horner) over the same tree the walk above follows:
//   node(E) = coef[rank(E)] + sum_{v >= last variable of E} x_v * node(E + e_v),   p = node(0),
// one multiply-add per non-constant monomial and output channel (3 * 125 for 126 coefficients)
// instead of a multiply for the monomial plus three multiply-adds (4 * 125).
template <class T, int NV, int LEFT, int V, int E0, int E1, int E2, int E3, int E4>
struct HornerNode {
  template <int v>
  static CURL_HD void children(const T* x, const Coef4* coef, T acc[3]) {
    if constexpr (v < NV) {
      constexpr int F0 = E0 + (v == 0), F1 = E1 + (v == 1), F2_ = E2 + (v == 2), F3 = E3 + (v == 3), F4 = E4 + (v == 4);
      T ch[3];
      HornerNode<T, NV, LEFT - 1, v, F0, F1, F2_, F3, F4>::eval(x, coef, ch);
      acc[0] = vfma(x[v], ch[0], acc[0]);
      acc[1] = vfma(x[v], ch[1], acc[1]);
      acc[2] = vfma(x[v], ch[2], acc[2]);
      children<v + 1>(x, coef, acc);
    }
  }
  static CURL_HD void eval(const T* x, const Coef4* coef, T acc[3]) {
    const Coef4 cf = coef[mono_rank(NV, E0, E1, E2, E3, E4)];
    acc[0] = T(cf.v[0]); acc[1] = T(cf.v[1]); acc[2] = T(cf.v[2]);
    if constexpr (LEFT > 0) children<V>(x, coef, acc);
  }
};

template <class T, int NV, int DEG>
CURL_HD void poly_eval(const T* x, const Coef4* coef, T out[3]) {
  HornerNode<T, NV, DEG, 0, 0, 0, 0, 0, 0>::eval(x, coef, out);
}

// Gradient of the polynomial w.r.t. its three colour variables, for the image gradient:
//   d[j] = d/dx_j  sum_c w[c] * P_c(x),  j = 0..2   (w = the gradient arriving at the three outputs).
// The three channels are combined first (q_E = w . coef[E], per pixel), then ONE scalar nested evaluation carries its
// three derivatives along (forward mode): with node(E) = q_E + sum_{v >= V} x_v node(E + e_v),
//   d_j node(E) = sum_{v >= V} x_v d_j node(E + e_v)  +  node(E + e_j)          (j >= V; 0 for j < V).
// A node of remaining degree 0 has no derivative and a subtree entered through x or y (v >= 3) contains no colour
// variable; both are known at compile time, so no accumulator is ever initialised with a zero or fed one.
template <class T, int NV, int LEFT, int V, int E0, int E1, int E2, int E3, int E4>
struct HornerGrad {
  template <int v>
  static CURL_HD void children(const T* x, const Coef4* coef, const T w[3], T& val, T d[3]) {
    if constexpr (v < NV) {
      constexpr int F0 = E0 + (v == 0), F1 = E1 + (v == 1), F2_ = E2 + (v == 2), F3 = E3 + (v == 3), F4 = E4 + (v == 4);
      constexpr bool CHILD_HAS_D = LEFT - 1 > 0 && v < 3;       // the child's d[j], j in [v, 2]
      T cv, cd[3];
      HornerGrad<T, NV, LEFT - 1, v, F0, F1, F2_, F3, F4>::eval(x, coef, w, cv, cd);
      val = vfma(x[v], cv, val);
#pragma unroll
      for (int j = 0; j < 3; ++j) {
        if (j < v) continue;                                      // the child's subtree has no x_j
        const bool direct = j == v;                               // d/dx_v (x_v * child) has the term `child`
        const bool first = LEFT - 1 > 0 ? v == V : v == j;        // nothing has been added to d[j] yet
        if (direct && CHILD_HAS_D) d[j] = first ? vfma(x[v], cd[j], cv) : vfma(x[v], cd[j], vadd(d[j], cv));
        else if (direct) d[j] = first ? cv : vadd(d[j], cv);
        else if (CHILD_HAS_D) d[j] = first ? vmul(x[v], cd[j]) : vfma(x[v], cd[j], d[j]);
      }
      children<v + 1>(x, coef, w, val, d);
    }
  }
  static CURL_HD void eval(const T* x, const Coef4* coef, const T w[3], T& val, T d[3]) {
    const Coef4 cf = coef[mono_rank(NV, E0, E1, E2, E3, E4)];
    val = vfma(w[2], T(cf.v[2]), vfma(w[1], T(cf.v[1]), vmul(w[0], T(cf.v[0]))));
    if constexpr (LEFT > 0) children<V>(x, coef, w, val, d);
  }
};

template <class T, int NV, int DEG>
CURL_HD void poly_grad_colour(const T* x, const Coef4* coef, const T w[3], T d[3]) {
  static_assert(DEG >= 1, "a constant has no gradient");
  T val;
  HornerGrad<T, NV, DEG, 0, 0, 0, 0, 0, 0>::eval(x, coef, w, val, d);
}

// sigmoid and its derivative from the value
template <class T>
CURL_HD T vsigmoid(T p) {
  const T e = vex2(vmul(p, T(-1.4426950408889634f)));        // exp(-p)
  return vrcp(vadd(e, T(1.0f)));
}

// ------------------------------------------------------------------------------------------------
// TriSpaceRegNet.generate_residual (+ generate_image when clamp)      (model.py:486-508)
//   rgb_res = sigmoid(poly(rgb | x, y ; R))
//   lab_res = lab2rgb(sigmoid(poly(rgb2lab(rgb) | x, y ; L)))
//   hsv_res = hsv2rgb(sigmoid(poly(rgb2hsv(rgb) | x, y ; H)))
//   residual = 2 (rgb_res - 0.5) + 2 (lab_res - 0.5) + 2 (hsv_res - 0.5)
//   out = clamp ? clamp(rgb + residual, 0, 1) : residual
// NV = 3 (spatial=False) or 5 (spatial=True: x = column / width, y = row / height appended, model.py:472-484).
// coef: [space R, L, H][NC] Coef4 records.
// ------------------------------------------------------------------------------------------------
// sink(space, channel, value) receives the nine sigmoid outputs, which is all the backward needs besides the
// image to skip re-evaluating the three polynomials (the forward kernel stores them as they appear).
// ROLLED: the loop over the three spaces is NOT unrolled on the device.  Fully unrolled, ptxas schedules the spaces'
// coefficient loads and colour conversions across each other; that overlap is worth 5-14 % where the registers
// hold it, but the 126-coefficient forward WITHOUT stores of the sigmoids in between (the inference path) then
// spills 132-224 bytes per thread whatever its register cap (4K x 8: 4.38 ms; rolled 3.33 ms, no spills, and the
// polynomial code exists once).  Same operations in the same order either way: bit-identical results.
struct NoSigSink { template <class T> CURL_HD void operator()(int, int, T) const {} };
template <class T, int NV, int DEG, bool ROLLED = false, class Sink = NoSigSink>
CURL_HD void trispace_fwd(const T rgb[3], T xs, T ys, const Coef4* coef, bool clamp, T out[3], Sink sink = Sink()) {
  constexpr int NC = poly_num_coeffs(NV, DEG);
  T res[3] = {T(0.0f), T(0.0f), T(0.0f)};
#if defined(__CUDA_ARCH__)
#pragma unroll(ROLLED ? 1 : 3)
#endif
  for (int sp = 0; sp < 3; ++sp) {
    T v[5], p[3], s[3], r[3];
    v[3] = xs; v[4] = ys;
    if (sp == 0) { v[0] = rgb[0]; v[1] = rgb[1]; v[2] = rgb[2]; }
    else if (sp == 1) rgb2lab_fwd<T, false>(rgb, v, nullptr);
    else rgb2hsv_fwd<T, false, false>(rgb, v, nullptr);
    poly_eval<T, NV, DEG>(v, coef + sp * NC, p);
#pragma unroll
    for (int k = 0; k < 3; ++k) { s[k] = vsigmoid(p[k]); sink(sp, k, s[k]); }              // saved plane 3 * sp + k
    if (sp == 0) { r[0] = s[0]; r[1] = s[1]; r[2] = s[2]; }                                // model.py:490
    else if (sp == 1) lab2rgb_fwd<T, false>(s, r, nullptr);                                // model.py:491
    else hsv2rgb_fwd<T, false, false>(s, r, nullptr);                                      // model.py:492
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const T term = vfma(r[k], T(2.0f), T(-1.0f));                                        // model.py:497-499
      res[k] = sp == 0 ? term : vadd(res[k], term);                                        // model.py:501
    }
  }
#pragma unroll
  for (int k = 0; k < 3; ++k) out[k] = clamp ? vadd_sat(rgb[k], res[k]) : res[k];          // model.py:508
}

// Backward, per-pixel part (the image is data -- TriSpaceRegNet.forward feeds the input or target
// image, model.py:517-519 -- so only coefficient gradients exist).  Recomputes the forward and
// returns, per space s, the three colour variables the polynomial saw (vars[s]) and the gradient
// arriving at the polynomial's three outputs (gp[s]).  The coefficient gradient is then
//   d coef[s][c][k] = sum over pixels of gp[s][c] * monomial_k(vars[s] | x, y),
// a reduction the caller organises (curl_trispace.cu).
// sig (nullable): the nine sigmoid outputs the forward saved; when given, no polynomial is evaluated here.
// IN_JAC (the image gradient, trispace_gimg): also returns the derivative records of the two input conversions
// (the values they produce are the same with and without) and the gradient that reaches the image directly through
// generate_image's  clamp(img + residual)  (model.py:508): gdirect = gout * [pass], or 0 without the clamp.
template <class T, int NV, int DEG, bool IN_JAC = false>
CURL_HD void trispace_gpoly(const T rgb[3], T xs, T ys, const Coef4* coef, bool clamp, const T gout[3],
                            T vars[3][3], T gp[3][3], const T* sig = nullptr, LabJac<T>* jin_lab = nullptr,
                            HsvJac<T>* jin_hsv = nullptr, T* gdirect = nullptr) {
  constexpr int NC = poly_num_coeffs(NV, DEG);
  T v[5], p[3], s[3][3], r_lab[3], r_hsv[3];
  RgbJac<T> jl; Hsv2RgbJac<T> jh;
  v[3] = xs; v[4] = ys;
  v[0] = vars[0][0] = rgb[0]; v[1] = vars[0][1] = rgb[1]; v[2] = vars[0][2] = rgb[2];
  rgb2lab_fwd<T, IN_JAC>(rgb, vars[1], IN_JAC ? jin_lab : nullptr);
  rgb2hsv_fwd<T, IN_JAC, false>(rgb, vars[2], IN_JAC ? jin_hsv : nullptr);
  if (sig) {
#pragma unroll
    for (int sp = 0; sp < 3; ++sp)
#pragma unroll
      for (int k = 0; k < 3; ++k) s[sp][k] = sig[sp * 3 + k];
  } else if (IN_JAC) {
    // the image-gradient kernel: the three evaluations in a rolled loop (see trispace_gimg), operands and results
    // moved with selects; same operations as below, bit-identical sigmoids
#if defined(__CUDA_ARCH__)
#pragma unroll 1
#endif
    for (int sp = 0; sp < 3; ++sp) {
#pragma unroll
      for (int k = 0; k < 3; ++k) v[k] = sp == 0 ? vars[0][k] : (sp == 1 ? vars[1][k] : vars[2][k]);
      poly_eval<T, NV, DEG>(v, coef + sp * NC, p);
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        const T sk = vsigmoid(p[k]);
        s[0][k] = sp == 0 ? sk : s[0][k];
        s[1][k] = sp == 1 ? sk : s[1][k];
        s[2][k] = sp == 2 ? sk : s[2][k];
      }
    }
  } else {
    poly_eval<T, NV, DEG>(v, coef, p);
#pragma unroll
    for (int k = 0; k < 3; ++k) s[0][k] = vsigmoid(p[k]);
    v[0] = vars[1][0]; v[1] = vars[1][1]; v[2] = vars[1][2];
    poly_eval<T, NV, DEG>(v, coef + NC, p);
#pragma unroll
    for (int k = 0; k < 3; ++k) s[1][k] = vsigmoid(p[k]);
    v[0] = vars[2][0]; v[1] = vars[2][1]; v[2] = vars[2][2];
    poly_eval<T, NV, DEG>(v, coef + 2 * NC, p);
#pragma unroll
    for (int k = 0; k < 3; ++k) s[2][k] = vsigmoid(p[k]);
  }
  lab2rgb_fwd<T, true>(s[1], r_lab, &jl);
  hsv2rgb_fwd<T, true, false>(s[2], r_hsv, &jh);
  T g2[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    T g = gout[k];
    if (clamp) {
      const T res = vadd(vadd(vfma(s[0][k], T(2.0f), T(-1.0f)), vfma(r_lab[k], T(2.0f), T(-1.0f))),
                         vfma(r_hsv[k], T(2.0f), T(-1.0f)));
      const T sum = vadd(rgb[k], res);
      g = vmul(g, vgate_eq(vsat(sum), sum));       // clamp passes gradient on [0,1] inclusive
    }
    if (IN_JAC) gdirect[k] = clamp ? g : T(0.0f);
    g2[k] = vmul(g, T(2.0f));
  }
  T ga[3][3];
  ga[0][0] = g2[0]; ga[0][1] = g2[1]; ga[0][2] = g2[2];
  lab2rgb_bwd(jl, g2, ga[1]);
  hsv2rgb_bwd<T, false>(jh, g2, ga[2]);
#pragma unroll
  for (int sp = 0; sp < 3; ++sp)
#pragma unroll
    for (int k = 0; k < 3; ++k) gp[sp][k] = vmul(ga[sp][k], vmul(s[sp][k], vsub(T(1.0f), s[sp][k])));   // d sigmoid
}

// Image gradient of generate_residual (+ generate_image).  The reference obtains it from autograd whenever the image
// handed to TriSpaceRegNet requires a gradient (model.py:486-508 is differentiable in `img`; its own scripts feed
// data, so they never ask).  x = column / W and y = row / H are constants (model.py:444-452, frozen parameters).
//   gimg = gdirect + dQ_R/d rgb + J_rgb2lab^T dQ_L/d lab + J_rgb2hsv^T dQ_H/d hsv,   Q_s = gp[s] . P_s
template <class T, int NV, int DEG>
CURL_HD void trispace_gimg(const T rgb[3], T xs, T ys, const Coef4* coef, bool clamp, const T gout[3], T gimg[3],
                           const T* sig = nullptr) {
  constexpr int NC = poly_num_coeffs(NV, DEG);
  T vars[3][3], gp[3][3], gd[3];
  LabJac<T> jlab; HsvJac<T> jhsv;
  trispace_gpoly<T, NV, DEG, true>(rgb, xs, ys, coef, clamp, gout, vars, gp, sig, &jlab, &jhsv, gd);
  // The loop over the spaces stays rolled on the device (as in trispace_fwd<.., ROLLED>: unrolled, ptxas hoists the
  // three polynomials' coefficient loads across each other -- 6.6 KB of spills per thread in the 126-coefficient
  // kernel); the space's operands are picked with selects, never by a run-time index (that would put them in local memory).
#if defined(__CUDA_ARCH__)
#pragma unroll 1
#endif
  for (int sp = 0; sp < 3; ++sp) {
    T v[5], w[3], d[3], gx[3];
    v[3] = xs; v[4] = ys;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      v[k] = sp == 0 ? vars[0][k] : (sp == 1 ? vars[1][k] : vars[2][k]);
      w[k] = sp == 0 ? gp[0][k] : (sp == 1 ? gp[1][k] : gp[2][k]);
    }
    poly_grad_colour<T, NV, DEG>(v, coef + sp * NC, w, d);
    if (sp == 0) { gx[0] = d[0]; gx[1] = d[1]; gx[2] = d[2]; }
    else if (sp == 1) rgb2lab_bwd(jlab, d, gx);
    else rgb2hsv_bwd(jhsv, d, gx);
#pragma unroll
    for (int k = 0; k < 3; ++k) gimg[k] = sp == 0 ? (clamp ? vadd(gd[k], gx[k]) : gx[k]) : vadd(gimg[k], gx[k]);
  }
}

// visitor that writes monomial k to dst[k * pitch]
template <class T>
struct MonoStore {
  T* dst;
  int pitch;
  template <int RANK>
  CURL_HD void operator()(T m) { dst[RANK * pitch] = m; }
};

}  // namespace curl
