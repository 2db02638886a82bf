"""Drop-in replacements for the reference's curves.py functions, backed by the sm_100a kernels.

Same names, arguments and return arity: adjust_*(img, P) -> (img, reg[B]); apply_curve(img, C,
slope_sqr_diff, channel_in, channel_out) -> (img, slope_sqr_diff).  As shipped the reference's
adjust_* raise TypeError (curves.py:56,111,155 start the accumulator at None); these implement the
evident intent (accumulator starts at zero), like the oracle.
"""
from . import _lib
from .functional import AdjustFn, ApplyCurveFn


def apply_curve(img, C, slope_sqr_diff, channel_in, channel_out):
    """reference curves.py:4-38.  C holds the knots themselves (already exp'd)."""
    out, reg = ApplyCurveFn.apply(img, C, int(channel_in), int(channel_out))
    if slope_sqr_diff is None:
        slope_sqr_diff = 0.0
    return out, slope_sqr_diff + reg


def adjust_lab(img, L):
    """reference curves.py:136-180"""
    return AdjustFn.apply(img, L, _lib.ADJUST_LAB)


def adjust_rgb(img, R):
    """reference curves.py:90-133"""
    return AdjustFn.apply(img, R, _lib.ADJUST_RGB)


def adjust_hsv(img, S):
    """reference curves.py:41-87"""
    return AdjustFn.apply(img, S, _lib.ADJUST_HSV)
