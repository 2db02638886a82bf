"""Drop-in replacement for the reference's model.py::CURLLayer (model.py:121-176), backed by one
fused sm_100a forward kernel and one fused recompute-backward kernel."""
import torch
import torch.nn as nn

from . import colors
from .functional import CurlLayerFn


class CURLLayer(nn.Module):
    """Same constructor, sub-module names (state-dict prefixes rgb2lab / lab2rgb / rgb2hsv /
    hsv2rgb) and forward(img, mask, L, R, H) -> (img, gradient_regulariser) as the reference.

    Only the leading num_lab_points / num_rgb_points / num_hsv_points columns of L / R / H are read
    (model.py:153,159,165); K = points / 3 (Lab, RGB) or / 4 (HSV) knots per curve.
    """

    def __init__(self, num_lab_points=48, num_rgb_points=48, num_hsv_points=64):
        super().__init__()
        if num_lab_points % 3 or num_rgb_points % 3 or num_hsv_points % 4:
            raise ValueError("num_lab_points and num_rgb_points must be multiples of 3, num_hsv_points of 4")
        if num_lab_points < 9 or num_rgb_points < 9 or num_hsv_points < 12:
            raise ValueError("each curve needs at least 3 knots")
        self.num_lab_points = num_lab_points
        self.num_rgb_points = num_rgb_points
        self.num_hsv_points = num_hsv_points
        self.rgb2lab = colors.RGB2LAB()
        self.lab2rgb = colors.LAB2RGB()
        self.rgb2hsv = colors.RGB2HSV()
        self.hsv2rgb = colors.HSV2RGB()

    def forward(self, img, mask, L=None, R=None, H=None):
        """forward(img, mask, L, R, H) as in the reference; forward(img, curve_params) -- the
        single-vector form BASELINE.json names -- splits curve_params like GCURLNet does
        (model.py:197-199) and uses an all-ones mask."""
        if L is None and R is None and H is None:
            curve_params, mask = mask, None
            b1 = self.num_lab_points
            b2 = b1 + self.num_rgb_points
            L, R, H = curve_params[:, :b1], curve_params[:, b1:b2], curve_params[:, b2:]
        for m in (self.rgb2lab, self.lab2rgb, self.rgb2hsv, self.hsv2rgb):
            m._check_constants()
        return CurlLayerFn.apply(img, mask, L, R, H, self.num_lab_points // 3, self.num_rgb_points // 3,
                                 self.num_hsv_points // 4)
