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


class GCURLNet(nn.Module):
    """The reference's only in-tree caller of CURLLayer (model.py:179-203): a backbone predicts the
    knot vector [L | R | H] for each image and the curve layer applies it to that same image.

    Same constructor arguments, attribute names (`backbone`, `curllayer`, `curve_break_1/2`) and
    forward(img, mask, L, R, H) as the reference; like there, the L/R/H arguments are placeholders that
    the predicted curves replace (model.py:196-199).  The split is three strided views of the backbone
    output -- no copy: the kernels take a row stride.

    The reference cannot be constructed as shipped (its classifier width reads the undefined
    `self.num_spaces/num_channels/num_coeffs`, model.py:190-191); the evident intent is one output
    per knot, which is what this class does.  `backbone` may be given explicitly: any module mapping
    [B,3,H,W] -> [B, >= num_lab_points + num_rgb_points + num_hsv_points].  Without it the
    reference's timm EfficientNetV2 is built (model.py:189) and an ImportError is raised if timm is
    missing.  The backbone stays on PyTorch/cuDNN (BASELINE.json north_star).
    """

    def __init__(self, num_lab_points=48, num_rgb_points=48, num_hsv_points=64, backbone=None, pretrained=True):
        super().__init__()
        self.num_lab_points = num_lab_points
        self.num_rgb_points = num_rgb_points
        self.num_hsv_points = num_hsv_points
        self.curve_break_1 = num_lab_points
        self.curve_break_2 = num_lab_points + num_rgb_points
        self.num_curve_params = num_lab_points + num_rgb_points + num_hsv_points
        if backbone is None:
            try:
                import timm
            except ImportError as e:
                raise ImportError("GCURLNet: timm is not installed; pass backbone=<nn.Module producing "
                                  "[B, %d] curve parameters>" % self.num_curve_params) from e
            backbone = timm.create_model("efficientnetv2_rw_s", pretrained=pretrained)        # model.py:189
            backbone.classifier = nn.Sequential(nn.Linear(in_features=1792, out_features=self.num_curve_params))
        self.backbone = backbone
        self.curllayer = CURLLayer(num_lab_points, num_rgb_points, num_hsv_points)

    def forward(self, img, mask, L=None, R=None, H=None):
        curves = self.backbone(img)
        if curves.dim() != 2 or curves.shape[1] < self.num_curve_params:
            raise ValueError("GCURLNet: backbone must return [B, >= %d], got %s"
                             % (self.num_curve_params, tuple(curves.shape)))
        L = curves[:, :self.curve_break_1]
        R = curves[:, self.curve_break_1:self.curve_break_2]
        H = curves[:, self.curve_break_2:self.num_curve_params]
        return self.curllayer(img, mask, L, R, H)


def _forward_u8(self, img_u8, mask, L, R, H, layout="CHW", out_layout=None):
    """Inference on 8-bit images: uint8 in -> x/255 -> the fused layer -> trunc(out*255) -> uint8 out,
    i.e. TF.to_tensor (data.py:152-158) and the `(out*255).astype('uint8')` of evaluate.py:63 folded
    into the kernel: 7 B/px of traffic instead of 25.  Returns (uint8 image, gradient_regulariser)."""
    from .functional import curl_layer_u8
    for m in (self.rgb2lab, self.lab2rgb, self.rgb2hsv, self.hsv2rgb):
        m._check_constants()
    return curl_layer_u8(img_u8, mask, L, R, H, self.num_lab_points // 3, self.num_rgb_points // 3,
                         self.num_hsv_points // 4, layout=layout, out_layout=out_layout)


CURLLayer.forward_u8 = _forward_u8
