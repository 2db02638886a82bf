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

    segment_clamp=True (not in the reference; default False) evaluates the curves as eq. (1) of the
    CURL paper defines them -- piecewise linear through all K knots, each segment term clamped to
    [0, 1] -- instead of the reference's unclamped sum (curves.py:29-32), through the
    curl_layer_seg_* kernels.
    """

    def __init__(self, num_lab_points=48, num_rgb_points=48, num_hsv_points=64, segment_clamp=False):
        super().__init__()
        self.segment_clamp = bool(segment_clamp)
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
                                 self.num_hsv_points // 4, self.segment_clamp)


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


class ChannelPolyLayer(nn.Module):
    """Holder of the reference's ChannelPolyLayer state (model.py:206-219): `degree`,
    `num_variables`, `num_out`, `num_coeffs` and the frozen `powers` parameter (state-dict key
    `polylayer.powers`), in the coefficient order of generate_powers (model.py:221-243).  On
    the path the reference's scripts run, the polynomial is evaluated fused inside the TriSpace kernels
    (curl_trispace_*: TriSpaceRegNet.generate_residual never calls this module's forward).  Called on its own --
    only the reference's PolyRegNet does that, which is outside this package's scope -- forward (model.py:296-333)
    runs on library element-wise ops, term by term, without the reference's [NC, B, V, H, W] power tensor."""

    def __init__(self, degree=3, num_variables=3, num_out=None):
        super().__init__()
        if not (isinstance(degree, int) and degree >= 0 and isinstance(num_variables, int) and num_variables >= 0):
            raise ValueError("`degree` and `num_variables` must be non-negative integers")
        self.degree = degree
        self.num_variables = num_variables
        self.num_out = num_variables if num_out is None else num_out
        powers = list(self.generate_powers(degree, num_variables))
        self.num_coeffs = len(powers)
        self.powers = nn.Parameter(torch.tensor(powers, dtype=torch.float32), requires_grad=False)

    @staticmethod
    def generate_powers(order, n_variables):
        """Exponent tuples by total degree, lexicographically descending inside a degree."""
        import itertools
        for d in range(order + 1):
            yield from sorted((t for t in itertools.product(range(d + 1), repeat=n_variables) if sum(t) == d),
                              reverse=True)

    def forward(self, img, coeffs):
        return _poly_forward(self, img, coeffs)


def _poly_forward(layer, img, coeffs):
    """ChannelPolyLayer.forward / Deg4MobilePolyLayer.forward (model.py:296-333, 345-408): img [B, V, H, W],
    coeffs [B, num_out, NC] -> [B, num_out, H, W] = sum_k coeffs[:, :, k] * prod_v img[:, v] ** powers[k, v].
    Stand-alone convenience on library ops (differentiable by autograd); not a kernel and not on the fused path."""
    if img.dim() != 4 or img.shape[1] != layer.num_variables:
        raise ValueError("There should be a polynomial variable per channel: expected [B, %d, H, W], got %s"
                         % (layer.num_variables, tuple(img.shape)))
    if coeffs.dim() != 3 or coeffs.shape[0] != img.shape[0] or coeffs.shape[1] != layer.num_out \
            or coeffs.shape[2] != layer.num_coeffs:
        raise ValueError("coeffs must have shape [%d, %d, %d] (degree %d, %d variables), got %s"
                         % (img.shape[0], layer.num_out, layer.num_coeffs, layer.degree, layer.num_variables,
                            tuple(coeffs.shape)))
    B = img.shape[0]
    pows = [[None, img[:, v:v + 1]] for v in range(layer.num_variables)]       # pows[v][p] = img[:, v] ** p
    for v in range(layer.num_variables):
        for _ in range(2, layer.degree + 1):
            pows[v].append(pows[v][-1] * pows[v][1])
    out = None
    for k, expo in enumerate(ChannelPolyLayer.generate_powers(layer.degree, layer.num_variables)):
        term = None
        for v, p in enumerate(expo):
            if p:
                term = pows[v][p] if term is None else term * pows[v][p]
        c = coeffs[:, :, k].reshape(B, layer.num_out, 1, 1)
        # (the constant term stays connected to img, as the reference's img ** 0 is: zero gradient, not "unused")
        contrib = c + img[:, 0:1] * 0.0 if term is None else c * term
        out = contrib if out is None else out + contrib
    return out


class Deg4MobilePolyLayer(nn.Module):
    """Holder of the reference's Deg4MobilePolyLayer state (model.py:336-343): `num_coeffs` = 126 and the
    frozen `powers` parameter of a degree-4, 5-variable polynomial, so that `infer.py:22-23` /
    `coreml_conversion.py:19-20` -- `TriSpaceRegNet(polynomial_order=4, spatial=True, is_train=False,
    polylayer=model.Deg4MobilePolyLayer())` -- construct unchanged and checkpoints written with either
    polynomial layer load into the other (the reference keeps `powers` for exactly that reason).  In the
    reference the class exists because Core ML cannot trace torch.pow; its products-written-out form and
    ChannelPolyLayer's differ by ~1e-5 in fp32 (tests/test_oracle_pins.py).  Here both run the same fused
    kernel (nested evaluation), which sits inside that spread of either."""

    degree = 4
    num_variables = 5
    num_out = 3

    def __init__(self):
        super().__init__()
        powers = list(ChannelPolyLayer.generate_powers(4, 5))
        self.num_coeffs = len(powers)           # 126
        self.powers = nn.Parameter(torch.tensor(powers, dtype=torch.float32), requires_grad=False)

    def forward(self, img, coeffs):
        return _poly_forward(self, img, coeffs)


class TriSpaceRegNet(nn.Module):
    """The reference's TriSpaceRegNet (model.py:411-520) with its per-pixel part -- colour
    conversions, three polynomials, sigmoids, inverse conversions, residual sum, final clamp -- as one
    fused sm_100a kernel forward and one backward.  Same constructor arguments, attribute and
    state-dict names (`polylayer`, `backbone`, `rgb2lab` .. `hsv2rgb`, `x`, `y`) and methods
    (generate_coefficients, generate_residual, generate_image, forward).  polynomial_order 1..4 (4 is the
    reference's default and what its scripts use).  `backbone`: any module mapping [B,3,H,W] -> [B, 9 * num_coeffs]; without
    it the reference's timm EfficientNetV2 + 4-layer head (model.py:427-435) is created, which needs
    timm.  The backbone stays on PyTorch/cuDNN."""

    def __init__(self, polynomial_order=4, spatial=False, max_resolution=10000, is_train=True,
                 use_sync_bn=False, polylayer=None, backbone=None, pretrained=True):
        super().__init__()
        self.num_channels = 3
        self.num_spaces = 3
        self.spatial = bool(spatial)
        self.num_in = self.num_channels + 2 * self.spatial
        self.order = polynomial_order
        self.is_train = is_train
        self.max_resolution = max_resolution
        self.polylayer = polylayer if polylayer is not None else ChannelPolyLayer(
            degree=self.order, num_variables=self.num_in, num_out=self.num_channels)
        self.num_coeffs = self.polylayer.num_coeffs
        want = len(list(ChannelPolyLayer.generate_powers(self.order, self.num_in)))
        if self.num_coeffs != want:
            raise ValueError("TriSpaceRegNet: polylayer has %d coefficients, a degree-%d polynomial in %d variables "
                             "(spatial=%s) has %d" % (self.num_coeffs, self.order, self.num_in, self.spatial, want))
        if backbone is None:
            try:
                import timm
            except ImportError as e:
                raise ImportError("TriSpaceRegNet: timm is not installed; pass backbone=<nn.Module producing "
                                  "[B, %d] coefficients>" % (9 * self.num_coeffs)) from e
            backbone = timm.create_model("efficientnetv2_rw_t", pretrained=pretrained)              # model.py:427
            if use_sync_bn:
                backbone = nn.SyncBatchNorm.convert_sync_batchnorm(backbone)
            backbone.classifier = nn.Sequential(nn.Linear(1024, 1024), nn.Linear(1024, 512), nn.Linear(512, 512),
                                                nn.Linear(512, 9 * self.num_coeffs))                # model.py:430-434
        self.backbone = backbone
        self.rgb2lab = colors.RGB2LAB()
        self.lab2rgb = colors.LAB2RGB()
        self.rgb2hsv = colors.RGB2HSV()
        self.hsv2rgb = colors.HSV2RGB()
        self.sigmoid = nn.Sigmoid()
        if not self.spatial:                                                                       # model.py:444-452
            x = torch.zeros(1, 0, 1, max_resolution)
            y = torch.zeros(1, 0, max_resolution, 1)
        else:
            x = torch.arange(0, max_resolution).reshape(1, 1, 1, max_resolution)
            y = torch.arange(0, max_resolution).reshape(1, 1, max_resolution, 1)
        self.x = nn.Parameter(x, requires_grad=False)
        self.y = nn.Parameter(y, requires_grad=False)
        # keep the nine sigmoid outputs (36 B/px) for the backward instead of recomputing the polynomials;
        # set False when memory is tighter than time (the reference keeps ~1.5 KB/px for autograd)
        self.save_activations = True

    def _fused(self, img, R, L, H, final_clamp):
        from .functional import TriSpaceFn
        if img.shape[2] > self.max_resolution or img.shape[3] > self.max_resolution:
            raise ValueError("img width and height must be less than `max_resolution` (%d)" % self.max_resolution)
        for m in (self.rgb2lab, self.lab2rgb, self.rgb2hsv, self.hsv2rgb):
            m._check_constants()
        return TriSpaceFn.apply(img, R, L, H, self.order, self.spatial, final_clamp, self.save_activations)

    def generate_residual(self, img, R, L, H):                                                     # model.py:486-503
        return self._fused(img, R, L, H, False)

    @staticmethod
    def generate_image(img, residual):                                                             # model.py:505-508
        return torch.clamp(img + residual, 0.0, 1.0)

    def generate_coefficients(self, img, mask):                                                    # model.py:510-515
        coeffs = self.backbone(img * mask).reshape(img.shape[0], self.num_spaces, self.num_channels, self.num_coeffs)
        return coeffs[:, 0], coeffs[:, 1], coeffs[:, 2]

    def forward(self, img, mask, target_img=None):                                                 # model.py:517-522
        R, L, H = self.generate_coefficients(img, mask)
        input_img = img if target_img is None else target_img
        # training: residual and generate_image in the same kernel; inference: the residual itself
        return self._fused(input_img, R, L, H, bool(self.is_train))


class CURLLoss(nn.Module):
    """Reference model.py:36-118, same constructor, attributes, state-dict keys and value:
        (L1_rgb + cosine_rgb + L1_lab + L1_hsv_cone + 10 * (1 - MS-SSIM(L))) / 5, masked.
    The colour terms -- masking, both RGB2LAB and both RGB2HSV conversions, the cone, the three L1
    sums and the cosine similarity -- are ONE kernel pass over (pred, target, mask) producing five
    sums per image and the two L planes; MS-SSIM of the L planes stays on the library convolution
    (metric.MSSSIMMetric).  The scalar arithmetic that combines the sums runs on the device (no
    host synchronisation), and autograd carries it back into the fused backward kernel."""

    def __init__(self, ssim_window_size=5, num_channel=1):
        super().__init__()
        from . import metric
        self.ssim_window_size = ssim_window_size
        self.num_channel = num_channel
        self.msssim_layer = metric.MSSSIMMetric(num_channel=num_channel)
        self.rgb2lab = colors.RGB2LAB()
        self.rgb2hsv = colors.RGB2HSV()

    def batch_lab_convert(self, img_batch):
        return torch.clamp(self.rgb2lab(img_batch), 0.0, 1.0)                       # model.py:55-58

    @staticmethod
    def batch_L_ssim_convert(lab_img_batch):
        return lab_img_batch[:, 0, :, :].unsqueeze(1)                               # model.py:60-62

    def batch_hsv_convert(self, img_batch):
        import math
        hsv = torch.clamp(self.rgb2hsv(img_batch), 0.0, 1.0)                        # model.py:64-79
        hue, sat, val = 2 * math.pi * hsv[:, 0], hsv[:, 1], hsv[:, 2]
        return torch.stack((val * sat * torch.cos(hue), val * sat * torch.sin(hue), val), 1)

    def terms(self, predicted_img_batch, target_img_batch, mask):
        """The five terms of model.py:91-108 as a dict of 0-d tensors (rgb, cosine, lab, hsv, ssim)."""
        from .functional import loss_terms
        for m in (self.rgb2lab, self.rgb2hsv):
            m._check_constants()
        B, C, H, W = predicted_img_batch.shape
        if isinstance(mask, torch.Tensor):
            # model.py:90 divides by `mask.sum()` of the tensor as passed: a broadcast [1,1,H,W] mask would make
            # the reference's terms B times larger than one mask row per image gives, so it is refused rather
            # than silently expanded; [B,H,W] is what data.py's loader yields before its unsqueeze
            if mask.dim() == 3:
                mask = mask.unsqueeze(1)
            if tuple(mask.shape) != (B, 1, H, W):
                raise ValueError("curl_b200: CURLLoss needs a mask of shape [%d, 1, %d, %d] (one row per image), got %s"
                                 % (B, H, W, tuple(mask.shape)))
        sums, L_pred, L_target = loss_terms(predicted_img_batch, target_img_batch, mask)
        tot = sums.sum(dim=0)
        n_px = float(B * H * W)
        if mask is None:
            n_masked = 0.0
        elif mask.dtype == torch.bool:
            n_masked = n_px - tot[4]
        else:
            n_masked = torch.logical_not(mask).expand(B, 1, H, W).sum()
        unmasked = C * tot[4]                                                       # model.py:90
        ssim = (1.0 - self.msssim_layer(L_pred, L_target)).mean()                   # model.py:104
        return {"rgb": tot[0] / unmasked,
                "cosine": 1.0 - (tot[1] + n_masked) / n_px,                         # model.py:95-96
                "lab": tot[2] / unmasked, "hsv": tot[3] / unmasked, "ssim": ssim}

    def forward(self, predicted_img_batch, target_img_batch, mask):
        t = self.terms(predicted_img_batch, target_img_batch, mask)
        return (t["rgb"] + t["cosine"] + t["lab"] + t["hsv"] + 10 * t["ssim"]) / 5  # model.py:110-116
