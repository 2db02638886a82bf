"""TEST INFRASTRUCTURE ONLY -- CPU oracle for the CURL curve-layer hot path.

A functional restatement (torch tensor ops on CPU, fp32 or fp64) of the reference's per-pixel
curve layer: colour chain (reference colors.py), knot curves (reference curves.py) and their
composition (reference model.py:137-176).  Gradients come from torch.autograd over this
restatement, which is how the reference itself differentiates the path.

Who may use it: tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs, as the *checker* and the *timed CPU baseline* only.  Nothing under curl_b200/ imports it; the
product path is the CUDA extension and fails loudly without it.

Parity pinning: the reference ships no tests or golden vectors for this path,
so the oracle is pinned by execution: tests/test_oracle_pins.py requires this file to be
BIT-IDENTICAL (fp32 and fp64, forward and backward) to the patched reference loaded from
/root/reference by oracle/ref_loader.py, and tests/golden/*.npz (generated from the reference
itself by oracle/make_golden.py) travel to the GPU box where the reference does not exist.
To stay bit-identical the operation ORDER of the reference is kept (both-branch * 0/1-mask
blends, einsum contractions, in-place channel writes), even where it is wasteful -- it is also
what makes this a fair stand-in when timed as the reference's CPU path.

The reference's curve code crashes as shipped (curves.py:56/111/155 initialise the regulariser
accumulator to None); the oracle, like oracle/ref_loader.py, starts it at 0.0.
"""
import torch

EPS = 6.0 / 29.0  # colors.py:43,106

# ---------------------------------------------------------------------------------------------
# Frozen colour constants, in the reference's registered shapes (colors.py:8-25, 69-86, 184).
# ---------------------------------------------------------------------------------------------

def _mat(rows, dtype):
    # reference builds [3,3] rows=input channel, then .transpose(1,0) and two unsqueeze(0)
    return torch.tensor(rows, dtype=torch.float).transpose(1, 0).unsqueeze(0).unsqueeze(0).to(dtype)


def constants(dtype=torch.float32):
    """The nine frozen parameters of CURLLayer().state_dict(), keyed like the reference."""
    white = torch.tensor([0.950456, 1.0, 1.088754], dtype=torch.float).reshape(1, 3, 1, 1).to(dtype)
    off = torch.tensor([16.0, 0.0, 0.0], dtype=torch.float).reshape(1, 3, 1, 1).to(dtype)
    return {
        # colors.py:8-15
        "rgb2lab.rgb_to_xyz": _mat([[0.412453, 0.212671, 0.019334],
                                    [0.357580, 0.715160, 0.119193],
                                    [0.180423, 0.072169, 0.950227]], dtype),
        # colors.py:16-23
        "rgb2lab.fxfyfz_to_lab": _mat([[0.0, 500.0, 0.0],
                                       [116.0, -500.0, 200.0],
                                       [0.0, 0.0, -200.0]], dtype),
        "rgb2lab.xyz_to_rgb_mult": white.clone(),          # colors.py:24
        "rgb2lab.lab_to_fxfyfz_offset": off.clone(),       # colors.py:25
        # colors.py:69-76
        "lab2rgb.xyz_to_rgb": _mat([[3.2404542, -0.9692660, 0.0556434],
                                    [-1.5371385, 1.8760108, -0.2040259],
                                    [-0.4985314, 0.0415560, 1.0572252]], dtype),
        # colors.py:77-84
        "lab2rgb.lab_to_fxfyfz": _mat([[1 / 116.0, 1 / 116.0, 1 / 116.0],
                                       [1 / 500.0, 0, 0],
                                       [0, 0, -1 / 200.0]], dtype),
        "lab2rgb.xyz_to_rgb_mult": white.clone(),          # colors.py:85
        "lab2rgb.lab_to_fxfyfz_offset": off.clone(),       # colors.py:86
        "rgb2hsv.comparison_zero": torch.tensor(0.0, dtype=torch.float).to(dtype),  # colors.py:184
    }


_CONST_CACHE = {}


def _c(dtype, device=None):
    key = (dtype, str(device))
    if key not in _CONST_CACHE:
        k = constants(dtype)
        _CONST_CACHE[key] = k if device is None else {n: v.to(device) for n, v in k.items()}
    return _CONST_CACHE[key]


def _mix(img, m):
    # colors.py:40,50,104,117 -- out[b,k,y,x] = sum_c img[b,c,y,x] * m[0,0,k,c]
    return torch.einsum('bcyx,bykc->bkyx', img, m)


# ---------------------------------------------------------------------------------------------
# Colour conversions
# ---------------------------------------------------------------------------------------------

def rgb2lab(img):
    """colors.py:27-62.  sRGB -> linear -> XYZ/white -> f(t) -> Lab, normalised to about [0,1]."""
    k = _c(img.dtype, img.device)
    x = img.contiguous()
    lo = (x / 12.92) * x.le(0.04045).float()                                       # colors.py:37
    hi = (((torch.clamp(x, min=0.0001) + 0.055) / 1.055) ** 2.4) * x.gt(0.04045).float()  # :37-38
    x = lo + hi
    x = _mix(x, k["rgb2lab.rgb_to_xyz"])                                           # colors.py:40
    x = torch.mul(x, 1 / k["rgb2lab.xyz_to_rgb_mult"])                             # colors.py:41
    lin = (x / (3.0 * EPS ** 2) + 4.0 / 29.0) * x.le(EPS ** 3).float()             # colors.py:45
    cub = torch.clamp(x, min=0.0001) ** (1.0 / 3.0) * x.gt(EPS ** 3).float()       # colors.py:46-47
    x = lin + cub
    x = _mix(x, k["rgb2lab.fxfyfz_to_lab"]) - k["rgb2lab.lab_to_fxfyfz_offset"]    # colors.py:50
    x[:, 0, :, :] = x[:, 0, :, :] / 100                                            # colors.py:57
    x[:, 1, :, :] = (x[:, 1, :, :] / 110 + 1) / 2                                  # colors.py:58
    x[:, 2, :, :] = (x[:, 2, :, :] / 110 + 1) / 2                                  # colors.py:59
    return x.contiguous()


def lab2rgb(img):
    """colors.py:88-123.  Inverse of rgb2lab; no clamp on the output."""
    k = _c(img.dtype, img.device)
    x = img.contiguous()
    c0 = torch.unsqueeze(x[:, 0, :, :] * 100, dim=1)                               # colors.py:97
    c1 = torch.unsqueeze(((x[:, 1, :, :] * 2) - 1) * 110, dim=1)                   # colors.py:98
    c2 = torch.unsqueeze(((x[:, 2, :, :] * 2) - 1) * 110, dim=1)                   # colors.py:99
    x = torch.cat([c0, c1, c2], dim=1)
    x = _mix(x + k["lab2rgb.lab_to_fxfyfz_offset"], k["lab2rgb.lab_to_fxfyfz"])    # colors.py:104-106
    lin = (3.0 * EPS ** 2 * (x - 4.0 / 29.0)) * x.le(EPS).float()                  # colors.py:110
    cub = (torch.clamp(x, min=0.0001) ** 3.0) * x.gt(EPS).float()                  # colors.py:111
    x = lin + cub
    x = torch.mul(x, k["lab2rgb.xyz_to_rgb_mult"])                                 # colors.py:114
    x = _mix(x, k["lab2rgb.xyz_to_rgb"])                                           # colors.py:117
    lo = x * 12.92 * x.le(0.0031308).float()                                       # colors.py:118
    hi = ((torch.clamp(x, min=0.0001) ** (1 / 2.4) * 1.055) - 0.055) * x.gt(0.0031308).float()  # :118-119
    return (lo + hi).contiguous()


def _inv_or_zero(t):
    # colors.py:186-193 -- 1/t where t != 0, else 0 (0*t keeps dtype/device, carries zero gradient)
    out = 0.0 * t
    nz = t != 0
    out[nz] = 1 / (t[nz])
    return out


def rgb2hsv(img):
    """colors.py:195-242.  Note the ADDITIVE hue selectors (ties add), colors.py:221-224."""
    k = _c(img.dtype, img.device)
    x = torch.clamp(img, 10 ** (-9), 1.0).contiguous()                              # colors.py:205-208
    mx = torch.max(x, 1)[0]                                                        # colors.py:211
    mn = torch.min(x, 1)[0]                                                        # colors.py:212
    df = torch.add(mx, -1.0 * mn)                                                  # colors.py:214
    r, g, b = x[:, 0, :, :], x[:, 1, :, :], x[:, 2, :, :]                          # colors.py:217
    dinv = _inv_or_zero(df)                                                        # colors.py:220
    x[:, 0, :, :] = torch.where(df == 0.0, k["rgb2hsv.comparison_zero"],
                                ((g - b) * dinv) * r.eq(mx).float()
                                + (2.0 + (b - r) * dinv) * g.eq(mx).float()
                                + (4.0 + (r - g) * dinv) * b.eq(mx).float())      # colors.py:221-224
    x[:, 0, :, :] = x[:, 0, :, :] * 60.0                                           # colors.py:225
    x[:, 0, :, :] = x[:, 0, :, :].lt(0.0).float() * (x[:, 0, :, :] + 360) \
        + x[:, 0, :, :].ge(0.0).float() * (x[:, 0, :, :])                          # colors.py:228-229
    x[:, 0, :, :] = x[:, 0, :, :] / 360                                            # colors.py:231
    minv = _inv_or_zero(mx)                                                        # colors.py:234
    x[:, 1, :, :] = torch.where(mx == 0.0, k["rgb2hsv.comparison_zero"],
                                mx.ne(0.0).float() * (df * minv) + mx.eq(0.0).float() * (0.0))  # :235-237
    x[:, 2, :, :] = mx                                                             # colors.py:238
    return torch.clamp(x, 10 ** (-9), 1.0)                                         # colors.py:240


def hsv2rgb(img):
    """colors.py:131-177.  Clamped-ramp HSV -> RGB; terms the reference multiplies by a literal 0
    (m1/m3/m5 = 0) are kept so that the float result is the same expression tree."""
    x = torch.clamp(img, 0.0, 1.0)                                                 # colors.py:141
    h, s, v = x[:, 0, :, :], x[:, 1, :, :], x[:, 2, :, :]

    def ramp(start, width):
        return torch.clamp(h * 360 - start, 0.0, width)

    m2 = (v * (1 - s) - v) / 60                                                    # colors.py:144
    m4 = -1 * m2                                                                   # colors.py:146
    r = v + ramp(0, 60.0) * 0 + ramp(60, 60.0) * m2 + ramp(120, 120.0) * 0 \
        + ramp(240, 60.0) * m4 + ramp(300, 60.0) * 0                               # colors.py:149-150
    m1 = (v - v * (1 - s)) / 60                                                    # colors.py:153
    m3 = -1 * m1                                                                   # colors.py:155
    g = v * (1 - s) + ramp(0, 60.0) * m1 + ramp(60, 120.0) * 0 + ramp(180, 60.0) * m3 \
        + ramp(240, 120.0) * 0                                                     # colors.py:158-159
    n2 = (v - v * (1 - s)) / 60                                                    # colors.py:163
    n4 = -1 * n2                                                                   # colors.py:165
    b = v * (1 - s) + ramp(0, 120.0) * 0 + ramp(120, 60.0) * n2 + ramp(180, 120.0) * 0 \
        + ramp(300, 60.0) * n4                                                     # colors.py:167-168
    out = torch.stack((r, g, b), 1).contiguous()                                   # colors.py:171-174
    return torch.clamp(out, 0.0, 1.0)                                              # colors.py:175


# ---------------------------------------------------------------------------------------------
# Knot curves
# ---------------------------------------------------------------------------------------------

def apply_curve(img, C, reg, channel_in, channel_out, segment_clamp=False):
    """curves.py:4-38.  NB: no per-segment clamp -- the scale is affine in the pixel.

    segment_clamp=True is NOT the reference: it is eq. (1) of the CURL paper (arXiv 1911.13175),
        scale = k_0 + sum_{i=0}^{K-2} (k_{i+1} - k_i) * clamp((K-1) x - i, 0, 1),
    the piecewise-linear curve through all K knots, which the reference code dropped (it sums the
    unclamped terms and leaves the last slope out).  PARITY UNPINNED for this flag: there is no
    reference code to execute; this restatement of the published formula is the only checker."""
    steps_n = C.shape[1] - 1                                                       # curves.py:14
    slope = C[:, 1:] - C[:, 0:-1]                                                  # curves.py:19
    reg = reg + ((slope[:, 1:] - slope[:, 0:-1]) ** 2).sum(1)                      # curves.py:24
    chan = torch.unsqueeze(img[:, channel_in, :, :], 1)                            # curves.py:30
    if segment_clamp:
        idx = torch.arange(0, slope.shape[1], device=img.device)
        scale = C[:, 0].reshape(-1, 1, 1) + (
            slope.reshape(slope.shape[0], slope.shape[1], 1, 1)
            * torch.clamp(steps_n * chan - idx.reshape(1, idx.shape[0], 1, 1), 0.0, 1.0)).sum(1)
    else:
        idx = torch.arange(0, slope.shape[1] - 1, device=img.device)               # curves.py:29
        scale = C[:, 0].reshape(-1, 1, 1) + (
            slope[:, :-1].reshape(slope.shape[0], slope.shape[1] - 1, 1, 1)
            * (steps_n * chan - idx.reshape(1, idx.shape[0], 1, 1))).sum(1)       # curves.py:31-32
    out = img.clone()                                                              # curves.py:34
    out[:, channel_out, :, :] = img[:, channel_out, :, :] * scale                  # curves.py:35
    return torch.clamp(out, 0.0, 1.0), reg                                         # curves.py:36-38


def _adjust(img, P, wiring, segment_clamp=False):
    img = img.contiguous()
    knots = [torch.exp(p) for p in torch.chunk(P, len(wiring), dim=1)]             # curves.py:53-54 etc.
    reg = 0.0
    for C, (cin, cout) in zip(knots, wiring):
        img, reg = apply_curve(img, C, reg, cin, cout, segment_clamp)
    return img.clone().contiguous(), reg


def adjust_lab(img, L, segment_clamp=False):
    """curves.py:136-180: three curves, channel c -> c."""
    return _adjust(img, L, ((0, 0), (1, 1), (2, 2)), segment_clamp)


def adjust_rgb(img, R, segment_clamp=False):
    """curves.py:90-133: three curves, channel c -> c."""
    return _adjust(img, R, ((0, 0), (1, 1), (2, 2)), segment_clamp)


def adjust_hsv(img, S, segment_clamp=False):
    """curves.py:41-87: hue<-f(hue), sat<-f(adjusted hue), sat<-f(sat), val<-f(val)."""
    return _adjust(img, S, ((0, 0), (0, 1), (1, 1), (2, 2)), segment_clamp)


# ---------------------------------------------------------------------------------------------
# The layer (model.py:137-176 without the three dead `feat` lines)
# ---------------------------------------------------------------------------------------------

def curl_layer(img, mask, L, R, H, num_lab_points=48, num_rgb_points=48, num_hsv_points=64, segment_clamp=False):
    lab = rgb2lab(img)                                                             # model.py:151
    lab, reg_lab = adjust_lab(lab, L[:, :num_lab_points], segment_clamp)           # model.py:153
    lab = lab * mask                                                               # model.py:154
    rgb = lab2rgb(lab)                                                             # model.py:157
    rgb, reg_rgb = adjust_rgb(rgb, R[:, :num_rgb_points], segment_clamp)           # model.py:159
    rgb = rgb * mask                                                               # model.py:160
    hsv = rgb2hsv(rgb)                                                             # model.py:163
    hsv, reg_hsv = adjust_hsv(hsv, H[:, :num_hsv_points], segment_clamp)           # model.py:165
    hsv = hsv * mask                                                               # model.py:166
    res = hsv2rgb(hsv)                                                             # model.py:169
    out = torch.clamp(img + res, 0.0, 1.0) * mask                                  # model.py:170
    return out, (reg_rgb + reg_lab + reg_hsv)                                      # model.py:172-176


def curl_layer_fwd_bwd(img, mask, L, R, H, gout, greg, **kw):
    """Forward + autograd backward; returns dict of detached tensors."""
    img = img.detach().clone().requires_grad_(True)
    L = L.detach().clone().requires_grad_(True)
    R = R.detach().clone().requires_grad_(True)
    H = H.detach().clone().requires_grad_(True)
    out, reg = curl_layer(img, mask, L, R, H, **kw)
    loss = (out * gout).sum() + (reg * greg).sum()
    gimg, gL, gR, gH = torch.autograd.grad(loss, (img, L, R, H))
    return {"out": out.detach(), "reg": reg.detach(), "gimg": gimg, "gL": gL, "gR": gR, "gH": gH}


# ---------------------------------------------------------------------------------------------
# 8-bit I/O around the layer: what the reference's data path does on either side of it
# ---------------------------------------------------------------------------------------------

def u8_to_unit(img_u8):
    """torchvision TF.to_tensor on a uint8 image (data.py:152-158, infer.py:38-41), minus its HWC->CHW
    permute: `img.to(float32).div(255)`.  Pinned against torchvision in tests/test_oracle_pins.py."""
    return img_u8.to(torch.float32).div(255)


def unit_to_u8(img):
    """evaluate.py:63 `(numpy_batch[i] * 255).astype('uint8')` == TF.to_pil_image's `.mul(255).byte()`
    (infer.py:47): multiply in fp32, truncate."""
    return img.mul(255).to(torch.uint8)


def curl_layer_u8(img_u8, mask, L, R, H, **kw):
    """uint8 CHW image in, uint8 CHW image out; returns (out_u8, reg, out_fp32)."""
    out, reg = curl_layer(u8_to_unit(img_u8), mask, L, R, H, **kw)
    return unit_to_u8(out), reg, out


# ---------------------------------------------------------------------------------------------
# Polynomial colour regression ("TriSpace" residual): ChannelPolyLayer (model.py:206-333) and
# TriSpaceRegNet.generate_residual / generate_image (model.py:472-508)
# ---------------------------------------------------------------------------------------------

def poly_powers(degree, num_variables):
    """The exponent tuples in the order ChannelPolyLayer.generate_powers (model.py:221-243) yields
    them: by total degree, and within a degree lexicographically descending.  Pinned against the
    reference generator in tests/test_oracle_pins.py."""
    import itertools
    out = []
    for d in range(degree + 1):
        tuples = [t for t in itertools.product(range(d + 1), repeat=num_variables) if sum(t) == d]
        out.extend(sorted(tuples, reverse=True))
    return out


def channel_poly(img, coeffs, degree):
    """ChannelPolyLayer.forward (model.py:319-333): img [B,V,H,W], coeffs [B,C,NC] -> [B,C,H,W]."""
    nv = img.shape[1]
    powers = torch.tensor(poly_powers(degree, nv), dtype=torch.float32).to(img.dtype).to(img.device)
    nc = powers.shape[0]
    img_us = torch.unsqueeze(img, dim=0)                                                       # model.py:319
    pwrs_rs = powers.reshape(nc, 1, nv, 1, 1)                                                  # model.py:320
    poly_terms = torch.permute(torch.pow(img_us, pwrs_rs), [1, 2, 3, 4, 0]).prod(dim=1)        # model.py:321
    return (coeffs.reshape(img.shape[0], coeffs.shape[1], 1, 1, nc) *
            torch.unsqueeze(poly_terms, dim=1)).sum(dim=-1)                                    # model.py:332-333


def cat_coords(img, spatial):
    """TriSpaceRegNet.cat_coords (model.py:472-484): appends x = column/width, y = row/height."""
    if not spatial:
        return img
    B, _, H, W = img.shape
    zeros = img[:, 0:1] * 0.0
    x = zeros + torch.arange(0, W, device=img.device).reshape(1, 1, 1, W) / W
    y = zeros + torch.arange(0, H, device=img.device).reshape(1, 1, H, 1) / H
    return torch.cat([img, x.to(img.dtype), y.to(img.dtype)], dim=1)


def trispace_residual(img, R, L, H, degree=4, spatial=False):
    """TriSpaceRegNet.generate_residual (model.py:486-503)."""
    img_rgb = cat_coords(img, spatial)
    img_lab = cat_coords(rgb2lab(img), spatial)
    img_hsv = cat_coords(rgb2hsv(img), spatial)
    rgb_res = torch.sigmoid(channel_poly(img_rgb, R, degree))
    lab_res = lab2rgb(torch.sigmoid(channel_poly(img_lab, L, degree)))
    hsv_res = hsv2rgb(torch.sigmoid(channel_poly(img_hsv, H, degree)))
    rgb_res = 2 * (rgb_res - 0.5)
    lab_res = 2 * (lab_res - 0.5)
    hsv_res = 2 * (hsv_res - 0.5)
    return rgb_res + lab_res + hsv_res


def trispace_image(img, residual):
    """TriSpaceRegNet.generate_image (model.py:505-508)."""
    return torch.clamp(img + residual, 0.0, 1.0)


def trispace_fwd_bwd(img, R, L, H, gout, degree=4, spatial=False, final_clamp=True):
    R, L, H = (t.detach().clone().requires_grad_(True) for t in (R, L, H))
    out = trispace_residual(img, R, L, H, degree, spatial)
    if final_clamp:
        out = trispace_image(img, out)
    gR, gL, gH = torch.autograd.grad((out * gout).sum(), (R, L, H))
    return {"out": out.detach(), "gR": gR, "gL": gL, "gH": gH}


# ---------------------------------------------------------------------------------------------
# CURLLoss (model.py:36-118) and the MS-SSIM it calls (metric.py:75-214).
# ---------------------------------------------------------------------------------------------
import math

MSSSIM_WEIGHTS = (0.0448, 0.2856, 0.3001, 0.2363, 0.1333)                                      # metric.py:78


def ssim_window(window_size=11, sigma=1.5, dtype=torch.float32, device=None):
    """MSSSIMMetric.gaussian / create_window (metric.py:88-117): fp32 weights, outer product."""
    g = torch.tensor([math.exp(-(x - window_size // 2) ** 2 / float(2 * sigma ** 2)) for x in range(window_size)],
                     dtype=torch.float32, device=device)
    g = (g / g.sum()).unsqueeze(1)
    return g.mm(g.t()).float().unsqueeze(0).unsqueeze(0).to(dtype)


def ssim(img1, img2, window):
    """MSSSIMMetric.compute_ssim (metric.py:119-160), one channel."""
    F = torch.nn.functional
    pad = window.shape[-1] // 2
    mu1 = F.conv2d(img1, window, padding=pad, groups=1)
    mu2 = F.conv2d(img2, window, padding=pad, groups=1)
    mu1_sq = mu1.pow(2)
    mu2_sq = mu2.pow(2)
    mu1_mu2 = mu1 * mu2
    sigma1_sq = F.conv2d(img1 * img1, window, padding=pad, groups=1) - mu1_sq
    sigma2_sq = F.conv2d(img2 * img2, window, padding=pad, groups=1) - mu2_sq
    sigma12 = F.conv2d(img1 * img2, window, padding=pad, groups=1) - mu1_mu2
    C1 = 0.01 ** 2
    C2 = 0.03 ** 2
    ssim_map1 = ((2 * mu1_mu2 + C1) * (2 * sigma12 + C2))
    ssim_map2 = ((mu1_sq + mu2_sq + C1) * (sigma1_sq + sigma2_sq + C2))
    ssim_map = ssim_map1 / ssim_map2
    v1 = 2.0 * sigma12 + C2
    v2 = sigma1_sq + sigma2_sq + C2
    cs = torch.mean(v1 / v2, dim=(1, 2, 3))
    return ssim_map.mean(dim=(1, 2, 3)), cs


def msssim(img1, img2):
    """MSSSIMMetric.compute_msssim (metric.py:162-211) for [B,1,H,W] inputs."""
    F = torch.nn.functional
    window = ssim_window(dtype=img1.dtype, device=img1.device)
    weights = torch.tensor(MSSSIM_WEIGHTS, dtype=torch.float32, device=img1.device).to(img1.dtype)
    ssims, mcs = [], []
    for _ in range(len(MSSSIM_WEIGHTS)):
        s, cs = ssim(img1, img2, window)
        ssims.append(s)
        mcs.append(cs)
        img1 = F.avg_pool2d(img1, (2, 2))
        img2 = F.avg_pool2d(img2, (2, 2))
    ssims = (torch.stack(ssims, dim=1) + 1) / 2
    mcs = (torch.stack(mcs, dim=1) + 1) / 2
    pow1 = mcs ** weights.reshape(1, -1)
    pow2 = ssims ** weights.reshape(1, -1)
    return torch.prod(pow1[:, :-1] * pow2[:, -1].reshape(-1, 1), dim=1)


def loss_lab(img):
    """CURLLoss.batch_lab_convert (model.py:55-58)."""
    return torch.clamp(rgb2lab(img), 0.0, 1.0)


def loss_hsv_cone(img):
    """CURLLoss.batch_hsv_convert (model.py:64-79)."""
    hsv = torch.clamp(rgb2hsv(img), 0.0, 1.0)
    hue = 2 * math.pi * hsv[:, 0, :, :]
    val = hsv[:, 2, :, :]
    sat = hsv[:, 1, :, :]
    c1 = val * sat * torch.cos(hue)
    c2 = val * sat * torch.sin(hue)
    return torch.stack((c1, c2, val), 1)


def curl_loss_parts(pred, target, mask):
    """CURLLoss.forward (model.py:81-118), returning every term and the L planes."""
    F = torch.nn.functional
    unmasked = pred.shape[1] * mask.sum()                                                      # model.py:90
    pred, target = pred * mask, target * mask
    rgb = F.l1_loss(pred, target, reduction='sum') / unmasked
    base_cos_sim = F.cosine_similarity(pred, target, dim=1)
    # model.py:96: [B,H,W] + [B,1,H,W] broadcasts to [B,B,H,W]; kept as the reference has it
    cosine = (1.0 - (base_cos_sim + torch.logical_not(mask)).mean(dim=(1, 2))).mean()
    target_lab = loss_lab(target)
    pred_lab = loss_lab(pred)
    lab = F.l1_loss(pred_lab, target_lab, reduction='sum') / unmasked
    L_target = target_lab[:, 0, :, :].unsqueeze(1)
    L_pred = pred_lab[:, 0, :, :].unsqueeze(1)
    ssim_term = (1.0 - msssim(L_pred, L_target)).mean()
    hsv = F.l1_loss(loss_hsv_cone(pred), loss_hsv_cone(target), reduction='sum') / unmasked
    total = (rgb + cosine + lab + hsv + 10 * ssim_term) / 5
    return {"loss": total, "rgb": rgb, "cosine": cosine, "lab": lab, "hsv": hsv, "ssim": ssim_term,
            "L_pred": L_pred, "L_target": L_target}


def curl_loss(pred, target, mask):
    return curl_loss_parts(pred, target, mask)["loss"]


def curl_loss_fwd_bwd(pred, target, mask):
    pred = pred.detach().clone().requires_grad_(True)
    parts = curl_loss_parts(pred, target, mask)
    (gpred,) = torch.autograd.grad(parts["loss"], (pred,))
    out = {k: v.detach() for k, v in parts.items()}
    out["gpred"] = gpred
    return out
