"""TEST / BASELINE INFRASTRUCTURE ONLY -- a torch.compile-friendly restatement of oracle/curl_oracle.py.

SURVEY.md 8(d) names "torch.compile of the reference on the B200" as the stronger stand-in a fused kernel
must beat.  The reference's op sequence does not survive dynamo as written: `non_nan_inv` (colors.py:186-193)
is a data-dependent boolean index_put (graph break), and the in-place channel writes of colors.py:57-59,
220-238 and curves.py:34-35 make the resumed graph fail in autograd.  This module changes ONLY those three
things -- `1/t where t != 0 else 0` as a pair of torch.where (with a safe denominator, so the untaken branch
carries no inf into the gradient), and channel tensors assembled with torch.stack instead of written in
place -- and reuses every other line of the oracle.  Everything else (both-branch blends, einsum
contractions, operation order) is the reference's, so eager CPU results equal the oracle's bit for bit on the
forward (tests/test_oracle_pins.py::test_compile_variant_equals_oracle).

Who may use it: tests/ and bench.py's `gpu_compile_baseline` leg (through dev/compile_baseline.py).
"""
import torch

from . import curl_oracle as O

EPS = O.EPS


def rgb2lab(img):
    k = O._c(img.dtype, img.device)
    x = img.contiguous()
    lo = (x / 12.92) * x.le(0.04045).float()
    hi = (((torch.clamp(x, min=0.0001) + 0.055) / 1.055) ** 2.4) * x.gt(0.04045).float()
    x = lo + hi
    x = O._mix(x, k["rgb2lab.rgb_to_xyz"])
    x = torch.mul(x, 1 / k["rgb2lab.xyz_to_rgb_mult"])
    lin = (x / (3.0 * EPS ** 2) + 4.0 / 29.0) * x.le(EPS ** 3).float()
    cub = torch.clamp(x, min=0.0001) ** (1.0 / 3.0) * x.gt(EPS ** 3).float()
    x = lin + cub
    x = O._mix(x, k["rgb2lab.fxfyfz_to_lab"]) - k["rgb2lab.lab_to_fxfyfz_offset"]
    return torch.stack((x[:, 0] / 100, (x[:, 1] / 110 + 1) / 2, (x[:, 2] / 110 + 1) / 2), 1).contiguous()   # colors.py:57-59


def _inv_or_zero(t):
    nz = t != 0                                                                    # colors.py:186-193
    return torch.where(nz, 1 / torch.where(nz, t, torch.ones_like(t)), torch.zeros_like(t))


def rgb2hsv(img):
    k = O._c(img.dtype, img.device)
    x = torch.clamp(img, 10 ** (-9), 1.0).contiguous()
    mx = torch.max(x, 1)[0]
    mn = torch.min(x, 1)[0]
    df = torch.add(mx, -1.0 * mn)
    r, g, b = x[:, 0, :, :], x[:, 1, :, :], x[:, 2, :, :]
    dinv = _inv_or_zero(df)
    h = torch.where(df == 0.0, k["rgb2hsv.comparison_zero"],
                    ((g - b) * dinv) * r.eq(mx).float()
                    + (2.0 + (b - r) * dinv) * g.eq(mx).float()
                    + (4.0 + (r - g) * dinv) * b.eq(mx).float())
    h = h * 60.0
    h = h.lt(0.0).float() * (h + 360) + h.ge(0.0).float() * h
    h = h / 360
    minv = _inv_or_zero(mx)
    s = torch.where(mx == 0.0, k["rgb2hsv.comparison_zero"],
                    mx.ne(0.0).float() * (df * minv) + mx.eq(0.0).float() * (0.0))
    return torch.clamp(torch.stack((h, s, mx), 1), 10 ** (-9), 1.0)


def apply_curve(img, C, reg, channel_in, channel_out):
    steps_n = C.shape[1] - 1
    slope = C[:, 1:] - C[:, 0:-1]
    reg = reg + ((slope[:, 1:] - slope[:, 0:-1]) ** 2).sum(1)
    chan = torch.unsqueeze(img[:, channel_in, :, :], 1)
    idx = torch.arange(0, slope.shape[1] - 1, device=img.device)
    scale = C[:, 0].reshape(-1, 1, 1) + (
        slope[:, :-1].reshape(slope.shape[0], slope.shape[1] - 1, 1, 1)
        * (steps_n * chan - idx.reshape(1, idx.shape[0], 1, 1))).sum(1)
    chans = [img[:, c] for c in range(3)]
    chans[channel_out] = img[:, channel_out, :, :] * scale                         # curves.py:34-35, out of place
    return torch.clamp(torch.stack(chans, 1), 0.0, 1.0), reg


def _adjust(img, P, wiring):
    knots = [torch.exp(p) for p in torch.chunk(P, len(wiring), dim=1)]
    reg = 0.0
    for C, (cin, cout) in zip(knots, wiring):
        img, reg = apply_curve(img, C, reg, cin, cout)
    return img, reg


def curl_layer(img, mask, L, R, H):
    lab = rgb2lab(img)
    lab, reg_lab = _adjust(lab, L[:, :48], ((0, 0), (1, 1), (2, 2)))
    lab = lab * mask
    rgb = O.lab2rgb(lab)
    rgb, reg_rgb = _adjust(rgb, R[:, :48], ((0, 0), (1, 1), (2, 2)))
    rgb = rgb * mask
    hsv = rgb2hsv(rgb)
    hsv, reg_hsv = _adjust(hsv, H[:, :64], ((0, 0), (0, 1), (1, 1), (2, 2)))
    hsv = hsv * mask
    res = O.hsv2rgb(hsv)
    out = torch.clamp(img + res, 0.0, 1.0) * mask
    return out, (reg_rgb + reg_lab + reg_hsv)
