"""TEST INFRASTRUCTURE: the CPU oracle evaluated in row chunks.

The layer is per-pixel independent given an image's knots, so a full-size image can be
checked against the oracle without holding autograd state for all of it (~3.6 kB per pixel in fp32):
rows are processed `rows` at a time, `out` / `gimg` are concatenated, the per-image knot gradients of the
chunks are added (in float64), and the pixel-free regulariser gradient (`greg`) is applied in the first
chunk only.  With one chunk this is exactly oracle.curl_layer_fwd_bwd.
"""
import numpy as np
import torch

from oracle import curl_oracle as O


def layer_fwd_bwd_chunked(img, mask, L, R, Hk, gout, greg, dtype=torch.float32, rows=128, **kw):
    """img [B,3,H,W], mask [B,1,H,W] (bool or float) or None, gout like img, greg [B]: CPU tensors.
    Returns numpy arrays out, gimg (dtype's precision), reg, and gL/gR/gH accumulated in float64."""
    B, _, H, W = img.shape
    if mask is None:
        mask = torch.ones(B, 1, H, W, dtype=torch.bool)
    conv = lambda t: t.to(dtype)
    Ld, Rd, Hd = conv(L), conv(R), conv(Hk)
    np_dtype = np.float32 if dtype == torch.float32 else np.float64
    out = np.empty((B, 3, H, W), np_dtype)
    gimg = np.empty((B, 3, H, W), np_dtype)
    acc = None
    reg = None
    for r0 in range(0, H, rows):
        r1 = min(H, r0 + rows)
        m = mask[:, :, r0:r1]
        m = m if m.dtype == torch.bool else conv(m)
        g = conv(greg) if r0 == 0 else torch.zeros(B, dtype=dtype)
        res = O.curl_layer_fwd_bwd(conv(img[:, :, r0:r1]).contiguous(), m.contiguous(), Ld, Rd, Hd,
                                   conv(gout[:, :, r0:r1]).contiguous(), g, **kw)
        out[:, :, r0:r1] = res["out"].numpy()
        gimg[:, :, r0:r1] = res["gimg"].numpy()
        if reg is None:
            reg = res["reg"].numpy().copy()
        part = [res[k].double().numpy() for k in ("gL", "gR", "gH")]
        acc = part if acc is None else [a + p for a, p in zip(acc, part)]
    return dict(out=out, gimg=gimg, reg=reg, gL=acc[0], gR=acc[1], gH=acc[2])


def trispace_fwd_bwd_chunked(img, R, L, Hc, gout, degree=4, spatial=False, final_clamp=True, dtype=torch.float32, rows=16):
    """TriSpaceRegNet.generate_residual (+ generate_image) of the oracle (O.trispace_residual / trispace_image,
    model.py:486-508) on row chunks of a full-size image: the reference materialises a [NC, B, V, H, W] power tensor
    per space (2.5 kB per pixel for 126 coefficients), so a megapixel image only fits in pieces.  Pixels are
    independent given the coefficients; the spatial variables are x = column / W and y = row / H of the WHOLE image
    (model.py:472-484, fp32 quotients), which is why a chunk cannot simply be handed to O.trispace_residual.
    Returns numpy out, gimg (dtype's precision) and gR / gL / gH accumulated in float64."""
    B, _, H, W = img.shape
    np_dtype = np.float32 if dtype == torch.float32 else np.float64
    out = np.empty((B, 3, H, W), np_dtype)
    gimg = np.empty((B, 3, H, W), np_dtype)
    acc = None
    for r0 in range(0, H, rows):
        r1 = min(H, r0 + rows)

        def cat(t):
            if not spatial:
                return t
            zeros = t[:, 0:1] * 0.0
            x = zeros + (torch.arange(0, W).reshape(1, 1, 1, W) / W).to(t.dtype)
            y = zeros + (torch.arange(r0, r1).reshape(1, 1, r1 - r0, 1) / H).to(t.dtype)
            return torch.cat([t, x, y], dim=1)

        xi = img[:, :, r0:r1].to(dtype).clone().requires_grad_(True)
        a = [t.to(dtype).clone().requires_grad_(True) for t in (R, L, Hc)]
        rgb_res = torch.sigmoid(O.channel_poly(cat(xi), a[0], degree))
        lab_res = O.lab2rgb(torch.sigmoid(O.channel_poly(cat(O.rgb2lab(xi)), a[1], degree)))
        hsv_res = O.hsv2rgb(torch.sigmoid(O.channel_poly(cat(O.rgb2hsv(xi)), a[2], degree)))
        res = 2 * (rgb_res - 0.5) + 2 * (lab_res - 0.5) + 2 * (hsv_res - 0.5)
        o = O.trispace_image(xi, res) if final_clamp else res
        g = torch.autograd.grad((o * gout[:, :, r0:r1].to(dtype)).sum(), a + [xi])
        out[:, :, r0:r1] = o.detach().numpy()
        gimg[:, :, r0:r1] = g[3].numpy()
        part = [t.double().numpy() for t in g[:3]]
        acc = part if acc is None else [p + q for p, q in zip(acc, part)]
    return dict(out=out, gimg=gimg, gR=acc[0], gL=acc[1], gH=acc[2])
