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
