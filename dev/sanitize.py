"""Drives every kernel of libcurl_b200.so once at small shapes, for compute-sanitizer:

    compute-sanitizer --tool memcheck|racecheck|synccheck|initcheck python dev/sanitize.py

Covers the vector and scalar (H*W % 4 != 0) paths, every mask kind, with and without image gradient,
the 8-bit paths, colour modules, curve groups, TriSpace (35 and 126 coefficients, saved / recomputed),
CURLLoss colour terms and the segment-clamped curves.  Prints 'sanitize: ok' when every call returned.
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import torch.nn as nn  # noqa: E402

import curl_b200  # noqa: E402
from curl_b200 import functional as F  # noqa: E402

dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(5)


def rnd(*s):
    return torch.rand(*s, generator=g, device=dev)


def layer_case(B, H, W, mask_kind, want_gimg, segment_clamp=False):
    img = rnd(B, 3, H, W).requires_grad_(want_gimg)
    mask = {"none": None, "bool": rnd(B, 1, H, W) > 0.2, "f32": (rnd(B, 1, H, W) > 0.2).float()}[mask_kind]
    L, R, Hk = ((0.1 * torch.randn(B, n, generator=g, device=dev)).requires_grad_(True) for n in (48, 48, 64))
    layer = curl_b200.CURLLayer(segment_clamp=segment_clamp).to(dev)
    out, reg = layer(img, mask if mask is not None else torch.ones(B, 1, H, W, device=dev, dtype=torch.bool), L, R, Hk)
    torch.autograd.backward([out, reg], [torch.randn_like(out), torch.ones_like(reg)])


for (H, W) in ((64, 96), (33, 37)):                 # TMA/vector path (>= one tile) and scalar path
    for mk in ("bool", "f32"):
        for gi in (True, False):
            layer_case(2, H, W, mk, gi)
layer_case(1, 128, 256, "bool", True)               # several tiles per block ring
layer_case(2, 64, 96, "bool", True, segment_clamp=True)
layer_case(2, 33, 37, "f32", False, segment_clamp=True)

# colour modules and curve groups, forward + backward
x = rnd(2, 3, 40, 52).requires_grad_(True)
for mod in (curl_b200.colors.RGB2LAB(), curl_b200.colors.LAB2RGB(), curl_b200.colors.RGB2HSV(), curl_b200.colors.HSV2RGB()):
    mod.to(dev)(x).sum().backward()
for fn, n in ((curl_b200.curves.adjust_lab, 48), (curl_b200.curves.adjust_rgb, 48), (curl_b200.curves.adjust_hsv, 64)):
    P = (0.1 * torch.randn(2, n, generator=g, device=dev)).requires_grad_(True)
    o, r = fn(x, P)
    (o.sum() + r.sum()).backward()

# 8-bit paths
u8 = (rnd(2, 3, 64, 96) * 255).to(torch.uint8)
layer = curl_b200.CURLLayer().to(dev)
Lk, Rk, Hk = (0.1 * torch.randn(2, n, generator=g, device=dev) for n in (48, 48, 64))
m = rnd(2, 1, 64, 96) > 0.2
with torch.no_grad():
    layer.forward_u8(u8, m, Lk, Rk, Hk)
    layer.forward_u8(u8.permute(0, 2, 3, 1).contiguous(), m, Lk, Rk, Hk, layout="HWC")
    F.to_tensor_u8(u8)
    F.to_tensor_u8(u8.permute(0, 2, 3, 1).contiguous(), layout="HWC")
    F.to_tensor_u8((rnd(1, 3, 33, 37) * 255).to(torch.uint8))

# TriSpace residual
for spatial in (True, False):
    for save in (True, False):
        tri = curl_b200.TriSpaceRegNet(spatial=spatial, backbone=nn.Identity()).to(dev)
        tri.save_activations = save
        nc = 126 if spatial else 35
        coefs = [(0.3 * torch.randn(2, 3, nc, generator=g, device=dev)).requires_grad_(True) for _ in range(3)]
        tri.generate_residual(rnd(2, 3, 48, 40), *coefs).sum().backward()
        tri.generate_residual(rnd(1, 3, 33, 37), *[c[:1] for c in coefs]).sum().backward()
        # image gradient too (curl_trispace_bwd_img_f32), fused clamp, vector and scalar paths
        tri._fused(rnd(2, 3, 48, 40).requires_grad_(True), *coefs, True).sum().backward()
        tri._fused(rnd(1, 3, 33, 37).requires_grad_(True), *[c[:1].detach() for c in coefs], True).sum().backward()
        with torch.no_grad():                                   # inference forward (no saved sigmoids: the rolled variant)
            tri._fused(rnd(2, 3, 48, 40), *coefs, True)
# lower polynomial degrees
for degree in (1, 2, 3):
    for spatial in (True, False):
        tri = curl_b200.TriSpaceRegNet(polynomial_order=degree, spatial=spatial, backbone=nn.Identity()).to(dev)
        coefs = [(0.3 * torch.randn(2, 3, tri.num_coeffs, generator=g, device=dev)).requires_grad_(True) for _ in range(3)]
        tri._fused(rnd(2, 3, 48, 40).requires_grad_(True), *coefs, True).sum().backward()

# CURLLoss colour terms + whole criterion
pred = rnd(2, 3, 64, 96).requires_grad_(True)
tgt = rnd(2, 3, 64, 96)
crit = curl_b200.CURLLoss().to(dev)
for mk in (m, m.float()):
    crit(pred, tgt, mk).backward()
torch.cuda.synchronize()
print("sanitize: ok")
