"""GPU: reference semantics at the edges of the module API that the fixtures do not reach:
uint8 masks multiply by VALUE like the reference, CURLLoss takes [B,H,W] masks, TriSpace's saved activations
survive retain_graph, and the net built around Deg4MobilePolyLayer (infer.py:22-23) matches its fixture."""
import numpy as np
import pytest
import torch

from conftest import rel_to_max
from oracle import curl_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def test_uint8_mask_multiplies_by_value_like_the_reference():
    import curl_b200
    gen = torch.Generator().manual_seed(8)
    B, H, W = 2, 24, 36
    img = torch.rand(B, 3, H, W, generator=gen)
    keep = torch.rand(B, 1, H, W, generator=gen) > 0.3
    L, R, Hk = (0.1 * torch.randn(B, n, generator=gen) for n in (48, 48, 64))
    layer = curl_b200.CURLLayer().to(DEV)
    dev = lambda t: t.to(DEV)
    # 0/1 bytes: same as the bool mask
    a, _ = layer(dev(img), dev(keep), dev(L), dev(R), dev(Hk))
    b, _ = layer(dev(img), dev(keep.to(torch.uint8)), dev(L), dev(R), dev(Hk))
    assert torch.equal(a, b)
    # 0/2 bytes: the reference multiplies every stage by 2 (model.py:154,160,166,170); so does the float path
    m2 = keep.to(torch.uint8) * 2
    ref, _ = O.curl_layer(img, m2.float(), L, R, Hk)
    got, _ = layer(dev(img), dev(m2), dev(L), dev(R), dev(Hk))
    same, _ = layer(dev(img), dev(m2.float()), dev(L), dev(R), dev(Hk))
    assert torch.equal(got, same)           # a byte mask takes the float path: value-exact
    err = (got.cpu() - ref).abs()           # values reach 2.0 and every stage is scaled by 2: a looser absolute bar
    assert float((err > 1e-4).float().mean()) <= 1e-2 and float(err.max()) <= 1e-2
    assert float(got.max()) > 1.5          # i.e. NOT read as zero / non-zero


def test_curlloss_accepts_3d_mask_and_rejects_broadcast():
    import curl_b200
    gen = torch.Generator().manual_seed(9)
    B, H, W = 3, 32, 40
    pred = torch.rand(B, 3, H, W, generator=gen)
    target = torch.rand(B, 3, H, W, generator=gen)
    mask = torch.rand(B, 1, H, W, generator=gen) > 0.2
    crit = curl_b200.CURLLoss().to(DEV)
    a = crit(pred.to(DEV), target.to(DEV), mask.to(DEV))
    b = crit(pred.to(DEV), target.to(DEV), mask[:, 0].to(DEV))                       # [B, H, W]
    c = crit(pred.to(DEV), target.to(DEV), mask[:, 0].float().to(DEV))               # float [B, H, W]
    assert torch.equal(a, b)
    assert abs(float(a) - float(c)) <= 1e-6
    assert abs(float(a) - float(O.curl_loss(pred, target, mask))) <= 1e-5
    with pytest.raises(ValueError, match="one row per image"):
        crit(pred.to(DEV), target.to(DEV), mask[:1].to(DEV))


def test_trispace_second_backward_with_retain_graph_is_identical():
    import curl_b200
    import torch.nn as nn
    gen = torch.Generator().manual_seed(10)
    net = curl_b200.TriSpaceRegNet(spatial=True, backbone=nn.Identity()).to(DEV)
    img = torch.rand(2, 3, 20, 28, generator=gen).to(DEV)
    co = [(0.3 * torch.randn(2, 3, 126, generator=gen)).to(DEV).requires_grad_(True) for _ in range(3)]
    out = net.generate_residual(img, *co)
    g = torch.randn(out.shape, generator=gen).to(DEV)
    first = torch.autograd.grad(out, co, g, retain_graph=True)
    second = torch.autograd.grad(out, co, g)
    for a, b in zip(first, second):
        assert torch.equal(a, b)


def test_trispace_with_deg4_mobile_polylayer_vs_fixture(trispace_golden):
    """infer.py:22-23: TriSpaceRegNet(polynomial_order=4, spatial=True, is_train=False,
    polylayer=model.Deg4MobilePolyLayer()); the `spatial_mobile` fixture was produced by the reference with
    exactly that layer (products written out instead of torch.pow)."""
    import curl_b200
    import torch.nn as nn
    g = trispace_golden
    net = curl_b200.model.TriSpaceRegNet(polynomial_order=4, spatial=True, is_train=False,
                                         polylayer=curl_b200.model.Deg4MobilePolyLayer(), backbone=nn.Identity()).to(DEV)
    img = torch.from_numpy(g["spatial_mobile.img"]).to(DEV)
    R, L, H = (torch.from_numpy(g[f"spatial_mobile.{k}"]).to(DEV) for k in ("R", "L", "H"))
    with torch.no_grad():
        out = net.generate_residual(img, R, L, H).cpu().numpy()
    r32, r64 = g["spatial_mobile.f32.c0.out"], g["spatial_mobile.f64.c0.out"]
    own = float(np.abs(r32 - r64).max())
    assert float(np.abs(out - r64).max()) <= max(2e-5, 3 * own)
