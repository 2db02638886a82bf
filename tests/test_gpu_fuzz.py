"""GPU: seeded differential sweep of the fused layer against the CPU oracle over shapes, knot counts,
curve strengths, mask kinds and input kinds (continuous, 8-bit grids with black / white / grey patches).
Judged ELEMENT-WISE against the oracle's fp32 run (conftest.frac_mismatch: a wrong decision at an exact
boundary -- L = 0 of black pixels, saturated channels, channel ties -- shows up here even when some
other element of the tensor is huge), with the reference's own fp32-vs-fp64 disagreement as the
allowance for ties that rounding decides."""
import numpy as np
import pytest
import torch

from conftest import frac_mismatch, rel_to_max
from oracle import curl_oracle as O

pytestmark = pytest.mark.gpu

CASES = []
_rng = np.random.RandomState(20261107)
for _i in range(18):
    B = int(_rng.randint(1, 4))
    H, W = int(_rng.randint(3, 70)), int(_rng.randint(3, 90))
    K = (int(_rng.randint(3, 22)), int(_rng.randint(3, 22)), int(_rng.randint(3, 22)))
    sigma = float(_rng.choice([0.03, 0.1, 0.25]))
    maskkind = str(_rng.choice(["none", "bool", "float"]))
    inkind = str(_rng.choice(["continuous", "uint8", "patches"]))
    CASES.append((B, H, W, K, sigma, maskkind, inkind, 1000 + _i))


def _inputs(B, H, W, K, sigma, maskkind, inkind, seed):
    gen = torch.Generator().manual_seed(seed)
    img = torch.rand(B, 3, H, W, generator=gen)
    if inkind != "continuous":
        img = (img * 255).round() / 255
    if inkind == "patches":          # black, white, mid-grey and two-channel ties, as real images have them
        h3 = max(1, H // 4)
        img[:, :, :h3, : W // 2] = 0.0
        img[:, :, :h3, W // 2:] = 1.0
        img[:, :, h3:2 * h3] = img[:, 0:1, h3:2 * h3]
        img[:, 1, 2 * h3:3 * h3] = img[:, 0, 2 * h3:3 * h3]
    if maskkind == "none":
        mask, omask = None, torch.ones(B, 1, H, W, dtype=torch.bool)
    elif maskkind == "bool":
        mask = torch.rand(B, 1, H, W, generator=gen) > 0.2
        omask = mask
    else:
        mask = (torch.rand(B, 1, H, W, generator=gen) > 0.2).float() * (0.5 + 0.5 * torch.rand(B, 1, H, W, generator=gen))
        omask = mask
    pts = (3 * K[0], 3 * K[1], 4 * K[2])
    L, R, Hk = (sigma * torch.randn(B, n, generator=gen) for n in pts)
    gout = torch.randn(B, 3, H, W, generator=gen)
    greg = torch.randn(B, generator=gen)
    return img, mask, omask, L, R, Hk, gout, greg, pts


@pytest.mark.parametrize("case", CASES, ids=lambda c: "B%dx%dx%d-K%s-s%g-%s-%s" % (c[0], c[1], c[2], "_".join(map(str, c[3])), c[4], c[5], c[6]))
def test_layer_fuzz_vs_oracle(case):
    import curl_b200
    dev = torch.device("cuda:0")
    img, mask, omask, L, R, Hk, gout, greg, pts = _inputs(*case)
    kw = dict(num_lab_points=pts[0], num_rgb_points=pts[1], num_hsv_points=pts[2])
    r32 = O.curl_layer_fwd_bwd(img, omask, L, R, Hk, gout, greg, **kw)
    m64 = omask if omask.dtype == torch.bool else omask.double()
    r64 = O.curl_layer_fwd_bwd(img.double(), m64, L.double(), R.double(), Hk.double(), gout.double(), greg.double(), **kw)
    layer = curl_b200.CURLLayer(*pts).to(dev)
    xi = img.to(dev).requires_grad_(True)
    kn = [t.to(dev).requires_grad_(True) for t in (L, R, Hk)]
    out, reg = layer(xi, None if mask is None else mask.to(dev), *kn)
    torch.autograd.backward([out, reg], [gout.to(dev), greg.to(dev)])
    got = dict(out=out.detach(), reg=reg.detach(), gimg=xi.grad, gL=kn[0].grad, gR=kn[1].grad, gH=kn[2].grad)
    got = {k: v.cpu().numpy() for k, v in got.items()}
    assert rel_to_max(got["reg"], r64["reg"].numpy()) <= 1e-5
    # forward: absolute, element-wise
    e32 = np.abs(got["out"] - r32["out"].numpy())
    own = np.abs(r32["out"].numpy() - r64["out"].numpy())
    assert (e32 > 1e-5).mean() <= max(2e-3, 2 * (own > 1e-5).mean()), ((e32 > 1e-5).mean(), (own > 1e-5).mean())
    # image gradient: element-wise against fp32, ties allowed at the reference's own flip rate
    fm = frac_mismatch(got["gimg"], r32["gimg"].numpy())
    own_fm = frac_mismatch(r64["gimg"].numpy(), r32["gimg"].numpy())
    assert fm <= max(3e-3, 2 * own_fm), (fm, own_fm)
    # knot gradients: sums over pixels; a flipped tie pixel moves them by one pixel's contribution
    for k in ("gL", "gR", "gH"):
        e = min(rel_to_max(got[k], r32[k].numpy()), rel_to_max(got[k], r64[k].numpy()))
        assert e <= max(1e-4, 4 * rel_to_max(r32[k].numpy(), r64[k].numpy())), (k, e)


LOSS_FUZZ = []
for _i in range(10):
    LOSS_FUZZ.append((int(_rng.randint(1, 4)), int(_rng.randint(32, 90)), int(_rng.randint(32, 100)),
                      str(_rng.choice(["none", "bool", "float"])), str(_rng.choice(["near", "far", "uint8", "patches"])), 2000 + _i))


@pytest.mark.parametrize("case", LOSS_FUZZ, ids=lambda c: "B%dx%dx%d-%s-%s" % c[:5])
def test_loss_fuzz_vs_oracle(case):
    """CURLLoss (fused colour terms + library MS-SSIM) against the oracle: value, and gradient element-wise
    against the oracle's fp32 run with the reference's own fp32-vs-fp64 disagreement as the allowance."""
    import curl_b200
    B, H, W, maskkind, inkind, seed = case
    dev = torch.device("cuda:0")
    gen = torch.Generator().manual_seed(seed)
    target = torch.rand(B, 3, H, W, generator=gen)
    pred = torch.rand(B, 3, H, W, generator=gen) if inkind == "far" else \
        (target + 0.05 * torch.randn(B, 3, H, W, generator=gen)).clamp(0, 1)
    if inkind in ("uint8", "patches"):
        pred, target = (pred * 255).round() / 255, (target * 255).round() / 255
    if inkind == "patches":
        h4 = H // 5
        pred[:, :, :h4, : W // 2] = 0.0
        pred[:, :, :h4, W // 2:] = 1.0
        pred[:, :, h4:2 * h4] = pred[:, 0:1, h4:2 * h4]
        target[:, :, 2 * h4:3 * h4, : W // 3] = 0.0
        target[:, :, 2 * h4:3 * h4, W // 3:] = 1.0
        pred[:, :, 3 * h4:4 * h4, : W // 4] = target[:, :, 3 * h4:4 * h4, : W // 4]
    if maskkind == "none":
        mask, omask = None, torch.ones(B, 1, H, W, dtype=torch.bool)
    elif maskkind == "bool":
        mask = torch.rand(B, 1, H, W, generator=gen) > 0.15
        omask = mask
    else:
        mask = (torch.rand(B, 1, H, W, generator=gen) > 0.15).float()
        omask = mask
    r32 = O.curl_loss_fwd_bwd(pred, target, omask)
    r64 = O.curl_loss_fwd_bwd(pred.double(), target.double(), omask)
    crit = curl_b200.CURLLoss().to(dev)
    p = pred.to(dev).requires_grad_(True)
    loss = crit(p, target.to(dev), None if mask is None else mask.to(dev))
    loss.backward()
    tol = max(1e-5, 3 * abs(float(r32["loss"]) - float(r64["loss"])))
    assert abs(float(loss.detach()) - float(r64["loss"])) <= tol, (float(loss.detach()), float(r32["loss"]), float(r64["loss"]))
    gp = p.grad.cpu().numpy()
    fm = frac_mismatch(gp, r32["gpred"].numpy())
    own = frac_mismatch(r64["gpred"].numpy(), r32["gpred"].numpy())
    assert fm <= max(2e-3, 2 * own), (fm, own)


TRI_FUZZ = [(int(_rng.randint(1, 4)), int(_rng.randint(2, 80)), int(_rng.randint(2, 100)), bool(_rng.randint(0, 2)),
             bool(_rng.randint(0, 2)), str(_rng.choice(["continuous", "uint8", "patches"])), bool(_rng.randint(0, 2)), 3000 + _i)
            for _i in range(10)]


@pytest.mark.parametrize("case", TRI_FUZZ, ids=lambda c: "B%dx%dx%d-sp%d-cl%d-%s-save%d" % c[:7])
def test_trispace_fuzz_vs_oracle(case):
    """Fused TriSpace residual / image and its coefficient gradients against the oracle over shapes
    (odd and tiny included), spatial on/off, with/without the final clamp, saved or recomputed backward."""
    import curl_b200
    import torch.nn as nn
    B, H, W, spatial, clamp, inkind, save, seed = case
    dev = torch.device("cuda:0")
    gen = torch.Generator().manual_seed(seed)
    img = torch.rand(B, 3, H, W, generator=gen)
    if inkind != "continuous":
        img = (img * 255).round() / 255
    if inkind == "patches":
        h3 = max(1, H // 3)
        img[:, :, :h3, : W // 2] = 0.0
        img[:, :, :h3, W // 2:] = 1.0
        img[:, :, h3:2 * h3] = img[:, 0:1, h3:2 * h3]
    nc = 126 if spatial else 35
    co = [0.3 * torch.randn(B, 3, nc, generator=gen) for _ in range(3)]
    gout = torch.randn(B, 3, H, W, generator=gen)
    r32 = O.trispace_fwd_bwd(img, *co, gout, 4, spatial, clamp)
    r64 = O.trispace_fwd_bwd(img.double(), *[c.double() for c in co], gout.double(), 4, spatial, clamp)
    net = curl_b200.TriSpaceRegNet(spatial=spatial, backbone=nn.Identity()).to(dev)
    net.save_activations = save
    cg = [c.to(dev).requires_grad_(True) for c in co]
    out = net._fused(img.to(dev), *cg, clamp)
    out.backward(gout.to(dev))
    o = out.detach().cpu().numpy()
    own = float(np.abs(r32["out"].numpy() - r64["out"].numpy()).max())
    assert np.abs(o - r64["out"].numpy()).max() <= max(3e-5, 3 * own)
    assert (np.abs(o - r32["out"].numpy()) > 2e-5).mean() <= 2e-3
    for k, c in zip(("gR", "gL", "gH"), cg):
        e = rel_to_max(c.grad.cpu().numpy(), r64[k].numpy())
        assert e <= max(1e-4, 4 * rel_to_max(r32[k].numpy(), r64[k].numpy())), (k, e)


TRI_FUZZ_DEG = [(int(_rng.randint(1, 3)), int(_rng.randint(2, 60)), int(_rng.randint(2, 80)), bool(_rng.randint(0, 2)),
                 bool(_rng.randint(0, 2)), int(_rng.randint(1, 5)), bool(_rng.randint(0, 2)), 4000 + _i) for _i in range(10)]


@pytest.mark.parametrize("case", TRI_FUZZ_DEG, ids=lambda c: "B%dx%dx%d-sp%d-cl%d-deg%d-save%d" % c[:7])
def test_trispace_fuzz_degrees_and_image_gradient(case):
    """Polynomial degrees 1-4 (ChannelPolyLayer is generic in the degree, model.py:206-333) and the image gradient
    (autograd of model.py:486-508 w.r.t. img) over random shapes, against the oracle."""
    import curl_b200
    import torch.nn as nn
    import parity
    from chunked import trispace_fwd_bwd_chunked
    B, H, W, spatial, clamp, degree, save, seed = case
    dev = torch.device("cuda:0")
    gen = torch.Generator().manual_seed(seed)
    img = torch.rand(B, 3, H, W, generator=gen)
    nc = len(O.poly_powers(degree, 5 if spatial else 3))
    co = [0.3 * torch.randn(B, 3, nc, generator=gen) for _ in range(3)]
    gout = torch.randn(B, 3, H, W, generator=gen)
    r32 = trispace_fwd_bwd_chunked(img, *co, gout, degree, spatial, clamp, dtype=torch.float32, rows=H)
    r64 = trispace_fwd_bwd_chunked(img, *co, gout, degree, spatial, clamp, dtype=torch.float64, rows=H)
    net = curl_b200.TriSpaceRegNet(polynomial_order=degree, spatial=spatial, backbone=nn.Identity()).to(dev)
    net.save_activations = save
    x = img.to(dev).requires_grad_(True)
    cg = [c.to(dev).requires_grad_(True) for c in co]
    out = net._fused(x, *cg, clamp)
    out.backward(gout.to(dev))
    o = out.detach().cpu().numpy()
    own = float(np.abs(r32["out"] - r64["out"]).max())
    assert np.abs(o - r64["out"]).max() <= max(3e-5, 3 * own)
    assert (np.abs(o - r32["out"]) > 2e-5 * np.maximum(1.0, np.abs(r64["out"]))).mean() <= 2e-3
    for k, c in zip(("gR", "gL", "gH"), cg):
        e = rel_to_max(c.grad.cpu().numpy(), r64[k])
        assert e <= max(1e-4, 4 * rel_to_max(r32[k], r64[k])), (k, e)
    if H * W >= 64:           # the fraction-based criterion needs a few pixels
        parity.check_trispace_gimg(x.grad.cpu().numpy(), r32["gimg"], r64["gimg"], case)
    else:
        scale = np.abs(r64["gimg"]).max()
        assert np.abs(x.grad.cpu().numpy() - r64["gimg"]).max() <= max(1e-4 * scale, 3 * np.abs(r32["gimg"] - r64["gimg"]).max())
