"""GPU parity at the sizes and loop structures the benchmark runs (VERDICT r01 "What's weak" #1).

The small fixtures give every block of the backward ONE 768-pixel tile and the forward one grid-stride
step, so the `it > 0` half of the backward loop (ring-stage reuse, mbarrier parity flip, refill behind the
barrier, mask prefetch across tiles) and the forward's later iterations were only ever checked against
themselves.  Here:

  * full-size images (1 x 4K, 2 x 1080p, a large shape with H*W % 4 != 0) are compared with the CPU oracle,
    fp32 and fp64, evaluated in row chunks (tests/chunked.py; pixels are independent) -- at B = 1 the 4K
    backward runs 3 552 blocks over 10 800 tiles, so every block takes >= 3 tiles and the forward
    grid-strides twice;
  * small seeded images are run with the launch capped to a few blocks per image
    (curl_test_set_grid_cap), which forces the same multi-tile paths (including a partial last tile and the
    scalar kernels) at sizes the oracle does in milliseconds, and must reproduce the uncapped result bit
    for bit on everything elementwise.

Statistics go to gpurun_out/parity_fullsize.json.
"""
import json
import os

import numpy as np
import pytest
import torch

import parity
from chunked import layer_fwd_bwd_chunked
from conftest import ROOT, frac_mismatch

pytestmark = pytest.mark.gpu

REPORT = {}


@pytest.fixture(scope="module", autouse=True)
def _write_report():
    yield
    out = os.path.join(ROOT, "gpurun_out")
    os.makedirs(out, exist_ok=True)
    with open(os.path.join(out, "parity_fullsize.json"), "w") as f:
        json.dump(REPORT, f, indent=1, sort_keys=True)


def _inputs(B, H, W, maskkind, seed, u8=False):
    gen = torch.Generator().manual_seed(seed)
    img = torch.rand(B, 3, H, W, generator=gen)
    if u8:
        img = (img * 255).round() / 255
    if maskkind == "none":
        mask = None
    elif maskkind == "bool":
        mask = torch.rand(B, 1, H, W, generator=gen) > 0.1
    else:
        mask = (torch.rand(B, 1, H, W, generator=gen) > 0.1).float()
    L, R, Hk = (0.1 * torch.randn(B, n, generator=gen) for n in (48, 48, 64))
    gout = torch.randn(B, 3, H, W, generator=gen)
    greg = torch.rand(B, generator=gen) + 0.5
    return img, mask, L, R, Hk, gout, greg


def _gpu(img, mask, L, R, Hk, gout, greg, want_gimg=True):
    import curl_b200
    dev = torch.device("cuda:0")
    layer = curl_b200.CURLLayer().to(dev)
    x = img.to(dev).requires_grad_(want_gimg)
    kn = [t.to(dev).requires_grad_(True) for t in (L, R, Hk)]
    out, reg = layer(x, None if mask is None else mask.to(dev), *kn)
    torch.autograd.backward([out, reg], [gout.to(dev), greg.to(dev)])
    res = dict(out=out.detach(), reg=reg.detach(), gL=kn[0].grad, gR=kn[1].grad, gH=kn[2].grad)
    if want_gimg:
        res["gimg"] = x.grad
    return {k: v.cpu().numpy() for k, v in res.items()}


def _strict(ours, r32, r64):
    """The plain north_star numbers, reported whether or not they pass (DESIGN.md section 2 quotes them)."""
    e = np.abs(ours.astype(np.float64) - r32)
    own = np.abs(r32.astype(np.float64) - r64)
    return dict(max_abs_vs_ref32=float(e.max()), frac_over_1e5_vs_ref32=float((e > 1e-5).mean()),
                max_abs_vs_ref64=float(np.abs(ours - r64).max()), ref32_vs_ref64_max=float(own.max()),
                ref32_vs_ref64_frac_over_1e5=float((own > 1e-5).mean()))


FULL = [("4k_b1_bool", (1, 2160, 3840), "bool", 96), ("1080p_b2_float", (2, 1080, 1920), "float", 120),
        ("odd_b1_1021x1531_bool", (1, 1021, 1531), "bool", 128), ("1080p_b1_nomask_u8grid", (1, 1080, 1920), "none", 120)]


@pytest.mark.parametrize("name,shape,maskkind,rows", FULL, ids=[c[0] for c in FULL])
def test_fullsize_vs_oracle(name, shape, maskkind, rows):
    B, H, W = shape
    args = _inputs(B, H, W, maskkind, 1234 + H, u8=name.endswith("u8grid"))
    r32 = layer_fwd_bwd_chunked(*args, dtype=torch.float32, rows=rows)
    r64 = layer_fwd_bwd_chunked(*args, dtype=torch.float64, rows=rows)
    ours = _gpu(*args, want_gimg=True)
    nog = _gpu(*args, want_gimg=False)
    rep = dict(shape=list(shape), mask=maskkind)
    rep["out"] = parity.check_image_stat(ours["out"], r32["out"], r64["out"])
    rep["out_strict"] = _strict(ours["out"], r32["out"], r64["out"])
    assert parity.rel_to_max(ours["reg"], r64["reg"]) <= 1e-5
    gscale = float(np.abs(r64["gimg"]).max())
    rep["gimg"] = parity.check_image_stat(ours["gimg"], r32["gimg"], r64["gimg"], scale=gscale, tol=parity.GRAD_TOL)
    rep["gimg_frac_mismatch_vs_ref32"] = frac_mismatch(ours["gimg"], r32["gimg"])
    rep["gimg_frac_mismatch_ref64_vs_ref32"] = frac_mismatch(r64["gimg"], r32["gimg"])
    assert rep["gimg_frac_mismatch_vs_ref32"] <= max(3e-3, 3 * rep["gimg_frac_mismatch_ref64_vs_ref32"]), rep
    for k in ("gL", "gR", "gH"):
        # (the chunked fp32 oracle adds its per-chunk fp32 sums in double; the unchunked reference would add 8 M
        # terms in fp32)
        rep[k] = parity.check_knots_stat(ours[k], r32[k], r64[k], n_pixels=H * W)
        # without the image gradient the same moments are reduced: bit-identical knot gradients
        assert np.array_equal(ours[k], nog[k]), k
    assert np.array_equal(ours["out"], nog["out"])
    REPORT[name] = rep


CAPPED = [
    # name, (B, H, W), mask, cap, seed          HW, tiles of 768 px
    ("aligned_bool_cap2", (2, 96, 128), "bool", 2, 1),          # 12288 px = 16 tiles -> 8 per block
    ("aligned_float_cap1", (3, 64, 100), "float", 1, 2),        # 6400 px: 9 tiles, the last one partial
    ("aligned_nomask_cap1_partial", (1, 40, 52), "none", 1, 3),  # 2080 px = 2 tiles + 544 px
    ("aligned_bool_cap3", (1, 120, 200), "bool", 3, 4),         # 24000 px: 32 tiles, uneven split over 3 blocks
    ("scalar_bool_cap3", (1, 50, 77), "bool", 3, 5),            # 3850 px, H*W % 4 == 2: scalar kernels loop
    ("scalar_float_cap2", (2, 37, 53), "float", 2, 6),          # odd H*W
    ("scalar_nomask_cap1", (1, 33, 65), "none", 1, 7),
]


@pytest.fixture
def grid_cap():
    from curl_b200 import _lib
    lib = _lib.load()
    yield lambda n: lib.curl_test_set_grid_cap(int(n))
    lib.curl_test_set_grid_cap(0)


@pytest.mark.parametrize("name,shape,maskkind,cap,seed", CAPPED, ids=[c[0] for c in CAPPED])
def test_capped_grid_vs_oracle(name, shape, maskkind, cap, seed, grid_cap):
    from oracle import curl_oracle as O
    B, H, W = shape
    img, mask, L, R, Hk, gout, greg = _inputs(B, H, W, maskkind, 777 + seed)
    omask = torch.ones(B, 1, H, W, dtype=torch.bool) if mask is None else mask
    m64 = omask if omask.dtype == torch.bool else omask.double()
    r32 = {k: v.numpy() for k, v in O.curl_layer_fwd_bwd(img, omask, L, R, Hk, gout, greg).items()}
    r64 = {k: v.numpy() for k, v in O.curl_layer_fwd_bwd(img.double(), m64, L.double(), R.double(), Hk.double(),
                                                        gout.double(), greg.double()).items()}
    free = _gpu(img, mask, L, R, Hk, gout, greg)
    assert grid_cap(cap) == 0
    capped = _gpu(img, mask, L, R, Hk, gout, greg)
    capped_nog = _gpu(img, mask, L, R, Hk, gout, greg, want_gimg=False)
    assert grid_cap(0) == cap
    rep = dict(shape=list(shape), mask=maskkind, cap=cap)
    # elementwise results do not depend on which block / tile / ring stage handled a pixel: the capped launch
    # (every block loops over many tiles) must reproduce the uncapped one (one tile per block) bit for bit ...
    assert np.array_equal(capped["out"], free["out"])
    assert np.array_equal(capped["gimg"], free["gimg"])
    assert np.array_equal(capped["reg"], free["reg"])
    for k in ("gL", "gR", "gH"):
        assert np.array_equal(capped[k], capped_nog[k]), k
        # ... and only the order in which the blocks' partial sums are folded differs for the knot gradients
        assert parity.rel_to_max(capped[k], free[k]) <= 2e-6, k
    # and both against the oracle
    rep["out"] = parity.check_image_stat(capped["out"], r32["out"], r64["out"])
    gscale = float(np.abs(r64["gimg"]).max())
    rep["gimg"] = parity.check_image_stat(capped["gimg"], r32["gimg"], r64["gimg"], scale=gscale, tol=parity.GRAD_TOL)
    for k in ("gL", "gR", "gH"):
        rep[k] = parity.check_knots_stat(capped[k], r32[k], r64[k], n_pixels=H * W)
    REPORT["capped." + name] = rep


def test_capped_grid_stage_kernels(grid_cap):
    """The stand-alone colour / curve-group kernels share the same loops: capped == uncapped, bit for bit."""
    import curl_b200.colors as C
    import curl_b200.curves as V
    dev = torch.device("cuda:0")
    gen = torch.Generator().manual_seed(31)
    x = torch.rand(2, 3, 72, 96, generator=gen).to(dev)
    gy = torch.randn(2, 3, 72, 96, generator=gen).to(dev)
    P = (0.2 * torch.randn(2, 64, generator=gen)).to(dev)

    def run():
        res = []
        for M in (C.RGB2LAB, C.LAB2RGB, C.RGB2HSV, C.HSV2RGB):
            xi = x.clone().requires_grad_(True)
            y = M().to(dev)(xi)
            y.backward(gy)
            res += [y.detach().clone(), xi.grad.clone()]
        for fn, n in ((V.adjust_lab, 48), (V.adjust_rgb, 48), (V.adjust_hsv, 64)):
            xi = x.clone().requires_grad_(True)
            Pi = P[:, :n].clone().requires_grad_(True)
            y, reg = fn(xi, Pi)
            torch.autograd.backward([y, reg], [gy, torch.ones_like(reg)])
            res += [y.detach().clone(), xi.grad.clone(), Pi.grad.clone()]
        return res

    free = run()
    grid_cap(2)
    capped = run()
    grid_cap(0)
    for i, (a, b) in enumerate(zip(capped, free)):
        if a.dim() == 2:      # knot gradients: fold order differs
            assert parity.rel_to_max(a.cpu().numpy(), b.cpu().numpy()) <= 2e-6, i
        else:
            assert torch.equal(a, b), i


# ---- TriSpace residual: the same two kinds of test ----------------------------------------------------------

def _tri_inputs(B, H, W, spatial, degree, seed):
    from oracle import curl_oracle as O
    gen = torch.Generator().manual_seed(seed)
    nc = len(O.poly_powers(degree, 5 if spatial else 3))
    img = torch.rand(B, 3, H, W, generator=gen)
    co = [0.3 * torch.randn(B, 3, nc, generator=gen) for _ in range(3)]
    gout = torch.randn(B, 3, H, W, generator=gen)
    return img, co, gout


def _tri_gpu(img, co, gout, spatial, degree, save=True, want_gimg=True):
    import curl_b200
    import torch.nn as nn
    dev = torch.device("cuda:0")
    net = curl_b200.TriSpaceRegNet(polynomial_order=degree, spatial=spatial, backbone=nn.Identity()).to(dev)
    net.save_activations = save
    x = img.to(dev).requires_grad_(want_gimg)
    a = [c.to(dev).requires_grad_(True) for c in co]
    out = net._fused(x, *a, True)
    out.backward(gout.to(dev))
    res = dict(out=out.detach(), gR=a[0].grad, gL=a[1].grad, gH=a[2].grad)
    if want_gimg:
        res["gimg"] = x.grad
    return {k: v.cpu().numpy() for k, v in res.items()}


def _tri_check(got, r32, r64, rep):
    own = np.abs(r32["out"] - r64["out"]).max()
    e64 = np.abs(got["out"] - r64["out"])
    rep["out_max_vs_ref64"] = float(e64.max()); rep["ref32_vs_ref64_max"] = float(own)
    assert e64.max() <= max(2e-5, 3 * own), rep
    rep["out_frac_over_1e5_vs_ref32"] = float((np.abs(got["out"] - r32["out"]) > 1e-5 * np.maximum(1.0, np.abs(r64["out"]))).mean())
    assert rep["out_frac_over_1e5_vs_ref32"] <= 2e-3, rep
    n_px = got["out"].shape[2] * got["out"].shape[3]
    for k in ("gR", "gL", "gH"):
        # a coefficient gradient is a random-walk sum over the image (gout is randn): one pixel whose final-clamp gate
        # flips between two roundings is worth ~ 1/sqrt(n) of it (7e-4 at 2 Mpx; seen in rgb_d4_1080p, where the
        # reference's own fp32 run is 3e-4 from its fp64 one) -- the same flip allowance as for the layer's knots
        rep[k] = parity.check_knots_stat(got[k], r32[k], r64[k], n_pixels=n_px)
    parity.check_trispace_gimg(got["gimg"], r32["gimg"], r64["gimg"], rep)


TRI_CAPPED = [("spatial_d4_cap3", (2, 72, 100), True, 4, 3), ("rgb_d4_cap2", (1, 96, 128), False, 4, 2),
              ("spatial_d2_odd_cap1", (1, 37, 53), True, 2, 1), ("rgb_d3_cap1", (2, 40, 52), False, 3, 1)]


@pytest.mark.parametrize("name,shape,spatial,degree,cap", TRI_CAPPED, ids=[c[0] for c in TRI_CAPPED])
def test_trispace_capped_grid_vs_oracle(name, shape, spatial, degree, cap, grid_cap):
    """Forward, coefficient-gradient and image-gradient kernels with a few blocks per image (every block takes many
    grid-stride steps / tiles): bit-identical to the uncapped launch where the result is element-wise, and against
    the oracle (whole image, fp32 and fp64)."""
    from chunked import trispace_fwd_bwd_chunked
    B, H, W = shape
    img, co, gout = _tri_inputs(B, H, W, spatial, degree, 900 + H)
    r32 = trispace_fwd_bwd_chunked(img, *co, gout, degree, spatial, True, dtype=torch.float32, rows=H)
    r64 = trispace_fwd_bwd_chunked(img, *co, gout, degree, spatial, True, dtype=torch.float64, rows=H)
    free = _tri_gpu(img, co, gout, spatial, degree)
    assert grid_cap(cap) == 0
    capped = _tri_gpu(img, co, gout, spatial, degree)
    capped_recompute = _tri_gpu(img, co, gout, spatial, degree, save=False)
    assert grid_cap(0) == cap
    for k in ("out", "gimg"):
        assert np.array_equal(capped[k], free[k]), k
        assert np.array_equal(capped_recompute[k], free[k]), k
    for k in ("gR", "gL", "gH"):
        assert np.array_equal(capped[k], capped_recompute[k]), k          # saved == recomputed sigmoids
        assert parity.rel_to_max(capped[k], free[k]) <= 2e-6, k           # only the fold order of the partials differs
    rep = dict(shape=list(shape), spatial=spatial, degree=degree, cap=cap)
    _tri_check(capped, r32, r64, rep)
    REPORT["trispace.capped." + name] = rep


TRI_FULL = [("rgb_d4_1080p", (1, 1080, 1920), False, 4, 32), ("spatial_d4_360p", (1, 360, 640), True, 4, 8)]


@pytest.mark.parametrize("name,shape,spatial,degree,rows", TRI_FULL, ids=[c[0] for c in TRI_FULL])
def test_trispace_fullsize_vs_oracle(name, shape, spatial, degree, rows):
    """Uncapped launches at image sizes where the forward grid-strides and the backward blocks take several tiles
    (2 Mpx at 35 coefficients), and where x = column / W, y = row / H run over realistic ranges (126 coefficients),
    against the oracle evaluated in row chunks (tests/chunked.py)."""
    from chunked import trispace_fwd_bwd_chunked
    B, H, W = shape
    img, co, gout = _tri_inputs(B, H, W, spatial, degree, 77 + H)
    r32 = trispace_fwd_bwd_chunked(img, *co, gout, degree, spatial, True, dtype=torch.float32, rows=rows)
    r64 = trispace_fwd_bwd_chunked(img, *co, gout, degree, spatial, True, dtype=torch.float64, rows=rows)
    got = _tri_gpu(img, co, gout, spatial, degree)
    rep = dict(shape=list(shape), spatial=spatial, degree=degree)
    _tri_check(got, r32, r64, rep)
    REPORT["trispace." + name] = rep
