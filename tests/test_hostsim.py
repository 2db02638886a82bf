"""CPU: the kernel's per-pixel templates (curl_b200/csrc/curl_pixel.cuh), instantiated on the host
with T = float by tests/hostsim, against the golden fixtures.  This checks the hand-derived
reverse-mode code, the clamp/tie/branch semantics and the moment -> knot-gradient expansion without
a GPU; the GPU tests then only have to establish that the MUFU approximations and the packed
arithmetic stay inside the same bounds."""
import numpy as np
import pytest
import torch

import hostsim
import parity
from conftest import LAYER_CASES, LOSS_CASES, STAGE_CASES, frac_mismatch, rel_to_max
from oracle import curl_oracle as O


@pytest.mark.parametrize("case", LAYER_CASES)
def test_hostsim_layer_vs_golden(case, layer_golden):
    g = layer_golden
    pts = tuple(int(v) for v in g[f"{case}.pts"])
    args = [g[f"{case}.{k}"] for k in ("img", "mask", "L", "R", "H")]
    fwd = hostsim.layer(*args, pts)
    bwd = hostsim.layer(*args, pts, gout=g[f"{case}.gout"], greg=g[f"{case}.greg"])
    r32 = {k: g[f"{case}.f32.{k}"] for k in ("out", "reg", "gimg", "gL", "gR", "gH")}
    r64 = {k: g[f"{case}.f64.{k}"] for k in ("out", "reg", "gimg", "gL", "gR", "gH")}
    parity.check_image(case, fwd["out"], r32["out"], r64["out"])
    assert parity.rel_to_max(fwd["reg"], r64["reg"]) <= 1e-5
    gscale = float(np.abs(r64["gimg"]).max())
    if np.isfinite(gscale) and gscale > 0:
        parity.check_image(case, bwd["gimg"], r32["gimg"], r64["gimg"], scale=gscale, tol=parity.GRAD_TOL)
    for k, n in (("gL", pts[0]), ("gR", pts[1]), ("gH", pts[2])):
        assert bwd[k].shape[1] == n
        parity.check_knots(case, bwd[k], r32[k][:, :n], r64[k][:, :n])
        # columns beyond num_*_points are sliced away by the layer: the reference gradient is zero there
        assert np.all(r64[k][:, n:] == 0)


@pytest.mark.parametrize("case", STAGE_CASES)
@pytest.mark.parametrize("conv,name", [(0, "rgb2lab"), (1, "lab2rgb"), (2, "rgb2hsv"), (3, "hsv2rgb")])
def test_hostsim_colors_vs_golden(case, conv, name, stage_golden):
    g = stage_golden
    x, gy = g[f"{case}.x"], g[f"{case}.gy"]
    y = hostsim.color(conv, x)
    gx = hostsim.color(conv, x, gy)
    y64 = g[f"{case}.{name}.f64.y"]
    scale = max(1.0, float(np.abs(y64).max()))
    kind = "smooth" if case == "unit" and name != "rgb2hsv" else "edges"
    parity.check_image(kind, y, g[f"{case}.{name}.f32.y"], y64, scale=scale)
    gx64 = g[f"{case}.{name}.f64.gx"]
    parity.check_image("edges", gx, g[f"{case}.{name}.f32.gx"], gx64, scale=float(np.abs(gx64).max()), tol=2e-5)


@pytest.mark.parametrize("case", STAGE_CASES)
@pytest.mark.parametrize("group,name,K", [(0, "adjust_lab", 16), (1, "adjust_rgb", 16), (2, "adjust_hsv", 16)])
def test_hostsim_adjust_vs_golden(case, group, name, K, stage_golden):
    g = stage_golden
    x, gy, P = g[f"{case}.x"], g[f"{case}.gy"], g[f"{case}.{name}.P"]
    f = hostsim.adjust(group, x, P, K)
    b = hostsim.adjust(group, x, P, K, gy=gy, greg=np.ones(x.shape[0], np.float32))
    pre = f"{case}.{name}"
    kind = "smooth" if case == "unit" else "edges"
    parity.check_image(kind, f["y"], g[f"{pre}.f32.y"], g[f"{pre}.f64.y"])
    assert parity.rel_to_max(f["reg"], g[f"{pre}.f64.reg"]) <= 1e-5
    gx64 = g[f"{pre}.f64.gx"]
    parity.check_image(kind, b["gx"], g[f"{pre}.f32.gx"], gx64, scale=float(np.abs(gx64).max()), tol=parity.GRAD_TOL)
    parity.check_knots(kind, b["gP"], g[f"{pre}.f32.gP"], g[f"{pre}.f64.gP"])


def test_hostsim_tie_semantics():
    """Hue selectors add on ties; grey pixels give h = s = 1e-9; max/min route to the first index."""
    px = lambda r, g, b: np.array([r, g, b], np.float32).reshape(1, 3, 1, 1)
    assert abs(hostsim.color(2, px(0.5, 0.5, 0.2))[0, 0, 0, 0] - 1 / 3) < 1e-6
    assert abs(hostsim.color(2, px(0.2, 0.5, 0.5))[0, 0, 0, 0] - 1.0) < 1e-6
    assert abs(hostsim.color(2, px(0.5, 0.2, 0.5))[0, 0, 0, 0] - 2 / 3) < 1e-6
    grey = hostsim.color(2, px(0.3, 0.3, 0.3))[0, :, 0, 0]
    assert grey[0] == np.float32(1e-9) and grey[1] == np.float32(1e-9) and grey[2] == np.float32(0.3)
    gx = hostsim.color(2, px(0.3, 0.3, 0.3), np.array([0, 0, 1], np.float32).reshape(1, 3, 1, 1))[0, :, 0, 0]
    assert gx.tolist() == [1.0, 0.0, 0.0]      # d v / d rgb goes to channel 0 on a three-way tie


def test_u8_conversions_are_exact():
    """curl_layer_fwd_u8's conversions: u8 -> x/255 must be the correctly rounded quotient for all 256
    inputs (the kernel uses a multiply + one Newton step, not a division), and the way out must be
    trunc(rn(y*255)) like numpy's astype('uint8') (evaluate.py:63)."""
    import torch
    from oracle import curl_oracle as O
    tab = hostsim.u8_to_unit_table()
    want = O.u8_to_unit(torch.arange(256, dtype=torch.uint8)).numpy()
    assert np.array_equal(tab.view(np.uint32), want.view(np.uint32))
    rng = np.random.default_rng(3)
    y = np.concatenate([rng.random(200000, dtype=np.float32), np.arange(256, dtype=np.float32) / np.float32(255),
                        np.nextafter(np.arange(256, dtype=np.float32) / np.float32(255), np.float32(0)),
                        np.array([0.0, 1.0, 0.999999, 1e-9], np.float32)]).astype(np.float32)
    got = hostsim.unit_to_u8(y)
    assert np.array_equal(got, O.unit_to_u8(torch.from_numpy(y)).numpy())
    assert np.array_equal(got, (y * np.float32(255)).astype("uint8"))


@pytest.mark.parametrize("case", ["rgb", "spatial", "spatial_mobile", "uint8", "edges",
                                  "deg1_spatial", "deg2_rgb", "deg2_spatial", "deg3_rgb", "deg3_spatial"])
def test_hostsim_trispace_vs_golden(case, trispace_golden):
    """Per-pixel TriSpace math of the kernels (curl_poly.cuh: compile-time monomial order, fused
    sigmoid / colour chain, hand-derived coefficient gradient) against the reference fixtures."""
    g = trispace_golden
    spatial, degree = bool(g[f"{case}.spatial"]), int(g[f"{case}.degree"])
    img, R, L, Hc, gout = (g[f"{case}.{k}"] for k in ("img", "R", "L", "H", "gout"))
    for clamp in (0, 1):
        out = hostsim.trispace(img, R, L, Hc, spatial, clamp, degree=degree)
        r32, r64 = g[f"{case}.f32.c{clamp}.out"], g[f"{case}.f64.c{clamp}.out"]
        own = np.abs(r32 - r64).max()
        assert np.abs(out - r64).max() <= max(2e-5, 3 * own), (case, clamp)
        # 1e-5 relative to max(1, |value|): the unclamped residual is not confined to [-3, 3] (lab2rgb of an arbitrary
        # sigmoid triple leaves the gamut -- up to 17.6 in deg2_spatial, where the reference's own fp32 result is
        # more than 1e-5 from its fp64 one at 0.26 % of the elements)
        assert (np.abs(out - r32) > 1e-5 * np.maximum(1.0, np.abs(r64))).mean() <= 2e-3, (case, clamp)
        gr = hostsim.trispace(img, R, L, Hc, spatial, clamp, gout, degree=degree)
        for k in ("gR", "gL", "gH"):
            assert rel_to_max(gr[k], g[f"{case}.f64.c{clamp}.{k}"]) <= 1e-4, (case, clamp, k)
        # image gradient (autograd of the reference through model.py:486-508)
        gi = hostsim.trispace(img, R, L, Hc, spatial, clamp, gout, degree=degree, want_gimg=True)
        parity.check_trispace_gimg(gi, g[f"{case}.f32.c{clamp}.gimg"], g[f"{case}.f64.c{clamp}.gimg"], (case, clamp))


def _loss_from_terms(sums, L_pred, L_target, mask, msssim):
    """model.py:90-117 from the five per-image sums and the L planes (what curl_b200.CURLLoss does)."""
    B, _, H, W = L_pred.shape
    tot = sums.sum(0)
    n = 3 * tot[4]
    n_masked = 0.0 if mask is None else float((np.asarray(mask) == 0).sum())
    return (tot[0] / n + 1.0 - (tot[1] + n_masked) / (B * H * W) + tot[2] / n + tot[3] / n
            + 10 * (1.0 - msssim(L_pred, L_target)).mean()) / 5


@pytest.mark.parametrize("case", LOSS_CASES)
def test_hostsim_loss_terms_vs_golden(case, loss_golden):
    """Per-pixel math of the fused CURLLoss colour terms (curl_loss.cuh: cosine similarity, clamped
    Lab, HSV cone and the hand-derived gradient) against the reference fixtures; MS-SSIM (library
    path) is taken from the oracle and differentiated by autograd down to the L plane."""
    g = loss_golden
    pred, target, mask = g[f"{case}.pred"], g[f"{case}.target"], g[f"{case}.mask"]
    r = hostsim.loss_terms(pred, target, mask)
    sums = torch.from_numpy(r["sums"]).requires_grad_(True)
    Lp = torch.from_numpy(r["L_pred"]).requires_grad_(True)
    loss = _loss_from_terms(sums, Lp, torch.from_numpy(r["L_target"]), mask, O.msssim)
    assert abs(float(loss.detach()) - float(g[f"{case}.f64.loss"])) <= 1e-5, case          # fp tolerance of north_star
    gs, gL = torch.autograd.grad(loss, (sums, Lp))
    gp = hostsim.loss_terms(pred, target, mask, w=gs[:, :4].numpy(), gL=gL.numpy())
    r32, r64 = g[f"{case}.f32.gpred"], g[f"{case}.f64.gpred"]
    assert rel_to_max(gp, r64) <= max(1e-4, 3 * rel_to_max(r32, r64)), case
    # element-wise against the reference's fp32 run: black, white and grey pixels sit exactly on clamp
    # boundaries and ties (L = 0, L = 1, df = 0) and must take the reference's side of each
    assert frac_mismatch(gp, r32) == 0.0, (case, frac_mismatch(gp, r32))


@pytest.mark.parametrize("shape,sigma,masked", [((2, 24, 32), 0.1, True), ((1, 17, 19), 0.1, False), ((2, 16, 16), 0.3, True)])
def test_hostsim_segment_clamp_vs_oracle(shape, sigma, masked):
    """Per-pixel math of the opt-in segment-clamped layer (curl_seg.cuh: segment lookup, tie
    semantics at knots, binned knot gradients) against the oracle's restatement of the paper's
    eq. (1) with autograd.  No reference code exists for this flag (parity unpinned)."""
    B, H, W = shape
    gen = torch.Generator().manual_seed(B * H + W)
    img = torch.rand(B, 3, H, W, generator=gen)
    mask = (torch.rand(B, 1, H, W, generator=gen) > 0.15) if masked else None
    L, R, Hk = (sigma * torch.randn(B, n, generator=gen) for n in (48, 48, 64))
    gout, greg = torch.randn(B, 3, H, W, generator=gen), torch.randn(B, generator=gen)
    om = mask if masked else torch.ones(B, 1, H, W, dtype=torch.bool)
    r64 = O.curl_layer_fwd_bwd(img.double(), om, L.double(), R.double(), Hk.double(), gout.double(), greg.double(), segment_clamp=True)
    r32 = O.curl_layer_fwd_bwd(img, om, L, R, Hk, gout, greg, segment_clamp=True)
    m_np = None if mask is None else mask.numpy()
    f = hostsim.layer_seg(img.numpy(), m_np, L.numpy(), R.numpy(), Hk.numpy())
    b = hostsim.layer_seg(img.numpy(), m_np, L.numpy(), R.numpy(), Hk.numpy(), gout=gout.numpy(), greg=greg.numpy())
    own = float((r32["out"].double() - r64["out"]).abs().max())
    assert np.abs(f["out"] - r64["out"].numpy()).max() <= max(2e-5, 3 * own)
    assert rel_to_max(f["reg"], r64["reg"].numpy()) <= 1e-5
    for k in ("gimg", "gL", "gR", "gH"):
        assert rel_to_max(b[k], r64[k].numpy()) <= max(1e-4, 3 * rel_to_max(r32[k].numpy(), r64[k].numpy())), k


def test_hsv2rgb_unclamped_output_stays_in_unit_range_and_matches_reference():
    """HSV->RGB is evaluated as one trapezoid per channel and without the reference's output clamp
    (colors.py:175), which is inert for that form: every channel is fl(p + w ramp) with 0 <= p <= v <= 1,
    0 <= w <= v, 0 <= ramp <= 1.  Dense random inputs, inputs outside [0, 1] and every ramp end (hue = n/6,
    their float neighbours, 0, 1, 1e-9; s, v in {0, 1e-9, 1/3, 1 - 2^-24, 1}) stay inside [0, 1] and agree with the
    reference's two-ramp form; at the ramp ends the hue gradient is the reference's (closed clamp intervals)."""
    import torch
    from oracle import curl_oracle as O
    rng = np.random.default_rng(3)
    hues = [0.0, 1e-9, 1.0]
    for n in range(1, 6):
        h = np.float32(n / 6.0)
        hues += [float(h), float(np.nextafter(h, np.float32(0))), float(np.nextafter(h, np.float32(1)))]
    sv = [0.0, 1e-9, 1.0 / 3.0, float(np.float32(1) - np.float32(2.0 ** -24)), 1.0]
    grid = np.array([(h, s, v) for h in hues for s in sv for v in sv], np.float32)          # [n, 3]
    dense = rng.random((200000, 3), dtype=np.float32)
    wild = (rng.random((20000, 3), dtype=np.float32) * 3.0 - 1.0)                           # outside [0, 1] too
    x = np.concatenate([grid, dense, wild]).T.reshape(1, 3, 1, -1).copy()
    y = hostsim.color(3, x)
    assert y.min() >= 0.0 and y.max() <= 1.0
    ref = O.hsv2rgb(torch.from_numpy(x).double()).numpy()
    assert np.abs(y - ref).max() <= 2e-6
    # gradients on the grid of ramp ends and on generic points against the fp64 reference
    xg = np.concatenate([grid, dense[:20000]]).T.reshape(1, 3, 1, -1).copy()
    gy = rng.standard_normal(xg.shape).astype(np.float32)
    gx = hostsim.color(3, xg, gy)
    t = torch.from_numpy(xg).double().requires_grad_(True)
    O.hsv2rgb(t).backward(torch.from_numpy(gy).double())
    gref = t.grad.numpy()
    bad = np.abs(gx - gref) > 1e-4 * max(1.0, np.abs(gref).max())
    # a hue within one float step of a ramp end may take the neighbouring slope in fp32 (the reference's own
    # fp32 run does too); exact ends and generic points must agree
    near_end = np.zeros(xg.shape[-1], bool)
    k6 = xg[0, 0, 0].astype(np.float64) * 6.0
    near_end |= (np.abs(k6 - np.round(k6)) < 1e-6) & (np.abs(k6 - np.round(k6)) > 0)
    assert not bad[:, :, :, ~near_end].any()
