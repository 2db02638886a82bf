"""GPU parity tests: the CUDA path (through the module API -> autograd.Function -> C ABI) against
the committed golden fixtures (generated from the reference itself) and against the CPU oracle on
seeded inputs.  Run on the B200 box: python -m pytest tests -m gpu."""
import json
import os

import numpy as np
import pytest
import torch

import parity
from conftest import LAYER_CASES, LOSS_CASES, STAGE_CASES, ROOT, frac_mismatch, frac_over, rel_to_max

pytestmark = pytest.mark.gpu

REPORT = {}

# the stand-in backbones run on cuDNN; keep them in true fp32 so that they can be compared with the CPU
torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False


def _dev():
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    return torch.device("cuda:0")


@pytest.fixture(scope="module", autouse=True)
def _write_report():
    yield
    out = os.path.join(ROOT, "gpurun_out")
    os.makedirs(out, exist_ok=True)
    with open(os.path.join(out, "parity_report.json"), "w") as f:
        json.dump(REPORT, f, indent=1, sort_keys=True)


def _run_layer(case, g, dev, want_gimg=True):
    import curl_b200
    pts = tuple(int(v) for v in g[f"{case}.pts"])
    layer = curl_b200.CURLLayer(*pts).to(dev)
    img = torch.from_numpy(g[f"{case}.img"]).to(dev).requires_grad_(want_gimg)
    mask = torch.from_numpy(g[f"{case}.mask"]).to(dev)
    L, R, H = (torch.from_numpy(g[f"{case}.{k}"]).to(dev).requires_grad_(True) for k in "LRH")
    out, reg = layer(img, mask, L, R, H)
    gout = torch.from_numpy(g[f"{case}.gout"]).to(dev)
    greg = torch.from_numpy(g[f"{case}.greg"]).to(dev)
    torch.autograd.backward([out, reg], [gout, greg])
    res = dict(out=out.detach(), reg=reg.detach(), gL=L.grad, gR=R.grad, gH=H.grad)
    if want_gimg:
        res["gimg"] = img.grad
    return {k: v.cpu().numpy() for k, v in res.items()}


@pytest.mark.parametrize("case", LAYER_CASES)
def test_layer_vs_golden(case, layer_golden):
    g = layer_golden
    ours = _run_layer(case, g, _dev())
    r32 = {k: g[f"{case}.f32.{k}"] for k in ("out", "reg", "gimg", "gL", "gR", "gH")}
    r64 = {k: g[f"{case}.f64.{k}"] for k in ("out", "reg", "gimg", "gL", "gR", "gH")}
    rep = {}
    rep["out"] = parity.check_image(case, ours["out"], r32["out"], r64["out"])
    assert parity.rel_to_max(ours["reg"], r64["reg"]) <= 1e-5
    gscale = float(np.abs(r64["gimg"]).max())
    rep["gimg"] = parity.check_image(case, ours["gimg"], r32["gimg"], r64["gimg"], scale=gscale, tol=parity.GRAD_TOL)
    for k in ("gL", "gR", "gH"):
        assert ours[k].shape == r32[k].shape
        # ('identity': every pixel sits on a tie / clamp bound and the reference's own fp32 and fp64 gradients differ
        # by 3 % of the max-norm; judged, like the other edge cases, by fp64 triangulation -- parity.check_knots)
        rep[k] = parity.check_knots(case, ours[k], r32[k], r64[k])
    # element-wise against the reference's fp32 run: pixels on exact boundaries (black, white, grey) must take
    # the reference's side.  Genuine ties that rounding decides are bounded by the reference's own
    # fp32-vs-fp64 flip rate (the 'identity' case is made of nothing else).
    own = frac_mismatch(r64["gimg"], r32["gimg"])
    rep["gimg_frac_mismatch_vs_ref32"] = frac_mismatch(ours["gimg"], r32["gimg"])
    rep["gimg_frac_mismatch_ref64_vs_ref32"] = own
    assert rep["gimg_frac_mismatch_vs_ref32"] <= max(2e-3, 2 * own), rep
    REPORT[f"layer.{case}"] = rep


def test_layer_no_gimg_matches(layer_golden):
    a = _run_layer("smooth", layer_golden, _dev(), want_gimg=True)
    b = _run_layer("smooth", layer_golden, _dev(), want_gimg=False)
    for k in ("out", "reg", "gL", "gR", "gH"):
        assert np.array_equal(a[k], b[k]), k   # deterministic reduction: bitwise equal


@pytest.mark.parametrize("case", STAGE_CASES)
@pytest.mark.parametrize("mod", ["rgb2lab", "lab2rgb", "rgb2hsv", "hsv2rgb"])
def test_color_modules_vs_golden(case, mod, stage_golden):
    import curl_b200.colors as C
    g = stage_golden
    dev = _dev()
    M = {"rgb2lab": C.RGB2LAB, "lab2rgb": C.LAB2RGB, "rgb2hsv": C.RGB2HSV, "hsv2rgb": C.HSV2RGB}[mod]().to(dev)
    x = torch.from_numpy(g[f"{case}.x"]).to(dev).requires_grad_(True)
    y = M(x)
    y.backward(torch.from_numpy(g[f"{case}.gy"]).to(dev))
    y32, y64 = g[f"{case}.{mod}.f32.y"], g[f"{case}.{mod}.f64.y"]
    gx32, gx64 = g[f"{case}.{mod}.f32.gx"], g[f"{case}.{mod}.f64.gx"]
    name = "smooth" if case == "unit" and mod != "rgb2hsv" else "edges"
    scale = max(1.0, float(np.abs(y64).max()))   # lab2rgb of out-of-gamut input reaches |20|
    rep = dict(y=parity.check_image(name, y.detach().cpu().numpy(), y32, y64, scale=scale))
    gscale = float(np.abs(gx64).max())
    rep["gx"] = parity.check_image("edges", x.grad.cpu().numpy(), gx32, gx64, scale=gscale, tol=2e-5)
    rep["gx_frac_mismatch_vs_ref32"] = frac_mismatch(x.grad.cpu().numpy(), gx32)
    assert rep["gx_frac_mismatch_vs_ref32"] <= max(2e-3, 2 * frac_mismatch(gx64, gx32)), rep
    REPORT[f"color.{mod}.{case}"] = rep


@pytest.mark.parametrize("case", STAGE_CASES)
@pytest.mark.parametrize("fn", ["adjust_lab", "adjust_rgb", "adjust_hsv"])
def test_adjust_vs_golden(case, fn, stage_golden):
    import curl_b200.curves as V
    g = stage_golden
    dev = _dev()
    x = torch.from_numpy(g[f"{case}.x"]).to(dev).requires_grad_(True)
    P = torch.from_numpy(g[f"{case}.{fn}.P"]).to(dev).requires_grad_(True)
    y, reg = getattr(V, fn)(x, P)
    torch.autograd.backward([y, reg], [torch.from_numpy(g[f"{case}.gy"]).to(dev), torch.ones_like(reg)])
    pre = f"{case}.{fn}"
    name = "smooth" if case == "unit" else "edges"
    rep = dict(y=parity.check_image(name, y.detach().cpu().numpy(), g[f"{pre}.f32.y"], g[f"{pre}.f64.y"]))
    assert parity.rel_to_max(reg.detach().cpu().numpy(), g[f"{pre}.f64.reg"]) <= 1e-5
    gscale = float(np.abs(g[f"{pre}.f64.gx"]).max())
    rep["gx"] = parity.check_image(name, x.grad.cpu().numpy(), g[f"{pre}.f32.gx"], g[f"{pre}.f64.gx"], scale=gscale)
    rep["gP"] = parity.check_knots(name, P.grad.cpu().numpy(), g[f"{pre}.f32.gP"], g[f"{pre}.f64.gP"])
    REPORT[f"{fn}.{case}"] = rep


def test_apply_curve_vs_oracle():
    import curl_b200.curves as V
    from oracle import curl_oracle as O
    dev = _dev()
    gen = torch.Generator().manual_seed(5)
    x = torch.rand(2, 3, 10, 12, generator=gen) * 1.2 - 0.1
    C = torch.exp(0.2 * torch.randn(2, 16, generator=gen))
    gy = torch.randn(2, 3, 10, 12, generator=gen)
    for cin, cout in ((0, 0), (0, 1), (2, 1), (2, 2)):
        xr = x.clone().requires_grad_(True); Cr = C.clone().requires_grad_(True)
        yr, rr = O.apply_curve(xr, Cr, 0.0, cin, cout)
        torch.autograd.backward([yr, rr], [gy, torch.ones(2)])
        xg = x.to(dev).requires_grad_(True); Cg = C.to(dev).requires_grad_(True)
        yg, rg = V.apply_curve(xg, Cg, None, cin, cout)
        torch.autograd.backward([yg, rg], [gy.to(dev), torch.ones(2, device=dev)])
        assert (yg.cpu() - yr).abs().max().item() <= 1e-6
        assert parity.rel_to_max(rg.detach().cpu().numpy(), rr.detach().numpy()) <= 1e-5
        assert parity.rel_to_max(xg.grad.cpu().numpy(), xr.grad.numpy()) <= 1e-5
        assert parity.rel_to_max(Cg.grad.cpu().numpy(), Cr.grad.numpy()) <= 1e-4


@pytest.mark.parametrize("shape,maskdtype", [((4, 3, 64, 96), torch.bool), ((1, 3, 37, 53), torch.bool),
                                             ((2, 3, 128, 128), None), ((2, 3, 40, 40), torch.float32)])
def test_layer_vs_oracle_seeded(shape, maskdtype):
    """Fresh seeded inputs (not in the fixtures), oracle evaluated on this box's CPU in fp32 and fp64."""
    import curl_b200
    from oracle import curl_oracle as O
    dev = _dev()
    gen = torch.Generator().manual_seed(1234 + shape[2])
    B, _, H, W = shape
    img = torch.rand(shape, generator=gen)
    if maskdtype is None:
        mask, omask = None, torch.ones(B, 1, H, W, dtype=torch.bool)
    elif maskdtype == torch.bool:
        mask = torch.rand(B, 1, H, W, generator=gen) > 0.1
        omask = mask
    else:
        mask = (torch.rand(B, 1, H, W, generator=gen) > 0.1).float()
        omask = mask
    L, R, Hk = (0.1 * torch.randn(B, n, generator=gen) for n in (48, 48, 64))
    gout = torch.randn(shape, generator=gen)
    greg = torch.rand(B, generator=gen)
    r32 = O.curl_layer_fwd_bwd(img, omask, L, R, Hk, gout, greg)
    m64 = omask if omask.dtype == torch.bool else omask.double()
    r64 = O.curl_layer_fwd_bwd(img.double(), m64, L.double(), R.double(), Hk.double(), gout.double(), greg.double())
    layer = curl_b200.CURLLayer().to(dev)
    ig = img.to(dev).requires_grad_(True)
    Lg, Rg, Hg = (t.to(dev).requires_grad_(True) for t in (L, R, Hk))
    out, reg = layer(ig, None if mask is None else mask.to(dev), Lg, Rg, Hg)
    torch.autograd.backward([out, reg], [gout.to(dev), greg.to(dev)])
    n = lambda t: t.detach().cpu().numpy()
    rep = dict(out=parity.check_image("smooth", n(out), n(r32["out"]), n(r64["out"])))
    gscale = float(r64["gimg"].abs().max())
    rep["gimg"] = parity.check_image("smooth", n(ig.grad), n(r32["gimg"]), n(r64["gimg"]), scale=gscale,
                                     tol=parity.GRAD_TOL)
    for k, t in (("gL", Lg), ("gR", Rg), ("gH", Hg)):
        rep[k] = parity.check_knots("smooth", n(t.grad), n(r32[k]), n(r64[k]))
    REPORT[f"seeded.{shape}.{maskdtype}"] = rep


def test_strided_knot_views_and_single_vector_form():
    """GCURLNet passes column slices of one [B,160] tensor (model.py:197-199): no copy, same result."""
    import curl_b200
    dev = _dev()
    gen = torch.Generator().manual_seed(9)
    img = torch.rand(3, 3, 32, 48, generator=gen).to(dev)
    cp = (0.1 * torch.randn(3, 160, generator=gen)).to(dev).requires_grad_(True)
    layer = curl_b200.CURLLayer().to(dev)
    out_a, reg_a = layer(img, cp)                       # forward(img, curve_params)
    (out_a.sum() + reg_a.sum()).backward()
    ga = cp.grad.clone()
    L, R, H = (cp.detach()[:, a:b].contiguous().requires_grad_(True) for a, b in ((0, 48), (48, 96), (96, 160)))
    out_b, reg_b = layer(img, torch.ones(3, 1, 32, 48, dtype=torch.bool, device=dev), L, R, H)
    (out_b.sum() + reg_b.sum()).backward()
    assert torch.equal(out_a, out_b) and torch.equal(reg_a, reg_b)
    assert torch.equal(ga, torch.cat([L.grad, R.grad, H.grad], 1))


def test_masked_pixels_are_zero_and_inputs_untouched():
    import curl_b200
    dev = _dev()
    gen = torch.Generator().manual_seed(3)
    img = torch.rand(2, 3, 64, 64, generator=gen).to(dev)
    keep = img.clone()
    mask = (torch.rand(2, 1, 64, 64, generator=gen) > 0.5).to(dev)
    L, R, H = ((0.1 * torch.randn(2, n, generator=gen)).to(dev) for n in (48, 48, 64))
    out, _ = curl_b200.CURLLayer().to(dev)(img, mask, L, R, H)
    assert torch.equal(img, keep)
    assert (out * (~mask)).abs().max().item() == 0.0
    assert out.min().item() >= 0.0 and out.max().item() <= 1.0


def test_backward_run_to_run_determinism():
    """The backward is bitwise reproducible run to run, with and without the image gradient, over many repeats
    at shapes that keep every SM's three resident blocks busy.  Guards the shared-memory ring of the backward
    kernel: a stage may only be refilled once every warp's reads of it have been consumed (a refill issued
    right behind a barrier that does not wait for outstanding LDS showed up as rare differences of whole
    warps' pixels)."""
    import curl_b200
    dev = torch.device("cuda:0")
    gen = torch.Generator(device=dev).manual_seed(11)
    layer = curl_b200.CURLLayer().to(dev)
    for B, H, W, reps in ((4, 2160, 3840, 12), (3, 1080, 1920, 40), (32, 512, 512, 40)):
        img0 = torch.rand(B, 3, H, W, generator=gen, device=dev)
        mask = torch.rand(B, 1, H, W, generator=gen, device=dev) > 0.1
        Ls = [(0.1 * torch.randn(B, n, generator=gen, device=dev)).requires_grad_(True) for n in (48, 48, 64)]
        gout = torch.randn(B, 3, H, W, generator=gen, device=dev)
        for want_gimg in (True, False):
            first = None
            for _ in range(reps):
                img = img0.clone().requires_grad_(want_gimg)
                for t in Ls:
                    t.grad = None
                o, r = layer(img, mask, *Ls)
                torch.autograd.backward([o, r], [gout, torch.ones_like(r)])
                cur = [t.grad.clone() for t in Ls] + ([img.grad] if want_gimg else [])
                if first is None:
                    first = cur
                else:
                    for a, b in zip(cur, first):
                        assert torch.equal(a, b)
        del img0, mask, gout, first, cur


def test_full_size_properties():
    """BASELINE config sizes (1080p batch 16 forward; 4K batch 4 fwd+bwd): size-independent checks --
    range, mask zeros, determinism, batch-shard equality (no cross-image coupling), knot-gradient
    additivity over pixels (gradient of two row-halves sums to the gradient of the whole)."""
    import curl_b200
    dev = _dev()
    gen = torch.Generator(device=dev).manual_seed(1234)
    layer = curl_b200.CURLLayer().to(dev)
    B, H, W = 16, 1080, 1920
    img = torch.rand(B, 3, H, W, generator=gen, device=dev)
    mask = torch.rand(B, 1, H, W, generator=gen, device=dev) > 0.1
    L, R, Hk = (0.1 * torch.randn(B, n, generator=gen, device=dev) for n in (48, 48, 64))
    out, reg = layer(img, mask, L, R, Hk)
    out2, reg2 = layer(img, mask, L, R, Hk)
    assert torch.equal(out, out2) and torch.equal(reg, reg2)
    assert out.min().item() >= 0.0 and out.max().item() <= 1.0
    assert out.masked_select(~mask.expand_as(out)).abs().max().item() == 0.0
    o_lo, r_lo = layer(img[:5], mask[:5], L[:5], R[:5], Hk[:5])      # shard == slice of the whole
    assert torch.equal(o_lo, out[:5]) and torch.equal(r_lo, reg[:5])
    del out2, o_lo

    B, H, W = 4, 2160, 3840
    img = torch.rand(B, 3, H, W, generator=gen, device=dev)
    mask = torch.rand(B, 1, H, W, generator=gen, device=dev) > 0.1
    Ls = [(0.1 * torch.randn(B, n, generator=gen, device=dev)).requires_grad_(True) for n in (48, 48, 64)]
    gout = torch.randn(B, 3, H, W, generator=gen, device=dev)

    def grads(sl):
        for t in Ls:
            t.grad = None
        o, r = layer(img[:, :, sl].contiguous(), mask[:, :, sl].contiguous(), *Ls)
        o.backward(gout[:, :, sl].contiguous())
        return [t.grad.clone() for t in Ls]

    whole = grads(slice(0, H))
    again = grads(slice(0, H))
    top, bot = grads(slice(0, H // 2)), grads(slice(H // 2, H))
    for w, a, t, b in zip(whole, again, top, bot):
        assert torch.equal(w, a)                                       # deterministic reduction
        assert parity.rel_to_max((t + b).cpu().numpy(), w.cpu().numpy()) <= 1e-5


def test_cuda_graph_capture():
    import curl_b200
    dev = _dev()
    gen = torch.Generator().manual_seed(11)
    layer = curl_b200.CURLLayer().to(dev)
    img = torch.rand(2, 3, 64, 64, generator=gen).to(dev)
    L, R, H = ((0.1 * torch.randn(2, n, generator=gen)).to(dev) for n in (48, 48, 64))
    mask = torch.ones(2, 1, 64, 64, dtype=torch.bool, device=dev)
    ref, _ = layer(img, mask, L, R, H)
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        layer(img, mask, L, R, H)
    torch.cuda.current_stream().wait_stream(s)
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        out, reg = layer(img, mask, L, R, H)
    out.zero_()
    graph.replay()
    torch.cuda.synchronize()
    assert torch.equal(out, ref)


def test_errors():
    import curl_b200
    dev = _dev()
    layer = curl_b200.CURLLayer().to(dev)
    img = torch.rand(1, 3, 8, 8, device=dev)
    z = lambda n: torch.zeros(1, n, device=dev)
    with pytest.raises(RuntimeError):
        layer(img.cpu(), None, z(48), z(48), z(64))
    with pytest.raises(TypeError):
        layer(img.double(), None, z(48), z(48), z(64))
    with pytest.raises(ValueError):
        layer(img, None, z(47), z(48), z(64))
    with pytest.raises(ValueError):
        layer(img[:, :2], None, z(48), z(48), z(64))
    with pytest.raises(ValueError):
        layer(img, torch.ones(1, 3, 8, 8, device=dev), z(48), z(48), z(64))
    # non-contiguous image is accepted (the reference calls .contiguous())
    nc = torch.rand(1, 8, 8, 3, device=dev).permute(0, 3, 1, 2)
    a, _ = layer(nc, None, z(48), z(48), z(64))
    b, _ = layer(nc.contiguous(), None, z(48), z(48), z(64))
    assert torch.equal(a, b)
    # a checkpoint with different frozen constants is refused, not ignored
    bad = curl_b200.CURLLayer().to(dev)
    with torch.no_grad():
        bad.rgb2lab.rgb_to_xyz.mul_(1.01)
    with pytest.raises(RuntimeError):
        bad(img, None, z(48), z(48), z(64))


def _standin_backbone(seed):
    """timm is not in the image: a small conv net with the GCURLNet head width (160 knots).  The
    final layer is scaled down so that the predicted curves stay near identity (|param| ~ 0.1)."""
    import torch.nn as nn
    torch.manual_seed(seed)
    net = nn.Sequential(nn.Conv2d(3, 8, 3, stride=2, padding=1), nn.SiLU(), nn.Conv2d(8, 16, 3, stride=2, padding=1),
                        nn.SiLU(), nn.AdaptiveAvgPool2d(4), nn.Flatten(), nn.Linear(256, 160))
    with torch.no_grad():
        net[-1].weight.mul_(0.3)
        net[-1].bias.normal_(0.0, 0.1)
    return net


@pytest.mark.gpu
def test_gcurlnet_training_step_vs_oracle():
    """BASELINE config 3 in miniature: backbone (PyTorch/cuDNN) -> knots -> fused layer -> L1 loss +
    regulariser, one optimiser-free training step; every backbone gradient against the CPU oracle
    driven by the same backbone weights (reference model.py:194-203, main.py:254-262)."""
    import copy
    import curl_b200
    from oracle import curl_oracle as O
    dev = _dev()
    gen = torch.Generator().manual_seed(77)
    B, H, W = 4, 64, 64
    img = torch.rand(B, 3, H, W, generator=gen)
    tgt = torch.rand(B, 3, H, W, generator=gen)
    mask = torch.rand(B, 1, H, W, generator=gen) > 0.1

    cpu_bb = _standin_backbone(5)
    gpu_net = curl_b200.GCURLNet(backbone=copy.deepcopy(cpu_bb)).to(dev)

    def loss_of(out, reg, target, m):
        return ((out - target).abs() * m).sum() / (3 * m.sum()) + 1e-6 * reg.mean()

    curves = cpu_bb(img)
    o_out, o_reg = O.curl_layer(img, mask, curves[:, :48], curves[:, 48:96], curves[:, 96:160])
    loss_of(o_out, o_reg, tgt, mask).backward()

    g_out, g_reg = gpu_net(img.to(dev), mask.to(dev))
    loss = loss_of(g_out, g_reg, tgt.to(dev), mask.to(dev))
    loss.backward()
    assert frac_over(g_out.detach().cpu().numpy(), o_out.detach().numpy(), 1e-5) <= 1e-3
    for (name, pc), pg in zip(cpu_bb.named_parameters(), gpu_net.backbone.parameters()):
        assert pg.grad is not None, name
        assert rel_to_max(pg.grad.cpu().numpy(), pc.grad.numpy()) <= 2e-4, name


@pytest.mark.parametrize("layout,out_layout", [("CHW", None), ("HWC", None), ("HWC", "CHW"), ("CHW", "HWC")])
@pytest.mark.parametrize("shape,masked", [((3, 64, 96), True), ((2, 37, 53), True), ((1, 128, 128), False)])
def test_layer_u8_vs_oracle(layout, out_layout, shape, masked):
    """curl_layer_fwd_u8 (8-bit in/out, reference data path folded into the kernel) against the
    oracle: the fp32 core is the same device code as curl_layer_fwd_f32, so the only differences
    allowed are the +-1 steps where the fp32 result times 255 sits within 1e-5*255 of an integer
    (the image tolerance of north_star, carried through the truncation)."""
    import curl_b200
    from oracle import curl_oracle as O
    dev = _dev()
    B, H, W = shape
    gen = torch.Generator().manual_seed(4321 + H)
    img8 = torch.randint(0, 256, (B, 3, H, W), generator=gen, dtype=torch.uint8)
    mask = (torch.rand(B, 1, H, W, generator=gen) > 0.1) if masked else None
    L, R, Hk = (0.1 * torch.randn(B, n, generator=gen) for n in (48, 48, 64))
    omask = mask if masked else torch.ones(B, 1, H, W, dtype=torch.bool)
    ref8, ref_reg, ref32 = O.curl_layer_u8(img8, omask, L, R, Hk)
    layer = curl_b200.CURLLayer().to(dev)
    src = img8 if layout == "CHW" else img8.permute(0, 2, 3, 1).contiguous()
    with torch.no_grad():
        out8, reg = layer.forward_u8(src.to(dev), None if mask is None else mask.to(dev), L.to(dev), R.to(dev),
                                     Hk.to(dev), layout=layout, out_layout=out_layout)
        # the fp32 entry point on the same pixels: the u8 kernel must agree with it exactly
        f32, _ = layer(O.u8_to_unit(img8).to(dev), None if mask is None else mask.to(dev), L.to(dev), R.to(dev), Hk.to(dev))
    eff = out_layout or layout
    assert out8.dtype == torch.uint8 and out8.shape == ((B, 3, H, W) if eff == "CHW" else (B, H, W, 3))
    got = out8.cpu() if eff == "CHW" else out8.cpu().permute(0, 3, 1, 2)
    assert torch.equal(got, O.unit_to_u8(f32.cpu())), "u8 kernel differs from quantised fp32 kernel"
    diff = (got.to(torch.int16) - ref8.to(torch.int16)).abs()
    assert int(diff.max()) <= 1
    v = ref32 * 255
    near = (v - v.round()).abs() <= 255 * 3e-5           # within the image tolerance of a truncation step
    assert bool(((diff == 0) | near).all()), "u8 output differs away from a quantisation boundary"
    assert float((diff != 0).float().mean()) <= 2e-3
    assert rel_to_max(reg.cpu().numpy(), ref_reg.numpy()) <= 1e-5


def test_layer_u8_errors():
    import curl_b200
    dev = _dev()
    layer = curl_b200.CURLLayer().to(dev)
    z = lambda n: torch.zeros(1, n, device=dev)
    with pytest.raises(TypeError, match="uint8"):
        layer.forward_u8(torch.zeros(1, 3, 8, 8, device=dev), None, z(48), z(48), z(64))
    with pytest.raises(ValueError, match="shape"):
        layer.forward_u8(torch.zeros(1, 8, 8, 3, dtype=torch.uint8, device=dev), None, z(48), z(48), z(64))
    with pytest.raises(ValueError, match="layout"):
        layer.forward_u8(torch.zeros(1, 3, 8, 8, dtype=torch.uint8, device=dev), None, z(48), z(48), z(64), layout="NHWC")
    with pytest.raises(RuntimeError, match="inference only"):
        layer.forward_u8(torch.zeros(1, 3, 8, 8, dtype=torch.uint8, device=dev), None, z(48).requires_grad_(True), z(48), z(64))


# ---- TriSpace polynomial residual (curl_trispace_*) -----------------------------------------------------

def _tri_check(out, r32, r64, grads, g64):
    own = np.abs(r32 - r64).max()
    assert np.abs(out - r64).max() <= max(2e-5, 3 * own)
    # relative to max(1, |value|): the unclamped residual leaves [-3, 3] (see test_hostsim_trispace_vs_golden)
    assert (np.abs(out - r32) > 1e-5 * np.maximum(1.0, np.abs(r64))).mean() <= 2e-3
    for k, v in grads.items():
        assert rel_to_max(v, g64[k]) <= 1e-4, k


@pytest.mark.parametrize("save", [True, False])
@pytest.mark.parametrize("clamp", [0, 1])
@pytest.mark.parametrize("case", ["rgb", "spatial", "spatial_mobile", "uint8", "edges",
                                  "deg1_spatial", "deg2_rgb", "deg2_spatial", "deg3_rgb", "deg3_spatial"])
def test_trispace_vs_golden(case, clamp, save, trispace_golden):
    """Fused generate_residual (+ generate_image) forward and coefficient gradients against the
    fixtures the reference itself produced (model.py:486-508 through oracle/ref_loader)."""
    import curl_b200
    import torch.nn as nn
    g = trispace_golden
    dev = _dev()
    spatial = bool(g[f"{case}.spatial"])
    degree = int(g[f"{case}.degree"])                # ChannelPolyLayer is generic in the degree (model.py:263-333)
    net = curl_b200.TriSpaceRegNet(polynomial_order=degree, spatial=spatial, backbone=nn.Identity()).to(dev)
    net.save_activations = save      # backward from the saved sigmoid outputs, or recomputing the polynomials
    img = torch.from_numpy(g[f"{case}.img"]).to(dev).requires_grad_(True)      # the image gradient as well
    R, L, Hc = (torch.from_numpy(g[f"{case}.{k}"]).to(dev).requires_grad_(True) for k in ("R", "L", "H"))
    out = net.generate_residual(img, R, L, Hc) if not clamp else net._fused(img, R, L, Hc, True)
    out.backward(torch.from_numpy(g[f"{case}.gout"]).to(dev))
    parity.check_trispace_gimg(img.grad.cpu().numpy(), g[f"{case}.f32.c{clamp}.gimg"], g[f"{case}.f64.c{clamp}.gimg"], (case, clamp, save))
    # coefficient gradients do not depend on whether the image gradient was asked for
    a = [t.detach().clone().requires_grad_(True) for t in (R, L, Hc)]
    o2 = net._fused(img.detach(), *a, bool(clamp))
    o2.backward(torch.from_numpy(g[f"{case}.gout"]).to(dev))
    assert torch.equal(o2, out.detach())
    for t, u in zip(a, (R, L, Hc)):
        assert torch.equal(t.grad, u.grad)
    img = img.detach()
    if clamp:       # the two-step public route gives the same image as the fused one
        two = net.generate_image(img, net.generate_residual(img, R.detach(), L.detach(), Hc.detach()))
        assert (two - out.detach()).abs().max().item() <= 1e-6
    _tri_check(out.detach().cpu().numpy(), g[f"{case}.f32.c{clamp}.out"], g[f"{case}.f64.c{clamp}.out"],
               dict(gR=R.grad.cpu().numpy(), gL=L.grad.cpu().numpy(), gH=Hc.grad.cpu().numpy()),
               {k: g[f"{case}.f64.c{clamp}.{k}"] for k in ("gR", "gL", "gH")})


@pytest.mark.parametrize("shape,spatial,degree", [((2, 3, 96, 128), True, 4), ((3, 3, 37, 53), True, 4), ((2, 3, 64, 64), False, 4),
                                                  ((1, 3, 255, 131), False, 4), ((2, 3, 80, 112), True, 3), ((2, 3, 45, 67), True, 2),
                                                  ((1, 3, 64, 96), False, 3), ((2, 3, 33, 35), False, 1)])
def test_trispace_training_step_vs_oracle(shape, spatial, degree):
    """TriSpaceRegNet.forward (is_train=True) with a stand-in backbone: output image and every
    backbone gradient against the oracle on CPU (fp32 and fp64); odd sizes take the scalar path."""
    import copy
    import curl_b200
    import torch.nn as nn
    from oracle import curl_oracle as O
    dev = _dev()
    B, _, H, W = shape
    gen = torch.Generator().manual_seed(100 + H)
    nc = len(list(curl_b200.ChannelPolyLayer.generate_powers(degree, 5 if spatial else 3)))
    assert degree != 4 or nc == (126 if spatial else 35)
    torch.manual_seed(3)
    bb = nn.Sequential(nn.Conv2d(3, 8, 3, stride=4, padding=1), nn.SiLU(), nn.AdaptiveAvgPool2d(2), nn.Flatten(),
                       nn.Linear(32, 9 * nc))
    with torch.no_grad():
        bb[-1].weight.mul_(0.5); bb[-1].bias.normal_(0.0, 0.3)
    img = torch.rand(shape, generator=gen)
    mask = (torch.rand(B, 1, H, W, generator=gen) > 0.1).float()
    gout = torch.randn(shape, generator=gen)
    net = curl_b200.TriSpaceRegNet(polynomial_order=degree, spatial=spatial, backbone=copy.deepcopy(bb)).to(dev)
    out = net(img.to(dev), mask.to(dev))
    out.backward(gout.to(dev))
    res = {}
    for dtype in (torch.float32, torch.float64):
        b2 = copy.deepcopy(bb).to(dtype)
        co = b2(img.to(dtype) * mask.to(dtype)).reshape(B, 3, 3, nc)
        o = O.trispace_image(img.to(dtype), O.trispace_residual(img.to(dtype), co[:, 0], co[:, 1], co[:, 2], degree, spatial))
        (o * gout.to(dtype)).sum().backward()
        res[dtype] = (o.detach().numpy(), [q.grad.numpy() for q in b2.parameters()])
    own = np.abs(res[torch.float32][0] - res[torch.float64][0]).max()
    o = out.detach().cpu().numpy()
    assert np.abs(o - res[torch.float64][0]).max() <= max(2e-5, 3 * own)
    assert frac_over(o, res[torch.float32][0], 1e-5) <= 2e-3
    for pg, g64 in zip(net.backbone.parameters(), res[torch.float64][1]):
        assert rel_to_max(pg.grad.cpu().numpy(), g64) <= 2e-4


def test_trispace_deterministic_and_errors():
    import curl_b200
    import torch.nn as nn
    dev = _dev()
    net = curl_b200.TriSpaceRegNet(spatial=True, backbone=nn.Identity()).to(dev)
    gen = torch.Generator().manual_seed(8)
    img = torch.rand(2, 3, 120, 200, generator=gen).to(dev)
    gout = torch.randn(2, 3, 120, 200, generator=gen).to(dev)
    co = [(0.3 * torch.randn(2, 3, 126, generator=gen)).to(dev) for _ in range(3)]
    runs = []
    for save in (True, True, False):
        net.save_activations = save
        a = [c.clone().requires_grad_(True) for c in co]
        net._fused(img, *a, True).backward(gout)
        runs.append([t.grad.clone() for t in a])
    for x, y, z in zip(*runs):
        assert torch.equal(x, y)                              # fixed-order reduction
        assert torch.equal(x, z)                              # saved sigmoids == recomputed sigmoids, bit for bit
    with pytest.raises(ValueError, match="shape"):
        net.generate_residual(img, co[0][:, :, :35], co[1], co[2])
    # the image gradient alone (frozen coefficients): same values as together with the coefficient gradients,
    # saved sigmoids == recomputed sigmoids bit for bit
    gi = []
    for save, with_coef in ((True, True), (True, False), (False, False)):
        net.save_activations = save
        xi = img.clone().requires_grad_(True)
        a = [c.clone().requires_grad_(with_coef) for c in co]
        net._fused(xi, *a, True).backward(gout)
        gi.append(xi.grad.clone())
    assert torch.equal(gi[0], gi[1]) and torch.equal(gi[0], gi[2])
    assert torch.isfinite(gi[0]).all() and gi[0].abs().max() > 0
    with pytest.raises(ValueError, match="variable per channel"):      # stand-alone polynomial: 5 variables expected
        net.polylayer(img, co[0])
    with pytest.raises(RuntimeError, match="bad argument"):      # degrees 1..4 are built
        curl_b200.TriSpaceRegNet(polynomial_order=5, spatial=False, backbone=nn.Identity()).to(dev).generate_residual(
            img, *[c[:, :, :56].contiguous() for c in co])


# ---- fused colour terms of CURLLoss (reference model.py:36-118) -------------------------------------
from oracle import curl_oracle as O  # noqa: E402

def _run_loss(pred, target, mask, dev):
    import curl_b200
    crit = curl_b200.CURLLoss().to(dev)
    p = torch.as_tensor(pred).to(dev).requires_grad_(True)
    m = None if mask is None else torch.as_tensor(mask).to(dev)
    loss = crit(p, torch.as_tensor(target).to(dev), m)
    loss.backward()
    return float(loss.detach().cpu()), p.grad.cpu().numpy()


@pytest.mark.parametrize("case", LOSS_CASES)
def test_loss_vs_golden(case, loss_golden):
    """curl_b200.CURLLoss (fused colour-term kernels + library MS-SSIM) against the fixtures generated
    by the reference's CURLLoss: loss within 1e-5, gradient within 1e-4 of its max-norm (or 3x the
    reference's own fp32 error)."""
    g = loss_golden
    loss, gp = _run_loss(g[f"{case}.pred"], g[f"{case}.target"], g[f"{case}.mask"], _dev())
    r32, r64 = g[f"{case}.f32.gpred"], g[f"{case}.f64.gpred"]
    REPORT[f"loss.{case}"] = dict(loss_err=abs(loss - float(g[f"{case}.f64.loss"])), gpred_rel=rel_to_max(gp, r64),
                                  ref32_gpred_rel=rel_to_max(r32, r64))
    assert abs(loss - float(g[f"{case}.f64.loss"])) <= 1e-5, case
    assert rel_to_max(gp, r64) <= max(1e-4, 3 * rel_to_max(r32, r64)), case
    # element-wise against the reference's fp32 run (black / white / grey pixels: exact clamp boundaries and
    # ties must fall on the reference's side; the max-norm above cannot see them)
    REPORT[f"loss.{case}"]["frac_mismatch_vs_ref32"] = frac_mismatch(gp, r32)
    assert frac_mismatch(gp, r32) <= 1e-3, (case, frac_mismatch(gp, r32))


@pytest.mark.parametrize("shape,masked", [((4, 3, 128, 160), True), ((2, 3, 67, 91), True), ((1, 3, 256, 256), False)])
def test_loss_vs_oracle_seeded(shape, masked):
    """Larger seeded batches (even and odd H*W: packed and scalar kernels) against the CPU oracle in
    fp64, every term separately, and the gradient."""
    import curl_b200
    dev = _dev()
    gen = torch.Generator().manual_seed(sum(shape))
    target = torch.rand(shape, generator=gen)
    pred = (target + 0.08 * torch.randn(shape, generator=gen)).clamp(0, 1)
    mask = (torch.rand(shape[0], 1, *shape[2:], generator=gen) > 0.1) if masked else torch.ones(shape[0], 1, *shape[2:], dtype=torch.bool)
    ref = O.curl_loss_fwd_bwd(pred.double(), target.double(), mask)
    ref32 = O.curl_loss_fwd_bwd(pred, target, mask)
    crit = curl_b200.CURLLoss().to(dev)
    p = pred.to(dev).requires_grad_(True)
    terms = crit.terms(p, target.to(dev), mask.to(dev) if masked else None)
    for k in ("rgb", "cosine", "lab", "hsv", "ssim"):
        assert abs(float(terms[k].detach()) - float(ref[k])) <= 1e-5, (k, float(terms[k].detach()), float(ref[k]))
    loss = crit(p, target.to(dev), mask.to(dev))
    loss.backward()
    assert abs(float(loss.detach()) - float(ref["loss"])) <= 1e-5
    gp, r64, r32 = p.grad.cpu().numpy(), ref["gpred"].numpy(), ref32["gpred"].numpy()
    REPORT[f"loss.seeded.{shape}"] = dict(gpred_rel=rel_to_max(gp, r64), ref32_gpred_rel=rel_to_max(r32, r64),
                                          frac_over_1e4=frac_over(gp / np.abs(r64).max(), r64 / np.abs(r64).max(), 1e-4))
    assert rel_to_max(gp, r64) <= max(1e-4, 3 * rel_to_max(r32, r64))
    # masked pixels get exactly zero gradient
    if masked:
        dead = ~mask.expand(-1, 3, -1, -1).numpy()
        assert np.all(gp[dead] == 0.0)


def test_loss_terms_abi_deterministic_and_errors():
    """The C-ABI entry points directly: sums and L planes against the oracle, run-to-run bit
    identity of the ticketed reduction, float mask == bool mask, argument errors."""
    import curl_b200
    from curl_b200 import _lib, functional
    dev = _dev()
    gen = torch.Generator().manual_seed(77)
    B, H, W = 3, 96, 112
    target = torch.rand(B, 3, H, W, generator=gen)
    pred = torch.rand(B, 3, H, W, generator=gen)
    mask = torch.rand(B, 1, H, W, generator=gen) > 0.25
    a = functional.loss_terms(pred.to(dev), target.to(dev), mask.to(dev))
    b = functional.loss_terms(pred.to(dev), target.to(dev), mask.to(dev))
    c = functional.loss_terms(pred.to(dev), target.to(dev), mask.float().to(dev))
    for x, y, z in zip(a, b, c):
        assert torch.equal(x, y) and torch.equal(x, z)
    sums = a[0].cpu().double()
    pm, tm = pred.double() * mask, target.double() * mask
    want = torch.stack([(pm - tm).abs().sum(dim=(1, 2, 3)),
                        torch.nn.functional.cosine_similarity(pm, tm, dim=1).sum(dim=(1, 2)),
                        (O.loss_lab(pm) - O.loss_lab(tm)).abs().sum(dim=(1, 2, 3)),
                        (O.loss_hsv_cone(pm) - O.loss_hsv_cone(tm)).abs().sum(dim=(1, 2, 3)),
                        mask.double().sum(dim=(1, 2, 3))], dim=1)
    assert torch.all((sums - want).abs() <= 2e-6 * want.abs().clamp(min=1.0)), (sums, want)
    assert torch.equal(sums[:, 4], want[:, 4])
    assert (a[1].cpu() - O.loss_lab(pm)[:, 0:1].float()).abs().max() <= 1e-5
    assert (a[2].cpu() - O.loss_lab(tm)[:, 0:1].float()).abs().max() <= 1e-5
    # raw ABI errors
    lib = _lib.load()
    p, t = pred.to(dev), target.to(dev)
    s = torch.empty(B, 5, device=dev)
    n = lib.curl_loss_terms_workspace_bytes(B, H, W)
    ws = torch.empty(n, dtype=torch.uint8, device=dev)
    ok = lambda *args: lib.curl_loss_terms_fwd_f32(*args)
    st = torch.cuda.current_stream(dev).cuda_stream
    assert ok(p.data_ptr(), t.data_ptr(), None, 0, s.data_ptr(), None, None, ws.data_ptr(), n, B, H, W, st) == 0
    assert ok(None, t.data_ptr(), None, 0, s.data_ptr(), None, None, ws.data_ptr(), n, B, H, W, st) == -1
    assert ok(p.data_ptr(), t.data_ptr(), None, 1, s.data_ptr(), None, None, ws.data_ptr(), n, B, H, W, st) == -1
    assert ok(p.data_ptr(), t.data_ptr(), None, 0, s.data_ptr(), None, None, ws.data_ptr(), n - 1, B, H, W, st) == -3
    assert ok(p.data_ptr(), t.data_ptr(), None, 0, s.data_ptr(), None, None, ws.data_ptr(), n, 0, H, W, st) == -1
    assert lib.curl_loss_terms_bwd_f32(p.data_ptr(), t.data_ptr(), None, 0, None, None, p.data_ptr(), B, H, W, st) == -1
    assert lib.curl_loss_terms_workspace_bytes(0, H, W) == 0
    torch.cuda.synchronize()
    # module-level errors
    crit = curl_b200.CURLLoss().to(dev)
    with pytest.raises(ValueError):
        crit(p, t[:, :, :-1], mask.to(dev))
    with pytest.raises(RuntimeError):
        crit(p, t.clone().requires_grad_(True), mask.to(dev)).backward()
    bad = curl_b200.CURLLoss().to(dev)
    with torch.no_grad():
        bad.rgb2lab.rgb_to_xyz.mul_(1.01)
    with pytest.raises(RuntimeError):
        bad(p, t, mask.to(dev))


def test_loss_training_step_through_layer():
    """criterion(net(img), target, mask).backward() with the curve layer underneath: gradients reach
    the knots through both fused backward kernels and match the oracle chain."""
    import curl_b200
    dev = _dev()
    gen = torch.Generator().manual_seed(123)
    B, H, W = 2, 64, 80
    img = torch.rand(B, 3, H, W, generator=gen)
    target = (img * 1.1 + 0.02).clamp(0, 1)
    mask = torch.rand(B, 1, H, W, generator=gen) > 0.1
    knots = [0.1 * torch.randn(B, n, generator=gen) for n in (48, 48, 64)]
    # oracle, fp64
    kd = [k.double().requires_grad_(True) for k in knots]
    out, reg = O.curl_layer(img.double(), mask, *kd)
    ref_loss = O.curl_loss(out, target.double(), mask) + 1e-6 * reg.sum()
    ref_g = torch.autograd.grad(ref_loss, kd)
    layer, crit = curl_b200.CURLLayer().to(dev), curl_b200.CURLLoss().to(dev)
    kg = [k.to(dev).requires_grad_(True) for k in knots]
    out, reg = layer(img.to(dev), mask.to(dev), *kg)
    loss = crit(out, target.to(dev), mask.to(dev)) + 1e-6 * reg.sum()
    loss.backward()
    assert abs(float(loss.detach()) - float(ref_loss.detach())) <= 1e-5
    for k, r in zip(kg, ref_g):
        assert rel_to_max(k.grad.cpu().numpy(), r.numpy()) <= 1e-4


def test_loss_full_size_properties():
    """BASELINE-size images (4K): properties that need no oracle.  pred == target makes the three L1
    sums exactly zero and the colour-term gradient exactly zero; per-image sums do not depend on
    what else is in the batch (to reduction-order tolerance); masked pixels never receive gradient."""
    from curl_b200 import functional
    dev = _dev()
    B, H, W = 3, 2160, 3840
    gen = torch.Generator(device=dev).manual_seed(4)
    target = torch.rand(B, 3, H, W, generator=gen, device=dev)
    pred = (target + 0.05 * torch.randn(B, 3, H, W, generator=gen, device=dev)).clamp_(0, 1)
    mask = torch.rand(B, 1, H, W, generator=gen, device=dev) > 0.1
    same, Lp, Lt = functional.loss_terms(target, target, mask)
    assert torch.all(same[:, [0, 2, 3]] == 0) and torch.equal(Lp, Lt)
    assert torch.equal(same[:, 4], mask.sum(dim=(1, 2, 3)).float())
    assert torch.all((same[:, 1] - same[:, 4]).abs() <= 1e-5 * same[:, 4])          # cos == 1 on unmasked (non-black) pixels
    t = target.clone().requires_grad_(True)
    s, lp, _ = functional.loss_terms(t, target, mask)
    torch.autograd.backward([s], [torch.tensor([[1.0, 0.0, 1.0, 1.0, 0.0]], device=dev).expand(B, 5).contiguous()])
    assert torch.count_nonzero(t.grad) == 0                                           # sign(0) == 0 everywhere
    whole, _, _ = functional.loss_terms(pred, target, mask)
    for b in range(B):
        one, _, _ = functional.loss_terms(pred[b:b + 1], target[b:b + 1], mask[b:b + 1])
        assert torch.all((one[0] - whole[b]).abs() <= 2e-6 * whole[b].abs().clamp(min=1.0)), b
    p = pred.clone().requires_grad_(True)
    s, lp, _ = functional.loss_terms(p, target, mask)
    (s.sum() + lp.sum()).backward()
    assert torch.count_nonzero(p.grad * (~mask)) == 0
    assert torch.isfinite(p.grad).all()


# ---- opt-in segment-clamped curves (CURL paper eq. 1; not the reference: parity unpinned) -----------

@pytest.mark.parametrize("shape,maskdtype", [((3, 3, 64, 96), torch.bool), ((1, 3, 37, 53), torch.float32), ((2, 3, 48, 48), None)])
def test_segment_clamp_layer_vs_oracle(shape, maskdtype):
    import curl_b200
    dev = _dev()
    B, _, H, W = shape
    gen = torch.Generator().manual_seed(H * W + B)
    img = torch.rand(shape, generator=gen)
    L, R, Hk = (0.1 * torch.randn(B, n, generator=gen) for n in (48, 48, 64))
    gout, greg = torch.randn(shape, generator=gen), torch.randn(B, generator=gen)
    if maskdtype is None:
        mask, om = None, torch.ones(B, 1, H, W, dtype=torch.bool)
    else:
        mask = (torch.rand(B, 1, H, W, generator=gen) > 0.1).to(maskdtype)
        om = mask
    r64 = O.curl_layer_fwd_bwd(img.double(), om if om.dtype == torch.bool else om.double(), L.double(), R.double(),
                               Hk.double(), gout.double(), greg.double(), segment_clamp=True)
    r32 = O.curl_layer_fwd_bwd(img, om, L, R, Hk, gout, greg, segment_clamp=True)
    layer = curl_b200.CURLLayer(segment_clamp=True).to(dev)
    xi = img.to(dev).requires_grad_(True)
    kn = [t.to(dev).requires_grad_(True) for t in (L, R, Hk)]
    out, reg = layer(xi, None if mask is None else mask.to(dev), *kn)
    torch.autograd.backward([out, reg], [gout.to(dev), greg.to(dev)])
    own = float((r32["out"].double() - r64["out"]).abs().max())
    assert (out.detach().cpu().double() - r64["out"]).abs().max().item() <= max(2e-5, 3 * own)
    assert frac_over(out.detach().cpu().numpy(), r32["out"].numpy(), 1e-5) <= 5e-3
    assert rel_to_max(reg.detach().cpu().numpy(), r64["reg"].numpy()) <= 1e-5
    got = dict(gimg=xi.grad, gL=kn[0].grad, gR=kn[1].grad, gH=kn[2].grad)
    rep = {}
    for k, v in got.items():
        rep[k] = rel_to_max(v.cpu().numpy(), r64[k].numpy())
        if k == "gimg":
            # d scale / dx jumps at every knot (15 per curve): a pixel within rounding of a knot may take the
            # neighbouring segment's slope -- the value is continuous there, the derivative is not.  Bound
            # the fraction of such pixels instead of the maximum.
            scale = np.abs(r64[k].numpy()).max()
            rep["gimg_frac_over_1e-4"] = frac_over(v.cpu().numpy() / scale, r64[k].numpy() / scale, 1e-4)
            assert rep["gimg_frac_over_1e-4"] <= 2e-3, rep
        else:
            # the same few pixels (expected: pixel-curve events x rounding difference in t, about 5 here) change
            # the knot sums by O(one pixel's contribution) each: measured 2e-3 of the max-norm with a random
            # gout on these small images; the CPU instantiation of the same code (libm instead of MUFU) holds
            # 1e-4 (tests/test_hostsim.py::test_hostsim_segment_clamp_vs_oracle)
            assert rep[k] <= 5e-3, (k, rep[k])
    REPORT[f"segclamp.{shape}"] = rep


@pytest.mark.parametrize("content", ["noise", "smooth"])
def test_segment_clamp_knot_gradients_are_reproducible(content):
    """The per-segment reduction adds without atomics in an order fixed by the lane numbers: three runs give the same
    bits.  `smooth` puts most lanes of a warp into one segment (the butterfly path of WarpBins::add), `noise`
    spreads them (the rank-rounds path); both against the oracle as well."""
    import curl_b200
    dev = _dev()
    B, H, W = 2, 96, 160
    gen = torch.Generator().manual_seed(5)
    if content == "noise":
        img = torch.rand(B, 3, H, W, generator=gen)
    else:
        yy = torch.linspace(0, 1, H).reshape(1, 1, H, 1); xx = torch.linspace(0, 1, W).reshape(1, 1, 1, W)
        img = (0.5 + 0.45 * torch.sin(2.0 * xx + 1.5 * yy + torch.arange(3.0).reshape(1, 3, 1, 1))).expand(B, 3, H, W).contiguous()
    mask = torch.rand(B, 1, H, W, generator=gen) > 0.1
    L, R, Hk = (0.1 * torch.randn(B, n, generator=gen) for n in (48, 48, 64))
    gout, greg = torch.randn(B, 3, H, W, generator=gen), torch.randn(B, generator=gen)
    layer = curl_b200.CURLLayer(segment_clamp=True).to(dev)
    runs = []
    for _ in range(3):
        xi = img.to(dev).requires_grad_(True)
        kn = [t.to(dev).requires_grad_(True) for t in (L, R, Hk)]
        out, reg = layer(xi, mask.to(dev), *kn)
        torch.autograd.backward([out, reg], [gout.to(dev), greg.to(dev)])
        runs.append([xi.grad.clone()] + [t.grad.clone() for t in kn])
    for other in runs[1:]:
        for a, b_ in zip(runs[0], other):
            assert torch.equal(a, b_)
    r64 = O.curl_layer_fwd_bwd(img.double(), mask, L.double(), R.double(), Hk.double(), gout.double(), greg.double(),
                               segment_clamp=True)
    for k, v in zip(("gL", "gR", "gH"), runs[0][1:]):
        assert rel_to_max(v.cpu().numpy(), r64[k].numpy()) <= 5e-3, k


def test_segment_clamp_properties_and_errors():
    """With all parameters zero every knot is 1, both curve definitions give scale == 1 and the two
    layers must agree exactly; collinear knots on [0, 1] likewise agree up to rounding; without an
    image gradient the knot gradients are unchanged; argument errors are reported."""
    import curl_b200
    dev = _dev()
    gen = torch.Generator().manual_seed(21)
    img = torch.rand(2, 3, 40, 56, generator=gen).to(dev)
    mask = (torch.rand(2, 1, 40, 56, generator=gen) > 0.2).to(dev)
    seg, ref = curl_b200.CURLLayer(segment_clamp=True).to(dev), curl_b200.CURLLayer().to(dev)
    z = [torch.zeros(2, n, device=dev) for n in (48, 48, 64)]
    a, ra = seg(img, mask, *z)
    b, rb = ref(img, mask, *z)
    assert (a - b).abs().max().item() <= 2e-6 and torch.equal(ra, rb)
    kn = [(0.1 * torch.randn(2, n, generator=gen)).to(dev).requires_grad_(True) for n in (48, 48, 64)]
    out, reg = seg(img, mask, *kn)
    (out.sum() + reg.sum()).backward()
    g1 = [k.grad.clone() for k in kn]
    for k in kn:
        k.grad = None
    xi = img.clone().requires_grad_(True)
    out, reg = seg(xi, mask, *kn)
    (out.sum() + reg.sum()).backward()
    for x, y in zip(g1, [k.grad for k in kn]):
        assert torch.equal(x, y)       # same bins with and without the image gradient, added in a fixed order
    assert torch.count_nonzero(xi.grad * (~mask)) == 0
    with pytest.raises(ValueError):
        seg(img, mask, kn[0][:, :47], kn[1], kn[2])
    lib = curl_b200._lib.load()
    assert lib.curl_layer_seg_bwd_workspace_bytes(0, 8, 8, 16, 16, 16) == 0
    assert lib.curl_layer_seg_fwd_f32(None, None, 0, None, 48, None, 48, None, 64, None, None, 1, 8, 8, 16, 16, 16, None) == -1


def test_widened_ops_capture_in_cuda_graphs_and_accept_views():
    """CURLLoss (fused terms + library MS-SSIM), the TriSpace residual and the segment-clamped layer
    issue no host synchronisation: forward + backward capture into a CUDA graph and replay to the
    same bits (the loss terms and TriSpace reductions are fixed-order).  Non-contiguous inputs are
    accepted like the reference accepts them."""
    import curl_b200
    import torch.nn as nn
    dev = _dev()
    gen = torch.Generator().manual_seed(31)
    B, H, W = 2, 64, 80
    target = torch.rand(B, 3, H, W, generator=gen).to(dev)
    pred = (target + 0.05 * torch.randn(B, 3, H, W, generator=gen).to(dev)).clamp(0, 1).requires_grad_(True)
    mask = (torch.rand(B, 1, H, W, generator=gen) > 0.1).to(dev)
    crit = curl_b200.CURLLoss().to(dev)
    tri = curl_b200.TriSpaceRegNet(spatial=True, backbone=nn.Identity()).to(dev)
    co = [(0.3 * torch.randn(B, 3, 126, generator=gen)).to(dev).requires_grad_(True) for _ in range(3)]

    def step():
        out = tri._fused(target, *co, True)
        loss = crit(pred, target, mask) + (out * target).mean()
        return torch.autograd.grad(loss, [pred] + co) + (loss.detach(),)

    eager = [t.clone() for t in step()]
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        step()
    torch.cuda.current_stream().wait_stream(s)
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        captured = step()
    for t in captured:
        t.zero_()
    graph.replay()
    torch.cuda.synchronize()
    for a, b in zip(eager, captured):
        assert torch.equal(a, b)
    # views: channels-last prediction, expanded mask
    nc = pred.detach().permute(0, 2, 3, 1).contiguous().permute(0, 3, 1, 2)
    assert not nc.is_contiguous()
    a = crit(nc, target, mask[:1].expand(B, 1, H, W) | mask)
    b = crit(pred.detach(), target, mask[:1] | mask)
    assert torch.equal(a, b)


@pytest.mark.parametrize("layout", ["CHW", "HWC"])
@pytest.mark.parametrize("shape", [(2, 64, 96), (1, 37, 53), (3, 2, 128)])
def test_to_tensor_u8_is_exact(layout, shape):
    """Device-side TF.to_tensor: bit-identical to u8 / 255 (torchvision, data.py:152-158) for every byte
    value, both layouts, vector and scalar paths; and a training step from its output equals the step from
    the host-converted tensor."""
    from curl_b200 import functional
    dev = _dev()
    B, H, W = shape
    gen = torch.Generator().manual_seed(H + W)
    u8 = torch.randint(0, 256, (B, 3, H, W), generator=gen, dtype=torch.uint8)
    u8.view(-1)[:256] = torch.arange(256, dtype=torch.uint8)
    src = u8 if layout == "CHW" else u8.permute(0, 2, 3, 1).contiguous()
    got = functional.to_tensor_u8(src.to(dev), layout).cpu()
    want = O.u8_to_unit(u8)
    assert got.dtype == torch.float32 and torch.equal(got.view(torch.int32), want.view(torch.int32))
    with pytest.raises(TypeError):
        functional.to_tensor_u8(src.float().to(dev), layout)
    with pytest.raises(ValueError):
        functional.to_tensor_u8(src.to(dev)[:, :2] if layout == "CHW" else src.to(dev)[..., :2], layout)
