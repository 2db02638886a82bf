"""CPU: the reference-facing module API (names, constructor arguments, state-dict keys/shapes,
error behaviour).  Compute needs a GPU and is covered by tests/test_gpu_parity.py."""
import pytest
import torch

import curl_b200
from curl_b200 import colors, curves
from oracle import curl_oracle as O
from oracle import ref_loader


def test_state_dict_matches_reference_layout():
    layer = curl_b200.CURLLayer()
    sd = layer.state_dict()
    ref = O.constants()          # names/shapes/values of the reference's CURLLayer().state_dict()
    assert list(sd) == ["rgb2lab.rgb_to_xyz", "rgb2lab.fxfyfz_to_lab", "rgb2lab.xyz_to_rgb_mult",
                        "rgb2lab.lab_to_fxfyfz_offset", "lab2rgb.xyz_to_rgb", "lab2rgb.lab_to_fxfyfz",
                        "lab2rgb.xyz_to_rgb_mult", "lab2rgb.lab_to_fxfyfz_offset", "rgb2hsv.comparison_zero"]
    for k, v in sd.items():
        assert v.shape == ref[k].shape and torch.equal(v, ref[k]), k
    assert all(not p.requires_grad for p in layer.parameters())
    assert (layer.num_lab_points, layer.num_rgb_points, layer.num_hsv_points) == (48, 48, 64)
    assert curl_b200.CURLLayer(30, 18, 40).num_hsv_points == 40
    # the opt-in paper curves are off by default and add no state
    assert layer.segment_clamp is False and curl_b200.CURLLayer(segment_clamp=True).segment_clamp is True
    assert list(curl_b200.CURLLayer(segment_clamp=True).state_dict()) == list(layer.state_dict())


def test_strict_load_and_legacy_matrix_conversion():
    """Strict load_state_dict (main.py:184,246; infer.py:29) and convert_state.py:11-13, which reshapes
    legacy 2-D colour matrices to [1,1,3,3] by key name, keep working."""
    layer = curl_b200.CURLLayer()
    sd = {("module." + k): v.clone() for k, v in O.constants().items()}
    for k in list(sd):
        if sd[k].dim() == 4 and sd[k].shape[-2:] == (3, 3):
            sd[k] = sd[k].reshape(3, 3)                      # a legacy checkpoint
    conv = {}
    for k, v in sd.items():                                  # what convert_state.py does
        k = k[len("module."):] if k.startswith("module.") else k
        if k.split(".")[-1] in ("rgb_to_xyz", "fxfyfz_to_lab", "xyz_to_rgb", "lab_to_fxfyfz") and v.dim() == 2:
            v = v.reshape(1, 1, 3, 3)
        conv[k] = v
    assert layer.load_state_dict(conv, strict=True).missing_keys == []


@pytest.mark.skipif(not ref_loader.available(), reason="reference tree not present on this box")
def test_load_reference_state_dict_directly():
    _, _, RefLayer = ref_loader.load()
    layer = curl_b200.CURLLayer()
    res = layer.load_state_dict(RefLayer().state_dict(), strict=True)
    assert not res.missing_keys and not res.unexpected_keys


def test_constructor_validation():
    with pytest.raises(ValueError):
        curl_b200.CURLLayer(47, 48, 64)
    with pytest.raises(ValueError):
        curl_b200.CURLLayer(48, 48, 8)


def test_no_cpu_fallback():
    layer = curl_b200.CURLLayer()
    img = torch.rand(1, 3, 4, 4)
    z = lambda n: torch.zeros(1, n)
    with pytest.raises(RuntimeError, match="CUDA"):
        layer(img, None, z(48), z(48), z(64))
    for M in (colors.RGB2LAB, colors.LAB2RGB, colors.RGB2HSV, colors.HSV2RGB):
        with pytest.raises(RuntimeError, match="CUDA"):
            M()(img)
    for fn, n in ((curves.adjust_lab, 48), (curves.adjust_rgb, 48), (curves.adjust_hsv, 64)):
        with pytest.raises(RuntimeError, match="CUDA"):
            fn(img, z(n))
    with pytest.raises(RuntimeError, match="CUDA"):
        curves.apply_curve(img, torch.ones(1, 16), None, 0, 0)


def test_changed_constants_are_rejected():
    m = colors.RGB2LAB()
    with torch.no_grad():
        m.rgb_to_xyz[0, 0, 0, 0] += 0.01
    with pytest.raises(RuntimeError, match="frozen colour constant"):
        m(torch.rand(1, 3, 4, 4))


def test_product_code_does_not_import_the_oracle():
    import os
    from conftest import ROOT
    for dirpath, _, files in os.walk(os.path.join(ROOT, "curl_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in src and "from oracle" not in src, f
                assert "import hostsim" not in src and "pixel_sim" not in src, f


def test_gcurlnet_api():
    """GCURLNet (reference model.py:179-203): attribute names, knot split, explicit backbone."""
    import torch.nn as nn
    net = curl_b200.GCURLNet(backbone=nn.Sequential(nn.AdaptiveAvgPool2d(1), nn.Flatten(), nn.Linear(3, 160)))
    assert (net.curve_break_1, net.curve_break_2, net.num_curve_params) == (48, 96, 160)
    assert isinstance(net.curllayer, curl_b200.CURLLayer)
    assert [k for k in net.state_dict() if k.startswith("curllayer.")][0] == "curllayer.rgb2lab.rgb_to_xyz"
    with pytest.raises(RuntimeError, match="CUDA"):           # the layer has no CPU path
        net(torch.rand(1, 3, 8, 8), None)
    narrow = curl_b200.GCURLNet(backbone=nn.Sequential(nn.AdaptiveAvgPool2d(1), nn.Flatten(), nn.Linear(3, 100)))
    with pytest.raises(ValueError, match="backbone must return"):
        narrow(torch.rand(1, 3, 8, 8), None)
    try:
        import timm  # noqa: F401
    except ImportError:
        with pytest.raises(ImportError, match="backbone="):
            curl_b200.GCURLNet()


def test_trispace_api():
    """TriSpaceRegNet / ChannelPolyLayer (reference model.py:206-333, 411-520): names, shapes, order."""
    import torch.nn as nn
    net = curl_b200.TriSpaceRegNet(spatial=True, backbone=nn.Identity())
    sd = net.state_dict()
    assert {"x", "y", "polylayer.powers", "rgb2lab.rgb_to_xyz", "hsv2rgb" not in sd and "rgb2hsv.comparison_zero"} <= set(sd)
    assert sd["polylayer.powers"].shape == (126, 5) and sd["x"].shape == (1, 1, 1, 10000) and sd["y"].shape == (1, 1, 10000, 1)
    assert (net.num_coeffs, net.num_in, net.num_spaces, net.order) == (126, 5, 3, 4)
    flat = curl_b200.TriSpaceRegNet(spatial=False, backbone=nn.Identity())
    assert flat.num_coeffs == 35 and flat.state_dict()["x"].shape == (1, 0, 1, 10000)
    assert [tuple(int(v) for v in r) for r in flat.polylayer.powers.tolist()] == O.poly_powers(4, 3)
    assert curl_b200.ChannelPolyLayer(3, 3).num_coeffs == 20
    with pytest.raises(RuntimeError, match="CUDA"):
        flat.generate_residual(torch.rand(1, 3, 4, 4), *[torch.zeros(1, 3, 35)] * 3)
    x = torch.rand(1, 3, 4, 4)
    assert torch.equal(flat.generate_image(x, x), torch.clamp(x + x, 0, 1))
    # infer.py:22-23 / coreml_conversion.py:19-20 build the net around model.Deg4MobilePolyLayer()
    mob = curl_b200.model.Deg4MobilePolyLayer()
    assert mob.num_coeffs == 126 and list(mob.state_dict()) == ["powers"]
    assert torch.equal(mob.powers, net.polylayer.powers)
    inf = curl_b200.model.TriSpaceRegNet(polynomial_order=4, spatial=True, is_train=False, polylayer=mob,
                                         backbone=nn.Identity())
    assert inf.polylayer is mob and inf.num_coeffs == 126
    assert inf.load_state_dict(net.state_dict(), strict=True).missing_keys == []      # checkpoints are interchangeable
    with pytest.raises(ValueError, match="coefficients"):
        curl_b200.TriSpaceRegNet(spatial=False, polylayer=mob, backbone=nn.Identity())


@pytest.mark.parametrize("degree,nv,nout", [(3, 3, 3), (4, 5, 3), (2, 2, None), (0, 3, 3), (1, 4, 2)])
def test_channel_poly_layer_standalone_forward(degree, nv, nout):
    """ChannelPolyLayer.forward on its own (model.py:296-333; only the reference's PolyRegNet uses it): value and
    autograd gradients against the reference's formula (oracle.channel_poly, pinned bit-identical to model.py's)."""
    gen = torch.Generator().manual_seed(100 * degree + nv)
    layer = curl_b200.ChannelPolyLayer(degree, nv, nout)
    img = torch.rand(2, nv, 6, 7, generator=gen, dtype=torch.float64).requires_grad_(True)
    co = torch.randn(2, layer.num_out, layer.num_coeffs, generator=gen, dtype=torch.float64).requires_grad_(True)
    out = layer(img, co)
    ref = O.channel_poly(img, co, degree)
    assert out.shape == ref.shape == (2, layer.num_out, 6, 7)
    assert torch.allclose(out, ref, rtol=1e-12, atol=1e-12)
    gout = torch.randn(out.shape, generator=gen, dtype=torch.float64)
    g = torch.autograd.grad((out * gout).sum(), (img, co))
    gr = torch.autograd.grad((ref * gout).sum(), (img, co))
    for a, b in zip(g, gr):
        assert torch.allclose(a, b, rtol=1e-10, atol=1e-12)
    with pytest.raises(ValueError, match="variable per channel"):
        layer(torch.rand(2, nv + 1, 6, 7), co)
    with pytest.raises(ValueError, match="coeffs"):
        layer(img, co[:, :, :-1] if layer.num_coeffs > 1 else co[:, :, :0])
    if (degree, nv) == (4, 5):
        mob = curl_b200.Deg4MobilePolyLayer()
        assert torch.equal(mob(img, co), out)


def test_curlloss_api():
    """CURLLoss / MSSSIMMetric (reference model.py:36-118, metric.py:75-214): constructor, attributes,
    state-dict keys; MS-SSIM (the library part of the loss) equals the oracle's on CPU; the fused
    colour terms have no CPU path."""
    crit = curl_b200.CURLLoss()
    assert (crit.ssim_window_size, crit.num_channel) == (5, 1)
    assert list(crit.state_dict()) == ["msssim_layer.msssim_weights", "rgb2lab.rgb_to_xyz", "rgb2lab.fxfyfz_to_lab",
                                       "rgb2lab.xyz_to_rgb_mult", "rgb2lab.lab_to_fxfyfz_offset",
                                       "rgb2hsv.comparison_zero"]
    ms = crit.msssim_layer
    assert isinstance(ms, curl_b200.metric.MSSSIMMetric)
    assert (ms.window_size, ms.num_channel, ms.levels) == (11, 1, 5) and ms.gaussian_window.shape == (1, 1, 11, 11)
    assert torch.equal(ms.gaussian_window, O.ssim_window())
    gen = torch.Generator().manual_seed(2)
    a = torch.rand(2, 1, 48, 40, generator=gen)
    b = (a + 0.1 * torch.randn(2, 1, 48, 40, generator=gen)).clamp(0, 1)
    assert torch.allclose(ms(a, b), O.msssim(a, b), rtol=0, atol=2e-6)
    a64 = a.double().requires_grad_(True)
    ga, = torch.autograd.grad(ms.double()(a64, b.double()).sum(), a64)
    a64o = a.double().requires_grad_(True)
    go, = torch.autograd.grad(O.msssim(a64o, b.double()).sum(), a64o)
    assert torch.allclose(ga, go, rtol=1e-9, atol=1e-12)
    with pytest.raises(RuntimeError, match="four dimensions"):
        ms(a[0], b[0])
    with pytest.raises(RuntimeError, match="CUDA"):
        crit(torch.rand(1, 3, 32, 32), torch.rand(1, 3, 32, 32), torch.ones(1, 1, 32, 32, dtype=torch.bool))
    # a mask that is not one row per image is refused before anything runs (model.py:90 sums the mask as passed)
    for bad in (torch.ones(1, 1, 32, 32, dtype=torch.bool), torch.ones(2, 3, 32, 32)):
        with pytest.raises(ValueError, match="one row per image"):
            crit(torch.rand(2, 3, 32, 32), torch.rand(2, 3, 32, 32), bad)
    assert torch.equal(crit.batch_L_ssim_convert(torch.arange(24.).reshape(1, 3, 2, 4)), torch.arange(8.).reshape(1, 1, 2, 4))
