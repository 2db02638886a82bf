"""Out-of-bounds and coverage checks of every output-writing C-ABI entry point (GPU).

compute-sanitizer is not available on the GPU pool, so the writes are checked directly: every output
and workspace buffer handed to the library is carved out of a larger allocation filled with a
sentinel bit pattern, with guard bands on both sides.  After the call (a) both guard bands must be
untouched -- nothing written outside [ptr, ptr + size) -- and (b) no sentinel may be left inside an
output -- every element the contract promises was written.  Shapes cover the vector/TMA paths, the
scalar fallbacks (H*W % 4 != 0) and base pointers that are only 4-byte aligned.
"""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu

GUARD = 2048                       # elements on either side
SENT32 = 0x7FC5A5A5                # a NaN payload no kernel produces
SENT8 = 0xA5

SHAPES = [(2, 64, 96), (2, 33, 37), (1, 128, 256), (3, 17, 19)]


@pytest.fixture(scope="module")
def lib():
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    from curl_b200 import _lib
    return _lib.load()


class Arena:
    def __init__(self, dev):
        self.dev = dev
        self.items = []

    def f32(self, *shape, misalign=0, full=True):
        n = int(math.prod(shape))
        buf = torch.full((2 * GUARD + n + 4,), SENT32, dtype=torch.int32, device=self.dev)
        off = GUARD + misalign
        self.items.append((buf, off, n, SENT32, full))
        return buf[off:off + n].view(torch.float32).view(*shape)

    def u8(self, *shape, misalign=0):
        n = int(math.prod(shape))
        buf = torch.full((2 * GUARD + n + 16,), SENT8, dtype=torch.uint8, device=self.dev)
        off = GUARD + misalign
        self.items.append((buf, off, n, SENT8, False))       # 0xA5 is a legal pixel value: guards only
        return buf[off:off + n].view(*shape)

    def check(self, what):
        torch.cuda.synchronize(self.dev)
        for k, (buf, off, n, sent, full) in enumerate(self.items):
            assert bool((buf[:off] == sent).all()), "%s: buffer %d written BEFORE its start" % (what, k)
            assert bool((buf[off + n:] == sent).all()), "%s: buffer %d written PAST its end" % (what, k)
            if full:
                left = int((buf[off:off + n] == sent).sum())
                assert left == 0, "%s: buffer %d has %d of %d elements never written" % (what, k, left, n)
        self.items = []


def _inputs(B, H, W, dev, seed=0):
    g = torch.Generator(device=dev).manual_seed(100 + seed)
    img = torch.rand(B, 3, H, W, generator=g, device=dev)
    gout = torch.randn(B, 3, H, W, generator=g, device=dev)
    mask = torch.rand(B, 1, H, W, generator=g, device=dev) > 0.2
    L, R, Hk = (0.1 * torch.randn(B, n, generator=g, device=dev) for n in (48, 48, 64))
    return img, gout, mask, L, R, Hk, torch.ones(B, device=dev)


def _masks(mask):
    return ((None, 0), (mask.contiguous(), 1), (mask.float().contiguous(), 2))


def _p(t):
    return None if t is None else t.data_ptr()


def _st(dev):
    return torch.cuda.current_stream(dev).cuda_stream


@pytest.mark.parametrize("shape", SHAPES)
@pytest.mark.parametrize("misalign", [0, 1])
@pytest.mark.parametrize("seg", [False, True])
def test_layer_writes_stay_inside(lib, shape, misalign, seg):
    B, H, W = shape
    dev = torch.device("cuda", 0)
    img, gout, mask, L, R, Hk, greg = _inputs(B, H, W, dev)
    fwd = lib.curl_layer_seg_fwd_f32 if seg else lib.curl_layer_fwd_f32
    bwd = lib.curl_layer_seg_bwd_f32 if seg else lib.curl_layer_bwd_f32
    nws = lib.curl_layer_seg_bwd_workspace_bytes(B, H, W, 16, 16, 16) if seg else lib.curl_bwd_workspace_bytes(B, H, W)
    for m, kind in _masks(mask):
        a = Arena(dev)
        out, reg = a.f32(B, 3, H, W, misalign=misalign), a.f32(B)
        rc = fwd(_p(img), _p(m), kind, _p(L), 48, _p(R), 48, _p(Hk), 64, _p(out), _p(reg), B, H, W, 16, 16, 16, _st(dev))
        assert rc == 0
        a.check("layer fwd seg=%s mask=%d" % (seg, kind))
        for want_gimg in (True, False):
            gimg = a.f32(B, 3, H, W, misalign=misalign) if want_gimg else None
            gL, gR, gH = a.f32(B, 48), a.f32(B, 48), a.f32(B, 64)
            ws = a.u8(max(nws, 1))
            rc = bwd(_p(img), _p(m), kind, _p(L), 48, _p(R), 48, _p(Hk), 64, _p(gout), _p(greg), _p(gimg), _p(gL),
                     _p(gR), _p(gH), _p(ws), nws, B, H, W, 16, 16, 16, _st(dev))
            assert rc == 0
            a.check("layer bwd seg=%s mask=%d gimg=%s" % (seg, kind, want_gimg))


@pytest.mark.parametrize("shape", SHAPES)
def test_u8_paths_write_inside(lib, shape):
    B, H, W = shape
    dev = torch.device("cuda", 0)
    img, _, mask, L, R, Hk, _ = _inputs(B, H, W, dev, 1)
    u8 = (img * 255).to(torch.uint8)
    hwc = u8.permute(0, 2, 3, 1).contiguous()
    for src, lin in ((u8, 0), (hwc, 1)):
        for lout in (0, 1):
            a = Arena(dev)
            out, reg = a.u8(B * 3 * H * W), a.f32(B)
            rc = lib.curl_layer_fwd_u8(_p(src), lin, _p(mask), 1, _p(L), 48, _p(R), 48, _p(Hk), 64, _p(out), lout, _p(reg),
                                       B, H, W, 16, 16, 16, _st(dev))
            assert rc == 0
            a.check("fwd_u8 in=%d out=%d" % (lin, lout))
        for mis in (0, 1):
            a = Arena(dev)
            f = a.f32(B, 3, H, W, misalign=mis)
            assert lib.curl_u8_to_unit_f32(_p(src), lin, _p(f), B, H, W, _st(dev)) == 0
            a.check("u8_to_unit layout=%d misalign=%d" % (lin, mis))


@pytest.mark.parametrize("shape", SHAPES)
@pytest.mark.parametrize("misalign", [0, 1])
def test_colour_and_curve_ops_write_inside(lib, shape, misalign):
    B, H, W = shape
    dev = torch.device("cuda", 0)
    img, gout, _, L, _, Hk, greg = _inputs(B, H, W, dev, 2)
    nws = lib.curl_bwd_workspace_bytes(B, H, W)
    for conv in range(4):
        a = Arena(dev)
        out = a.f32(B, 3, H, W, misalign=misalign)
        assert lib.curl_color_fwd_f32(conv, _p(img), _p(out), B, H, W, _st(dev)) == 0
        a.check("color fwd %d" % conv)
        gin = a.f32(B, 3, H, W, misalign=misalign)
        assert lib.curl_color_bwd_f32(conv, _p(img), _p(gout), _p(gin), B, H, W, _st(dev)) == 0
        a.check("color bwd %d" % conv)
    for group, P, nc in ((0, L, 3), (1, L, 3), (2, Hk, 4)):
        a = Arena(dev)
        out, reg = a.f32(B, 3, H, W, misalign=misalign), a.f32(B)
        assert lib.curl_adjust_fwd_f32(group, _p(img), _p(P), P.shape[1], _p(out), _p(reg), B, H, W, 16, _st(dev)) == 0
        a.check("adjust fwd %d" % group)
        for want_gin in (True, False):
            gin = a.f32(B, 3, H, W, misalign=misalign) if want_gin else None
            gP, ws = a.f32(B, nc * 16), a.u8(max(nws, 1))
            rc = lib.curl_adjust_bwd_f32(group, _p(img), _p(P), P.shape[1], _p(gout), _p(greg), _p(gin), _p(gP), _p(ws), nws,
                                         B, H, W, 16, _st(dev))
            assert rc == 0
            a.check("adjust bwd %d gin=%s" % (group, want_gin))
    C = torch.exp(L[:, :16]).contiguous()
    for cin, cout in ((0, 0), (0, 1), (2, 2)):
        a = Arena(dev)
        out, reg = a.f32(B, 3, H, W, misalign=misalign), a.f32(B)
        assert lib.curl_apply_curve_fwd_f32(_p(img), _p(C), 16, cin, cout, _p(out), _p(reg), B, H, W, 16, _st(dev)) == 0
        a.check("apply_curve fwd %d->%d" % (cin, cout))
        gin, gC, ws = a.f32(B, 3, H, W, misalign=misalign), a.f32(B, 16), a.u8(max(nws, 1))
        rc = lib.curl_apply_curve_bwd_f32(_p(img), _p(C), 16, cin, cout, _p(gout), _p(greg), _p(gin), _p(gC), _p(ws), nws,
                                          B, H, W, 16, _st(dev))
        assert rc == 0
        a.check("apply_curve bwd %d->%d" % (cin, cout))


@pytest.mark.parametrize("shape", SHAPES)
@pytest.mark.parametrize("spatial", [0, 1])
def test_trispace_writes_stay_inside(lib, shape, spatial):
    B, H, W = shape
    dev = torch.device("cuda", 0)
    img, gout, *_ = _inputs(B, H, W, dev, 3)
    nc = lib.curl_trispace_num_coeffs(4, spatial)
    g = torch.Generator(device=dev).manual_seed(9)
    R, L, Hc = (0.3 * torch.randn(B, 3, nc, generator=g, device=dev) for _ in range(3))
    nws = lib.curl_trispace_bwd_workspace_bytes(B, H, W, 4, spatial)
    for clamp in (0, 1):
        for save in (True, False):
            a = Arena(dev)
            out = a.f32(B, 3, H, W)
            saved = a.f32(B, 9, H, W) if save else None
            rc = lib.curl_trispace_fwd_f32(_p(img), _p(R), _p(L), _p(Hc), 4, spatial, clamp, _p(out), _p(saved), B, H, W, _st(dev))
            assert rc == 0
            keep = a.items[1][0] if save else None            # the saved activations feed the backward
            a.check("trispace fwd spatial=%d clamp=%d save=%s" % (spatial, clamp, save))
            gR, gL, gH, ws = a.f32(B, 3, nc), a.f32(B, 3, nc), a.f32(B, 3, nc), a.u8(max(nws, 1))
            rc = lib.curl_trispace_bwd_f32(_p(img), _p(R), _p(L), _p(Hc), 4, spatial, clamp, _p(gout), _p(saved),
                                           _p(gR), _p(gL), _p(gH), _p(ws), nws, B, H, W, _st(dev))
            assert rc == 0
            a.check("trispace bwd spatial=%d clamp=%d save=%s" % (spatial, clamp, save))
            gimg = a.f32(B, 3, H, W)
            rc = lib.curl_trispace_bwd_img_f32(_p(img), _p(R), _p(L), _p(Hc), 4, spatial, clamp, _p(gout), _p(saved), _p(gimg),
                                               B, H, W, _st(dev))
            assert rc == 0
            a.check("trispace bwd_img spatial=%d clamp=%d save=%s" % (spatial, clamp, save))
            del keep


@pytest.mark.parametrize("shape", SHAPES)
def test_loss_terms_write_inside(lib, shape):
    B, H, W = shape
    dev = torch.device("cuda", 0)
    pred, gLp3, mask, *_ = _inputs(B, H, W, dev, 4)
    target = (pred * 0.8 + 0.1).contiguous()
    gLp = gLp3[:, :1].contiguous()
    w = torch.full((B, 4), 0.01, device=dev)
    nws = lib.curl_loss_terms_workspace_bytes(B, H, W)
    for m, kind in _masks(mask):
        a = Arena(dev)
        sums, Lp, Lt, ws = a.f32(B, 5), a.f32(B, 1, H, W), a.f32(B, 1, H, W), a.u8(max(nws, 1))
        rc = lib.curl_loss_terms_fwd_f32(_p(pred), _p(target), _p(m), kind, _p(sums), _p(Lp), _p(Lt), _p(ws), nws, B, H, W, _st(dev))
        assert rc == 0
        a.check("loss terms fwd mask=%d" % kind)
        for mis in (0, 1):
            gp = a.f32(B, 3, H, W, misalign=mis)
            rc = lib.curl_loss_terms_bwd_f32(_p(pred), _p(target), _p(m), kind, _p(w), _p(gLp), _p(gp), B, H, W, _st(dev))
            assert rc == 0
            a.check("loss terms bwd mask=%d misalign=%d" % (kind, mis))
