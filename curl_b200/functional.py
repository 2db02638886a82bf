"""torch.autograd.Function wrappers over the C ABI (include/curl_b200.h).

Call chain: module/function API (colors.py, curves.py, model.py) -> *Fn.apply -> ctypes ->
libcurl_b200.so -> sm_100a kernels on torch's current stream.  Backward recomputes the forward
inside the kernel, so the only tensors saved for backward are the op's inputs.
"""
import torch

from . import _lib


def _stream(dev):
    return torch.cuda.current_stream(dev).cuda_stream


def _ptr(t):
    return None if t is None else t.data_ptr()


def _image(x, name):
    """Validates an NCHW fp32 CUDA image and returns it contiguous (the reference calls
    .contiguous() itself: colors.py:35,95,208; curves.py:50,99,146)."""
    if not isinstance(x, torch.Tensor):
        raise TypeError("%s must be a torch.Tensor" % name)
    if not x.is_cuda:
        raise RuntimeError("curl_b200: %s must be a CUDA tensor (this path has no CPU implementation)" % name)
    if x.dtype != torch.float32:
        raise TypeError("curl_b200: %s must be float32, got %s" % (name, x.dtype))
    if x.dim() != 4 or x.shape[1] != 3:
        raise ValueError("curl_b200: %s must have shape [B, 3, H, W], got %s" % (name, tuple(x.shape)))
    if x.numel() == 0:
        raise ValueError("curl_b200: %s is empty" % name)
    return x.contiguous()


def _knots(P, n, name, dev):
    """Validates a [B, >= n] fp32 knot tensor; returns (tensor, row_stride) with unit column stride."""
    if not isinstance(P, torch.Tensor) or P.dim() != 2:
        raise ValueError("curl_b200: %s must be a 2-D tensor [B, >= %d]" % (name, n))
    if P.device != dev:
        raise RuntimeError("curl_b200: %s is on %s, image is on %s" % (name, P.device, dev))
    if P.dtype != torch.float32:
        raise TypeError("curl_b200: %s must be float32, got %s" % (name, P.dtype))
    if P.shape[1] < n:
        raise ValueError("curl_b200: %s has %d columns, needs at least %d" % (name, P.shape[1], n))
    if P.stride(1) != 1 or (P.shape[0] > 1 and P.stride(0) < n):
        P = P.contiguous()
    return P, (P.stride(0) if P.shape[0] > 1 else max(P.stride(0), n))


def _mask(mask, img):
    """Returns (mask_tensor_or_None, mask_kind).  torch.bool -> 1 byte per pixel (zero / non-zero); every
    other dtype -- torch.uint8 included: the reference MULTIPLIES by the mask's value, so a 0/255 byte mask
    scales by 255 there -- is converted to fp32 and multiplied in at every stage like the reference
    (model.py:154,160,166,170)."""
    if mask is None:
        return None, _lib.MASK_NONE
    B, _, H, W = img.shape
    if not isinstance(mask, torch.Tensor):
        raise TypeError("curl_b200: mask must be a torch.Tensor or None")
    if mask.device != img.device:
        raise RuntimeError("curl_b200: mask is on %s, image is on %s" % (mask.device, img.device))
    if mask.requires_grad:
        raise RuntimeError("curl_b200: gradients w.r.t. the mask are not supported")
    if mask.dim() == 3:
        mask = mask.unsqueeze(1)
    if mask.dim() != 4 or mask.shape[1] != 1 or mask.shape[2:] != (H, W) or mask.shape[0] not in (1, B):
        raise ValueError("curl_b200: mask must have shape [B, 1, H, W], got %s" % (tuple(mask.shape),))
    if mask.shape[0] != B:
        mask = mask.expand(B, 1, H, W)
    if mask.dtype == torch.bool:
        return mask.contiguous(), _lib.MASK_U8
    return mask.to(torch.float32).contiguous(), _lib.MASK_F32


def _workspace(B, H, W, dev):
    n = _lib.load().curl_bwd_workspace_bytes(B, H, W)
    return torch.empty(n, dtype=torch.uint8, device=dev), n


class CurlLayerFn(torch.autograd.Function):
    """Fused CURLLayer.forward (reference model.py:137-176) and its recompute backward.
    segment_clamp=True selects the opt-in curves of the CURL paper's eq. (1) (curl_layer_seg_*)."""

    @staticmethod
    def forward(ctx, img, mask, L, R, H, k_lab, k_rgb, k_hsv, segment_clamp=False):
        lib = _lib.load()
        img = _image(img, "img")
        dev = img.device
        L, ldL = _knots(L, 3 * k_lab, "L", dev)
        R, ldR = _knots(R, 3 * k_rgb, "R", dev)
        H, ldH = _knots(H, 4 * k_hsv, "H", dev)
        B, _, h, w = img.shape
        for P, nm in ((L, "L"), (R, "R"), (H, "H")):
            if P.shape[0] != B:
                raise ValueError("curl_b200: %s has batch %d, image has %d" % (nm, P.shape[0], B))
        mask, kind = _mask(mask, img)
        out = torch.empty_like(img)
        reg = torch.empty(B, dtype=torch.float32, device=dev)
        fwd = lib.curl_layer_seg_fwd_f32 if segment_clamp else lib.curl_layer_fwd_f32
        with torch.cuda.device(dev):
            rc = fwd(_ptr(img), _ptr(mask), kind, _ptr(L), ldL, _ptr(R), ldR, _ptr(H), ldH,
                     _ptr(out), _ptr(reg), B, h, w, k_lab, k_rgb, k_hsv, _stream(dev))
        _lib.check(rc, "curl_layer_seg_fwd_f32" if segment_clamp else "curl_layer_fwd_f32")
        _lib.count_launch()
        ctx.save_for_backward(img, mask, L, R, H)
        ctx.cfg = (kind, ldL, ldR, ldH, k_lab, k_rgb, k_hsv)
        ctx.segment_clamp = bool(segment_clamp)
        return out, reg

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, gout, greg):
        lib = _lib.load()
        img, mask, L, R, H = ctx.saved_tensors
        kind, ldL, ldR, ldH, k_lab, k_rgb, k_hsv = ctx.cfg
        dev = img.device
        B, _, h, w = img.shape
        gout = gout.contiguous() if gout is not None else torch.zeros_like(img)
        greg = greg.contiguous() if greg is not None else None
        want_gimg = ctx.needs_input_grad[0]
        gimg = torch.empty_like(img) if want_gimg else None
        gL = torch.empty(B, 3 * k_lab, dtype=torch.float32, device=dev)
        gR = torch.empty(B, 3 * k_rgb, dtype=torch.float32, device=dev)
        gH = torch.empty(B, 4 * k_hsv, dtype=torch.float32, device=dev)
        if ctx.segment_clamp:
            ws_bytes = lib.curl_layer_seg_bwd_workspace_bytes(B, h, w, k_lab, k_rgb, k_hsv)
            ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
            bwd, what = lib.curl_layer_seg_bwd_f32, "curl_layer_seg_bwd_f32"
        else:
            ws, ws_bytes = _workspace(B, h, w, dev)
            bwd, what = lib.curl_layer_bwd_f32, "curl_layer_bwd_f32"
        with torch.cuda.device(dev):
            rc = bwd(_ptr(img), _ptr(mask), kind, _ptr(L), ldL, _ptr(R), ldR, _ptr(H), ldH,
                     _ptr(gout), _ptr(greg), _ptr(gimg), _ptr(gL), _ptr(gR), _ptr(gH),
                     _ptr(ws), ws_bytes, B, h, w, k_lab, k_rgb, k_hsv, _stream(dev))
        _lib.check(rc, what)
        _lib.count_launch()

        def widen(g, P):
            # columns beyond num_*_points were sliced away by the layer: zero gradient
            if g.shape[1] == P.shape[1]:
                return g
            full = torch.zeros_like(P, memory_format=torch.contiguous_format)
            full[:, :g.shape[1]] = g
            return full

        return gimg, None, widen(gL, L), widen(gR, R), widen(gH, H), None, None, None, None


class ColorFn(torch.autograd.Function):
    """Stand-alone colour conversion (reference colors.py modules)."""

    @staticmethod
    def forward(ctx, img, conversion):
        lib = _lib.load()
        img = _image(img, "img")
        B, _, h, w = img.shape
        out = torch.empty_like(img)
        with torch.cuda.device(img.device):
            rc = lib.curl_color_fwd_f32(conversion, _ptr(img), _ptr(out), B, h, w, _stream(img.device))
        _lib.check(rc, "curl_color_fwd_f32")
        _lib.count_launch()
        ctx.save_for_backward(img)
        ctx.conversion = conversion
        return out

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, gout):
        lib = _lib.load()
        (img,) = ctx.saved_tensors
        B, _, h, w = img.shape
        gout = gout.contiguous()
        gin = torch.empty_like(img)
        with torch.cuda.device(img.device):
            rc = lib.curl_color_bwd_f32(ctx.conversion, _ptr(img), _ptr(gout), _ptr(gin), B, h, w, _stream(img.device))
        _lib.check(rc, "curl_color_bwd_f32")
        _lib.count_launch()
        return gin, None


class AdjustFn(torch.autograd.Function):
    """Stand-alone curve group (reference curves.adjust_lab / adjust_rgb / adjust_hsv)."""

    @staticmethod
    def forward(ctx, img, P, group):
        lib = _lib.load()
        img = _image(img, "img")
        nc = 4 if group == _lib.ADJUST_HSV else 3
        if P.dim() != 2 or P.shape[1] < 3 * nc:
            raise ValueError("curl_b200: curve parameters must be [B, %d*K] with K >= 3" % nc)
        K = -(-P.shape[1] // nc)          # torch.chunk(P, nc, dim=1) chunk size (curves.py:53,105,152)
        if K * nc != P.shape[1]:
            raise ValueError("curl_b200: parameter width %d is not a multiple of %d curves" % (P.shape[1], nc))
        P, ldP = _knots(P, nc * K, "P", img.device)
        B, _, h, w = img.shape
        if P.shape[0] != B:
            raise ValueError("curl_b200: P has batch %d, image has %d" % (P.shape[0], B))
        out = torch.empty_like(img)
        reg = torch.empty(B, dtype=torch.float32, device=img.device)
        with torch.cuda.device(img.device):
            rc = lib.curl_adjust_fwd_f32(group, _ptr(img), _ptr(P), ldP, _ptr(out), _ptr(reg), B, h, w, K,
                                         _stream(img.device))
        _lib.check(rc, "curl_adjust_fwd_f32")
        _lib.count_launch()
        ctx.save_for_backward(img, P)
        ctx.cfg = (group, ldP, K, nc)
        return out, reg

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, gout, greg):
        lib = _lib.load()
        img, P = ctx.saved_tensors
        group, ldP, K, nc = ctx.cfg
        dev = img.device
        B, _, h, w = img.shape
        gout = gout.contiguous() if gout is not None else torch.zeros_like(img)
        greg = greg.contiguous() if greg is not None else None
        gin = torch.empty_like(img) if ctx.needs_input_grad[0] else None
        gP = torch.empty(B, nc * K, dtype=torch.float32, device=dev)
        ws, ws_bytes = _workspace(B, h, w, dev)
        with torch.cuda.device(dev):
            rc = lib.curl_adjust_bwd_f32(group, _ptr(img), _ptr(P), ldP, _ptr(gout), _ptr(greg), _ptr(gin), _ptr(gP),
                                         _ptr(ws), ws_bytes, B, h, w, K, _stream(dev))
        _lib.check(rc, "curl_adjust_bwd_f32")
        _lib.count_launch()
        return gin, gP, None


class ApplyCurveFn(torch.autograd.Function):
    """One curve with knots C given directly (reference curves.apply_curve, curves.py:4-38)."""

    @staticmethod
    def forward(ctx, img, C, channel_in, channel_out):
        lib = _lib.load()
        img = _image(img, "img")
        if C.dim() != 2 or C.shape[1] < 3:
            raise ValueError("curl_b200: C must be [B, K] with K >= 3")
        K = C.shape[1]
        C, ldC = _knots(C, K, "C", img.device)
        B, _, h, w = img.shape
        if C.shape[0] != B:
            raise ValueError("curl_b200: C has batch %d, image has %d" % (C.shape[0], B))
        out = torch.empty_like(img)
        reg = torch.empty(B, dtype=torch.float32, device=img.device)
        with torch.cuda.device(img.device):
            rc = lib.curl_apply_curve_fwd_f32(_ptr(img), _ptr(C), ldC, channel_in, channel_out, _ptr(out), _ptr(reg),
                                              B, h, w, K, _stream(img.device))
        _lib.check(rc, "curl_apply_curve_fwd_f32")
        _lib.count_launch()
        ctx.save_for_backward(img, C)
        ctx.cfg = (ldC, K, channel_in, channel_out)
        return out, reg

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, gout, greg):
        lib = _lib.load()
        img, C = ctx.saved_tensors
        ldC, K, cin, cout = ctx.cfg
        dev = img.device
        B, _, h, w = img.shape
        gout = gout.contiguous() if gout is not None else torch.zeros_like(img)
        greg = greg.contiguous() if greg is not None else None
        gin = torch.empty_like(img) if ctx.needs_input_grad[0] else None
        gC = torch.empty(B, K, dtype=torch.float32, device=dev)
        ws, ws_bytes = _workspace(B, h, w, dev)
        with torch.cuda.device(dev):
            rc = lib.curl_apply_curve_bwd_f32(_ptr(img), _ptr(C), ldC, cin, cout, _ptr(gout), _ptr(greg), _ptr(gin),
                                              _ptr(gC), _ptr(ws), ws_bytes, B, h, w, K, _stream(dev))
        _lib.check(rc, "curl_apply_curve_bwd_f32")
        _lib.count_launch()
        return gin, gC, None, None


def curl_layer_u8(img, mask, L, R, H, k_lab, k_rgb, k_hsv, layout="CHW", out_layout=None):
    """Inference on 8-bit images (curl_layer_fwd_u8): img is uint8 [B,3,H,W] (layout "CHW", what
    TF.to_tensor starts from after its permute) or [B,H,W,3] ("HWC", what PIL / plt.imsave hold);
    returns (uint8 image in `out_layout` (default: same as the input), reg[B]).  x = u8/255 on the way
    in (data.py:152-158), trunc(out*255) on the way out (evaluate.py:63).  No autograd."""
    lib = _lib.load()
    layouts = {"CHW": _lib.LAYOUT_CHW, "HWC": _lib.LAYOUT_HWC}
    if layout not in layouts or (out_layout is not None and out_layout not in layouts):
        raise ValueError("curl_b200: layout must be 'CHW' or 'HWC'")
    out_layout = layout if out_layout is None else out_layout
    if not isinstance(img, torch.Tensor) or not img.is_cuda:
        raise RuntimeError("curl_b200: img must be a CUDA tensor (this path has no CPU implementation)")
    if img.dtype != torch.uint8:
        raise TypeError("curl_b200: forward_u8 takes a uint8 image, got %s" % img.dtype)
    cdim = 1 if layout == "CHW" else 3
    if img.dim() != 4 or img.shape[cdim] != 3 or img.numel() == 0:
        raise ValueError("curl_b200: img must have shape %s, got %s"
                         % ("[B,3,H,W]" if layout == "CHW" else "[B,H,W,3]", tuple(img.shape)))
    img = img.contiguous()
    dev = img.device
    B = img.shape[0]
    h, w = (img.shape[2], img.shape[3]) if layout == "CHW" else (img.shape[1], img.shape[2])
    L, ldL = _knots(L, 3 * k_lab, "L", dev)
    R, ldR = _knots(R, 3 * k_rgb, "R", dev)
    H, ldH = _knots(H, 4 * k_hsv, "H", dev)
    for P, nm in ((L, "L"), (R, "R"), (H, "H")):
        if P.shape[0] != B:
            raise ValueError("curl_b200: %s has batch %d, image has %d" % (nm, P.shape[0], B))
    if any(t.requires_grad for t in (L, R, H)) and torch.is_grad_enabled():
        raise RuntimeError("curl_b200: forward_u8 is inference only; call it under torch.no_grad()")
    shape_like = torch.empty(B, 3, h, w, device="meta")
    mask, kind = _mask(mask, _MetaShape(shape_like, dev))
    out = torch.empty((B, 3, h, w) if out_layout == "CHW" else (B, h, w, 3), dtype=torch.uint8, device=dev)
    reg = torch.empty(B, dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        rc = lib.curl_layer_fwd_u8(_ptr(img), layouts[layout], _ptr(mask), kind, _ptr(L), ldL, _ptr(R), ldR,
                                   _ptr(H), ldH, _ptr(out), layouts[out_layout], _ptr(reg), B, h, w,
                                   k_lab, k_rgb, k_hsv, _stream(dev))
    _lib.check(rc, "curl_layer_fwd_u8")
    _lib.count_launch()
    return out, reg


class _MetaShape:
    """What _mask() needs to know about the image: its NCHW shape and device."""

    def __init__(self, t, dev):
        self.shape = t.shape
        self.device = dev


class TriSpaceFn(torch.autograd.Function):
    """Fused TriSpaceRegNet.generate_residual (+ generate_image) (reference model.py:486-508) and its
    recompute backward: coefficient gradients (curl_trispace_bwd_f32) and, only when the image requires one, the
    image gradient (curl_trispace_bwd_img_f32 -- the reference's scripts feed data there, model.py:517-519, but
    its model.py:486-508 is differentiable in the image and autograd would provide it)."""

    @staticmethod
    def forward(ctx, img, R, L, H, degree, spatial, final_clamp, save_activations=True):
        lib = _lib.load()
        img = _image(img, "img")
        dev = img.device
        B, _, h, w = img.shape
        nc = lib.curl_trispace_num_coeffs(degree, int(spatial))
        coefs = []
        for P, nm in ((R, "R"), (L, "L"), (H, "H")):
            if not isinstance(P, torch.Tensor) or P.device != dev or P.dtype != torch.float32:
                raise TypeError("curl_b200: %s must be a float32 tensor on %s" % (nm, dev))
            if tuple(P.shape) != (B, 3, nc):
                raise ValueError("curl_b200: %s must have shape [%d, 3, %d] (degree %d, spatial=%s), got %s"
                                 % (nm, B, nc, degree, bool(spatial), tuple(P.shape)))
            coefs.append(P.contiguous())
        out = torch.empty_like(img)
        # the nine sigmoid outputs (36 B/px): with them the backward evaluates no polynomial
        saved = None
        if save_activations and any(ctx.needs_input_grad[0:4]):
            saved = torch.empty(B, 9, h, w, dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            rc = lib.curl_trispace_fwd_f32(_ptr(img), _ptr(coefs[0]), _ptr(coefs[1]), _ptr(coefs[2]), degree,
                                           int(spatial), int(final_clamp), _ptr(out), _ptr(saved), B, h, w, _stream(dev))
        _lib.check(rc, "curl_trispace_fwd_f32")
        _lib.count_launch()
        # `saved` goes through save_for_backward like the inputs: autograd frees it with the graph, a second
        # backward under retain_graph=True sees the same tensor, and saved-tensor hooks (checkpointing,
        # CPU offload) and the in-place/version checks apply to it
        ctx.save_for_backward(img, *coefs, saved)
        ctx.cfg = (degree, int(spatial), int(final_clamp))
        return out

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, gout):
        lib = _lib.load()
        img, R, L, H, saved = ctx.saved_tensors
        degree, spatial, final_clamp = ctx.cfg
        dev = img.device
        B, _, h, w = img.shape
        gout = gout.contiguous()
        gimg = gR = gL = gH = None
        with torch.cuda.device(dev):
            if any(ctx.needs_input_grad[1:4]):
                gR, gL, gH = torch.empty_like(R), torch.empty_like(L), torch.empty_like(H)
                n = lib.curl_trispace_bwd_workspace_bytes(B, h, w, degree, spatial)
                ws = torch.empty(n, dtype=torch.uint8, device=dev)
                rc = lib.curl_trispace_bwd_f32(_ptr(img), _ptr(R), _ptr(L), _ptr(H), degree, spatial, final_clamp,
                                               _ptr(gout), _ptr(saved), _ptr(gR), _ptr(gL), _ptr(gH), _ptr(ws), n, B, h, w,
                                               _stream(dev))
                _lib.check(rc, "curl_trispace_bwd_f32")
                _lib.count_launch(2)
            if ctx.needs_input_grad[0]:
                gimg = torch.empty_like(img)
                rc = lib.curl_trispace_bwd_img_f32(_ptr(img), _ptr(R), _ptr(L), _ptr(H), degree, spatial, final_clamp,
                                                   _ptr(gout), _ptr(saved), _ptr(gimg), B, h, w, _stream(dev))
                _lib.check(rc, "curl_trispace_bwd_img_f32")
                _lib.count_launch()
        return gimg, gR, gL, gH, None, None, None, None


class LossTermsFn(torch.autograd.Function):
    """Colour terms of CURLLoss in one pass (reference model.py:89-108): returns
    (sums[B, 5], L_pred[B, 1, H, W], L_target[B, 1, H, W]) -- see include/curl_b200.h for the five
    sums.  Differentiable w.r.t. pred only (the target is data; its L plane carries no gradient)."""

    @staticmethod
    def forward(ctx, pred, target, mask):
        lib = _lib.load()
        pred = _image(pred, "predicted_img_batch")
        target = _image(target, "target_img_batch")
        if target.shape != pred.shape or target.device != pred.device:
            raise ValueError("curl_b200: predicted and target batches differ: %s on %s vs %s on %s"
                             % (tuple(pred.shape), pred.device, tuple(target.shape), target.device))
        m, kind = _mask(mask, pred)
        dev = pred.device
        B, _, h, w = pred.shape
        sums = torch.empty(B, 5, dtype=torch.float32, device=dev)
        L_pred = torch.empty(B, 1, h, w, dtype=torch.float32, device=dev)
        L_target = torch.empty(B, 1, h, w, dtype=torch.float32, device=dev)
        n = lib.curl_loss_terms_workspace_bytes(B, h, w)
        ws = torch.empty(n, dtype=torch.uint8, device=dev)
        with torch.cuda.device(dev):
            rc = lib.curl_loss_terms_fwd_f32(_ptr(pred), _ptr(target), _ptr(m), kind, _ptr(sums), _ptr(L_pred),
                                             _ptr(L_target), _ptr(ws), n, B, h, w, _stream(dev))
        _lib.check(rc, "curl_loss_terms_fwd_f32")
        _lib.count_launch()
        ctx.save_for_backward(pred, target, m)
        ctx.kind = kind
        ctx.mark_non_differentiable(L_target)
        return sums, L_pred, L_target

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, gsums, gL_pred, _gL_target):
        if ctx.needs_input_grad[1]:
            raise RuntimeError("curl_b200: the loss terms have no gradient w.r.t. the target batch; detach it")
        lib = _lib.load()
        pred, target, m = ctx.saved_tensors
        dev = pred.device
        B, _, h, w = pred.shape
        wts = gsums[:, :4].to(torch.float32).contiguous()
        gL = None if gL_pred is None else gL_pred.to(torch.float32).contiguous()
        gpred = torch.empty_like(pred)
        with torch.cuda.device(dev):
            rc = lib.curl_loss_terms_bwd_f32(_ptr(pred), _ptr(target), _ptr(m), ctx.kind, _ptr(wts), _ptr(gL),
                                             _ptr(gpred), B, h, w, _stream(dev))
        _lib.check(rc, "curl_loss_terms_bwd_f32")
        _lib.count_launch()
        return gpred, None, None


def loss_terms(pred, target, mask):
    """(sums[B, 5], L_pred, L_target) of the fused CURLLoss colour terms; see LossTermsFn."""
    return LossTermsFn.apply(pred, target, mask)


def to_tensor_u8(img, layout="CHW"):
    """torchvision's TF.to_tensor for 8-bit images, on the device (curl_u8_to_unit_f32): uint8
    [B,3,H,W] ("CHW") or [B,H,W,3] ("HWC", what PIL holds) -> float32 [B,3,H,W] = u8 / 255, exactly the
    value the reference's loader produces (data.py:152-158).  The result is an ordinary tensor: feed it to
    CURLLayer / TriSpaceRegNet / CURLLoss for training, having moved 3 B/px over PCIe instead of 12."""
    lib = _lib.load()
    layouts = {"CHW": _lib.LAYOUT_CHW, "HWC": _lib.LAYOUT_HWC}
    if layout not in layouts:
        raise ValueError("curl_b200: layout must be 'CHW' or 'HWC'")
    if not isinstance(img, torch.Tensor) or not img.is_cuda:
        raise RuntimeError("curl_b200: img must be a CUDA tensor (this path has no CPU implementation)")
    if img.dtype != torch.uint8:
        raise TypeError("curl_b200: to_tensor_u8 takes a uint8 image, got %s" % img.dtype)
    cdim = 1 if layout == "CHW" else 3
    if img.dim() != 4 or img.shape[cdim] != 3 or img.numel() == 0:
        raise ValueError("curl_b200: img must have shape %s, got %s"
                         % ("[B,3,H,W]" if layout == "CHW" else "[B,H,W,3]", tuple(img.shape)))
    img = img.contiguous()
    B = img.shape[0]
    h, w = (img.shape[2], img.shape[3]) if layout == "CHW" else (img.shape[1], img.shape[2])
    out = torch.empty(B, 3, h, w, dtype=torch.float32, device=img.device)
    with torch.cuda.device(img.device):
        rc = lib.curl_u8_to_unit_f32(_ptr(img), layouts[layout], _ptr(out), B, h, w, _stream(img.device))
    _lib.check(rc, "curl_u8_to_unit_f32")
    _lib.count_launch()
    return out
