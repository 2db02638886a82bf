"""TEST INFRASTRUCTURE ONLY -- generates tests/golden/*.npz from the REAL reference.

Run in the build container (needs /root/reference):   python -m oracle.make_golden

Every fixture is produced by executing the reference's own code (oracle/ref_loader.py: colors.py
unmodified, curves.py / model.py::CURLLayer with the documented minimal patch) on CPU in fp32 and,
on the same fp32 inputs, in fp64.  Inputs are seeded and stored in the file, so the fixtures
regenerate bit-identically on the same torch build.  The GPU box has no reference tree; it checks
the CUDA path (and the oracle) against these files.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from oracle import ref_loader  # noqa: E402

GOLDEN_DIR = os.path.join(os.path.dirname(HERE), "tests", "golden")


def _edge_image():
    """1x3xHxW image made of every RGB triple from a list of branch-threshold / tie values."""
    vals = [0.0, 1e-9, 0.0031308, 0.04045, 0.2, 0.5, 0.5000001, 1.0]
    trip = torch.tensor([[a, b, c] for a in vals for b in vals for c in vals], dtype=torch.float32)  # 512 px
    return trip.t().reshape(1, 3, 16, 32).contiguous()


def layer_cases():
    g = torch.Generator().manual_seed(20261104)
    cases = {}

    def knots(B, nl, nr, nh, sigma):
        return (sigma * torch.randn(B, nl, generator=g), sigma * torch.randn(B, nr, generator=g),
                sigma * torch.randn(B, nh, generator=g))

    B, H, W = 2, 12, 16
    img = torch.rand(B, 3, H, W, generator=g)
    mask = torch.rand(B, 1, H, W, generator=g) > 0.1
    cases["smooth"] = dict(img=img, mask=mask, knots=knots(B, 48, 48, 64, 0.1), pts=(48, 48, 64))

    img = (torch.rand(B, 3, H, W, generator=g) * 255).round() / 255
    mask = torch.rand(B, 1, H, W, generator=g) > 0.1
    cases["uint8"] = dict(img=img, mask=mask, knots=knots(B, 48, 48, 64, 0.1), pts=(48, 48, 64))

    img = _edge_image()
    cases["edges"] = dict(img=img, mask=torch.ones(1, 1, 16, 32, dtype=torch.bool),
                          knots=knots(1, 48, 48, 64, 0.1), pts=(48, 48, 64))

    # identity curves (all knots 0 -> C == 1 -> scale == 1): exercises saturation/tie paths cleanly
    cases["identity"] = dict(img=_edge_image(), mask=torch.ones(1, 1, 16, 32, dtype=torch.bool),
                             knots=(torch.zeros(1, 48), torch.zeros(1, 48), torch.zeros(1, 64)),
                             pts=(48, 48, 64))

    # strong curves: heavy clamping everywhere
    img = torch.rand(B, 3, H, W, generator=g)
    mask = torch.rand(B, 1, H, W, generator=g) > 0.1
    cases["strong"] = dict(img=img, mask=mask, knots=knots(B, 48, 48, 64, 0.5), pts=(48, 48, 64))

    # non-default knot counts, and knot tensors wider than num_*_points (model.py:153,159,165 slices)
    img = torch.rand(3, 3, 5, 7, generator=g)  # H*W = 35: odd, exercises the unaligned tail
    mask = torch.rand(3, 1, 5, 7, generator=g) > 0.2
    cases["widek"] = dict(img=img, mask=mask, knots=knots(3, 40, 33, 50, 0.1), pts=(30, 18, 40))

    # float mask with fractional values (infer.py:41 passes a float mask)
    img = torch.rand(B, 3, H, W, generator=g)
    fmask = torch.rand(B, 1, H, W, generator=g)
    fmask = torch.where(fmask < 0.2, torch.zeros_like(fmask), torch.where(fmask > 0.6, torch.ones_like(fmask), fmask))
    cases["floatmask"] = dict(img=img, mask=fmask, knots=knots(B, 48, 48, 64, 0.1), pts=(48, 48, 64))

    for c in cases.values():
        Bc = c["img"].shape[0]
        c["gout"] = torch.randn(c["img"].shape, generator=g)
        c["greg"] = torch.rand(Bc, generator=g) + 0.5
    return cases


def run_layer(CURLLayer, case, dtype):
    nl, nr, nh = case["pts"]
    layer = CURLLayer(nl, nr, nh).to(dtype)
    img = case["img"].to(dtype).clone().requires_grad_(True)
    L, R, Hk = (k.to(dtype).clone().requires_grad_(True) for k in case["knots"])
    mask = case["mask"] if case["mask"].dtype == torch.bool else case["mask"].to(dtype)
    out, reg = layer(img, mask, L, R, Hk)
    loss = (out * case["gout"].to(dtype)).sum() + (reg * case["greg"].to(dtype)).sum()
    gimg, gL, gR, gH = torch.autograd.grad(loss, (img, L, R, Hk))
    return dict(out=out.detach(), reg=reg.detach(), gimg=gimg, gL=gL, gR=gR, gH=gH)


def stage_cases():
    g = torch.Generator().manual_seed(4242)
    x = torch.rand(2, 3, 9, 11, generator=g) * 1.3 - 0.15      # out-of-range values on purpose
    e = _edge_image()
    return {"wide": x, "edges": e, "unit": torch.rand(2, 3, 9, 11, generator=g)}


def main():
    os.makedirs(GOLDEN_DIR, exist_ok=True)
    colors, curves, CURLLayer = ref_loader.load()

    blob = {}
    for name, case in layer_cases().items():
        blob[f"{name}.img"] = case["img"].numpy()
        blob[f"{name}.mask"] = case["mask"].numpy()
        for k, t in zip("LRH", case["knots"]):
            blob[f"{name}.{k}"] = t.numpy()
        blob[f"{name}.pts"] = np.array(case["pts"], dtype=np.int64)
        blob[f"{name}.gout"] = case["gout"].numpy()
        blob[f"{name}.greg"] = case["greg"].numpy()
        for tag, dtype in (("f32", torch.float32), ("f64", torch.float64)):
            res = run_layer(CURLLayer, case, dtype)
            for k, t in res.items():
                blob[f"{name}.{tag}.{k}"] = t.numpy()
    np.savez_compressed(os.path.join(GOLDEN_DIR, "layer.npz"), **blob)

    blob = {}
    mods = {"rgb2lab": colors.RGB2LAB, "lab2rgb": colors.LAB2RGB, "rgb2hsv": colors.RGB2HSV, "hsv2rgb": colors.HSV2RGB}
    g = torch.Generator().manual_seed(777)
    for cname, x in stage_cases().items():
        blob[f"{cname}.x"] = x.numpy()
        gy = torch.randn(x.shape, generator=g)
        blob[f"{cname}.gy"] = gy.numpy()
        for mname, M in mods.items():
            for tag, dtype in (("f32", torch.float32), ("f64", torch.float64)):
                xi = x.to(dtype).clone().requires_grad_(True)
                y = M().to(dtype)(xi * 1.0)       # *1.0: the modules write in place into their input
                (gx,) = torch.autograd.grad((y * gy.to(dtype)).sum(), xi)
                blob[f"{cname}.{mname}.{tag}.y"] = y.detach().numpy()
                blob[f"{cname}.{mname}.{tag}.gx"] = gx.numpy()
        for aname, fn, width in (("adjust_lab", curves.adjust_lab, 48), ("adjust_rgb", curves.adjust_rgb, 48),
                                 ("adjust_hsv", curves.adjust_hsv, 64)):
            P = 0.2 * torch.randn(x.shape[0], width, generator=g)
            blob[f"{cname}.{aname}.P"] = P.numpy()
            for tag, dtype in (("f32", torch.float32), ("f64", torch.float64)):
                xi = x.to(dtype).clone().requires_grad_(True)
                Pi = P.to(dtype).clone().requires_grad_(True)
                y, reg = fn(xi, Pi)
                gx, gP = torch.autograd.grad((y * gy.to(dtype)).sum() + reg.sum(), (xi, Pi))
                blob[f"{cname}.{aname}.{tag}.y"] = y.detach().numpy()
                blob[f"{cname}.{aname}.{tag}.reg"] = reg.detach().numpy()
                blob[f"{cname}.{aname}.{tag}.gx"] = gx.numpy()
                blob[f"{cname}.{aname}.{tag}.gP"] = gP.numpy()
    np.savez_compressed(os.path.join(GOLDEN_DIR, "stages.npz"), **blob)
    trispace_golden()
    loss_golden()
    real_golden()
    for f in ("layer.npz", "stages.npz", "trispace.npz", "loss.npz", "real.npz"):
        print(f, os.path.getsize(os.path.join(GOLDEN_DIR, f)), "bytes")


TRISPACE_CASES = ["rgb", "spatial", "spatial_mobile", "uint8", "edges"]
# ChannelPolyLayer(degree, ...) is generic in the degree (model.py:263-333); appended after the degree-4 cases so
# that those regenerate bit for bit (one seeded generator runs through the list)
TRISPACE_DEG_CASES = {"deg1_spatial": 1, "deg2_rgb": 2, "deg2_spatial": 2, "deg3_rgb": 3, "deg3_spatial": 3}


def trispace_golden():
    """tests/golden/trispace.npz: the reference's TriSpaceRegNet.generate_residual / generate_image
    (model.py:486-508) executed through oracle/ref_loader.load_trispace, fp32 and fp64, with the
    coefficient gradients of loss = sum(out * gout) by autograd."""
    g = torch.Generator().manual_seed(20261105)
    blob = {}
    for case in TRISPACE_CASES + list(TRISPACE_DEG_CASES):
        spatial = "spatial" in case
        degree = TRISPACE_DEG_CASES.get(case, 4)
        B, H, W = (1, 16, 32) if case == "edges" else (2, 14, 18)
        if case == "edges":
            img = _edge_image()
        else:
            img = torch.rand(B, 3, H, W, generator=g)
            if case == "uint8":
                img = (img * 255).round() / 255
        _, net = ref_loader.load_trispace(degree, spatial, mobile=(case == "spatial_mobile"))
        nc = net.polylayer.num_coeffs
        R, L, Hc = (0.3 * torch.randn(B, 3, nc, generator=g) for _ in range(3))
        gout = torch.randn(B, 3, H, W, generator=g)
        blob.update({f"{case}.img": img.numpy(), f"{case}.R": R.numpy(), f"{case}.L": L.numpy(),
                     f"{case}.H": Hc.numpy(), f"{case}.gout": gout.numpy(),
                     f"{case}.spatial": np.array(int(spatial)), f"{case}.degree": np.array(degree)})
        for tag, dtype in (("f32", torch.float32), ("f64", torch.float64)):
            net_d = net.to(dtype) if dtype == torch.float64 else net
            for clamp in (0, 1):
                a = [t.to(dtype).clone().requires_grad_(True) for t in (R, L, Hc)]
                xi = img.to(dtype).clone().requires_grad_(True)      # the image gradient too (autograd of the reference)
                if dtype == torch.float64:                    # x / y coordinate parameters are int64: keep them
                    net_d.x.data = net_d.x.data.to(torch.int64) if spatial else net_d.x.data
                    net_d.y.data = net_d.y.data.to(torch.int64) if spatial else net_d.y.data
                out = net_d.generate_residual(xi, a[0], a[1], a[2])
                if clamp:
                    out = net_d.generate_image(xi, out)
                gr = torch.autograd.grad((out * gout.to(dtype)).sum(), a + [xi])
                blob[f"{case}.{tag}.c{clamp}.out"] = out.detach().numpy()
                for k, v in zip(("gR", "gL", "gH", "gimg"), gr):
                    blob[f"{case}.{tag}.c{clamp}.{k}"] = v.numpy()
    np.savez_compressed(os.path.join(GOLDEN_DIR, "trispace.npz"), **blob)


LOSS_CASES = ["near", "random", "uint8", "odd", "allmasked_image", "floatmask", "saturated"]


def loss_golden():
    """tests/golden/loss.npz: the reference's CURLLoss.forward (model.py:81-118, with the reference's
    MSSSIMMetric inside) executed through oracle/ref_loader.load_loss, fp32 and fp64, with the
    gradient w.r.t. the predicted batch by autograd."""
    g = torch.Generator().manual_seed(20261106)
    blob = {}
    for case in LOSS_CASES:
        B, H, W = (1, 37, 53) if case == "odd" else (2, 32, 36)
        target = torch.rand(B, 3, H, W, generator=g)
        if case == "random":
            pred = torch.rand(B, 3, H, W, generator=g)
        else:
            pred = (target + 0.05 * torch.randn(B, 3, H, W, generator=g)).clamp(0.0, 1.0)
        if case == "saturated":                  # what a clamped prediction looks like: exact 0 / 1, whites, greys, channel ties
            pred = ((target - 0.5) * 2.2 + 0.5 + 0.05 * torch.randn(B, 3, H, W, generator=g)).clamp(0.0, 1.0)
            pred[:, :, 4:6] = pred[:, 0:1, 4:6]          # grey rows
            pred[:, 1, 8:10] = pred[:, 0, 8:10]          # r == g rows
        if case == "uint8":                      # quantised: exact ties pred == target (sign(0) = 0), black pixels
            pred, target = (pred * 255).round() / 255, (target * 255).round() / 255
            pred[:, :, :2] = 0.0
            target[:, :, 1:3] = 0.0
        mask = torch.rand(B, 1, H, W, generator=g) > 0.15
        if case == "allmasked_image":
            mask[1] = False                      # one image contributes nothing
        if case == "floatmask":
            mask = mask.to(torch.float32)
        blob.update({f"{case}.pred": pred.numpy(), f"{case}.target": target.numpy(), f"{case}.mask": mask.numpy()})
        for tag, dtype in (("f32", torch.float32), ("f64", torch.float64)):
            crit = ref_loader.load_loss().to(dtype)
            p = pred.to(dtype).clone().requires_grad_(True)
            loss = crit(p, target.to(dtype), mask)
            (gp,) = torch.autograd.grad(loss, p)
            blob[f"{case}.{tag}.loss"] = loss.detach().numpy()
            blob[f"{case}.{tag}.gpred"] = gp.numpy()
    np.savez_compressed(os.path.join(GOLDEN_DIR, "loss.npz"), **blob)


REAL_DIR = "/root/reference/adobe5k_dpe/curl_example_test_input"
REAL_CROP = (48, 64)      # rows, columns of the window whose reference outputs are stored


def real_images():
    """The nine 8-bit example photographs the reference ships (512 x 340/341 RGBA, alpha == 255), as
    uint8 [H, W, 3] arrays in file-name order -- BASELINE.md section 4(i)'s substitute for config 1."""
    from PIL import Image
    out = []
    for name in sorted(os.listdir(REAL_DIR)):
        if not name.lower().endswith(".png"):
            continue
        a = np.array(Image.open(os.path.join(REAL_DIR, name)))
        assert a.dtype == np.uint8 and a.ndim == 3 and a.shape[2] == 4 and (a[..., 3] == 255).all(), name
        out.append((name, np.ascontiguousarray(a[..., :3])))
    assert len(out) == 9
    return out


def real_golden():
    """tests/golden/real.npz: the nine real photographs (whole, uint8) and, for one window of each, the
    reference CURLLayer's forward + backward in fp32 and fp64 (x = u8 / 255 as TF.to_tensor does,
    data.py:152-158; knots 0.1 * randn seed 1234; mask == 1 on even images, a random bool mask on odd
    ones).  The whole images travel so that the GPU box can (a) run the CUDA path on real saturation /
    smooth-region statistics against the oracle and (b) time the CPU arm on them one at a time."""
    _, _, CURLLayer = ref_loader.load()
    g = torch.Generator().manual_seed(1234)
    blob = {"n": np.array(9), "crop": np.array(REAL_CROP)}
    ch, cw = REAL_CROP
    for i, (name, a) in enumerate(real_images()):
        H, W, _ = a.shape
        blob[f"{i}.name"] = np.array(name)
        blob[f"{i}.u8"] = a
        L, R, Hk = (0.1 * torch.randn(1, n, generator=g) for n in (48, 48, 64))
        for k, t in zip("LRH", (L, R, Hk)):
            blob[f"{i}.{k}"] = t.numpy()
        # the window whose reference OUTPUT is closest to 30 % saturated (exact 0 / 1: the clamp ties of
        # SURVEY.md 8a) -- a window that is all white or all black carries no gradient information
        full = torch.from_numpy(a).permute(2, 0, 1)[None].float() / 255.0
        with torch.no_grad():
            o_full, _ = CURLLayer(48, 48, 64)(full.contiguous(), torch.ones(1, 1, H, W, dtype=torch.bool), L, R, Hk)
        sat = ((o_full[0] == 0) | (o_full[0] == 1)).float().mean(0).numpy()
        best, by, bx = 9.0, (H - ch) // 2, (W - cw) // 2
        for y0 in range(0, H - ch + 1, ch // 2):
            for x0 in range(0, W - cw + 1, cw // 2):
                v = abs(float(sat[y0:y0 + ch, x0:x0 + cw].mean()) - 0.3)
                if v < best:
                    best, by, bx = v, y0, x0
        blob[f"{i}.win"] = np.array([by, bx])
        x = torch.from_numpy(a[by:by + ch, bx:bx + cw]).permute(2, 0, 1)[None].float() / 255.0   # TF.to_tensor
        mask = torch.ones(1, 1, ch, cw, dtype=torch.bool) if i % 2 == 0 else torch.rand(1, 1, ch, cw, generator=g) > 0.1
        case = dict(img=x.contiguous(), mask=mask, knots=(L, R, Hk), pts=(48, 48, 64),
                    gout=torch.randn(1, 3, ch, cw, generator=g), greg=torch.ones(1))
        blob[f"{i}.mask"] = mask.numpy()
        blob[f"{i}.gout"] = case["gout"].numpy()
        for tag, dtype in (("f32", torch.float32), ("f64", torch.float64)):
            for k, t in run_layer(CURLLayer, case, dtype).items():
                blob[f"{i}.{tag}.{k}"] = t.numpy()
    np.savez_compressed(os.path.join(GOLDEN_DIR, "real.npz"), **blob)


if __name__ == "__main__":
    if "--loss-only" in sys.argv:
        loss_golden()
    elif "--trispace-only" in sys.argv:
        trispace_golden()
    elif "--real-only" in sys.argv:
        real_golden()
        print("real.npz", os.path.getsize(os.path.join(GOLDEN_DIR, "real.npz")), "bytes")
    else:
        main()
