"""Dev: times the segment-clamped layer kernels next to the default ones (CUDA events)."""
import os, sys, json
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import curl_b200
dev = torch.device("cuda:0")


def timeit(fn, reps=5):
    for _ in range(2): fn()
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


B, H, W = 8, 2160, 3840
g = torch.Generator(device=dev).manual_seed(1)
for kind in ("random", "smooth"):
    if kind == "random":
        img = torch.rand(B, 3, H, W, generator=g, device=dev)
    else:   # smooth image: neighbouring pixels share segments -> same-bin conflicts inside a warp
        yy = torch.linspace(0, 1, H, device=dev).reshape(1, 1, H, 1); xx = torch.linspace(0, 1, W, device=dev).reshape(1, 1, 1, W)
        img = (0.5 + 0.5 * torch.sin(6.0 * xx + 4.0 * yy + torch.arange(3, device=dev).reshape(1, 3, 1, 1))).expand(B, 3, H, W).contiguous()
    mask = torch.rand(B, 1, H, W, generator=g, device=dev) > 0.1
    gout = torch.randn(B, 3, H, W, generator=g, device=dev)
    res = {"image": kind, "shape": [B, H, W]}
    for name, seg in (("default", False), ("segment_clamp", True)):
        layer = curl_b200.CURLLayer(segment_clamp=seg).to(dev)
        kn = [(0.1 * torch.randn(B, n, generator=g, device=dev)).requires_grad_(True) for n in (48, 48, 64)]
        xi = img.clone().requires_grad_(True)
        with torch.no_grad():
            f = timeit(lambda: layer(img, mask, *kn))
        def step():
            out, reg = layer(xi, mask, *kn)
            torch.autograd.backward([out, reg], [gout, torch.ones(B, device=dev)])
        s = timeit(step)
        res[name] = {"fwd_ms": round(f, 3), "fwd_bwd_ms": round(s, 3), "fwd_Gpx": round(B * H * W / f / 1e6, 1), "step_Gpx": round(B * H * W / s / 1e6, 1)}
    print(json.dumps(res), flush=True)
