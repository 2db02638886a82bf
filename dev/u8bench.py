"""Dev: times CURLLayer.forward_u8 (8-bit in / out, planar) with CUDA events."""
import os, sys, json
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import curl_b200
dev = torch.device("cuda:0")
B, H, W = int(os.environ.get("KB_B", "64")), 2160, 3840
g = torch.Generator(device=dev).manual_seed(1)
img = (torch.rand(B, 3, H, W, generator=g, device=dev) * 255).to(torch.uint8)
mask = torch.rand(B, 1, H, W, generator=g, device=dev) > 0.1
kn = [0.1 * torch.randn(B, n, generator=g, device=dev) for n in (48, 48, 64)]
layer = curl_b200.CURLLayer().to(dev)
with torch.no_grad():
    for _ in range(3): layer.forward_u8(img, mask, *kn)
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): layer.forward_u8(img, mask, *kn)
    e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 10
print(json.dumps({"shape": [B, H, W], "u8_fwd_ms": round(ms, 4), "Gpx": round(B * H * W / ms / 1e6, 1)}))
