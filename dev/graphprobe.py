import os, sys, torch, torch.nn as nn
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import curl_b200
from curl_b200 import functional
dev = torch.device("cuda:0")
gen = torch.Generator().manual_seed(31)
B, H, W = 2, 64, 80
target = torch.rand(B, 3, H, W, generator=gen).to(dev)
pred = (target + 0.05 * torch.randn(B, 3, H, W, generator=gen).to(dev)).clamp(0, 1).requires_grad_(True)
mask = (torch.rand(B, 1, H, W, generator=gen) > 0.1).to(dev)
crit = curl_b200.CURLLoss().to(dev)
tri = curl_b200.TriSpaceRegNet(spatial=True, backbone=nn.Identity()).to(dev)
co = [(0.3 * torch.randn(B, 3, 126, generator=gen)).to(dev).requires_grad_(True) for _ in range(3)]
seg = curl_b200.CURLLayer(segment_clamp=True).to(dev)
kn = [(0.1 * torch.randn(B, n, generator=gen)).to(dev).requires_grad_(True) for n in (48, 48, 64)]
def f_terms():
    s, lp, lt = functional.loss_terms(pred, target, mask)
    return torch.autograd.grad(s.sum() + lp.sum(), [pred])
def f_msssim():
    lp = pred[:, :1]
    return torch.autograd.grad(crit.msssim_layer(lp, target[:, :1]).sum(), [pred])
def f_crit():
    return torch.autograd.grad(crit(pred, target, mask), [pred])
def f_tri_fwd():
    with torch.no_grad():
        return tri._fused(target, *co, True)
def f_tri():
    return torch.autograd.grad((tri._fused(target, *co, True) * target).mean(), co)
def f_seg():
    out, reg = seg(target, mask, *kn)
    return torch.autograd.grad(out.mean() + reg.sum(), kn)
for name, fn in (("terms", f_terms), ("msssim", f_msssim), ("crit", f_crit), ("tri_fwd", f_tri_fwd), ("tri", f_tri), ("seg", f_seg)):
    try:
        s = torch.cuda.Stream(); s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            fn(); fn()
        torch.cuda.current_stream().wait_stream(s); torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            fn()
        g.replay(); torch.cuda.synchronize()
        print(name, "captured ok", flush=True)
    except Exception as e:
        print(name, "FAILED:", str(e).splitlines()[0][:160], flush=True)
        torch.cuda.synchronize()
