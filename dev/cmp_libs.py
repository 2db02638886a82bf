"""Dev check: two builds of the library against the fp64 oracle on one seeded batch (knot gradients relative to
max-norm, image gradient relative to max-norm, output max-abs).  python dev/cmp_libs.py a.so b.so"""
import ctypes, os, sys, json
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from curl_b200 import _lib as L
from oracle import curl_oracle as O

def bind(path):
    lib = ctypes.CDLL(path)
    for name, (res, args) in L.SIGNATURES.items():
        if hasattr(lib, name):
            fn = getattr(lib, name); fn.restype = res; fn.argtypes = args
    return lib

B, H, W = int(os.environ.get("KB_B", "2")), int(os.environ.get("KB_H", "1080")), int(os.environ.get("KB_W", "1920"))
dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(7)
img = torch.rand(B, 3, H, W, generator=g); mask = torch.rand(B, 1, H, W, generator=g) > 0.1
Lk, Rk, Hk = (0.1 * torch.randn(B, n, generator=g) for n in (48, 48, 64))
gout = torch.randn(B, 3, H, W, generator=g); greg = torch.ones(B)
ref = O.curl_layer_fwd_bwd(img.double(), mask, Lk.double(), Rk.double(), Hk.double(), gout.double(), greg.double())
ref32 = O.curl_layer_fwd_bwd(img, mask, Lk, Rk, Hk, gout, greg)
def rel(a, b): return float((a.double() - b).abs().max() / b.abs().max())
print("ref32 vs ref64:", {k: rel(ref32[k], ref[k]) for k in ("out", "gimg", "gL", "gR", "gH")})
d = {k: v.to(dev) for k, v in dict(img=img, mask=mask, L=Lk, R=Rk, H=Hk, gout=gout, greg=greg).items()}
st = torch.cuda.current_stream().cuda_stream
res = []
for path in sys.argv[1:]:
    lib = bind(path)
    out = torch.empty_like(d["img"]); reg = torch.empty(B, device=dev); gimg = torch.empty_like(d["img"])
    gL = torch.empty_like(d["L"]); gR = torch.empty_like(d["R"]); gH = torch.empty_like(d["H"])
    wsb = lib.curl_bwd_workspace_bytes(B, H, W); ws = torch.empty(wsb, dtype=torch.uint8, device=dev)
    a = (d["img"].data_ptr(), d["mask"].data_ptr(), 1, d["L"].data_ptr(), 48, d["R"].data_ptr(), 48, d["H"].data_ptr(), 64)
    assert lib.curl_layer_fwd_f32(*a, out.data_ptr(), reg.data_ptr(), B, H, W, 16, 16, 16, st) == 0
    assert lib.curl_layer_bwd_f32(*a, d["gout"].data_ptr(), d["greg"].data_ptr(), gimg.data_ptr(), gL.data_ptr(), gR.data_ptr(),
                                  gH.data_ptr(), ws.data_ptr(), wsb, B, H, W, 16, 16, 16, st) == 0
    torch.cuda.synchronize()
    cur = dict(out=out.cpu(), gimg=gimg.cpu(), gL=gL.cpu(), gR=gR.cpu(), gH=gH.cpu())
    res.append(cur)
    print(os.path.basename(path), "vs ref64:", {k: rel(cur[k], ref[k]) for k in cur}, "out max-abs", float((cur["out"].double() - ref["out"]).abs().max()))
if len(res) == 2:
    print("a vs b:", {k: rel(res[0][k], res[1][k].double()) for k in res[0]})
    dg = (res[0]["gimg"] - res[1]["gimg"]).abs()
    print("gimg pixels differing by > 1e-3 of max:", int((dg > 1e-3 * ref["gimg"].abs().max()).sum()), "of", dg.numel())
