"""Dev: forward accuracy of one or more builds of the library (dev/variants/*.so) against the CPU oracle, fp32 and
fp64, on seeded continuous and 8-bit inputs.   python dev/fwd_accuracy.py lib1.so [lib2.so ...]"""
import ctypes, os, sys, json
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from curl_b200 import _lib as L
from oracle import curl_oracle as O

def bind(path):
    lib = ctypes.CDLL(os.path.abspath(path))
    for name, (res, args) in L.SIGNATURES.items():
        if hasattr(lib, name):
            fn = getattr(lib, name); fn.restype = res; fn.argtypes = args
    return lib

dev = torch.device("cuda:0")
cases = {}
for name, u8 in (("continuous", False), ("u8grid", True)):
    g = torch.Generator().manual_seed(77)
    B, H, W = 4, 384, 512
    img = torch.rand(B, 3, H, W, generator=g)
    if u8: img = (img * 255).round() / 255
    mask = torch.rand(B, 1, H, W, generator=g) > 0.1
    kn = [0.1 * torch.randn(B, n, generator=g) for n in (48, 48, 64)]
    r32, _ = O.curl_layer(img, mask, *kn)
    r64, _ = O.curl_layer(img.double(), mask, *[k.double() for k in kn])
    cases[name] = (img, mask, kn, r32, r64)
st = torch.cuda.current_stream().cuda_stream
for path in sys.argv[1:]:
    lib = bind(path)
    res = {"lib": os.path.basename(path)}
    for name, (img, mask, kn, r32, r64) in cases.items():
        B, _, H, W = img.shape
        di, dm = img.to(dev), mask.to(dev)
        dk = [k.to(dev) for k in kn]
        out = torch.empty_like(di); reg = torch.empty(B, device=dev)
        rc = lib.curl_layer_fwd_f32(di.data_ptr(), dm.data_ptr(), 1, dk[0].data_ptr(), 48, dk[1].data_ptr(), 48, dk[2].data_ptr(), 64,
                                    out.data_ptr(), reg.data_ptr(), B, H, W, 16, 16, 16, st)
        assert rc == 0
        o = out.cpu()
        e32 = (o - r32).abs(); e64 = (o.double() - r64).abs(); own = (r32.double() - r64).abs()
        res[name] = {"max_vs_ref32": float(e32.max()), "frac_over_1e5_vs_ref32": float((e32 > 1e-5).float().mean()),
                     "frac_over_3e6_vs_ref64": float((e64 > 3e-6).float().mean()), "mean_abs_vs_ref64": float(e64.mean()),
                     "max_vs_ref64": float(e64.max()), "ref_own_max": float(own.max()), "ref_own_mean": float(own.mean()),
                     "ref_frac_over_3e6": float((own > 3e-6).float().mean())}
    print(json.dumps(res), flush=True)
