"""Dev check: bitwise run-to-run determinism of the layer backward (knot gradients, image gradient) of one or more
builds.   python dev/determinism.py a.so [b.so ...]   (KB_B, KB_H, KB_W, KB_REPS)"""
import ctypes, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from curl_b200 import _lib as L

def bind(path):
    lib = ctypes.CDLL(path)
    for name, (res, args) in L.SIGNATURES.items():
        if hasattr(lib, name):
            fn = getattr(lib, name); fn.restype = res; fn.argtypes = args
    return lib

B = int(os.environ.get("KB_B", "4")); H = int(os.environ.get("KB_H", "2160")); W = int(os.environ.get("KB_W", "3840"))
reps = int(os.environ.get("KB_REPS", "20"))
dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(1)
img = torch.rand(B, 3, H, W, generator=g, device=dev); mask = torch.rand(B, 1, H, W, generator=g, device=dev) > 0.1
Lk, Rk, Hk = (0.1 * torch.randn(B, n, generator=g, device=dev) for n in (48, 48, 64))
gout = torch.randn(B, 3, H, W, generator=g, device=dev); greg = torch.ones(B, device=dev)
st = torch.cuda.current_stream().cuda_stream
for path in sys.argv[1:]:
    lib = bind(path)
    for want in (True, False):
        first = None; bad = 0; worst = 0.0
        for r in range(reps):
            gimg = torch.empty_like(img); gL = torch.empty_like(Lk); gR = torch.empty_like(Rk); gH = torch.empty_like(Hk)
            wsb = lib.curl_bwd_workspace_bytes(B, H, W); ws = torch.empty(wsb, dtype=torch.uint8, device=dev)
            rc = lib.curl_layer_bwd_f32(img.data_ptr(), mask.data_ptr(), 1, Lk.data_ptr(), 48, Rk.data_ptr(), 48, Hk.data_ptr(), 64,
                                        gout.data_ptr(), greg.data_ptr(), gimg.data_ptr() if want else None, gL.data_ptr(), gR.data_ptr(),
                                        gH.data_ptr(), ws.data_ptr(), wsb, B, H, W, 16, 16, 16, st)
            assert rc == 0, rc
            torch.cuda.synchronize()
            cur = [gL.clone(), gR.clone(), gH.clone()] + ([gimg.clone()] if want else [])
            if first is None: first = cur
            else:
                same = all(torch.equal(a, b) for a, b in zip(cur, first))
                if not same:
                    bad += 1
                    if bad <= 2:
                        names = ["gL", "gR", "gH", "gimg"]
                        for nm, a, b in zip(names, cur, first):
                            d = (a != b)
                            if d.any():
                                idx = d.nonzero()
                                print("   ", nm, "differs in", int(d.sum()), "elements; first", idx[0].tolist(), "last", idx[-1].tolist(), flush=True)
                    worst = max(worst, max(float(((a - b).abs().max() / b.abs().max())) for a, b in zip(cur, first)))
        print(os.path.basename(path), "gimg" if want else "nogimg", "runs differing from the first: %d of %d" % (bad, reps - 1), "worst rel %.3g" % worst, flush=True)
