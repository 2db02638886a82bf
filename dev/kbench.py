"""Dev micro-benchmark: times the layer forward / backward kernels of one or more builds of the
library through the C ABI (CUDA events on the launching stream) and diffs their results against the
first build.   python dev/kbench.py lib1.so [lib2.so ...]   (B=16 4K by default: 1.6 GB per tensor)"""
import ctypes, os, sys, json
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from curl_b200 import _lib as L

def bind(path):
    lib = ctypes.CDLL(os.path.abspath(path))
    for name, (res, args) in L.SIGNATURES.items():
        if not hasattr(lib, name):
            continue        # a layer-only experiment build
        fn = getattr(lib, name); fn.restype = res; fn.argtypes = args
    return lib

def main():
    libs = [a for a in sys.argv[1:] if a.endswith(".so")]
    B = int(os.environ.get("KB_B", "16")); H = int(os.environ.get("KB_H", "2160")); W = int(os.environ.get("KB_W", "3840"))
    reps = int(os.environ.get("KB_REPS", "10"))
    mk = int(os.environ.get("KB_MASK", "1"))
    dev = torch.device("cuda:0")
    g = torch.Generator(device=dev).manual_seed(1234)
    img = torch.rand(B, 3, H, W, generator=g, device=dev)
    mask = (torch.rand(B, 1, H, W, generator=g, device=dev) > 0.1)
    if mk == 2: mask = mask.float()
    Lk, Rk, Hk = (0.1 * torch.randn(B, n, generator=g, device=dev) for n in (48, 48, 64))
    gout = torch.randn(B, 3, H, W, generator=g, device=dev)
    greg = torch.ones(B, device=dev)
    npx = B * H * W
    ref = None
    st = torch.cuda.current_stream().cuda_stream
    for path in libs:
        lib = bind(path)
        out = torch.empty_like(img); reg = torch.empty(B, device=dev)
        gimg = torch.empty_like(img); gL = torch.empty_like(Lk); gR = torch.empty_like(Rk); gH = torch.empty_like(Hk)
        wsb = lib.curl_bwd_workspace_bytes(B, H, W)
        ws = torch.empty(wsb, dtype=torch.uint8, device=dev)
        mp = mask.data_ptr() if mk else None
        def fwd():
            rc = lib.curl_layer_fwd_f32(img.data_ptr(), mp, mk, Lk.data_ptr(), 48, Rk.data_ptr(), 48, Hk.data_ptr(), 64,
                                        out.data_ptr(), reg.data_ptr(), B, H, W, 16, 16, 16, st)
            assert rc == 0, rc
        def bwd(want=True):
            rc = lib.curl_layer_bwd_f32(img.data_ptr(), mp, mk, Lk.data_ptr(), 48, Rk.data_ptr(), 48, Hk.data_ptr(), 64,
                                        gout.data_ptr(), greg.data_ptr(), gimg.data_ptr() if want else None, gL.data_ptr(), gR.data_ptr(),
                                        gH.data_ptr(), ws.data_ptr(), wsb, B, H, W, 16, 16, 16, st)
            assert rc == 0, rc
        res = {"lib": os.path.basename(path)}
        for name, fn, bpp in (("fwd", fwd, 25 if mk == 1 else (24 if mk == 0 else 28)), ("bwd", bwd, 37 if mk == 1 else (36 if mk == 0 else 40)),
                              ("bwd_nogimg", lambda: bwd(False), 25)):
            for _ in range(3): fn()
            torch.cuda.synchronize()
            e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(reps): fn()
            e1.record(); torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / reps
            res[name + "_ms"] = round(ms, 4); res[name + "_GBs"] = round(npx * bpp / ms / 1e6, 1)
            res[name + "_Gpx"] = round(npx / ms / 1e6, 2)
        bwd(True); torch.cuda.synchronize()
        cur = [t.clone() for t in (out, reg, gimg, gL, gR, gH)]
        if ref is None: ref = cur
        else:
            res["maxdiff_vs_first"] = [float((a - b).abs().max()) for a, b in zip(cur, ref)]
            res["rel_knot"] = [float((a - b).abs().max() / b.abs().max()) for a, b in zip(cur[3:], ref[3:])]
        print(json.dumps(res), flush=True)

main()
