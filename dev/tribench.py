"""Dev: times the TriSpace forward / backward kernels through the C ABI (CUDA events)."""
import ctypes, os, sys, json
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from curl_b200 import _lib as L
lib = ctypes.CDLL(sys.argv[1] if len(sys.argv) > 1 else L.LIB_PATH)
for name, (res, args) in L.SIGNATURES.items():
    if not hasattr(lib, name):
        continue        # a trispace-only experiment build
    fn = getattr(lib, name); fn.restype = res; fn.argtypes = args
dev = torch.device("cuda:0")
st = torch.cuda.current_stream().cuda_stream
SHAPES = ((8, 2160, 3840),) if os.environ.get('TB_4K') else ((32, 512, 512), (8, 2160, 3840))
for (B, H, W) in SHAPES:
    for spatial in (1, 0):
        nc = lib.curl_trispace_num_coeffs(4, spatial)
        g = torch.Generator(device=dev).manual_seed(1)
        img = torch.rand(B, 3, H, W, generator=g, device=dev)
        co = [0.3 * torch.randn(B, 3, nc, generator=g, device=dev) for _ in range(3)]
        gout = torch.randn(B, 3, H, W, generator=g, device=dev)
        out = torch.empty_like(img); gc = [torch.empty_like(c) for c in co]
        n = lib.curl_trispace_bwd_workspace_bytes(B, H, W, 4, spatial); ws = torch.empty(n, dtype=torch.uint8, device=dev)
        sig = torch.empty(B, 9, H, W, device=dev) if os.environ.get('TB_SAVE', '1') == '1' else None
        sig_ptr = None if sig is None else sig.data_ptr()
        def fwd():
            assert lib.curl_trispace_fwd_f32(img.data_ptr(), co[0].data_ptr(), co[1].data_ptr(), co[2].data_ptr(), 4, spatial, 1, out.data_ptr(), sig_ptr, B, H, W, st) == 0
        def bwd():
            assert lib.curl_trispace_bwd_f32(img.data_ptr(), co[0].data_ptr(), co[1].data_ptr(), co[2].data_ptr(), 4, spatial, 1, gout.data_ptr(), sig_ptr,
                                             gc[0].data_ptr(), gc[1].data_ptr(), gc[2].data_ptr(), ws.data_ptr(), n, B, H, W, st) == 0
        gimg = torch.empty_like(img)
        def bwd_img():
            assert lib.curl_trispace_bwd_img_f32(img.data_ptr(), co[0].data_ptr(), co[1].data_ptr(), co[2].data_ptr(), 4, spatial, 1, gout.data_ptr(), sig_ptr,
                                                 gimg.data_ptr(), B, H, W, st) == 0
        res = {"shape": [B, H, W], "spatial": spatial, "nc": nc, "saved": sig is not None}
        legs = (("fwd", fwd), ("bwd", bwd)) + ((("bwd_img", bwd_img),) if hasattr(lib, "curl_trispace_bwd_img_f32") else ())
        for name, fn in legs:
            for _ in range(2): fn()
            torch.cuda.synchronize()
            e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(5): fn()
            e1.record(); torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / 5
            npx = B * H * W
            flops = npx * (3 * 2 * 3 * (nc - 1)) if name == "fwd" else npx * 3 * (nc + 2 * 3 * nc)   # nested evaluation: one FMA per coefficient and channel; backward: monomials + the gradient product
            res[name + "_ms"] = round(ms, 3); res[name + "_Gpx"] = round(npx / ms / 1e6, 2); res[name + "_TFLOPs"] = round(flops / ms / 1e9, 2)
        print(json.dumps(res), flush=True)
