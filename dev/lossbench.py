"""Dev: times the CURLLoss colour-term kernels through the C ABI (CUDA events), next to the eager
PyTorch colour terms on the same GPU and the library MS-SSIM."""
import ctypes, os, sys, json
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from curl_b200 import _lib as L
lib = ctypes.CDLL(sys.argv[1] if len(sys.argv) > 1 and sys.argv[1].endswith(".so") else L.LIB_PATH)
for name, (res, args) in L.SIGNATURES.items():
    if not hasattr(lib, name):
        continue        # a loss-only experiment build
    fn = getattr(lib, name); fn.restype = res; fn.argtypes = args
dev = torch.device("cuda:0")
st = torch.cuda.current_stream().cuda_stream


def timeit(fn, reps=5):
    for _ in range(2): fn()
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


for (B, H, W) in ((8, 2160, 3840), (32, 512, 512)):
    g = torch.Generator(device=dev).manual_seed(1)
    target = torch.rand(B, 3, H, W, generator=g, device=dev)
    pred = (target + 0.05 * torch.randn(B, 3, H, W, generator=g, device=dev)).clamp(0, 1)
    mask = torch.rand(B, 1, H, W, generator=g, device=dev) > 0.1
    sums = torch.empty(B, 5, device=dev); Lp = torch.empty(B, 1, H, W, device=dev); Lt = torch.empty_like(Lp)
    w = torch.full((B, 4), 1e-6, device=dev); gL = torch.randn(B, 1, H, W, generator=g, device=dev) * 1e-6
    gp = torch.empty_like(pred)
    n = lib.curl_loss_terms_workspace_bytes(B, H, W); ws = torch.empty(n, dtype=torch.uint8, device=dev)
    def fwd():
        assert lib.curl_loss_terms_fwd_f32(pred.data_ptr(), target.data_ptr(), mask.data_ptr(), 1, sums.data_ptr(), Lp.data_ptr(), Lt.data_ptr(), ws.data_ptr(), n, B, H, W, st) == 0
    def bwd():
        assert lib.curl_loss_terms_bwd_f32(pred.data_ptr(), target.data_ptr(), mask.data_ptr(), 1, w.data_ptr(), gL.data_ptr(), gp.data_ptr(), B, H, W, st) == 0
    npx = B * H * W
    res = {"shape": [B, H, W]}
    for name, fn, bpp in (("fwd", fwd, 33), ("bwd", bwd, 41)):
        ms = timeit(fn)
        res[name + "_ms"] = round(ms, 3); res[name + "_Gpx"] = round(npx / ms / 1e6, 2); res[name + "_GBs"] = round(npx * bpp / ms / 1e6, 1)
    if "--eager" in sys.argv:
        from oracle import curl_oracle as O
        import curl_b200
        ms_layer = curl_b200.metric.MSSSIMMetric(num_channel=1).to(dev)
        def eager_colour():
            p, t = pred * mask, target * mask
            a = (p - t).abs().sum(); c = torch.nn.functional.cosine_similarity(p, t, dim=1).sum()
            lp, lt = O.loss_lab(p), O.loss_lab(t)
            b_ = (lp - lt).abs().sum(); d = (O.loss_hsv_cone(p) - O.loss_hsv_cone(t)).abs().sum()
            return a + c + b_ + d
        res["eager_colour_fwd_ms"] = round(timeit(eager_colour, 3), 3)
        res["msssim_fwd_ms"] = round(timeit(lambda: ms_layer(Lp, Lt), 3), 3)
        crit = curl_b200.CURLLoss().to(dev)
        pr = pred.clone().requires_grad_(True)
        def ours_step():
            pr.grad = None
            crit(pr, target, mask).backward()
        res["ours_loss_fwd_bwd_ms"] = round(timeit(ours_step, 3), 3)
        def eager_step():
            pr.grad = None
            O.curl_loss(pr, target, mask).backward()
        if B * H * W <= 32 * 512 * 512:
            res["eager_loss_fwd_bwd_ms"] = round(timeit(eager_step, 3), 3)
    print(json.dumps(res), flush=True)
