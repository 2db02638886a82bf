"""One-off baseline (not part of bench.py, not a product path): the reference's op sequence (oracle/curl_oracle.py)
under torch.compile (inductor) on the B200 -- "the reference's Blackwell path" of SURVEY.md 8(d) -- beside
this library, forward + backward with image gradient.  Prints one JSON line.

    python dev/compile_baseline.py [B H W]
"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from oracle import curl_oracle as O  # noqa: E402
import curl_b200  # noqa: E402


def timed(fn, reps):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


def main():
    B, H, W = (int(v) for v in sys.argv[1:4]) if len(sys.argv) >= 4 else (4, 1080, 1920)
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev).manual_seed(1234)
    img = torch.rand(B, 3, H, W, generator=g, device=dev)
    mask = torch.rand(B, 1, H, W, generator=g, device=dev) > 0.1
    L, R, Hk = (0.1 * torch.randn(B, n, generator=g, device=dev) for n in (48, 48, 64))
    gout = torch.randn(B, 3, H, W, generator=g, device=dev)
    greg = torch.ones(B, device=dev)
    mpx = B * H * W / 1e6
    res = {"shape": [B, H, W]}

    def run(layer_fn):
        def f():
            x = img.detach().requires_grad_(True)
            l, r, h = (t.detach().requires_grad_(True) for t in (L, R, Hk))
            o, rg = layer_fn(x, mask, l, r, h)
            torch.autograd.backward([o, rg], [gout, greg])
            return x.grad
        return f

    ours = curl_b200.CURLLayer().to(dev)
    ms = timed(run(ours), 10)
    res["curl_b200"] = {"ms": ms, "mpx_per_s": mpx / (ms / 1e3)}
    g_ours = run(ours)().clone()

    ms = timed(run(O.curl_layer), 3)
    res["eager"] = {"ms": ms, "mpx_per_s": mpx / (ms / 1e3)}

    t0 = time.time()
    try:
        comp = torch.compile(O.curl_layer)
        fc = run(comp)
        g_c = fc().clone()
        torch.cuda.synchronize()
        res["compile_s"] = time.time() - t0
        ms = timed(fc, 5)
        res["torch_compile"] = {"ms": ms, "mpx_per_s": mpx / (ms / 1e3),
                                "max_abs_gimg_vs_ours": float((g_c - g_ours).abs().max())}
    except Exception as e:
        res["torch_compile"] = {"unavailable": "%s: %s" % (type(e).__name__, str(e)[:300])}
    print(json.dumps(res))


if __name__ == "__main__":
    main()
