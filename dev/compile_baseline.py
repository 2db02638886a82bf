"""`torch.compile` of the reference's op sequence on the B200 -- "the reference's Blackwell path" of SURVEY.md 8(d),
the strongest stand-in the fused kernels must beat -- beside eager PyTorch and this library.  bench.py runs it
as a subprocess (gpu_compile_baseline; a hung or failed compile must not take the headline line with it) and
embeds the one JSON line it prints.

    python dev/compile_baseline.py [B H W]

What is compiled: oracle/curl_oracle_compile.py -- the oracle with `non_nan_inv` written as torch.where and the
in-place channel writes as torch.stack, the only changes dynamo needs; pinned to the oracle on CPU in
tests/test_oracle_pins.py::test_compile_variant_equals_oracle.  fullgraph=True (one graph, no breaks),
inductor defaults (Triton kernels).  Not a product path.
"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from oracle import curl_oracle as O  # noqa: E402
from oracle import curl_oracle_compile as OC  # noqa: E402
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
    res = {"shape": [B, H, W], "mpx_per_step": mpx, "torch": torch.__version__}

    def fwd_bwd(layer_fn):
        def f():
            x = img.detach().requires_grad_(True)
            l, r, h = (t.detach().requires_grad_(True) for t in (L, R, Hk))
            o, rg = layer_fn(x, mask, l, r, h)
            torch.autograd.backward([o, rg], [gout, greg])
            return x.grad, l.grad, o.detach()
        return f

    def fwd(layer_fn):
        def f():
            with torch.no_grad():
                return layer_fn(img, mask, L, R, Hk)[0]
        return f

    def entry(ms_f, ms_fb):
        return {"fwd_ms": ms_f, "fwd_mpx_per_s": mpx / (ms_f / 1e3), "fwd_bwd_ms": ms_fb,
                "fwd_bwd_mpx_per_s": mpx / (ms_fb / 1e3)}

    ours = curl_b200.CURLLayer().to(dev)
    res["curl_b200"] = entry(timed(fwd(ours), 20), timed(fwd_bwd(ours), 10))
    gi_o, gl_o, out_o = (t.clone() for t in fwd_bwd(ours)())

    torch.cuda.reset_peak_memory_stats(dev)
    res["eager"] = entry(timed(fwd(O.curl_layer), 3), timed(fwd_bwd(O.curl_layer), 3))
    res["eager"]["peak_mem_gb"] = torch.cuda.max_memory_allocated(dev) / 1e9

    t0 = time.time()
    try:
        comp = torch.compile(OC.curl_layer, fullgraph=True)
        fb = fwd_bwd(comp)
        gi_c, gl_c, out_c = (t.clone() for t in fb())
        ff = fwd(comp)
        ff()
        torch.cuda.synchronize()
        res["compile_s"] = time.time() - t0
        torch.cuda.reset_peak_memory_stats(dev)
        e = entry(timed(ff, 10), timed(fb, 5))
        e["peak_mem_gb"] = torch.cuda.max_memory_allocated(dev) / 1e9
        e["frac_out_over_1e5_vs_curl_b200"] = float(((out_c - out_o).abs() > 1e-5).float().mean())
        e["rel_gL_vs_curl_b200"] = float((gl_c - gl_o).abs().max() / gl_o.abs().max())
        e["what"] = ("torch.compile(fullgraph=True, inductor) of oracle/curl_oracle_compile.py = the reference's op "
                     "sequence with non_nan_inv as torch.where and channel writes as torch.stack")
        res["torch_compile"] = e
    except Exception as e:
        res["torch_compile"] = {"unavailable": "%s: %s" % (type(e).__name__, str(e)[:300])}
    print(json.dumps(res))


if __name__ == "__main__":
    main()
