"""GPU: the allocation-free backward entry point, curl_layer_bwd_noalloc_f32 (no counterpart in the reference),
called through the C ABI and compared with curl_layer_bwd_f32 on the same
inputs -- two calls in a row on different inputs and shapes, so that the shared static scratch array (which keeps the
previous call's contents) is seen to be cleared where it must start at zero.  The two entry points run the same
kernel with the same grid, and the reduction is deterministic, so the results are expected to be bit-identical; the
assertion uses the suite's tolerances (parity.KNOT_TOL / GRAD_TOL relative to the max-norm).

Collected last on purpose (file name): a failure here must not hide other tests.
Runs alone in a fresh process as well as after the rest of the suite (it owns no state but the library's)."""
import numpy as np
import pytest
import torch

import parity

pytestmark = pytest.mark.gpu

SCRATCH_BYTES = 1 << 20          # CURL_NOALLOC_SCRATCH_BYTES (include/curl_b200.h)
ERR_TOO_LARGE = -6


def _case(B, H, W, maskkind, seed, dev):
    gen = torch.Generator().manual_seed(seed)
    img = torch.rand(B, 3, H, W, generator=gen).to(dev)
    if maskkind == "bool":
        mask, kind = (torch.rand(B, 1, H, W, generator=gen) > 0.1).to(dev), 1
    elif maskkind == "float":
        mask, kind = (torch.rand(B, 1, H, W, generator=gen) > 0.1).float().to(dev), 2
    else:
        mask, kind = None, 0
    L, R, Hk = ((0.1 * torch.randn(B, n, generator=gen)).to(dev) for n in (48, 48, 64))
    gout = torch.randn(B, 3, H, W, generator=gen).to(dev)
    greg = (torch.rand(B, generator=gen) + 0.5).to(dev)
    return img, mask, kind, L, R, Hk, gout, greg


def _run(lib, case, noalloc, want_gimg=True):
    img, mask, kind, L, R, Hk, gout, greg = case
    B, _, H, W = img.shape
    dev = img.device
    st = torch.cuda.current_stream(dev).cuda_stream
    gimg = torch.full_like(img, float("nan")) if want_gimg else None
    gL = torch.full((B, 48), float("nan"), device=dev)
    gR = torch.full((B, 48), float("nan"), device=dev)
    gH = torch.full((B, 64), float("nan"), device=dev)
    mp = None if mask is None else mask.data_ptr()
    gp = None if gimg is None else gimg.data_ptr()
    head = (img.data_ptr(), mp, kind, L.data_ptr(), 48, R.data_ptr(), 48, Hk.data_ptr(), 64, gout.data_ptr(), greg.data_ptr(),
            gp, gL.data_ptr(), gR.data_ptr(), gH.data_ptr())
    tail = (B, H, W, 16, 16, 16, st)
    if noalloc:
        rc = lib.curl_layer_bwd_noalloc_f32(*head, *tail)
    else:
        n = lib.curl_bwd_workspace_bytes(B, H, W)
        ws = torch.empty(n, dtype=torch.uint8, device=dev)
        rc = lib.curl_layer_bwd_f32(*head, ws.data_ptr(), n, *tail)
    assert rc == 0, lib.curl_strerror(rc)
    torch.cuda.synchronize(dev)
    out = dict(gL=gL, gR=gR, gH=gH)
    if want_gimg:
        out["gimg"] = gimg
    return {k: v.cpu().numpy() for k, v in out.items()}


def _compare(a, b):
    for k in a:
        assert np.isfinite(a[k]).all(), k                      # every promised element written
        tol = parity.GRAD_TOL if k == "gimg" else parity.KNOT_TOL
        assert parity.rel_to_max(a[k], b[k]) <= tol, (k, parity.rel_to_max(a[k], b[k]))
    return all(np.array_equal(a[k], b[k]) for k in a)


def test_noalloc_backward_matches_workspace_backward_twice_in_a_row():
    from curl_b200 import _lib
    lib = _lib.load()
    dev = torch.device("cuda:0")
    # first call: multi-tile aligned path (TMA ring, tensor-memory parking), byte mask, several blocks per image
    c1 = _case(3, 120, 200, "bool", 101, dev)
    # second call, straight after: other inputs, other shape (scalar path: H*W odd), float mask, fewer tickets -- the
    # scratch still holds the first call's tickets and partial rows when it starts
    c2 = _case(2, 37, 53, "float", 202, dev)
    assert lib.curl_bwd_workspace_bytes(3, 120, 200) <= SCRATCH_BYTES
    ref1, ref2 = _run(lib, c1, noalloc=False), _run(lib, c2, noalloc=False)
    got1 = _run(lib, c1, noalloc=True)
    got2 = _run(lib, c2, noalloc=True)
    exact1, exact2 = _compare(got1, ref1), _compare(got2, ref2)
    # and a third, repeating the first shape without the image gradient, after the scratch was used by another shape
    got3 = _run(lib, c1, noalloc=True, want_gimg=False)
    _compare(got3, {k: ref1[k] for k in got3})
    assert exact1 and exact2, "same kernel, same grid, deterministic fold: expected bit-identical results"


def test_noalloc_backward_at_a_benchmark_shape():
    """1080p x 4 (the torch.compile baseline's shape): whole waves of resident blocks, every block several tiles."""
    from curl_b200 import _lib
    lib = _lib.load()
    dev = torch.device("cuda:0")
    c = _case(4, 1080, 1920, "bool", 303, dev)
    assert lib.curl_bwd_workspace_bytes(4, 1080, 1920) <= SCRATCH_BYTES
    assert lib.curl_bwd_workspace_bytes(64, 2160, 3840) <= SCRATCH_BYTES      # the headline shape fits too
    _compare(_run(lib, c, noalloc=True), _run(lib, c, noalloc=False))


def test_noalloc_backward_rejects_shapes_beyond_its_scratch():
    from curl_b200 import _lib
    lib = _lib.load()
    dev = torch.device("cuda:0")
    B, H, W = 20000, 4, 4                      # 20000 images x one block x 80 B of partials + tickets > 1 MiB
    assert lib.curl_bwd_workspace_bytes(B, H, W) > SCRATCH_BYTES
    img = torch.rand(B, 3, H, W, device=dev)
    L, R, Hk = (torch.zeros(B, n, device=dev) for n in (48, 48, 64))
    g = torch.zeros(B, 48, device=dev); gh = torch.zeros(B, 64, device=dev)
    rc = lib.curl_layer_bwd_noalloc_f32(img.data_ptr(), None, 0, L.data_ptr(), 48, R.data_ptr(), 48, Hk.data_ptr(), 64,
                                        img.data_ptr(), None, None, g.data_ptr(), g.data_ptr(), gh.data_ptr(), B, H, W, 16, 16, 16,
                                        torch.cuda.current_stream(dev).cuda_stream)
    assert rc == ERR_TOO_LARGE
    assert b"1 MiB" in lib.curl_strerror(rc)
    # the ordinary argument checks still come first
    assert lib.curl_layer_bwd_noalloc_f32(None, None, 0, L.data_ptr(), 48, R.data_ptr(), 48, Hk.data_ptr(), 64, img.data_ptr(),
                                          None, None, g.data_ptr(), g.data_ptr(), gh.data_ptr(), 1, H, W, 16, 16, 16, None) == -1
