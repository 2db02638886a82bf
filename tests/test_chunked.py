"""CPU: the row-chunked oracle used by the full-size GPU parity tests (tests/chunked.py) equals the plain
oracle -- pixels are independent, knot gradients add over rows, the regulariser gradient is applied once.
Not bit for bit: torch's vectorised CPU kernels (pow, exp) round the SIMD body and the scalar tail of a
tensor differently, so a pixel's last bit depends on where it sits in the chunk; the differences are a few
ulp (measured below 1e-6 on outputs in fp32), far inside the 1e-5 bar the chunked oracle is used for."""
import numpy as np
import pytest
import torch

from chunked import layer_fwd_bwd_chunked
from oracle import curl_oracle as O


@pytest.mark.parametrize("maskkind", ["none", "bool", "float"])
@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
def test_chunked_oracle_equals_plain(maskkind, dtype):
    gen = torch.Generator().manual_seed(3)
    B, H, W = 2, 37, 29
    img = torch.rand(B, 3, H, W, generator=gen)
    mask = None if maskkind == "none" else (torch.rand(B, 1, H, W, generator=gen) > 0.2)
    if maskkind == "float":
        mask = mask.float() * 0.75
    L, R, Hk = (0.1 * torch.randn(B, n, generator=gen) for n in (48, 48, 64))
    gout = torch.randn(B, 3, H, W, generator=gen)
    greg = torch.rand(B, generator=gen) + 0.5
    om = torch.ones(B, 1, H, W, dtype=torch.bool) if mask is None else mask
    om = om if om.dtype == torch.bool else om.to(dtype)
    ref = O.curl_layer_fwd_bwd(img.to(dtype), om, L.to(dtype), R.to(dtype), Hk.to(dtype), gout.to(dtype), greg.to(dtype))
    for rows in (5, 16, 64):
        got = layer_fwd_bwd_chunked(img, mask, L, R, Hk, gout, greg, dtype=dtype, rows=rows)
        tol = 2e-6 if dtype == torch.float32 else 1e-13     # fp32: ulp-level, and a different summation order over pixels
        assert np.abs(got["out"] - ref["out"].numpy()).max() <= tol
        # an ulp on a clamp bound or tie flips a gate: isolated pixels may differ, as between any two fp32 runs
        g, r = got["gimg"], ref["gimg"].numpy()
        assert (np.abs(g - r) > 10 * tol * np.abs(r).max()).mean() <= 2e-3
        assert np.array_equal(got["reg"], ref["reg"].numpy())
        for k in ("gL", "gR", "gH"):
            r = ref[k].double().numpy()
            assert np.abs(got[k] - r).max() <= 20 * tol * np.abs(r).max(), (k, rows)


@pytest.mark.parametrize("spatial,degree", [(True, 4), (False, 4), (True, 2)])
@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
def test_chunked_trispace_oracle_equals_plain(spatial, degree, dtype):
    """The row-chunked TriSpace oracle (its own coordinate planes: y = row / H of the whole image) against
    O.trispace_fwd_bwd on the whole image."""
    from chunked import trispace_fwd_bwd_chunked
    gen = torch.Generator().manual_seed(11)
    B, H, W = 2, 23, 19
    nc = len(O.poly_powers(degree, 5 if spatial else 3))
    img = torch.rand(B, 3, H, W, generator=gen)
    R, L, Hc = (0.3 * torch.randn(B, 3, nc, generator=gen) for _ in range(3))
    gout = torch.randn(B, 3, H, W, generator=gen)
    for clamp in (True, False):
        ref = O.trispace_fwd_bwd(img.to(dtype), R.to(dtype), L.to(dtype), Hc.to(dtype), gout.to(dtype), degree, spatial, clamp)
        for rows in (4, 23):
            got = trispace_fwd_bwd_chunked(img, R, L, Hc, gout, degree, spatial, clamp, dtype=dtype, rows=rows)
            tol = 3e-6 if dtype == torch.float32 else 1e-12
            r = ref["out"].numpy()
            assert np.abs(got["out"] - r).max() <= tol * max(1.0, np.abs(r).max()), (clamp, rows)
            for k in ("gR", "gL", "gH"):
                r = ref[k].double().numpy()
                assert np.abs(got[k] - r).max() <= 20 * tol * np.abs(r).max(), (k, clamp, rows)
