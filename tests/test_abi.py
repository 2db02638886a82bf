"""CPU: the C-ABI library loads, exports every symbol include/curl_b200.h declares, and validates
its arguments (no compute calls: there is no GPU here)."""
import ctypes
import os
import re

import pytest

from conftest import ROOT
from curl_b200 import _lib, build


def header_functions():
    src = open(os.path.join(ROOT, "include", "curl_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(curl_\w+)\s*\(", src)))


@pytest.fixture(scope="module")
def lib():
    build.build()
    return _lib.load()


def test_header_and_binding_agree(lib):
    names = header_functions()
    assert len(names) >= 11
    assert sorted(_lib.SIGNATURES) == names
    for n in names:
        assert hasattr(lib, n), "symbol %s declared in the header but not exported" % n


def test_version_and_strerror(lib):
    assert lib.curl_abi_version() == 10
    assert lib.curl_strerror(0) == b"success"
    for code in (-1, -2, -3, -4, -5, -6, -99):
        assert len(lib.curl_strerror(code)) > 0


def test_workspace_query(lib):
    assert lib.curl_bwd_workspace_bytes(0, 8, 8) == 0
    a = lib.curl_bwd_workspace_bytes(1, 8, 8)
    b = lib.curl_bwd_workspace_bytes(64, 2160, 3840)
    assert 0 < a <= b < (64 << 20)


def test_argument_validation_returns_codes(lib):
    P = ctypes.c_void_p
    fake = P(0x1000)     # never dereferenced: validation rejects the call first
    n = None
    # null image
    assert lib.curl_layer_fwd_f32(n, n, 0, fake, 48, fake, 48, fake, 64, fake, fake, 1, 8, 8, 16, 16, 16, n) == -1
    # K < 3
    assert lib.curl_layer_fwd_f32(fake, n, 0, fake, 48, fake, 48, fake, 64, fake, fake, 1, 8, 8, 2, 16, 16, n) == -1
    # row stride smaller than the knots read
    assert lib.curl_layer_fwd_f32(fake, n, 0, fake, 47, fake, 48, fake, 64, fake, fake, 1, 8, 8, 16, 16, 16, n) == -1
    # mask kind / pointer mismatch
    assert lib.curl_layer_fwd_f32(fake, n, 1, fake, 48, fake, 48, fake, 64, fake, fake, 1, 8, 8, 16, 16, 16, n) == -1
    # too many knots
    assert lib.curl_layer_fwd_f32(fake, n, 0, fake, 3000, fake, 3000, fake, 4000, fake, fake, 1, 8, 8, 400, 400, 400, n) == -2
    # misaligned image pointer
    assert lib.curl_layer_fwd_f32(P(0x1001), n, 0, fake, 48, fake, 48, fake, 64, fake, fake, 1, 8, 8, 16, 16, 16, n) == -5
    # non-positive shape, unknown conversion / group
    assert lib.curl_color_fwd_f32(0, fake, fake, 0, 8, 8, n) == -1
    assert lib.curl_color_fwd_f32(7, fake, fake, 1, 8, 8, n) == -1
    assert lib.curl_adjust_fwd_f32(5, fake, fake, 48, fake, fake, 1, 8, 8, 16, n) == -1
    assert lib.curl_apply_curve_fwd_f32(fake, fake, 16, 3, 0, fake, fake, 1, 8, 8, 16, n) == -1
    # backward without workspace
    assert lib.curl_layer_bwd_f32(fake, n, 0, fake, 48, fake, 48, fake, 64, fake, n, fake, fake, fake, fake, n, 0,
                                  1, 8, 8, 16, 16, 16, n) == -3


def test_missing_library_fails_loudly(monkeypatch):
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", os.path.join(ROOT, "curl_b200", "no_such_library.so"))
    with pytest.raises(RuntimeError, match="no CPU or PyTorch fallback"):
        _lib.load()
