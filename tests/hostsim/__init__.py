"""TEST INFRASTRUCTURE ONLY: builds and binds tests/hostsim/pixel_sim.cpp (host instantiation of the
kernel's per-pixel templates).  See the header of pixel_sim.cpp."""
import ctypes
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "pixel_sim.cpp")
HDR = os.path.join(HERE, "..", "..", "curl_b200", "csrc", "curl_pixel.cuh")
HDR2 = os.path.join(HERE, "..", "..", "curl_b200", "csrc", "curl_vec.cuh")
HDR3 = os.path.join(HERE, "..", "..", "curl_b200", "csrc", "curl_poly.cuh")
HDR4 = os.path.join(HERE, "..", "..", "curl_b200", "csrc", "curl_loss.cuh")
HDR5 = os.path.join(HERE, "..", "..", "curl_b200", "csrc", "curl_seg.cuh")
OUT = os.path.join(HERE, "_build", "libpixel_sim.so")

_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    newest = max(os.path.getmtime(SRC), os.path.getmtime(HDR), os.path.getmtime(HDR2), os.path.getmtime(HDR3), os.path.getmtime(HDR4), os.path.getmtime(HDR5))
    if not os.path.exists(OUT) or os.path.getmtime(OUT) < newest:
        os.makedirs(os.path.dirname(OUT), exist_ok=True)
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-shared", "-fPIC", "-o", OUT, SRC])
    _lib = ctypes.CDLL(OUT)
    return _lib


def _f(a):
    return None if a is None else a.ctypes.data_as(ctypes.c_void_p)


def _c(a, dtype=np.float32):
    return None if a is None else np.ascontiguousarray(a, dtype=dtype)


def _mask(mask):
    if mask is None:
        return None, 0
    if mask.dtype == np.bool_ or mask.dtype == np.uint8:
        return np.ascontiguousarray(mask).view(np.uint8), 1
    return np.ascontiguousarray(mask, dtype=np.float32), 2


def layer(img, mask, L, R, H, pts=(48, 48, 64), gout=None, greg=None):
    """Returns dict(out, reg) or, when gout is given, dict(gimg, gL, gR, gH)."""
    img = _c(img); L = _c(L); R = _c(R); H = _c(H)
    B, _, h, w = img.shape
    m, kind = _mask(mask)
    Kl, Kr, Kh = pts[0] // 3, pts[1] // 3, pts[2] // 4
    i64 = ctypes.c_int64
    if gout is None:
        out = np.empty_like(img); reg = np.empty(B, np.float32)
        rc = lib().sim_layer(0, _f(img), _f(m), kind, _f(L), i64(L.shape[1]), _f(R), i64(R.shape[1]), _f(H),
                             i64(H.shape[1]), _f(out), _f(reg), None, None, None, None, None, None, B, h * w, Kl, Kr, Kh)
        assert rc == 0
        return dict(out=out, reg=reg)
    gout = _c(gout); greg = _c(greg)
    gimg = np.empty_like(img)
    gL = np.zeros((B, 3 * Kl), np.float32); gR = np.zeros((B, 3 * Kr), np.float32); gH = np.zeros((B, 4 * Kh), np.float32)
    rc = lib().sim_layer(1, _f(img), _f(m), kind, _f(L), i64(L.shape[1]), _f(R), i64(R.shape[1]), _f(H), i64(H.shape[1]),
                         None, None, _f(gout), _f(greg), _f(gimg), _f(gL), _f(gR), _f(gH), B, h * w, Kl, Kr, Kh)
    assert rc == 0
    return dict(gimg=gimg, gL=gL, gR=gR, gH=gH)


def color(conv, x, gy=None):
    x = _c(x)
    B, _, h, w = x.shape
    out = np.empty_like(x)
    gy = _c(gy)
    rc = lib().sim_color(0 if gy is None else 1, conv, _f(x), _f(gy), _f(out), B, h * w)
    assert rc == 0
    return out


def adjust(group, x, P, K, gy=None, greg=None):
    x = _c(x); P = _c(P)
    B, _, h, w = x.shape
    nc = 4 if group == 2 else 3
    i64 = ctypes.c_int64
    if gy is None:
        out = np.empty_like(x); reg = np.empty(B, np.float32)
        rc = lib().sim_adjust(0, group, _f(x), _f(P), i64(P.shape[1]), _f(out), _f(reg), None, None, None, None, B, h * w, K)
        assert rc == 0
        return dict(y=out, reg=reg)
    gy = _c(gy); greg = _c(greg)
    gx = np.empty_like(x); gP = np.zeros((B, nc * K), np.float32)
    rc = lib().sim_adjust(1, group, _f(x), _f(P), i64(P.shape[1]), None, None, _f(gy), _f(greg), _f(gx), _f(gP), B, h * w, K)
    assert rc == 0
    return dict(gx=gx, gP=gP)


def u8_to_unit_table():
    out = np.empty(256, np.float32)
    lib().sim_u8_to_unit_table(_f(out))
    return out


def unit_to_u8(y):
    y = _c(y)
    out = np.empty(y.shape, np.uint8)
    lib().sim_unit_to_u8(_f(y), _f(out), ctypes.c_longlong(y.size))
    return out


def trispace(img, R, L, Hc, spatial, clamp, gout=None, degree=4, want_gimg=False):
    """Returns out[B,3,H,W] or, when gout is given, dict(gR, gL, gH) (want_gimg: the image gradient instead)."""
    img = _c(img); R = _c(R); L = _c(L); Hc = _c(Hc)
    B, _, h, w = img.shape
    if gout is None:
        out = np.empty_like(img)
        rc = lib().sim_trispace_deg(0, _f(img), _f(R), _f(L), _f(Hc), int(spatial), int(degree), int(clamp), None, _f(out),
                                    None, None, None, B, h, w)
        assert rc == 0
        return out
    gout = _c(gout)
    if want_gimg:
        gimg = np.empty_like(img)
        rc = lib().sim_trispace_deg(2, _f(img), _f(R), _f(L), _f(Hc), int(spatial), int(degree), int(clamp), _f(gout), _f(gimg),
                                    None, None, None, B, h, w)
        assert rc == 0
        return gimg
    gR = np.empty_like(R); gL = np.empty_like(L); gH = np.empty_like(Hc)
    rc = lib().sim_trispace_deg(1, _f(img), _f(R), _f(L), _f(Hc), int(spatial), int(degree), int(clamp), _f(gout), None,
                                _f(gR), _f(gL), _f(gH), B, h, w)
    assert rc == 0
    return dict(gR=gR, gL=gL, gH=gH)


def loss_terms(pred, target, mask, w=None, gL=None):
    """Returns dict(sums[B,5], L_pred, L_target) or, when w[B,4] is given, gpred[B,3,H,W]."""
    pred = _c(pred); target = _c(target)
    B, _, h, wd = pred.shape
    m = None if mask is None else _c(np.broadcast_to(np.asarray(mask, dtype=np.float32), (B, 1, h, wd)))
    if w is None:
        sums = np.empty((B, 5), np.float32)
        Lp = np.empty((B, 1, h, wd), np.float32); Lt = np.empty((B, 1, h, wd), np.float32)
        rc = lib().sim_loss_terms(0, _f(pred), _f(target), _f(m), _f(sums), _f(Lp), _f(Lt), None, None, None, B, h * wd)
        assert rc == 0
        return dict(sums=sums, L_pred=Lp, L_target=Lt)
    w = _c(w); gL = _c(gL)
    gpred = np.empty_like(pred)
    rc = lib().sim_loss_terms(1, _f(pred), _f(target), _f(m), None, None, None, _f(w), _f(gL), _f(gpred), B, h * wd)
    assert rc == 0
    return gpred


def layer_seg(img, mask, L, R, H, pts=(48, 48, 64), gout=None, greg=None):
    """Segment-clamped layer.  Returns dict(out, reg) or, when gout is given, dict(gimg, gL, gR, gH)."""
    img = _c(img)
    B, _, h, w = img.shape
    Kl, Kr, Kh = pts[0] // 3, pts[1] // 3, pts[2] // 4
    knots = _c(np.concatenate([np.asarray(L)[:, :pts[0]], np.asarray(R)[:, :pts[1]], np.asarray(H)[:, :pts[2]]], axis=1))
    m = None if mask is None else _c(np.broadcast_to(np.asarray(mask, dtype=np.float32), (B, 1, h, w)))
    if gout is None:
        out = np.empty_like(img); reg = np.empty(B, np.float32)
        rc = lib().sim_layer_seg(0, _f(img), _f(m), _f(knots), Kl, Kr, Kh, _f(out), _f(reg), None, None, None, None, B, h * w)
        assert rc == 0
        return dict(out=out, reg=reg)
    gout = _c(gout); greg = _c(greg)
    gimg = np.empty_like(img); gk = np.empty_like(knots)
    rc = lib().sim_layer_seg(1, _f(img), _f(m), _f(knots), Kl, Kr, Kh, None, None, _f(gout), _f(greg), _f(gimg), _f(gk), B, h * w)
    assert rc == 0
    return dict(gimg=gimg, gL=gk[:, :pts[0]], gR=gk[:, pts[0]:pts[0] + pts[1]], gH=gk[:, pts[0] + pts[1]:])
