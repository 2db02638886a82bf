"""curl_b200 -- B200-native (sm_100a) implementation of the CURL per-pixel curve layer.

Mirrors the reference's module API for this path:
    curl_b200.colors  : RGB2LAB, LAB2RGB, RGB2HSV, HSV2RGB        (reference colors.py)
    curl_b200.curves  : apply_curve, adjust_lab, adjust_rgb, adjust_hsv   (reference curves.py)
    curl_b200.model   : CURLLayer                                 (reference model.py:121-176)
                        GCURLNet (backbone -> knots -> layer)     (reference model.py:179-203)
                        TriSpaceRegNet, ChannelPolyLayer, Deg4MobilePolyLayer (fused polynomial residual, model.py:206-333, 411-520)
All of them run hand-written CUDA kernels through the C ABI in include/curl_b200.h; there is no
CPU or PyTorch fallback.
"""
from . import colors, curves, metric, model  # noqa: F401
from .model import CURLLayer, CURLLoss, GCURLNet, ChannelPolyLayer, Deg4MobilePolyLayer, TriSpaceRegNet  # noqa: F401

__all__ = ["colors", "curves", "metric", "model", "CURLLayer", "CURLLoss", "GCURLNet", "ChannelPolyLayer",
           "Deg4MobilePolyLayer", "TriSpaceRegNet"]
