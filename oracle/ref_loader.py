"""TEST INFRASTRUCTURE ONLY -- loads the *real* reference curve layer for pinning the oracle.

Only usable where /root/reference exists (the build container; never the GPU box).  Nothing from
the reference is copied into this repo: the three source files are read where they lie and patched
in memory, exactly as SURVEY.md section 8(c) documents:

  * colors.py            -- imported unmodified.
  * curves.py            -- the three `slope_sqr_diff = None` initialisers (curves.py:56,111,155)
                            make `apply_curve` raise TypeError at curves.py:24; they are replaced by
                            `slope_sqr_diff = 0.0` (0.0 += Tensor[B] yields the [B] tensor).
  * model.py::CURLLayer  -- model.py:121-176 is extracted with `ast`; the three dead
                            `torch.cat((feat, ...))` lines (model.py:152,158,164; `feat` is
                            undefined, results unused) are dropped.  model.py itself cannot be
                            imported (matplotlib/timm absent).

Used by oracle/make_golden.py (to generate tests/golden/*.npz) and by tests/test_oracle_pins.py
(to prove oracle/curl_oracle.py is bit-identical to the reference on CPU).
"""
import ast
import importlib.util
import os
import sys
import types

REFERENCE_ROOT = os.environ.get("CURL_REFERENCE_ROOT", "/root/reference")


def available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "curves.py"))


def load():
    """Returns (colors_module, curves_module_patched, CURLLayer_class_patched)."""
    import torch
    import torch.nn as nn

    if not available():
        raise RuntimeError("reference tree not present at %s" % REFERENCE_ROOT)

    spec = importlib.util.spec_from_file_location("_ref_colors", os.path.join(REFERENCE_ROOT, "colors.py"))
    colors = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(colors)

    with open(os.path.join(REFERENCE_ROOT, "curves.py")) as f:
        src = f.read()
    assert src.count("slope_sqr_diff = None") == 3, "reference curves.py changed; re-derive the patch"
    src = src.replace("slope_sqr_diff = None", "slope_sqr_diff = 0.0")
    curves = types.ModuleType("_ref_curves")
    exec(compile(src, os.path.join(REFERENCE_ROOT, "curves.py"), "exec"), curves.__dict__)

    with open(os.path.join(REFERENCE_ROOT, "model.py")) as f:
        msrc = f.read()
    tree = ast.parse(msrc)
    cls = [n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == "CURLLayer"]
    assert len(cls) == 1
    lines = msrc.splitlines()[cls[0].lineno - 1: cls[0].end_lineno]
    kept = [ln for ln in lines if "torch.cat((feat" not in ln]
    assert len(lines) - len(kept) == 3, "reference model.py changed; re-derive the patch"
    ns = {"torch": torch, "nn": nn, "colors": colors, "curves": curves}
    exec(compile("\n".join(kept), os.path.join(REFERENCE_ROOT, "model.py"), "exec"), ns)
    return colors, curves, ns["CURLLayer"]


def load_trispace(degree=4, spatial=False, mobile=False):
    """Returns (ChannelPolyLayer_class, net) where `net` is a reference TriSpaceRegNet carrying
    everything its generate_residual / generate_image need, WITHOUT its backbone: __init__
    (model.py:413-457) calls timm.create_model, which is not installed, so the object is made with
    __new__ and given the attributes __init__ would set (polylayer, the four colour modules, sigmoid,
    the x / y coordinate parameters of model.py:444-452).  The classes themselves are the reference's
    source, extracted from model.py with ast and executed unmodified.
    mobile=True uses the reference's Deg4MobilePolyLayer (model.py:336-408) as the polylayer."""
    import functools
    import operator
    import torch
    import torch.nn as nn

    colors, _, _ = load()
    with open(os.path.join(REFERENCE_ROOT, "model.py")) as f:
        msrc = f.read()
    tree = ast.parse(msrc)
    ns = {"torch": torch, "nn": nn, "colors": colors, "op": operator, "reduce": functools.reduce}
    for name in ("ChannelPolyLayer", "Deg4MobilePolyLayer", "TriSpaceRegNet"):
        cls = [n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == name]
        assert len(cls) == 1, name
        src = "\n".join(msrc.splitlines()[cls[0].lineno - 1: cls[0].end_lineno])
        exec(compile(src, os.path.join(REFERENCE_ROOT, "model.py"), "exec"), ns)
    Net = ns["TriSpaceRegNet"]
    net = Net.__new__(Net)
    nn.Module.__init__(net)
    net.num_channels, net.num_spaces = 3, 3
    net.num_in = 3 + 2 * spatial
    net.order = degree
    net.max_resolution = 10000
    if mobile:
        assert degree == 4 and spatial
        net.polylayer = ns["Deg4MobilePolyLayer"]()
    else:
        net.polylayer = ns["ChannelPolyLayer"](degree=degree, num_variables=net.num_in, num_out=3)
    net.rgb2lab, net.lab2rgb = colors.RGB2LAB(), colors.LAB2RGB()
    net.rgb2hsv, net.hsv2rgb = colors.RGB2HSV(), colors.HSV2RGB()
    net.sigmoid = nn.Sigmoid()
    if not spatial:                                                            # model.py:444-452
        x = torch.zeros(1, 0, 1, net.max_resolution)
        y = torch.zeros(1, 0, net.max_resolution, 1)
    else:
        x = torch.arange(0, net.max_resolution).reshape(1, 1, 1, net.max_resolution)
        y = torch.arange(0, net.max_resolution).reshape(1, 1, net.max_resolution, 1)
    net.x = torch.nn.Parameter(x, requires_grad=False)
    net.y = torch.nn.Parameter(y, requires_grad=False)
    return ns["ChannelPolyLayer"], net


def load_loss():
    """Returns a reference CURLLoss instance (model.py:36-118) with the reference's own MSSSIMMetric
    (metric.py:75-214) inside.  Both classes are the reference's source, extracted with ast; one
    patch: metric.py hard-codes `.cuda()` on intermediates (metric.py:116,150-157), which cannot run
    in this GPU-less container, so every `.cuda()` is dropped (a no-op for CPU tensors)."""
    import math
    import torch
    import torch.nn as nn
    import torch.nn.functional as F
    from torch.autograd import Variable

    colors, _, _ = load()
    with open(os.path.join(REFERENCE_ROOT, "metric.py")) as f:
        src = f.read()
    tree = ast.parse(src)
    cls = [n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == "MSSSIMMetric"]
    assert len(cls) == 1
    csrc = "\n".join(src.splitlines()[cls[0].lineno - 1: cls[0].end_lineno])
    assert csrc.count(".cuda()") == 10, "reference metric.py changed; re-derive the patch"
    metric = types.ModuleType("_ref_metric")
    metric.__dict__.update({"torch": torch, "nn": nn, "F": F, "Variable": Variable, "exp": math.exp})
    exec(compile(csrc.replace(".cuda()", ""), os.path.join(REFERENCE_ROOT, "metric.py"), "exec"), metric.__dict__)

    with open(os.path.join(REFERENCE_ROOT, "model.py")) as f:
        msrc = f.read()
    tree = ast.parse(msrc)
    cls = [n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == "CURLLoss"]
    assert len(cls) == 1
    lsrc = "\n".join(msrc.splitlines()[cls[0].lineno - 1: cls[0].end_lineno])
    ns = {"torch": torch, "nn": nn, "F": F, "math": math, "colors": colors, "metric": metric}
    exec(compile(lsrc, os.path.join(REFERENCE_ROOT, "model.py"), "exec"), ns)
    return ns["CURLLoss"]()
