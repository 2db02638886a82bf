"""ctypes binding of libcurl_b200.so (include/curl_b200.h).  There is no fallback: if the library
is missing or a call fails, a RuntimeError is raised."""
import ctypes
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libcurl_b200.so")

ABI_VERSION = 10

MASK_NONE, MASK_U8, MASK_F32 = 0, 1, 2
RGB2LAB, LAB2RGB, RGB2HSV, HSV2RGB = 0, 1, 2, 3
ADJUST_LAB, ADJUST_RGB, ADJUST_HSV = 0, 1, 2
LAYOUT_CHW, LAYOUT_HWC = 0, 1

_p = ctypes.c_void_p
_i = ctypes.c_int
_i64 = ctypes.c_int64
_sz = ctypes.c_size_t

# name -> (restype, argtypes); must list every symbol include/curl_b200.h declares
SIGNATURES = {
    "curl_abi_version": (_i, []),
    "curl_strerror": (ctypes.c_char_p, [_i]),
    "curl_test_set_grid_cap": (_i, [_i]),
    "curl_layer_fwd_f32": (_i, [_p, _p, _i, _p, _i64, _p, _i64, _p, _i64, _p, _p, _i, _i, _i, _i, _i, _i, _p]),
    "curl_layer_bwd_f32": (_i, [_p, _p, _i, _p, _i64, _p, _i64, _p, _i64, _p, _p, _p, _p, _p, _p, _p, _sz,
                                _i, _i, _i, _i, _i, _i, _p]),
    "curl_layer_bwd_noalloc_f32": (_i, [_p, _p, _i, _p, _i64, _p, _i64, _p, _i64, _p, _p, _p, _p, _p, _p,
                                        _i, _i, _i, _i, _i, _i, _p]),
    "curl_layer_fwd_u8": (_i, [_p, _i, _p, _i, _p, _i64, _p, _i64, _p, _i64, _p, _i, _p, _i, _i, _i, _i, _i, _i, _p]),
    "curl_bwd_workspace_bytes": (_sz, [_i, _i, _i]),
    "curl_color_fwd_f32": (_i, [_i, _p, _p, _i, _i, _i, _p]),
    "curl_color_bwd_f32": (_i, [_i, _p, _p, _p, _i, _i, _i, _p]),
    "curl_adjust_fwd_f32": (_i, [_i, _p, _p, _i64, _p, _p, _i, _i, _i, _i, _p]),
    "curl_adjust_bwd_f32": (_i, [_i, _p, _p, _i64, _p, _p, _p, _p, _p, _sz, _i, _i, _i, _i, _p]),
    "curl_apply_curve_fwd_f32": (_i, [_p, _p, _i64, _i, _i, _p, _p, _i, _i, _i, _i, _p]),
    "curl_trispace_num_coeffs": (_i, [_i, _i]),
    "curl_trispace_fwd_f32": (_i, [_p, _p, _p, _p, _i, _i, _i, _p, _p, _i, _i, _i, _p]),
    "curl_trispace_bwd_workspace_bytes": (_sz, [_i, _i, _i, _i, _i]),
    "curl_trispace_bwd_f32": (_i, [_p, _p, _p, _p, _i, _i, _i, _p, _p, _p, _p, _p, _p, _sz, _i, _i, _i, _p]),
    "curl_trispace_bwd_img_f32": (_i, [_p, _p, _p, _p, _i, _i, _i, _p, _p, _p, _i, _i, _i, _p]),
    "curl_layer_seg_fwd_f32": (_i, [_p, _p, _i, _p, _i64, _p, _i64, _p, _i64, _p, _p, _i, _i, _i, _i, _i, _i, _p]),
    "curl_layer_seg_bwd_f32": (_i, [_p, _p, _i, _p, _i64, _p, _i64, _p, _i64, _p, _p, _p, _p, _p, _p, _p, _sz,
                                    _i, _i, _i, _i, _i, _i, _p]),
    "curl_layer_seg_bwd_workspace_bytes": (_sz, [_i, _i, _i, _i, _i, _i]),
    "curl_u8_to_unit_f32": (_i, [_p, _i, _p, _i, _i, _i, _p]),
    "curl_loss_terms_workspace_bytes": (_sz, [_i, _i, _i]),
    "curl_loss_terms_fwd_f32": (_i, [_p, _p, _p, _i, _p, _p, _p, _p, _sz, _i, _i, _i, _p]),
    "curl_loss_terms_bwd_f32": (_i, [_p, _p, _p, _i, _p, _p, _p, _i, _i, _i, _p]),
    "curl_apply_curve_bwd_f32": (_i, [_p, _p, _i64, _i, _i, _p, _p, _p, _p, _p, _sz, _i, _i, _i, _i, _p]),
}

_lib = None
launches = 0   # kernels this process has enqueued through the ABI (bench.py reports it)


def load():
    """Loads the library once; raises RuntimeError (never falls back) if it cannot."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            "curl_b200: %s is missing. Build it with `python -m curl_b200.build` (needs nvcc); "
            "there is no CPU or PyTorch fallback for this path." % LIB_PATH)
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)   # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    v = lib.curl_abi_version()
    if v != ABI_VERSION:
        raise RuntimeError("curl_b200: ABI version mismatch (library %d, binding %d); rebuild" % (v, ABI_VERSION))
    _lib = lib
    return lib


def check(rc, what):
    if rc != 0:
        msg = load().curl_strerror(rc).decode()
        raise RuntimeError("curl_b200: %s failed: %s (code %d)" % (what, msg, rc))


def count_launch(n=1):
    global launches
    launches += n
