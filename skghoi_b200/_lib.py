"""ctypes binding of libskghoi_b200.so (the C-ABI declared in include/skghoi_b200.h).

There is NO fallback: if the shared library is missing or a call fails, an exception is raised.
PyTorch is used only to own device memory and streams; tensors cross the boundary as raw
``data_ptr()`` addresses plus explicit sizes.
"""
import ctypes
import os
from ctypes import c_char_p, c_float, c_int, c_int64, c_void_p

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libskghoi_b200.so")

P, I, L, F = c_void_p, c_int, c_int64, c_float

_SIGNATURES = {
    "skg_last_error": (c_char_p, []),
    "skg_abi_version": (I, []),
    "skg_num_sms": (I, []),
    "skg_pairs_spatial_encode": (I, [P, P, P, P, P, P, P, I, L, I, F, I, P, P, P, P, P, P, P, P, P, P, P, P]),
    "skg_spatial_encode_rows": (I, [P, P, P, P, I, L, I, F, P, P, P]),
    "skg_transh_forward": (I, [P, P, P, I, I, I, P, P, P, L, I, I, P, P, P, P, P, P]),
    "skg_transh_forward_grid": (I, [P, P, P, I, I, I, L, P, L, I, I, P, P, P, P, P, P]),
    "skg_transh_backward": (I, [P, P, P, I, I, I, P, P, P, L, I, I, P, P, P, P, P]),
    "skg_margin_loss": (I, [P, P, L, L, F, F, I, P, P, P, F, P]),
    "skg_focal_loss": (I, [P, P, L, F, F, F, P, P, P, P]),
}

_lib = None


class SkgError(RuntimeError):
    pass


def declared_symbols():
    return sorted(_SIGNATURES)


def register(name, restype, argtypes):
    _SIGNATURES[name] = (restype, argtypes)
    if _lib is not None:
        fn = getattr(_lib, name)
        fn.restype, fn.argtypes = restype, argtypes


def lib():
    """Loads the library (once). Raises if it has not been built -- there is no CPU path."""
    global _lib
    if _lib is None:
        if not os.path.isfile(LIB_PATH):
            raise SkgError(
                "libskghoi_b200.so is not built (%s). Run `python -c 'import __graft_entry__ as g; g.build()'`"
                " or `python skghoi_b200/build.py`; there is no CPU fallback." % LIB_PATH)
        l = ctypes.CDLL(LIB_PATH)
        for name, (restype, argtypes) in _SIGNATURES.items():
            fn = getattr(l, name)          # AttributeError if the .so lacks a declared symbol
            fn.restype, fn.argtypes = restype, argtypes
        _lib = l
    return _lib


def ptr(t):
    """Device/host address of a tensor (None -> NULL)."""
    if t is None:
        return None
    return ctypes.c_void_p(t.data_ptr())


def stream_ptr(device=None):
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def call(name, *args):
    l = lib()
    rc = getattr(l, name)(*args)
    if rc != 0:
        raise SkgError("%s failed (%d): %s" % (name, rc, l.skg_last_error().decode(errors="replace")))


def require_cuda(*tensors):
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise SkgError("skghoi_b200 runs on CUDA tensors only (got a %s tensor); there is no CPU fallback"
                           % t.device.type)
