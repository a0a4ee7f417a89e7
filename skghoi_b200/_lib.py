"""ctypes binding of libskghoi_b200.so (the C-ABI declared in include/skghoi_b200.h).

There is NO fallback: if the shared library is missing or a call fails, an exception is raised.
PyTorch is used only to own device memory and streams; tensors cross the boundary as raw
``data_ptr()`` addresses plus explicit sizes.
"""
import contextlib
import ctypes
import os
from ctypes import c_char_p, c_float, c_int, c_int64, c_void_p

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("SKG_LIB_PATH") or os.path.join(_HERE, "libskghoi_b200.so")   # (override: kernel experiments only)

P, I, L, F = c_void_p, c_int, c_int64, c_float

_SIGNATURES = {
    "skg_last_error": (c_char_p, []),
    "skg_abi_version": (I, []),
    "skg_num_sms": (I, []),
    "skg_launch_count": (ctypes.c_longlong, []),
    "skg_pairs_spatial_encode": (I, [P, P, P, P, P, P, P, I, L, I, F, I, P, P, P, P, P, P, P, P, P, P, P, P]),
    "skg_spatial_encode_rows": (I, [P, P, P, P, I, L, I, F, P, P, P]),
    "skg_transh_forward": (I, [P, P, P, I, I, I, P, P, P, L, I, I, P, P, P, P, P, P]),
    "skg_transh_forward_grid": (I, [P, P, P, I, I, I, L, P, L, I, I, P, P, P, P, P, P]),
    "skg_transh_backward": (I, [P, P, P, I, I, I, P, P, P, L, I, I, P, P, P, P, P]),
    "skg_margin_loss": (I, [P, P, L, L, F, F, I, P, P, P, F, P]),
    "skg_focal_loss": (I, [P, P, L, F, F, F, P, P, P, P]),
    "skg_cast_pad_bf16": (I, [P, L, P, I, L, I, P, L, I, P]),
    "skg_gemm_bf16_wrapk": (I, [P, L, I, P, L, I, I, I, P, I, I, P, L, P]),
    "skg_spatial_fc1": (I, [P, P, P, L, I, I, P, P]),
    "skg_node_inputs": (I, [P, I, P, I, I, I, P, P, I, I, I, P, P, I, P]),
    "skg_global_pool": (I, [P, L, I, P, P, P]),
    "skg_fc_bias_fold": (I, [P, P, P, I, I, I, P, P]),
    "skg_node_prep": (I, [P, I, I, P, P, P, I, I, P, I, I, P, P]),
    "skg_message_aggregate": (I, [P, I, F, P, P, L, P, P, P, I, I, I, I, I, P, P, P, P]),
    "skg_layernorm": (I, [P, L, L, I, P, P, P, P, F, P, P, P]),
    "skg_hadamard_cells": (I, [P, L, P, L, P, L, P, P, L, I, P, L, P]),
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


def raw_stream(device=None) -> int:
    """cudaStream_t of torch's current stream on ``device`` as an integer (one C call, no Stream object)."""
    idx = device.index if isinstance(device, torch.device) else device
    if idx is None:
        idx = torch.cuda.current_device()
    return torch._C._cuda_getCurrentRawStream(idx)


_NULL_CTX = contextlib.nullcontext()


def on_device(device):
    """``torch.cuda.device(device)`` only when ``device`` is not already current (the context manager costs ~10 us)."""
    return _NULL_CTX if torch.cuda.current_device() == device.index else torch.cuda.device(device)


def stream_ptr(device=None):
    return ctypes.c_void_p(raw_stream(device))


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


class EpilogueGroup(ctypes.Structure):
    """Mirror of ``skg_epilogue_group`` (include/skghoi_b200.h)."""
    _fields_ = [
        ("n_begin", ctypes.c_int32), ("flags", ctypes.c_int32),
        ("bias", c_void_p),
        ("mul_p", c_void_p), ("row_p", c_void_p), ("ld_p", ctypes.c_int32),
        ("mul_q", c_void_p), ("row_q", c_void_p), ("ld_q", ctypes.c_int32),
        ("residual", c_void_p), ("row_r", c_void_p), ("ld_r", ctypes.c_int32),
        ("dot_w", c_void_p), ("dot_out", c_void_p), ("ld_dot", ctypes.c_int32),
        ("out_bf16", c_void_p), ("ld_ob", ctypes.c_int32),
        ("out_f32", c_void_p), ("ld_of", ctypes.c_int32),
        ("raw_bf16", c_void_p), ("ld_raw", ctypes.c_int32),
        ("out_row", c_void_p),
        ("lo_bf16", c_void_p), ("ld_lo", ctypes.c_int32),
        ("mask_bf16", c_void_p), ("ld_mask", ctypes.c_int32),
    ]


class PostTables(ctypes.Structure):
    """Mirror of ``skg_post_tables``."""
    _fields_ = [(k, c_void_p) for k in ("verb_off", "verb_idx", "pair2cell", "cell_hrow", "cell_orow", "cell_img",
                                        "hum_off", "box_off", "pair_off", "labels", "scores", "box_prefix", "box_cnt",
                                        "hum_base", "logit_row")]


_SIGNATURES["skg_multiscale_roi_align"] = (I, [P, P, P, P, I, I, I, P, P, L, I, I, I, I, F, I, P, P, L, P])
_SIGNATURES["skg_postprocess"] = (I, [P, I, I, ctypes.POINTER(PostTables), L, F, P, L, P, P, P, P, P, P, P, P, P])
_SIGNATURES["skg_head_post"] = (I, [P, I, I, ctypes.POINTER(PostTables), P, I, I, I, L, F, P, L, P, P, P, P, P, P, P, P, P, P, P, P, P])
_SIGNATURES["skg_prior_dense"] = (I, [ctypes.POINTER(PostTables), L, I, F, P, P])
_SIGNATURES["skg_post_offsets"] = (I, [P, P, P, P, P, I, P, P, P, P, P, P])
_SIGNATURES["skg_gather_pair_boxes"] = (I, [P, ctypes.POINTER(PostTables), L, P, P, P])
_SIGNATURES["skg_associate_gt"] = (I, [P, ctypes.POINTER(PostTables), L, I, P, P, P, P, F, P, P])

EPI_RELU = 1
_SIGNATURES["skg_gemm_bf16"] = (I, [P, L, P, L, I, I, I, ctypes.POINTER(EpilogueGroup), I, I, P])


def epilogue_group(n_begin=0, relu=False, bias=None, mul_p=None, row_p=None, mul_q=None, row_q=None,
                   residual=None, row_r=None, dot_w=None, dot_out=None, out_bf16=None, out_f32=None,
                   raw_bf16=None, out_row=None, lo_bf16=None, mask_bf16=None):
    """Builds one epilogue group from tensors (leading dimensions are taken from ``stride(0)``)."""
    def a(t):
        return t.data_ptr() if t is not None else None

    def ld(t):
        return int(t.stride(0)) if t is not None and t.dim() >= 2 else 0
    g = EpilogueGroup()
    g.n_begin = int(n_begin); g.flags = EPI_RELU if relu else 0
    g.bias = a(bias)
    g.mul_p, g.row_p, g.ld_p = a(mul_p), a(row_p), ld(mul_p)
    g.mul_q, g.row_q, g.ld_q = a(mul_q), a(row_q), ld(mul_q)
    g.residual, g.row_r, g.ld_r = a(residual), a(row_r), ld(residual)
    g.dot_w, g.dot_out, g.ld_dot = a(dot_w), a(dot_out), ld(dot_out)
    g.out_bf16, g.ld_ob = a(out_bf16), ld(out_bf16)
    g.out_f32, g.ld_of = a(out_f32), ld(out_f32)
    g.raw_bf16, g.ld_raw = a(raw_bf16), ld(raw_bf16)
    g.out_row = a(out_row)
    g.lo_bf16, g.ld_lo = a(lo_bf16), ld(lo_bf16)
    g.mask_bf16, g.ld_mask = a(mask_bf16), ld(mask_bf16)
    return g


_SIGNATURES["skg_gemm_bf16_splitk"] = (I, [P, L, P, L, I, I, I, ctypes.POINTER(EpilogueGroup), I, I, P, L, P])
_SIGNATURES["skg_gemm_splitk_scratch_bytes"] = (c_int64, [I, I])
_SIGNATURES["skg_gemm_bf16_splitk_static"] = (I, [P, L, P, L, I, I, I, ctypes.POINTER(EpilogueGroup), I, I, P])
_SIGNATURES["skg_gemm_static_scratch_bytes"] = (c_int64, [])


def gemm_bf16(A, B, groups, M=None, N=None, K=None, block_n=0, scratch=None):
    """C = epilogue(A[M,K] @ B[N,K]^T); A, B bf16 CUDA tensors with unit inner stride.  ``scratch`` (a zeroed
    uint8 CUDA tensor of skg_gemm_splitk_scratch_bytes) enables split-K for few-tile shapes."""
    require_cuda(A, B)
    assert A.dtype == torch.bfloat16 and B.dtype == torch.bfloat16 and A.stride(1) == 1 and B.stride(1) == 1
    M = A.shape[0] if M is None else M
    N = B.shape[0] if N is None else N
    K = A.shape[1] if K is None else K
    arr = (EpilogueGroup * len(groups))(*groups)
    with torch.cuda.device(A.device):
        if scratch is not None:
            call("skg_gemm_bf16_splitk", ptr(A), A.stride(0), ptr(B), B.stride(0), M, N, K, arr, len(groups), block_n,
                 ptr(scratch), scratch.numel(), stream_ptr(A.device))
        else:
            call("skg_gemm_bf16", ptr(A), A.stride(0), ptr(B), B.stride(0), M, N, K, arr, len(groups), block_n,
                 stream_ptr(A.device))


# ---- whole-path forward (csrc/forward.cu) -------------------------------------------------------
_HW_INTS = ("R", "E", "K0", "pool_k", "fc_k", "Kg", "Cg", "num_cls", "n_logit", "n_logit_pad", "human_idx", "ent_tot",
            "ent_dim", "sp_in")


class HeadWeights(ctypes.Structure):
    """Mirror of ``skg_head_weights``: raw device pointers of the packed operands."""
    _fields_ = [(k, ctypes.c_int32) for k in _HW_INTS] + [
        ("Wb1", c_void_p), ("bb1", c_void_p), ("Wb2", c_void_p), ("bb2", c_void_p),
        ("Wsp0", c_void_p), ("bsp0", c_void_p), ("Wsp1", c_void_p), ("bsp1", c_void_p), ("Wsp2", c_void_p), ("bsp2", c_void_p),
        ("Wfh", c_void_p), ("bfh", c_void_p), ("Wft", c_void_p), ("bft", c_void_p),
        ("Wfc", c_void_p), ("bfc", c_void_p),
        ("W_hproj", c_void_p), ("b_hproj", c_void_p), ("W_oproj", c_void_p), ("b_oproj", c_void_p),
        ("Wg1", c_void_p), ("bg1", c_void_p), ("W2cat", c_void_p), ("b2cat", c_void_p),
        ("W3att", c_void_p), ("b3att", c_void_p), ("W3os", c_void_p), ("b3os", c_void_p),
        ("W3so", c_void_p), ("b3so", c_void_p), ("W3g", c_void_p), ("b3g", c_void_p),
        ("w_adj", c_void_p), ("b_adj", c_float),
        ("ln_h_w", c_void_p), ("ln_h_b", c_void_p), ("ln_o_w", c_void_p), ("ln_o_b", c_void_p), ("ln_eps", c_float),
        ("Wpred", c_void_p), ("bpred", c_void_p), ("Wpred_hh", c_void_p), ("Wpred_lo", c_void_p),
        ("ent", c_void_p), ("Wpred_hhl", c_void_p), ("Wfh_ent", c_void_p)]


class HeadBatch(ctypes.Structure):
    """Mirror of ``skg_head_batch``."""
    _fields_ = [(k, ctypes.c_int32) for k in ("num_images", "total_boxes", "total_humans", "max_boxes_per_image", "feat_hw")] + \
               [("total_cells", c_int64), ("total_pairs", c_int64), ("seg_bytes", c_int64)] + \
               [(k, c_void_p) for k in ("features3", "box_features", "boxes", "box_off", "n_h", "hum_off", "cell_off",
                                        "pair_off", "hw", "box_features_bf16", "maps_ready_event")]


HEAD_NUM_BUFFERS = 96


class HeadPlan(ctypes.Structure):
    """Mirror of ``skg_head_plan_t``."""
    _fields_ = [("total_bytes", c_int64), ("num_buffers", ctypes.c_int32),
                ("offset", c_int64 * HEAD_NUM_BUFFERS), ("rows", c_int64 * HEAD_NUM_BUFFERS),
                ("cols", c_int64 * HEAD_NUM_BUFFERS), ("elem_size", ctypes.c_int32 * HEAD_NUM_BUFFERS)]


HEAD_PRECISE_LOGITS = 1
HEAD_WANT_TAPS = 2
HEAD_PROFILE = 4
HEAD_GRAPH = 8
HEAD_TRAIN = 16
HEAD_CAPACITY = 32
_SIGNATURES["skg_head_capacity"] = (I, [ctypes.POINTER(HeadWeights), ctypes.POINTER(HeadBatch), ctypes.POINTER(HeadBatch)])


class SessionConfig(ctypes.Structure):
    """Mirror of ``skg_session_config``."""
    _fields_ = [("weights", HeadWeights), ("verb_len", c_void_p), ("verb_off_dev", c_void_p), ("verb_idx_dev", c_void_p),
                ("num_obj_cls", ctypes.c_int32), ("max_images", ctypes.c_int32), ("flags", ctypes.c_int32), ("prior_power", c_float)]


class StepSizes(ctypes.Structure):
    """Mirror of ``skg_step_sizes``."""
    _fields_ = [(k, c_int64) for k in ("boxes", "humans", "cells", "pairs", "entries", "cap_boxes", "cap_humans", "cap_cells",
                                       "workspace_bytes", "out_i64", "out_f32")]


class StepIO(ctypes.Structure):
    """Mirror of ``skg_step_io``."""
    _fields_ = [(k, c_void_p) for k in ("features3", "box_features", "box_features_bf16", "boxes", "labels", "scores", "ent",
                                        "out_i64", "out_f32")]


_SIGNATURES["skg_session_create"] = (I, [ctypes.POINTER(SessionConfig), ctypes.POINTER(c_void_p)])
_SIGNATURES["skg_session_destroy"] = (None, [c_void_p])
_SIGNATURES["skg_session_prepare"] = (I, [c_void_p, I, P, P, P, P, I, ctypes.POINTER(StepSizes), P, P])
_SIGNATURES["skg_session_run"] = (I, [c_void_p, ctypes.POINTER(StepIO), P, L, P])
_SIGNATURES["skg_session_plans"] = (I, [c_void_p, ctypes.POINTER(HeadPlan), ctypes.POINTER(HeadPlan)])
_SIGNATURES["skg_tensor_map_cache_stats"] = (None, [ctypes.POINTER(ctypes.c_longlong)] * 2)
_SIGNATURES["skg_head_profile_get"] = (I, [I, ctypes.POINTER(c_char_p), ctypes.POINTER(I), ctypes.POINTER(c_float)])
_SIGNATURES["skg_head_buffer_name"] = (c_char_p, [I])
_SIGNATURES["skg_head_plan"] = (I, [ctypes.POINTER(HeadWeights), ctypes.POINTER(HeadBatch), I, ctypes.POINTER(HeadPlan)])
_SIGNATURES["skg_head_forward"] = (I, [ctypes.POINTER(HeadWeights), ctypes.POINTER(HeadBatch), I, P, L, P])
_SIGNATURES["skg_head_graph_stats"] = (None, [ctypes.POINTER(ctypes.c_longlong)] * 3)
_SIGNATURES["skg_segments_pack"] = (c_int64, [P, P, P, I, P, L, P, P])
_SIGNATURES["skg_post_count_host"] = (c_int64, [P, I, P, P, P, P, I, P])

_SIGNATURES["skg_preprocess_detections"] = (I, [P, P, P, P, I, I, F, F, I, I, I, P, P, P])

_SIGNATURES["skg_kg_sample"] = (I, [P, P, P, P, P, P, P, P, L, L, P, P, I, L, I, I, P, P, P, P, P])

class KgTrainArgs(ctypes.Structure):
    """Mirror of ``skg_kg_train_args``."""
    _fields_ = [(k, c_void_p) for k in ("head_hrt", "tail_hrt", "lef_head", "rig_head", "lef_tail", "rig_tail", "left_mean", "right_mean")] + \
               [("train_total", c_int64), ("ent_total", c_int64), ("rel_total", c_int64), ("seeds", c_void_p)] + \
               [(k, ctypes.c_int32) for k in ("work_threads", "neg_rate", "mode", "dim", "p_norm", "norm_flag", "num_steps", "pad_")] + \
               [("batch_size", c_int64), ("ent", c_void_p), ("rel", c_void_p), ("nrm", c_void_p), ("margin", c_float), ("lr", c_float),
                ("loss_out", c_void_p), ("scratch", c_void_p), ("scratch_bytes", c_int64)]


_SIGNATURES["skg_kg_train_scratch_bytes"] = (c_int64, [L, I, L, L, I])
_SIGNATURES["skg_kg_train_steps"] = (I, [ctypes.POINTER(KgTrainArgs), P])

# ---- training step (csrc/trainops.cu, csrc/backward.cu); struct pointers travel as void* (ctypes.byref) -------------
_SIGNATURES["skg_block_copy"] = (I, [P, P, I, P])
_SIGNATURES["skg_head_backward"] = (I, [P, P, P, P, P, I, P, L, P])
_SIGNATURES["skg_head_backward_graph_stats"] = (None, [ctypes.POINTER(ctypes.c_longlong)] * 3)
_SIGNATURES["skg_head_backward_staged"] = (I, [P, P, P, P, P, I, P, L, P, P, I])
_SIGNATURES["skg_transpose_bf16"] = (I, [P, L, L, I, P, P, L, P, P])
_SIGNATURES["skg_select_count"] = (I, [P, P, I, I, P, P])
_SIGNATURES["skg_select_fill"] = (I, [P, P, I, I, P, ctypes.c_ulonglong, P, P, P])
_SIGNATURES["skg_loss_normalisers"] = (I, [P, P, L, I, P, P])
_SIGNATURES["skg_hum2box"] = (I, [P, I, P, P])
_SIGNATURES["skg_head_loss_backward"] = (I, [P, I, I, P, P, L, F, P, P, P, P, P, I, P])
_SIGNATURES["skg_hadamard_backward"] = (I, [P, I, I, P])
_SIGNATURES["skg_message_aggregate_backward"] = (I, [P, P, P, P, P, L, P, P, P, I, I, I, I, I, P, P, L, P, P, P])
_SIGNATURES["skg_adjacency_backward"] = (I, [P, P, P, L, P, L, I, P, L, P, P, P])
_SIGNATURES["skg_layernorm_backward"] = (I, [P, P, P, P, L, L, I, P, P, F, P, P, P, P, P, P, P])
_SIGNATURES["skg_spatial_fc1_backward"] = (I, [P, P, L, I, I, P, P, P])

_buffer_names = None


def head_buffer_names():
    global _buffer_names
    if _buffer_names is None:
        l = lib()
        names, i = [], 0
        while True:
            s = l.skg_head_buffer_name(i)
            if s is None:
                break
            names.append(s.decode()); i += 1
        _buffer_names = names
    return _buffer_names


def graph_stats():
    """(replays, captures, plain launches) of the skg_head_forward graph cache on this thread."""
    a, b, c = ctypes.c_longlong(), ctypes.c_longlong(), ctypes.c_longlong()
    lib().skg_head_graph_stats(ctypes.byref(a), ctypes.byref(b), ctypes.byref(c))
    return a.value, b.value, c.value
